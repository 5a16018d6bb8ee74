# final measurements of round 2 on one B200 (after tests are green): smoke, default bench, reference arm, variants
set -x
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo smoke rc=$?
python bench.py > gpurun_out/r02_bench_c2_n1.json 2> gpurun_out/bench_n1.err; echo bench rc=$?
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r02_bench_reference_arm.json 2> gpurun_out/bench_ref.err; echo ref rc=$?
python bench.py --workload c4 --steps 3 --warmup 3 --no-cpu-baseline --no-dropin > gpurun_out/r02_bench_c4_n1.json 2> gpurun_out/bench_c4.err; echo c4 rc=$?
python bench.py --workload c2_errors --steps 3 --warmup 2 --no-cpu-baseline --no-dropin > gpurun_out/r02_bench_c2_errors.json 2>/dev/null; echo errors rc=$?
python bench.py --sweep --steps 2 --warmup 2 --no-cpu-baseline --no-e2e --no-dropin > gpurun_out/r02_bench_c2_sweep.json 2>/dev/null; echo sweep rc=$?
