set -x
python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/noalloc_alone_final.log 2>&1; echo alone rc=$?
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo smoke rc=$?
python bench.py > gpurun_out/r02_bench_c2_n1.json 2> gpurun_out/bench_n1.err; echo bench rc=$?
python bench.py --workload c4 --steps 3 --warmup 3 --no-cpu-baseline --no-dropin > gpurun_out/r02_bench_c4_n1.json 2> gpurun_out/bench_c4.err; echo c4 rc=$?
