timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_full_size.py -x -q -m gpu > gpurun_out/parity24.log 2>&1; echo rc=$?
timeout 200 python bench.py --steps 4 --warmup 2 --no-cpu-baseline --no-e2e --no-dropin > gpurun_out/cw.json 2>/dev/null; echo rc=$?
timeout 300 python bench.py --workload c4 --steps 2 --warmup 2 --no-cpu-baseline --no-e2e --no-dropin > gpurun_out/cw_c4.json 2>/dev/null; echo rc=$?
