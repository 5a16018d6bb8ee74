python -m pytest tests/test_gpu_parity.py tests/test_gpu_full_size.py -x -q -m gpu > gpurun_out/parity15.log 2>&1
python bench.py --steps 4 --warmup 2 --no-cpu-baseline --no-e2e --no-dropin > gpurun_out/dense.json 2>/dev/null
python bench.py --workload c4 --steps 2 --warmup 1 --no-cpu-baseline --no-e2e --no-dropin > gpurun_out/dense_c4.json 2>/dev/null
