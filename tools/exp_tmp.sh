python -m pytest tests/test_gpu_parity.py -x -q -m gpu > gpurun_out/parity12.log 2>&1
python bench.py --steps 4 --warmup 2 --no-cpu-baseline --no-e2e --no-dropin > gpurun_out/sel.json 2>/dev/null
