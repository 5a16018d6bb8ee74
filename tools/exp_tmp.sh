python -m pytest tests/test_gpu_parity.py tests/test_gpu_full_size.py -x -q -m gpu > gpurun_out/parity10.log 2>&1
MCAAT_DLS_VERBOSE=1 python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-dropin > /dev/null 2> gpurun_out/dlsv.err
python bench.py --steps 4 --warmup 2 --no-cpu-baseline --no-e2e --no-dropin > gpurun_out/dcut.json 2>/dev/null
