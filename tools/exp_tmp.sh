CMD="python bench.py --steps 4 --warmup 2 --no-cpu-baseline --no-e2e --no-dropin"
python -m pytest tests/test_gpu_parity.py tests/test_gpu_full_size.py -x -q -m gpu > gpurun_out/parity7.log 2>&1
$CMD > gpurun_out/sterm.json 2>/dev/null
