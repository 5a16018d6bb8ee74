CMD="python bench.py --steps 4 --warmup 2 --no-cpu-baseline --no-e2e --no-dropin"
python -m pytest tests/test_gpu_parity.py tests/test_gpu_full_size.py -x -q -m gpu > gpurun_out/parity8.log 2>&1
MCAAT_PEEL_VERBOSE=1 $CMD > gpurun_out/branch.json 2> gpurun_out/branch.err
