python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "dump or empty or ingest" > gpurun_out/parity14.log 2>&1
BENCH_DEBUG=1 python bench.py --steps 4 --warmup 2 --no-cpu-baseline --no-dropin > gpurun_out/e2e.json 2> gpurun_out/e2e.err
