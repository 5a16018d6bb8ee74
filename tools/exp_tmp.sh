CMD="python bench.py --workload c4 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-dropin"
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:^k_ --csv --log-file gpurun_out/r02_launches_c4.csv $CMD > /dev/null 2>&1
