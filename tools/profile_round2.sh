set -x
CMD="python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e --no-dropin"
$CMD > gpurun_out/r02_plain.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:^k_ --csv --log-file gpurun_out/r02_launches_c2.csv $CMD > gpurun_out/r02_ncu_launch.log 2>&1
for k in k_enum_pipeline k_split_walk k_dls_ball k_chain_init k_chain_resolve k_peel_final; do
  ncu --set full --clock-control none --import-source on -k regex:^$k\$ -s 1 -c 1 -o gpurun_out/r02_ncufull_$k -f $CMD > gpurun_out/r02_ncufull_$k.log 2>&1
done
MCAAT_ENUM_TRACE=gpurun_out/r02_enum_trace_c2.csv $CMD > gpurun_out/r02_trace_run.json 2>/dev/null
ls -la gpurun_out/r02_ncufull_* | head -20
