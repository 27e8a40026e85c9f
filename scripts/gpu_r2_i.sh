# launch list (gpu__time_duration) of the distinct-tuples path of configs[4]
mkdir -p gpurun_out
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/i_launches.csv \
  python scripts/union_probe.py --words 4000000 union_dedup=1 > gpurun_out/i_ncu.log 2>&1; echo "ncu exit $?"
tail -2 gpurun_out/i_ncu.log
