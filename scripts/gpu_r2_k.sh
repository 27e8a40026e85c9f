# launch list (gpu__time_duration) of the final distinct-tuples path, configs[4] at one GPU
mkdir -p gpurun_out
timeout 65 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/k_launches.csv \
  python scripts/union_probe.py --words 4000000 union_dedup=1 > gpurun_out/k_ncu.log 2>&1; echo "ncu exit $?"
