# distinct-tuples path of configs[4]: plain timing of both claim kernels, then the launch list (gpu__time_duration)
mkdir -p gpurun_out
timeout 400 python scripts/union_probe.py --words 4000000 union_dedup_keys=0 union_dedup_keys=1 2>&1 | grep setting | cut -c1-200
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/i_launches.csv \
  python scripts/union_probe.py --words 4000000 union_dedup_keys=0 union_dedup_keys=1 > gpurun_out/i_ncu.log 2>&1; echo "ncu exit $?"
