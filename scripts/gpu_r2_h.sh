# ncu --set full of the copy-out kernel of the distinct-tuples path (full_gather_kernel over the temporary), configs[4]
mkdir -p gpurun_out
timeout 400 python scripts/union_probe.py --words 4000000 union_dedup=1 > gpurun_out/h_plain.log 2>&1; echo "plain exit $?"
grep setting gpurun_out/h_plain.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:full_gather_kernel -s 1 -c 1 -f -o gpurun_out/r02_copy_out \
  python scripts/union_probe.py --words 4000000 union_dedup=1 > gpurun_out/h_ncu.log 2>&1; echo "ncu exit $?"
tail -3 gpurun_out/h_ncu.log
