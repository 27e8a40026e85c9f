# ncu --set full of the union average's second launch (SYM = 3) at 4 M Zipf words, every query decoded
mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"trained_decode_kernel" -s 13 -c 1 -f -o gpurun_out/r02_union_sym3 python scripts/union_probe.py --words 4000000 union_dedup=1 > gpurun_out/ncu_sym3.log 2>&1
echo "ncu exit $?"
