# ncu --set full captures for profiles/ (each after the same command ran clean without ncu)
# 1) headline: ring kernel, 1 M random ids, 4-bit table
python bench.py --quick --steps 3 --warmup 3 > gpurun_out/ncu_ring.plain.log 2>&1 &&
timeout 300 ncu --set full --clock-control none --import-source on -k regex:trained_decode_kernel -s 3 -c 1 -f -o gpurun_out/r02_k1_ring python bench.py --quick --steps 3 --warmup 3 > gpurun_out/ncu_ring.log 2>&1
echo "ncu ring exit $?"
# 2) single-copy 12-bit LUT (8-bit table) and 3) second-level table (16-bit codes): random ids
python scripts/sweep_probe.py --bits 8 --deep > gpurun_out/ncu_sweep.plain.log 2>&1 &&
timeout 400 ncu --set full --clock-control none --import-source on -k regex:trained_decode_kernel -c 8 -f -o gpurun_out/r02_k1_sweep python scripts/sweep_probe.py --bits 8 --deep > gpurun_out/ncu_sweep.log 2>&1
echo "ncu sweep exit $?"
# 4) launch list of the quick bench
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches.csv python bench.py --quick --steps 3 --warmup 3 > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches exit $?"
