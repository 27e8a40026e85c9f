# ncu --set full of trained_rows_kernel: single reader (random ids) and the one-launch union (average)
python bench.py --quick --steps 3 --warmup 3 --option trained_rows_kernel=2 > gpurun_out/ncu_rows1.plain.log 2>&1 &&
timeout 300 ncu --set full --clock-control none --import-source on -k regex:trained_rows_kernel -s 3 -c 1 -f -o gpurun_out/r02_rows_single python bench.py --quick --steps 3 --warmup 3 --option trained_rows_kernel=2 > gpurun_out/ncu_rows1.log 2>&1
echo "ncu rows single exit $?"
python scripts/union_probe.py trained_rows_kernel=1 > gpurun_out/ncu_rows2.plain.log 2>&1 &&
timeout 400 ncu --set full --clock-control none --import-source on -k regex:trained_rows_kernel -s 2 -c 1 -f -o gpurun_out/r02_rows_union_average python scripts/union_probe.py trained_rows_kernel=1 > gpurun_out/ncu_rows2.log 2>&1
echo "ncu rows union exit $?"
