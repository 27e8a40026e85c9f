# round 2, final: parity suite, the full bench line, then (each after the same command exited 0 without ncu) the launch
# list, one --set full capture of the headline kernel, and one of the union average's three kernels (symbol-word
# decode, merged flush on the per-bank pair table, copy-out of the distinct path)
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q --timeout 900 > gpurun_out/gpu_tests.log 2>&1; echo "pytest exit $?" >> gpurun_out/gpu_tests.log
tail -3 gpurun_out/gpu_tests.log
/usr/bin/time -v python bench.py --steps 20 --warmup 5 > gpurun_out/bench_full.json 2> gpurun_out/bench_full.err; echo "bench exit $?"
grep -E "Elapsed|Maximum resident" gpurun_out/bench_full.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; echo "reference arm exit $?"
python bench.py --quick --steps 3 --warmup 3 > gpurun_out/quick.plain.log 2>&1 &&
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches.csv python bench.py --quick --steps 3 --warmup 3 > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches exit $?"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:trained_decode_kernel -s 3 -c 1 -f -o gpurun_out/r02_k1_ring python bench.py --quick --steps 3 --warmup 3 > gpurun_out/ncu_ring.log 2>&1
echo "ncu ring exit $?"
python scripts/union_probe.py --words 4000000 union_dedup=1 > gpurun_out/union.plain.log 2>&1 &&
timeout 400 ncu --set full --clock-control none --import-source on -k regex:"trained_decode_kernel|full_gather_kernel|dedup_claim64" -s 12 -c 4 -f -o gpurun_out/r02_union_average python scripts/union_probe.py --words 4000000 union_dedup=1 > gpurun_out/ncu_union.log 2>&1
echo "ncu union exit $?"
