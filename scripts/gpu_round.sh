# One GPU call: parity tests, full bench, then the ncu launch list and one --set full capture
# of the decode kernel (each ncu run only after the same command exited 0 without ncu).
python -m pytest tests -m gpu -x -q --timeout 900 > gpurun_out/gpu_tests.log 2>&1; echo "pytest exit $?" >> gpurun_out/gpu_tests.log
tail -3 gpurun_out/gpu_tests.log
python bench.py > gpurun_out/bench_full.json 2> gpurun_out/bench_full.err; echo "bench exit $?"
tail -2 gpurun_out/bench_full.err
python bench.py --steps 3 --warmup 3 --no-extra > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 3 --warmup 3 --no-extra > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches exit $?"
ncu --set full --clock-control none --import-source on -k regex:trained_decode_kernel -s 3 -c 2 \
    -f -o gpurun_out/prof python bench.py --steps 3 --warmup 3 --no-extra > gpurun_out/ncu_full.log 2>&1
echo "ncu full exit $?"
