# round 2: GPU tests + the full default bench line at one GPU
python -m pytest tests -m gpu -x -q --timeout 900 > gpurun_out/r2b_tests.log 2>&1; echo "pytest exit $?" >> gpurun_out/r2b_tests.log
tail -5 gpurun_out/r2b_tests.log
/usr/bin/time -v python bench.py --steps 20 --warmup 5 > gpurun_out/r2b_bench.json 2> gpurun_out/r2b_bench.err; echo "bench exit $?"
tail -5 gpurun_out/r2b_bench.err
