# round 2: GPU tests, whole-row kernel vs ring kernel (quick), then the full default bench line at one GPU
python -m pytest tests -m gpu -x -q --timeout 900 > gpurun_out/r2b_tests.log 2>&1; echo "pytest exit $?" >> gpurun_out/r2b_tests.log
tail -5 gpurun_out/r2b_tests.log
rm -f gpurun_out/r2b_quick.jsonl
for opt in "trained_rows_kernel=2" "trained_rows_kernel=0" "trained_rows_warps=8"; do
  for ids in random sequential; do
    python bench.py --quick --steps 50 --ids $ids --option trained_rows_kernel=2 --option $opt 2>> gpurun_out/r2b_quick.err | tee -a gpurun_out/r2b_quick.jsonl | cut -c1-170
  done
done
/usr/bin/time -v python bench.py --steps 20 --warmup 5 > gpurun_out/r2b_bench.json 2> gpurun_out/r2b_bench.err; echo "bench exit $?"
tail -5 gpurun_out/r2b_bench.err | cut -c1-300
