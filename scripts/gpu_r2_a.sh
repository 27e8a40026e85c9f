# round 2: full GPU test suite, then the whole-row kernel against the ring kernel
python -m pytest tests -m gpu -x -q --timeout 900 > gpurun_out/r2a_tests.log 2>&1; echo "pytest exit $?" >> gpurun_out/r2a_tests.log
tail -5 gpurun_out/r2a_tests.log
rm -f gpurun_out/r2a_bench.jsonl
for opt in "trained_rows_kernel=1" "trained_rows_kernel=0" "trained_rows_warps=8" "trained_rows_warps=6"; do
  for ids in random sequential; do
    python bench.py --quick --steps 50 --ids $ids --option $opt 2>> gpurun_out/r2a_bench.err | tee -a gpurun_out/r2a_bench.jsonl | cut -c1-200
  done
done
