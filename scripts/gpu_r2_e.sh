# round 2, late: (1) GPU parity suite on the in-tree build, (2) union average with the pair table as one copy
# (union_pair_table=2) against one copy per bank (=3), (3) K1 with the MEMB_BOUNDARY_STORE variants of scripts/ab.sh
mkdir -p gpurun_out
cp memb_b200/libmemb_b200.so /tmp/libmemb_b200.default.so
timeout 900 python -m pytest tests -m gpu -x -q --timeout 600 > gpurun_out/e_tests.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/e_tests.log
tail -3 gpurun_out/e_tests.log
timeout 400 python scripts/union_probe.py --words 4000000 union_pair_table=2 union_pair_table=3 union_pair_table=2 union_pair_table=3 > gpurun_out/e_union.log 2>&1; echo "union exit $?"
grep setting gpurun_out/e_union.log
timeout 600 bash scripts/ab_run.sh > gpurun_out/e_ab.log 2>&1; echo "ab exit $?"
cat gpurun_out/e_ab.log
cp /tmp/libmemb_b200.default.so memb_b200/libmemb_b200.so
