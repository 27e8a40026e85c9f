# round 2, late: K1 with the MEMB_BOUNDARY_STORE variants of scripts/ab.sh, then the union average per variant
mkdir -p gpurun_out
cp memb_b200/libmemb_b200.so /tmp/libmemb_b200.default.so
timeout 900 bash scripts/ab_run.sh > gpurun_out/f_ab.log 2>&1; echo "ab exit $?"
cat gpurun_out/f_ab.log
for d in memb_b200/variants/*/; do
  cp $d/libmemb_b200.so memb_b200/libmemb_b200.so
  echo "== $(cat $d/defs.txt)"
  timeout 300 python scripts/union_probe.py --words 4000000 union_pair_table=3 2>&1 | grep setting
done 2>&1 | tee gpurun_out/f_union.log
cp /tmp/libmemb_b200.default.so memb_b200/libmemb_b200.so
