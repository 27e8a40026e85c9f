# runs the GPU parity tests (set AB_TESTS=1) and bench.py once per variant built by scripts/ab.sh
# (the variant's library replaces the in-tree one)
for d in memb_b200/variants/*/; do
  cp $d/libmemb_b200.so memb_b200/libmemb_b200.so
  if [ -n "$AB_TESTS" ]; then python -m pytest tests -m gpu -q -x --timeout 900 2>&1 | tail -1; fi
  for mode in random sequential; do
    python bench.py --steps 100 --warmup 5 --no-extra --ids $mode "$@" 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('$(cat $d/defs.txt) | $mode: ms %.4f rows/s %.4g parity %s' % (d['ms_per_step'], d['value'], d['parity_vs_oracle']))"
  done
done
