# interleaved repeat of the scripts/ab.sh variants with clocks, to separate box noise from the effect
mkdir -p gpurun_out
cp memb_b200/libmemb_b200.so /tmp/libmemb_b200.default.so
nvidia-smi --query-gpu=name,serial,uuid,clocks.sm,clocks.mem,power.draw,temperature.gpu,temperature.memory --format=csv
for rep in 1 2; do
for v in ${AB_VARIANTS:-a b d}; do
  d=memb_b200/variants/$v
  cp $d/libmemb_b200.so memb_b200/libmemb_b200.so
  for mode in ${AB_MODES:-random sequential}; do
    python bench.py --steps 200 --warmup 10 --no-extra --ids $mode 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('$(cat $d/defs.txt) | $mode: ms %.4f rows/s %.4g parity %s clocks %s' % (d['ms_per_step'], d['value'], d['parity_vs_oracle'], d['clocks']))"
  done
done
done 2>&1 | tee gpurun_out/g_ab.log
cp /tmp/libmemb_b200.default.so memb_b200/libmemb_b200.so
