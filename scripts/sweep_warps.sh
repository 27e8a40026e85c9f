# development aid: decode rate as a function of resident warps per SM
for w in 6 8 9 10; do
  python bench.py --steps 50 --warmup 5 --no-extra --trained-warps $w > gpurun_out/sw.json 2>/dev/null
  python -c "import json; d=json.load(open('gpurun_out/sw.json')); print('warps $w rows/s %.4g ms %.4f' % (d['value'], d['ms_per_step']))"
done
