# quick GPU check: parity tests + a short bench (summary on stdout, JSON in gpurun_out/bench_q.json)
python -m pytest tests -m gpu -q -x --timeout 900 2>&1 | tail -4
python bench.py --steps 100 --warmup 5 > gpurun_out/bench_q.json 2> gpurun_out/bench_q.err
python -c "
import json; d=json.load(open('gpurun_out/bench_q.json')); print('rows/s %.4g ms %.4f frac %.4f parity %s launches %d clocks %s' % (d['value'], d['ms_per_step'], d['roofline']['frac'], d['parity_vs_oracle'], d['gpu_launches'], d['clocks'])); print(d.get('extra'))"
tail -3 gpurun_out/bench_q.err
