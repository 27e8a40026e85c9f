# usage: bash scripts/gpu_n.sh N [extra bench args]  -> gpurun_out/bench_nN.json
N=$1; shift
SECONDS=0
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 20 --warmup 5 "$@" > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err
echo "bench n=$N exit $? after $SECONDS s"
grep -v "^\*\|OMP_NUM\|^$" gpurun_out/bench_n$N.err | tail -6 | cut -c1-300
python -c "
import json; d=json.load(open('gpurun_out/bench_n$N.json')); print(json.dumps(d['strong_scaling'], indent=0)); print(d['parity'])"
