# one ncu --set full capture of the decode kernel (after the same command exited 0 without ncu)
# usage: bash scripts/gpu_ncu.sh [extra bench.py args]   -> gpurun_out/prof.ncu-rep
python bench.py --steps 3 --warmup 3 --no-extra "$@" > gpurun_out/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:trained_decode_kernel -s 3 -c 1 \
    -f -o gpurun_out/prof python bench.py --steps 3 --warmup 3 --no-extra "$@" > gpurun_out/ncu_full.log 2>&1
echo "ncu full exit $?"
