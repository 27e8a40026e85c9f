# ncu --set full of trained_rows_kernel (after the same command ran clean without ncu)
# usage: bash scripts/gpu_ncu_rows.sh <tag> [bench args]
tag=$1; shift
python bench.py --quick --steps 3 --warmup 3 "$@" > gpurun_out/ncu_$tag.plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:trained_rows_kernel -s 3 -c 1 \
    -f -o gpurun_out/rows_$tag python bench.py --quick --steps 3 --warmup 3 "$@" > gpurun_out/ncu_$tag.log 2>&1
echo "ncu $tag exit $?"
