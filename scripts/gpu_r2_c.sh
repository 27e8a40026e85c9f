# round 2: the full default bench line at one GPU
SECONDS=0
python bench.py --steps 20 --warmup 5 > gpurun_out/r2c_bench.json 2> gpurun_out/r2c_bench.err; echo "bench exit $? after $SECONDS s"
tail -5 gpurun_out/r2c_bench.err | cut -c1-300
