python -m pytest tests -m gpu -x -q --timeout 900 > gpurun_out/r2d_tests.log 2>&1; echo "pytest exit $?" >> gpurun_out/r2d_tests.log
tail -4 gpurun_out/r2d_tests.log
rm -f gpurun_out/r2d_quick.jsonl
for opt in "trained_tma=1" "trained_tma=0"; do
    python bench.py --quick --steps 50 --ids random --option $opt 2>> gpurun_out/r2d_quick.err | tee -a gpurun_out/r2d_quick.jsonl | cut -c1-300
done
python scripts/sweep_probe.py --bits 8 --deep 2>>gpurun_out/r2d_quick.err
python scripts/sweep_probe.py --bits 8 --deep --option trained_tma=0 2>>gpurun_out/r2d_quick.err
tail -3 gpurun_out/r2d_quick.err
