set -x
cd $GRAFT_REPO_ROOT
timeout 400 python -m pytest tests/test_gpu_zline.py -x -q -s 2>&1 | tail -25 > gpurun_out/r2_16_pytest_zline.log; echo "pytest rc=$?"; tail -8 gpurun_out/r2_16_pytest_zline.log
for z in 0 1; do
timeout 200 python bench.py --steps 5 --warmup 3 --no-cpu --no-parity --workload tgv3d --param zline=$z > gpurun_out/r2_16_bench1_zline$z.json 2> gpurun_out/r2_16_bench1_zline$z.err; echo "bench rc=$?"
python tools/sumbench.py gpurun_out/r2_16_bench1_zline$z.json | head -9
done
