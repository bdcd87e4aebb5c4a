set -x
cd $GRAFT_REPO_ROOT
nvidia-smi -L | head -2
timeout 300 python -m pytest tests/test_gpu_fast.py -x -q -s > gpurun_out/r2_01_pytest_fast.log 2>&1; echo "rc=$?" >> gpurun_out/r2_01_pytest_fast.log
tail -25 gpurun_out/r2_01_pytest_fast.log
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu > gpurun_out/r2_01_bench_smag_fast.json 2> gpurun_out/r2_01_bench_smag_fast.err; echo "rc=$?"
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu --no-parity --param exact=1 > gpurun_out/r2_01_bench_smag_exact.json 2> gpurun_out/r2_01_bench_smag_exact.err; echo "rc=$?"
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu --workload tgv3d > gpurun_out/r2_01_bench_tgv_fast.json 2> gpurun_out/r2_01_bench_tgv_fast.err; echo "rc=$?"
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu --no-parity --workload tgv3d --param exact=1 > gpurun_out/r2_01_bench_tgv_exact.json 2> gpurun_out/r2_01_bench_tgv_exact.err; echo "rc=$?"
for f in gpurun_out/r2_01_bench_*.json; do python tools/sumbench.py $f 2>/dev/null || head -c 600 $f; done
tail -5 gpurun_out/r2_01_bench_smag_fast.err
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_stage3d_fast --launch-skip 14 -c 4 -o gpurun_out/r2_01_ncu_fast -f python bench.py --steps 2 --warmup 3 --no-cpu --no-parity --modes-init > gpurun_out/r2_01_ncu.log 2>&1; echo "ncu rc=$?"
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2_01_pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/r2_01_pytest_gpu.log
tail -15 gpurun_out/r2_01_pytest_gpu.log
