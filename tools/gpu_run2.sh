set -x
cd $GRAFT_REPO_ROOT
timeout 300 python -m pytest tests/test_gpu_fast.py -x -q -s > gpurun_out/r2_02_pytest_fast.log 2>&1; echo "rc=$?" >> gpurun_out/r2_02_pytest_fast.log
tail -12 gpurun_out/r2_02_pytest_fast.log
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu > gpurun_out/r2_02_bench_smag_fast.json 2> gpurun_out/r2_02_bench_smag_fast.err; echo "rc=$?"
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu --no-parity --param exact=1 > gpurun_out/r2_02_bench_smag_exact.json 2> gpurun_out/r2_02_bench_smag_exact.err; echo "rc=$?"
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu --workload tgv3d > gpurun_out/r2_02_bench_tgv_fast.json 2> gpurun_out/r2_02_bench_tgv_fast.err; echo "rc=$?"
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu --no-parity --workload tgv3d --param exact=1 > gpurun_out/r2_02_bench_tgv_exact.json 2> gpurun_out/r2_02_bench_tgv_exact.err; echo "rc=$?"
for f in gpurun_out/r2_02_bench_*.json; do python tools/sumbench.py $f; done
tail -3 gpurun_out/r2_02_bench_smag_fast.err
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_stage3d_fast --launch-skip 14 -c 4 -o gpurun_out/r2_02_ncu_fast -f python bench.py --steps 2 --warmup 3 --no-cpu --no-parity --modes-init > gpurun_out/r2_02_ncu.log 2>&1; echo "ncu rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_stage3d_fast --launch-skip 14 -c 2 -o gpurun_out/r2_02_ncu_fast_tgv -f python bench.py --steps 2 --warmup 3 --no-cpu --no-parity --modes-init --workload tgv3d > gpurun_out/r2_02_ncu_tgv.log 2>&1; echo "ncu rc=$?"
