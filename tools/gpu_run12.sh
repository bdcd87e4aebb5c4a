set -x
cd $GRAFT_REPO_ROOT
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 240 $TR --master-port 29511 tests/mgpu_parity.py > gpurun_out/r2_12_mgpu2.log 2>&1 || { echo "PARITY FAILED rc=$?"; tail -30 gpurun_out/r2_12_mgpu2.log; exit 1; }
grep -E "mgpu" gpurun_out/r2_12_mgpu2.log
MGPU_TMODE=1 MGPU_CASES=1,4 timeout 200 $TR --master-port 29512 tests/mgpu_parity.py 2>&1 | grep mgpu
MGPU_TMODE=0 MGPU_CASES=1,4 MGPU_BIG=0 timeout 200 $TR --master-port 29513 tests/mgpu_parity.py 2>&1 | grep mgpu
timeout 200 $TR --master-port 29514 tools/sweep_mgpu.py --steps=5 "transpose_mode=2" "transpose_mode=1" "transpose_mode=0" "transpose_mode=2,fft_pipe=2" "transpose_mode=2,fft_pipe=8" > gpurun_out/r2_12_sweep2.log 2>&1; grep sweep gpurun_out/r2_12_sweep2.log
BENCH_PARITY_SMALL=1 timeout 200 $TR --master-port 29515 bench.py --gpus 2 --steps 200 --warmup 5 --no-cpu --workload kolmo2d > gpurun_out/r2_12_bench2_kolmo.json 2> gpurun_out/r2_12_bench2_kolmo.err || { echo "KOLMO2 FAILED rc=$?"; tail -20 gpurun_out/r2_12_bench2_kolmo.err; }
python tools/sumbench.py gpurun_out/r2_12_bench2_kolmo.json
timeout 900 python -m pytest tests -m gpu -q -s -x > gpurun_out/r2_12_pytest_gpu.log 2>&1; echo "suite rc=$?"; tail -4 gpurun_out/r2_12_pytest_gpu.log; grep "tolerance" gpurun_out/r2_12_pytest_gpu.log | sort | uniq | head -60
