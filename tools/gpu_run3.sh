set -x
cd $GRAFT_REPO_ROOT
nvidia-smi -L
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 400 $TR --master-port 29511 tests/mgpu_parity.py > gpurun_out/r2_03_mgpu2_default.log 2>&1; echo "rc=$?" >> gpurun_out/r2_03_mgpu2_default.log
grep -E "mgpu|rc=|Error|error" gpurun_out/r2_03_mgpu2_default.log | tail -12
MGPU_BIG=0 MGPU_FASTBAR=0 timeout 300 $TR --master-port 29512 tests/mgpu_parity.py > gpurun_out/r2_03_mgpu2_ncclbar.log 2>&1; echo "rc=$?" >> gpurun_out/r2_03_mgpu2_ncclbar.log
grep -E "mgpu|rc=" gpurun_out/r2_03_mgpu2_ncclbar.log | tail -8
MGPU_BIG=0 MGPU_P2P=0 timeout 300 $TR --master-port 29513 tests/mgpu_parity.py > gpurun_out/r2_03_mgpu2_nccl.log 2>&1; echo "rc=$?" >> gpurun_out/r2_03_mgpu2_nccl.log
grep -E "mgpu|rc=" gpurun_out/r2_03_mgpu2_nccl.log | tail -8
BENCH_RANK_TIMERS=1 timeout 400 $TR --master-port 29514 bench.py --gpus 2 --steps 10 --warmup 3 --no-cpu > gpurun_out/r2_03_bench2_smag.json 2> gpurun_out/r2_03_bench2_smag.err; echo "rc=$?"
python tools/sumbench.py gpurun_out/r2_03_bench2_smag.json
grep "rank" gpurun_out/r2_03_bench2_smag.err | tail -4
timeout 400 $TR --master-port 29515 bench.py --gpus 2 --steps 10 --warmup 3 --no-cpu --no-parity --param fast_barrier=0 > gpurun_out/r2_03_bench2_smag_ncclbar.json 2> gpurun_out/r2_03_bench2_smag_ncclbar.err; echo "rc=$?"
python tools/sumbench.py gpurun_out/r2_03_bench2_smag_ncclbar.json
timeout 400 $TR --master-port 29516 bench.py --gpus 2 --steps 10 --warmup 3 --no-cpu --no-parity --param fft_pipe_min_ranks=2 > gpurun_out/r2_03_bench2_smag_pipe.json 2> gpurun_out/r2_03_bench2_smag_pipe.err; echo "rc=$?"
python tools/sumbench.py gpurun_out/r2_03_bench2_smag_pipe.json
timeout 400 $TR --master-port 29517 bench.py --gpus 2 --steps 10 --warmup 3 --no-cpu --workload tgv3d > gpurun_out/r2_03_bench2_tgv.json 2> gpurun_out/r2_03_bench2_tgv.err; echo "rc=$?"
python tools/sumbench.py gpurun_out/r2_03_bench2_tgv.json
tail -3 gpurun_out/r2_03_bench2_smag.err
