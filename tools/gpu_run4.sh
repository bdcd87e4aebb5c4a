set -x
cd $GRAFT_REPO_ROOT
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
BENCH_RANK_TIMERS=1 timeout 240 $TR --master-port 29514 bench.py --gpus 2 --steps 10 --warmup 3 --no-cpu > gpurun_out/r2_04_bench2_smag.json 2> gpurun_out/r2_04_bench2_smag.err || { echo "FAILED rc=$?"; tail -20 gpurun_out/r2_04_bench2_smag.err; exit 1; }
python tools/sumbench.py gpurun_out/r2_04_bench2_smag.json
grep "rank" gpurun_out/r2_04_bench2_smag.err | tail -4
timeout 150 $TR --master-port 29515 bench.py --gpus 2 --steps 10 --warmup 3 --no-cpu --no-parity --param fast_barrier=0 > gpurun_out/r2_04_bench2_smag_ncclbar.json 2> gpurun_out/r2_04_bench2_smag_ncclbar.err || { echo "FAILED rc=$?"; exit 1; }
python tools/sumbench.py gpurun_out/r2_04_bench2_smag_ncclbar.json
timeout 150 $TR --master-port 29516 bench.py --gpus 2 --steps 10 --warmup 3 --no-cpu --no-parity --param fft_pipe_min_ranks=2 > gpurun_out/r2_04_bench2_smag_pipe.json 2> gpurun_out/r2_04_bench2_smag_pipe.err || { echo "FAILED rc=$?"; exit 1; }
python tools/sumbench.py gpurun_out/r2_04_bench2_smag_pipe.json
timeout 200 $TR --master-port 29517 bench.py --gpus 2 --steps 10 --warmup 3 --no-cpu --workload tgv3d > gpurun_out/r2_04_bench2_tgv.json 2> gpurun_out/r2_04_bench2_tgv.err || { echo "FAILED rc=$?"; exit 1; }
python tools/sumbench.py gpurun_out/r2_04_bench2_tgv.json
