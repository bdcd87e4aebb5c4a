set -x
cd $GRAFT_REPO_ROOT
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
MGPU_BIG=1 MGPU_CASES=0,1,2,4 timeout 240 $TR --master-port 29511 tests/mgpu_parity.py > gpurun_out/r2_08_mgpu2.log 2>&1 || { echo "PARITY FAILED rc=$?"; tail -30 gpurun_out/r2_08_mgpu2.log; exit 1; }
grep -E "mgpu" gpurun_out/r2_08_mgpu2.log
MGPU_BIG=1 MGPU_CASES=4 timeout 240 $TR --master-port 29518 tests/mgpu_parity.py > gpurun_out/r2_08_mgpu2b.log 2>&1 || { echo "PARITY(b) FAILED rc=$?"; tail -30 gpurun_out/r2_08_mgpu2b.log; exit 1; }
BENCH_PARITY_SMALL=1 BENCH_RANK_TIMERS=1 timeout 300 $TR --master-port 29514 bench.py --gpus 2 --steps 10 --warmup 3 --no-cpu --also-tgv > gpurun_out/r2_08_bench2.json 2> gpurun_out/r2_08_bench2.err || { echo "FAILED rc=$?"; tail -20 gpurun_out/r2_08_bench2.err; exit 1; }
python tools/sumbench.py gpurun_out/r2_08_bench2.json
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/r2_08_bench2.json') if l.startswith('{')][-1])
print(json.dumps(d['parity_check'])); c5=d['config5']; print('config5', c5['ms_per_step'], c5['step_roofline']['frac'], c5['parity_check']['ok'], c5['nvlink']['achieved'])
PY
timeout 200 $TR --master-port 29512 tools/sweep_mgpu.py "transpose_ce=1" "transpose_ce=0" "transpose_ce=0,scatter_blocks=64" > gpurun_out/r2_08_sweep2.log 2>&1; grep sweep gpurun_out/r2_08_sweep2.log
