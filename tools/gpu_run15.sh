set -x
cd $GRAFT_REPO_ROOT
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1"
MGPU_BIG=1 MGPU_CASES=0,4,5 timeout 200 $TR --master-port 29511 tests/mgpu_parity.py > gpurun_out/r2_15_mgpu4.log 2>&1; echo "parity rc=$?"
grep -E "mgpu|Error" gpurun_out/r2_15_mgpu4.log | head
BENCH_PARITY_SMALL=1 BENCH_RANK_TIMERS=1 timeout 240 $TR --master-port 29514 bench.py --gpus 4 --steps 10 --warmup 3 --no-cpu --also-tgv > gpurun_out/r2_15_bench4.json 2> gpurun_out/r2_15_bench4.err; echo "bench rc=$?"
python tools/sumbench.py gpurun_out/r2_15_bench4.json | head -12
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/r2_15_bench4.json') if l.startswith('{')][-1])
c5=d['config5']; print('tgv4 ms/step', c5['ms_per_step'], 'frac', c5['step_roofline']['frac'], 'parity', d['parity_check']['ok'], c5['parity_check']['ok'])
PY
BENCH_PARITY_SMALL=1 timeout 200 $TR --master-port 29515 bench.py --gpus 4 --steps 200 --warmup 5 --no-cpu --workload kolmo2d > gpurun_out/r2_15_bench4_kolmo.json 2> gpurun_out/r2_15_bench4_kolmo.err; echo "kolmo rc=$?"
python tools/sumbench.py gpurun_out/r2_15_bench4_kolmo.json | head -5
