set -x
cd $GRAFT_REPO_ROOT
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
MGPU_BIG=1 timeout 300 $TR --master-port 29511 tests/mgpu_parity.py > gpurun_out/r2_17_mgpu2.log 2>&1; echo "parity rc=$?"
grep -E "mgpu|Error|error" gpurun_out/r2_17_mgpu2.log | head -12
MGPU_BIG=0 MGPU_ZLCHUNKS=2 timeout 200 $TR --master-port 29512 tests/mgpu_parity.py > gpurun_out/r2_17_mgpu2_chunks.log 2>&1; echo "parity chunks rc=$?"
grep -E "mgpu|Error|error" gpurun_out/r2_17_mgpu2_chunks.log | head -12
BENCH_RANK_TIMERS=1 timeout 300 $TR --master-port 29514 bench.py --gpus 2 --steps 10 --warmup 3 --no-cpu --also-tgv > gpurun_out/r2_17_bench2.json 2> gpurun_out/r2_17_bench2.err; echo "bench rc=$?"
python tools/sumbench.py gpurun_out/r2_17_bench2.json 2>/dev/null | head -14
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/r2_17_bench2.json') if l.startswith('{')][-1])
c5=d['config5']; print('tgv2 ms/step', c5['ms_per_step'], 'frac', c5['step_roofline']['frac'], 'parity', d['parity_check']['ok'], c5['parity_check']['ok'])
print({k: round(v['ms_per_step'],3) for k,v in c5['kernels'].items()})
PY
tail -5 gpurun_out/r2_17_bench2.err
