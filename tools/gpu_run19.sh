set -x
cd $GRAFT_REPO_ROOT
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
MGPU_BIG=1 MGPU_CASES=0,4 timeout 150 $TR --master-port 29511 tests/mgpu_parity.py > gpurun_out/r2_19_mgpu8.log 2>&1; echo "parity rc=$?"
grep -E "mgpu|Error|error" gpurun_out/r2_19_mgpu8.log | head
timeout 120 $TR --master-port 29512 tools/sweep_mgpu.py --steps=4 "zline=1" "zline_chunks=1" "zline_chunks=4" > gpurun_out/r2_19_sweep8.log 2>&1; echo "sweep rc=$?"
grep -E "sweep|Error|error" gpurun_out/r2_19_sweep8.log | head -12
BENCH_PARITY_SMALL=1 BENCH_RANK_TIMERS=1 timeout 240 $TR --master-port 29514 bench.py --gpus 8 --steps 10 --warmup 3 --no-cpu > gpurun_out/r2_19_bench8.json 2> gpurun_out/r2_19_bench8.err; echo "bench rc=$?"
python tools/sumbench.py gpurun_out/r2_19_bench8.json 2>/dev/null | head -16
python - <<'PY'
import json
try:
    d=json.loads([l for l in open('gpurun_out/r2_19_bench8.json') if l.startswith('{')][-1])
    print(json.dumps(d['parity_check'])); print(json.dumps(d['nvlink']))
    c5=d['config5']; print('config5 ms/step', c5['ms_per_step'], 'frac', c5['step_roofline']['frac'], 'parity', c5['parity_check']['ok'])
    for k,v in sorted(c5['kernels'].items(), key=lambda kv:-kv[1]['ms_per_step'])[:12]: print('   ', k, round(v['ms_per_step'],3))
except Exception as e: print('ERR', e)
PY
grep "rank" gpurun_out/r2_19_bench8.err | cut -c1-600 | tail -16 > gpurun_out/r2_19_rank_timers.txt
tail -3 gpurun_out/r2_19_bench8.err | cut -c1-300
