set -x
cd $GRAFT_REPO_ROOT
timeout 300 python -m pytest tests/test_zz_no_alloc.py -x -q > gpurun_out/r2_09_noalloc_alone.log 2>&1; echo "alone rc=$?"; tail -5 gpurun_out/r2_09_noalloc_alone.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
BENCH_PARITY_SMALL=1 timeout 300 $TR --master-port 29514 bench.py --gpus 2 --steps 10 --warmup 3 --no-cpu --also-tgv > gpurun_out/r2_09_bench2.json 2> gpurun_out/r2_09_bench2.err || { echo "BENCH2 FAILED rc=$?"; tail -20 gpurun_out/r2_09_bench2.err; }
python tools/sumbench.py gpurun_out/r2_09_bench2.json
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/r2_09_bench2.json') if l.startswith('{')][-1])
print(json.dumps(d['parity_check'])); c5=d['config5']; print('config5', c5['ms_per_step'], c5['step_roofline']['frac'], c5['parity_check']['ok'], c5['nvlink']['achieved'])
PY
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_09_pytest_gpu.log 2>&1; echo "suite rc=$?"; tail -8 gpurun_out/r2_09_pytest_gpu.log
