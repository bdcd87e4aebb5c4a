set -x
cd $GRAFT_REPO_ROOT
python tools/bench_zsolve2.py 256 256 1024 > gpurun_out/r2_14_zsolve_1024.txt 2>&1; cat gpurun_out/r2_14_zsolve_1024.txt | tail -12
python tools/bench_zsolve2.py 512 512 512 > gpurun_out/r2_14_zsolve_512.txt 2>&1; grep -E "zcols=auto" gpurun_out/r2_14_zsolve_512.txt
for wl in dt2d_f64 tgv3d_f64 kolmo2d; do
  st=400; [ $wl = kolmo2d ] && st=100
  timeout 300 python bench.py --steps $st --warmup 10 --cpu-steps 3 --workload $wl > gpurun_out/r2_14_bench_$wl.json 2> gpurun_out/r2_14_bench_$wl.err; echo "$wl rc=$?"
  python tools/sumbench.py gpurun_out/r2_14_bench_$wl.json | head -6
  python - <<PY
import json
d=json.loads([l for l in open('gpurun_out/r2_14_bench_$wl.json') if l.startswith('{')][-1])
print('   cpu', d['cpu_baseline']['value'], d['cpu_baseline']['cores'], d['cpu_baseline']['ms_per_step'], 'gpu value', d['value'], 'step frac', d['step_roofline']['frac'], 'clocks', d['clocks'], 'parity', d['parity_check']['ok'], d['parity_check']['rel_l2_u'], 'graphs', d['graph_replays'])
PY
done
timeout 300 python bench.py --steps 20 --warmup 5 --workload tgv3d --no-cpu > gpurun_out/r2_14_bench_tgv3d.json 2> gpurun_out/r2_14_bench_tgv3d.err; python tools/sumbench.py gpurun_out/r2_14_bench_tgv3d.json | head -8
timeout 600 python -m pytest tests -m gpu -q -x > gpurun_out/r2_14_pytest_gpu.log 2>&1; echo "suite rc=$?"; tail -4 gpurun_out/r2_14_pytest_gpu.log
