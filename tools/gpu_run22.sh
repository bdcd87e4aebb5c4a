cd $GRAFT_REPO_ROOT
timeout 100 python -m pytest tests/test_gpu_zline.py -x -q 2>&1 | tail -3
timeout 60 python bench.py --steps 5 --warmup 3 --no-cpu --no-parity --workload tgv3d --param zline=1 2>/dev/null | python -c "import sys,json; d=json.loads([l for l in sys.stdin if l.startswith('{')][-1]); print(d['ms_per_step'], {k:round(v['ms_per_step'],3) for k,v in d['kernels'].items() if 'poisson' in k})"
