set -x
cd $GRAFT_REPO_ROOT
timeout 400 python -m pytest tests/test_gpu_zline.py -x -q 2>&1 | tail -5 > gpurun_out/r2_18_pytest_zline.log; cat gpurun_out/r2_18_pytest_zline.log
timeout 200 python bench.py --steps 5 --warmup 3 --no-cpu --no-parity --workload tgv3d --param zline=1 > gpurun_out/r2_18_bench1_zline1.json 2> gpurun_out/r2_18_bench1_zline1.err; echo "bench rc=$?"
python tools/sumbench.py gpurun_out/r2_18_bench1_zline1.json 2>/dev/null | head -9
timeout 300 ncu --set full --import-source on --clock-control none -k regex:k_zline --launch-skip 6 -c 3 -o gpurun_out/r2_18_zline python bench.py --steps 1 --warmup 1 --no-cpu --no-parity --workload tgv3d --param zline=1 > gpurun_out/r2_18_ncu.log 2>&1; echo "ncu rc=$?"
ncu -i gpurun_out/r2_18_zline.ncu-rep --page raw --csv > gpurun_out/r2_18_zline_raw.csv 2>/dev/null
ls -la gpurun_out/r2_18_zline*
