set -x
cd $GRAFT_REPO_ROOT
timeout 300 python -m pytest tests/test_zz_no_alloc.py -x -q > gpurun_out/r2_10_noalloc_alone.log 2>&1; echo "alone rc=$?"; tail -5 gpurun_out/r2_10_noalloc_alone.log
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_10_pytest_gpu.log 2>&1; echo "suite rc=$?"; tail -8 gpurun_out/r2_10_pytest_gpu.log
timeout 300 python bench.py --steps 10 --warmup 3 --cpu-steps 1 > gpurun_out/r2_10_bench1.json 2> gpurun_out/r2_10_bench1.err; echo "bench rc=$?"; python tools/sumbench.py gpurun_out/r2_10_bench1.json; tail -3 gpurun_out/r2_10_bench1.err
