cd $GRAFT_REPO_ROOT
timeout 150 python tools/zline_long.py --workload tgv3d --blocks 12 --steps 10 "zline=0" "zline=1,zline_chunks=2" > gpurun_out/r2_20_long1.log 2>&1; echo "rc=$?"
grep -E "block|Error" gpurun_out/r2_20_long1.log | cut -c1-200
