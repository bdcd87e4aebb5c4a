cd $GRAFT_REPO_ROOT
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 100 $TR --master-port 29521 tools/zline_long.py --workload tgv3d --blocks 30 --steps 20 "zline=1" > gpurun_out/r2_21_long2.log 2>&1; echo "rc=$?"
grep -E "block|Error" gpurun_out/r2_21_long2.log | grep -E "r0\]" | awk 'NR%3==0' | cut -c1-200
grep -c "nan\|inf" gpurun_out/r2_21_long2.log
