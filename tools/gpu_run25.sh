cd $GRAFT_REPO_ROOT
timeout 14 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 16 python -m pytest tests/test_gpu_fast.py -x -q 2>&1 | tail -2
