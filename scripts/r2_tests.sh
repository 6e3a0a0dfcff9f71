cd $GRAFT_REPO_ROOT
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest_full.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest_full.log
tail -n 30 gpurun_out/r2_pytest_full.log | cut -c1-250
python __graft_entry__.py smoke 2>&1 | tail -3
