cd $GRAFT_REPO_ROOT
python -m pytest tests -m gpu -x -q --timeout 900 2>&1 | tail -12 > gpurun_out/r2k_tests.log
cat gpurun_out/r2k_tests.log
