cd $GRAFT_REPO_ROOT
python -m pytest tests -m gpu -x -q --timeout 900 2>&1 | tail -6 > gpurun_out/r2o_tests.log
python bench.py > gpurun_out/r2o_bench.json 2> gpurun_out/r2o_bench.err
cat gpurun_out/r2o_tests.log; head -c 300 gpurun_out/r2o_bench.json
