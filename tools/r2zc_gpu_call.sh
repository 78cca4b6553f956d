cd $GRAFT_REPO_ROOT
V=$GRAFT_REPO_ROOT/rl-cattle-herding_b200/variants
( python tools/ab_bench.py head c3 mid c4 big 2>&1 | grep -v Warning | sed 's/^/base /'
  CHS_LIB=$V/libchs_w5.so python tools/ab_bench.py head c3 mid c4 big 2>&1 | grep -v Warning | sed 's/^/w5   /' ) > gpurun_out/r2zc_ab.txt
cat gpurun_out/r2zc_ab.txt
CHS_LIB=$V/libchs_w5.so timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
