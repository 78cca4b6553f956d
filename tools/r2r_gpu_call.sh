cd $GRAFT_REPO_ROOT
V=$GRAFT_REPO_ROOT/rl-cattle-herding_b200/variants
( python tools/ab_bench.py head c3 mid c4 big 2>&1 | grep -v Warning | sed 's/^/d1   /'
  CHS_LIB=$V/libchs_d2.so python tools/ab_bench.py head c3 mid c4 big 2>&1 | grep -v Warning | sed 's/^/d2   /'
) > gpurun_out/r2r_ab.txt
cat gpurun_out/r2r_ab.txt
CHS_LIB=$V/libchs_d2.so python tools/kernel_times.py 262144 split > gpurun_out/r2r_ktimes.txt 2>&1; tail -8 gpurun_out/r2r_ktimes.txt
CHS_LIB=$V/libchs_d2.so python tools/kernel_times.py 65536 fused >> gpurun_out/r2r_ktimes.txt 2>&1; tail -4 gpurun_out/r2r_ktimes.txt
CHS_LIB=$V/libchs_d2.so timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
