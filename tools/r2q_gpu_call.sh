cd $GRAFT_REPO_ROOT
V=$GRAFT_REPO_ROOT/rl-cattle-herding_b200/variants
( python tools/ab_bench.py head c3 mid c4 big 2>&1 | grep -v Warning | sed 's/^/base /'
  CHS_LIB=$V/libchs_d1.so python tools/ab_bench.py head c3 mid c4 big 2>&1 | grep -v Warning | sed 's/^/d1   /'
) > gpurun_out/r2q_ab.txt
cat gpurun_out/r2q_ab.txt
python tools/host_path_times.py 4096 > gpurun_out/r2q_host.txt 2>&1; tail -30 gpurun_out/r2q_host.txt
