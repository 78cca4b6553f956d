cd $GRAFT_REPO_ROOT
V=$GRAFT_REPO_ROOT/rl-cattle-herding_b200/variants
( python tools/ab_bench.py head c3 mid c4 big 2>&1 | grep -v Warning | sed 's/^/d2   /'
  CHS_LIB=$V/libchs_i1.so python tools/ab_bench.py head c3 mid c4 big 2>&1 | grep -v Warning | sed 's/^/i1   /'
) > gpurun_out/r2u_ab.txt
cat gpurun_out/r2u_ab.txt
CHS_LIB=$V/libchs_i1.so timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
NCU="ncu --set full --clock-control none --import-source on"
timeout 900 $NCU -k regex:"drone_kernel|step_kernel|reset_list_kernel" --launch-skip 183 --launch-count 3 -f -o gpurun_out/prof_r2s_split_262144 python tools/ncu_target.py 262144 3 20 split 32 256 > gpurun_out/ncu_s_c.log 2>&1; tail -3 gpurun_out/ncu_s_c.log
