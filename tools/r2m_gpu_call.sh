cd $GRAFT_REPO_ROOT
V=$GRAFT_REPO_ROOT/rl-cattle-herding_b200/variants
for v in e0 e2 e4 e6; do
  CHS_LIB=$V/libchs_$v.so python tools/ab_bench.py head c3 mid 2>&1 | grep -v Warning
done > gpurun_out/r2m_ab.txt
CHS_LIB=$V/libchs_e2.so python tools/phase_timing.py 1024 4 256 8 100 > gpurun_out/r2m_phase_c3_e2.txt 2>&1
CHS_LIB=$V/libchs_e2.so python tools/phase_timing.py 4096 16 256 > gpurun_out/r2m_phase_4096_e2.txt 2>&1
NCU="ncu --set full --clock-control none --import-source on --warp-sampling-interval 0"
CHS_LIB=$V/libchs_e0.so timeout 600 $NCU -k regex:step_kernel --launch-skip 60 --launch-count 1 -f -o gpurun_out/prof_r2m_c3_e0 python tools/ncu_target.py 1024 8 100 fused 4 256 > gpurun_out/ncu_r2m_a.log 2>&1
CHS_LIB=$V/libchs_e0.so timeout 600 $NCU -k regex:step_kernel --launch-skip 60 --launch-count 1 -f -o gpurun_out/prof_r2m_4096_e0 python tools/ncu_target.py 4096 3 20 fused 16 256 > gpurun_out/ncu_r2m_b.log 2>&1
cat gpurun_out/r2m_ab.txt; grep -E "p2 |sum" gpurun_out/r2m_phase_c3_e2.txt gpurun_out/r2m_phase_4096_e2.txt; ls -la gpurun_out/*.ncu-rep
