cd $GRAFT_REPO_ROOT
V=$GRAFT_REPO_ROOT/rl-cattle-herding_b200/variants
for v in e0 e1; do
  CHS_LIB=$V/libchs_$v.so python tools/ab_bench.py head c3 mid c4 2>&1 | grep -v Warning
done > gpurun_out/r2l_ab.txt
CHS_LIB=$V/libchs_e0.so python tools/phase_timing.py 1024 4 256 8 100 > gpurun_out/r2l_phase_c3_e0.txt 2>&1
CHS_LIB=$V/libchs_e0.so NO_FLUSH=1 python tools/phase_timing.py 1024 4 256 8 100 > gpurun_out/r2l_phase_c3_e0_noflush.txt 2>&1
CHS_LIB=$V/libchs_e1.so python tools/phase_timing.py 1024 4 256 8 100 > gpurun_out/r2l_phase_c3_e1.txt 2>&1
CHS_LIB=$V/libchs_e1.so python tools/phase_timing.py 4096 16 256 > gpurun_out/r2l_phase_4096_e1.txt 2>&1
cat gpurun_out/r2l_ab.txt gpurun_out/r2l_phase_c3_e0.txt gpurun_out/r2l_phase_c3_e0_noflush.txt gpurun_out/r2l_phase_c3_e1.txt gpurun_out/r2l_phase_4096_e1.txt
