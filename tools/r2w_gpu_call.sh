cd $GRAFT_REPO_ROOT
export CHS_LIB=$GRAFT_REPO_ROOT/rl-cattle-herding_b200/variants/libchs_prof.so
for cfg in "4096 16 256 3 20" "1024 4 256 8 100"; do
  echo "=== lean stamps, L2 flush: $cfg"; python tools/phase_timing.py $cfg 2>&1 | grep -v Warn
  echo "=== lean stamps, no flush: $cfg"; NO_FLUSH=1 python tools/phase_timing.py $cfg 2>&1 | grep -v Warn
done > gpurun_out/r2w_phase_lean.txt
cat gpurun_out/r2w_phase_lean.txt
