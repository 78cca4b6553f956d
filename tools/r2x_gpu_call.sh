cd $GRAFT_REPO_ROOT
V=$GRAFT_REPO_ROOT/rl-cattle-herding_b200/variants
for cfg in "4096 16 256 3 20" "1024 4 256 8 100"; do
  echo "=== prof2 lean stamps, no flush: $cfg"; CHS_LIB=$V/libchs_prof2.so NO_FLUSH=1 python tools/phase_timing.py $cfg 2>&1 | grep -E "p2 |sum|p3 reduce|p3 detail|CTA total"
done > gpurun_out/r2x_phase.txt
cat gpurun_out/r2x_phase.txt
( CHS_LIB=$V/libchs_prof.so python tools/ab_bench.py head c3 2>&1 | grep -v Warning | sed 's/^/prof  /'
  CHS_LIB=$V/libchs_prof2.so python tools/ab_bench.py head c3 mid 2>&1 | grep -v Warning | sed 's/^/prof2 /' ) > gpurun_out/r2x_ab.txt
cat gpurun_out/r2x_ab.txt
