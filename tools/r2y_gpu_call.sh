cd $GRAFT_REPO_ROOT
V=$GRAFT_REPO_ROOT/rl-cattle-herding_b200/variants
( python tools/ab_bench.py head c3 mid c4 big 2>&1 | grep -v Warning | sed 's/^/i1   /'
  CHS_LIB=$V/libchs_n1.so python tools/ab_bench.py head c3 mid c4 big 2>&1 | grep -v Warning | sed 's/^/n1   /' ) > gpurun_out/r2y_ab.txt
cat gpurun_out/r2y_ab.txt
for cfg in "4096 16 256 3 20" "1024 4 256 8 100"; do
  echo "=== prof3 lean stamps, no flush: $cfg"; CHS_LIB=$V/libchs_prof3.so NO_FLUSH=1 python tools/phase_timing.py $cfg 2>&1 | grep -E "p0 issue|p1 thread0|p2 |sum|p3 reduce|p3 reward|p3 stores|p5 reset|p3 detail|CTA total|global timer"
done > gpurun_out/r2y_phase.txt
cat gpurun_out/r2y_phase.txt
CHS_LIB=$V/libchs_n1.so timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
