cd $GRAFT_REPO_ROOT
V=$GRAFT_REPO_ROOT/rl-cattle-herding_b200/variants
for cfg in "4096 16 256 3 20" "1024 4 256 8 100"; do
  CHS_LIB=$V/libchs_y1.so NO_FLUSH=1 python tools/phase_timing_exp.py $cfg 2>&1 | grep -E "p2 detail|p2 neigh|ranking loop stamps|since stamp 4"
done
