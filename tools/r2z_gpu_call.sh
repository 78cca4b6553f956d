cd $GRAFT_REPO_ROOT
V=$GRAFT_REPO_ROOT/rl-cattle-herding_b200/variants
timeout 600 python -m pytest tests/test_cuda_contract.py -m gpu -x -q -k knobs 2>&1 | tail -3
for v in x0 x16 x32 x48; do
for cfg in "4096 16 256 3 20" "1024 4 256 8 100"; do
  echo "=== $v (16: workers skip herd+winding, 32: workers skip cattle pieces), no flush: $cfg"; CHS_LIB=$V/libchs_$v.so NO_FLUSH=1 python tools/phase_timing.py $cfg 2>&1 | grep -E "p1 thread0|p2 thread0|p2 detail|p2 neigh|p3 reduce|p3 detail: bar|sum"
done; done > gpurun_out/r2z_phase.txt
cat gpurun_out/r2z_phase.txt
