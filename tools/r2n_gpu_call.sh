cd $GRAFT_REPO_ROOT
V=$GRAFT_REPO_ROOT/rl-cattle-herding_b200/variants
( CHS_LIB=$V/libchs_p0.so python tools/ab_bench.py head c3 mid c4 big 2>&1 | grep -v Warning | sed 's/^/PDL   /'
  CHS_NO_PDL=1 CHS_LIB=$V/libchs_p0.so python tools/ab_bench.py head c3 mid c4 big 2>&1 | grep -v Warning | sed 's/^/noPDL /'
  CHS_LIB=$V/libchs_p8.so python tools/ab_bench.py head c3 mid 2>&1 | grep -v Warning | sed 's/^/PDL   /'
  CHS_NO_PDL=1 CHS_LIB=$V/libchs_p8.so python tools/ab_bench.py head c3 mid 2>&1 | grep -v Warning | sed 's/^/noPDL /'
) > gpurun_out/r2n_ab.txt
cat gpurun_out/r2n_ab.txt
CHS_LIB=$V/libchs_p8.so timeout 900 python -m pytest tests/test_cuda_env.py tests/test_cuda_shapes.py tests/test_cuda_parity.py tests/test_cuda_contract.py tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -5 > gpurun_out/r2n_tests.log
cat gpurun_out/r2n_tests.log
