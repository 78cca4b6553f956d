cd $GRAFT_REPO_ROOT
V=$GRAFT_REPO_ROOT/rl-cattle-herding_b200/variants
export CHS_LIB=$V/libchs_h1.so
timeout 900 python -m pytest tests/test_cuda_env.py tests/test_cuda_adapters.py tests/test_cuda_contract.py tests/test_cuda_marl.py -m gpu -x -q 2>&1 | tail -4
python tools/host_path_times.py 4096 2>&1 | grep -v Warn | tail -5
CHS_HOST_COPY=1 python tools/host_path_times.py 4096 2>&1 | grep -v Warn | tail -5
CHS_NO_DEAD_SKIP=1 python tools/host_path_times.py 4096 2>&1 | grep -v Warn | tail -5
python tools/r2_probe.py host 2>&1 | tail -3
NCU="ncu --set full --clock-control none --import-source on"
unset CHS_LIB
timeout 900 $NCU -k regex:chs:: --launch-skip 183 --launch-count 3 -f -o gpurun_out/prof_r2s_split_262144 python tools/ncu_target.py 262144 3 20 split 32 256 > gpurun_out/ncu_s_c.log 2>&1; tail -3 gpurun_out/ncu_s_c.log
