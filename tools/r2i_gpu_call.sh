cd $GRAFT_REPO_ROOT
python -m pytest tests/test_cuda_contract.py tests/test_cuda_env.py tests/test_cuda_parity.py tests/test_cuda_shapes.py tests/test_cuda_adapters.py tests/test_zz_no_alloc.py -m gpu -x -q --timeout 900 2>&1 | tail -12 > gpurun_out/r2i_tests.log
python tools/r2_probe.py host big > gpurun_out/r2i_probe.txt 2>&1
python tools/host_path_times.py 4096 > gpurun_out/r2i_host.txt 2>&1
cat gpurun_out/r2i_tests.log gpurun_out/r2i_probe.txt gpurun_out/r2i_host.txt
