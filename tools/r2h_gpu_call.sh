cd $GRAFT_REPO_ROOT
python -m pytest tests/test_cuda_parity.py tests/test_cuda_shapes.py tests/test_cuda_env.py tests/test_cuda_marl.py tests/test_cuda_edge.py tests/test_ppo.py -m gpu -x -q --timeout 900 2>&1 | tail -6 > gpurun_out/r2h_tests.log
python tools/r2_probe.py c3 rot > gpurun_out/r2h_probe.txt 2>&1
python tools/host_path_times.py 4096 > gpurun_out/r2h_host.txt 2>&1
CHS_NO_DIRECT=1 python tools/host_path_times.py 4096 >> gpurun_out/r2h_host.txt 2>&1
cat gpurun_out/r2h_tests.log gpurun_out/r2h_probe.txt gpurun_out/r2h_host.txt
