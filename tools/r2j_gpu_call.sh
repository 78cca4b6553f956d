cd $GRAFT_REPO_ROOT
python -m pytest tests/test_telemetry.py tests/test_cuda_parity.py tests/test_cuda_env.py tests/test_cuda_marl.py -m gpu -x -q --timeout 900 2>&1 | tail -12 > gpurun_out/r2j_tests.log
python bench.py --steps 200 --warmup 20 --no-cpu-baseline > gpurun_out/r2j_bench.json 2> gpurun_out/r2j_bench.err
python tools/phase_timing.py 4096 16 256 > gpurun_out/r2j_phase_4096.txt 2>&1
python tools/phase_timing.py 1024 4 256 8 100 > gpurun_out/r2j_phase_c3.txt 2>&1
cat gpurun_out/r2j_tests.log gpurun_out/r2j_phase_4096.txt gpurun_out/r2j_phase_c3.txt
