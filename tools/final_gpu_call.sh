# GPU box: what the driver runs at round end, in the same order -- gpu tests, smoke, both bench arms (short reference arm)
cd $GRAFT_REPO_ROOT
python -m pytest tests -m gpu -x -q --timeout 900 2>&1 | tail -4
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
python bench.py --impl reference --steps 3 --warmup 1 --cpu-seconds 8 > gpurun_out/final_ref_arm.json 2> gpurun_out/final_ref_arm.err; tail -c 400 gpurun_out/final_ref_arm.json; echo
python bench.py > gpurun_out/final_bench.json 2> gpurun_out/final_bench.err; head -c 300 gpurun_out/final_bench.json; echo
