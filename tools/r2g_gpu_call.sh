set -x
cd $GRAFT_REPO_ROOT
python bench.py --impl reference --steps 3 --warmup 1 --cpu-seconds 8 > gpurun_out/r2g_ref_arm.json 2> gpurun_out/r2g_ref_arm.err
for tool in memcheck racecheck; do
  timeout 900 compute-sanitizer --tool $tool --print-limit 20 python tools/sanitize_run.py > gpurun_out/r2g_sanitizer_$tool.log 2>&1
done
timeout 600 compute-sanitizer --tool synccheck --print-limit 20 python tools/sanitize_run.py fused split > gpurun_out/r2g_sanitizer_synccheck.log 2>&1
NCU="ncu --set full --clock-control none --import-source on"
timeout 600 $NCU -k regex:step_kernel --launch-skip 60 --launch-count 2 -f -o gpurun_out/prof_r2g_4096_fused python tools/ncu_target.py 4096 3 20 fused 16 256 > gpurun_out/ncu_a.log 2>&1
timeout 600 $NCU -k regex:step_kernel --launch-skip 60 --launch-count 2 -f -o gpurun_out/prof_r2g_c3 python tools/ncu_target.py 1024 8 100 fused 4 256 > gpurun_out/ncu_b.log 2>&1
timeout 900 $NCU -k regex:chs:: --launch-skip 181 --launch-count 3 -f -o gpurun_out/prof_r2g_split_262144 python tools/ncu_target.py 262144 3 20 split 32 256 > gpurun_out/ncu_c.log 2>&1
ls -la gpurun_out/*.ncu-rep
tail -3 gpurun_out/r2g_sanitizer_*.log
tail -c 1500 gpurun_out/r2g_ref_arm.json
