set -x
cd $GRAFT_REPO_ROOT
NCU="ncu --set full --clock-control none --import-source on"
timeout 900 $NCU -k regex:drone_kernel\|step_kernel --launch-skip 120 --launch-count 2 -f -o gpurun_out/prof_r2z_split_262144 python tools/ncu_target.py 262144 3 20 split 32 256 > gpurun_out/ncu_z_c.log 2>&1
timeout 900 $NCU -k regex:drone_kernel\|step_kernel --launch-skip 120 --launch-count 2 -f -o gpurun_out/prof_r2z_split_65536 python tools/ncu_target.py 65536 3 20 split 32 256 > gpurun_out/ncu_z_d.log 2>&1
ls -la gpurun_out/*.ncu-rep
