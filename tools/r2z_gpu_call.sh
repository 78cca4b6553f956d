# GPU box: the ncu evidence of the end-of-round kernels (profiles/r2z_*): launch list of the bench command, then one
# `--set full` capture each of the fused headline kernel, config 3 and the two split-mode kernels at 65 536 / 262 144 envs
set -x
cd $GRAFT_REPO_ROOT
python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-sweep --no-configs > gpurun_out/r2z_b.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/r2z_launches_4096env.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-sweep --no-configs > gpurun_out/r2z_ncu_list.log 2>&1
NCU="ncu --set full --clock-control none --import-source on"
timeout 600 $NCU -k regex:step_kernel --launch-skip 60 --launch-count 2 -f -o gpurun_out/prof_r2z_4096_fused python tools/ncu_target.py 4096 3 20 fused 28 512 > gpurun_out/ncu_z_a.log 2>&1
timeout 600 $NCU -k regex:step_kernel --launch-skip 60 --launch-count 2 -f -o gpurun_out/prof_r2z_c3 python tools/ncu_target.py 1024 8 100 fused 4 256 > gpurun_out/ncu_z_b.log 2>&1
timeout 900 $NCU -k regex:drone_kernel\|step_kernel --launch-skip 120 --launch-count 2 -f -o gpurun_out/prof_r2z_split_262144 python tools/ncu_target.py 262144 3 20 split 32 256 > gpurun_out/ncu_z_c.log 2>&1
timeout 900 $NCU -k regex:drone_kernel\|step_kernel --launch-skip 120 --launch-count 2 -f -o gpurun_out/prof_r2z_split_65536 python tools/ncu_target.py 65536 3 20 split 32 256 > gpurun_out/ncu_z_d.log 2>&1
ls -la gpurun_out/*.ncu-rep
