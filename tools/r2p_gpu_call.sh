cd $GRAFT_REPO_ROOT
export NCCL_DEBUG=WARN
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
( echo "== ctde graphs"; timeout 300 $RUN --master-port 29511 -m rl_cattle_herding_b200.train --mode ctde --envs 8192 --iters 4 2>&1 | grep -E "^\{|rror|Traceback" | cut -c1-700
  echo "== ctde eager"; timeout 300 $RUN --master-port 29512 -m rl_cattle_herding_b200.train --mode ctde --envs 8192 --iters 4 --no-graphs 2>&1 | grep -E "^\{|rror|Traceback" | cut -c1-700
  echo "== dtde graphs"; timeout 300 $RUN --master-port 29513 -m rl_cattle_herding_b200.train --mode dtde --envs 8192 --iters 3 2>&1 | grep -E "^\{|rror|Traceback" | cut -c1-700
  echo "== bench 2 gpus"; timeout 600 $RUN --master-port 29514 bench.py --gpus 2 --steps 200 --warmup 20 2>&1 | grep -E "^\{|rror|Traceback" | cut -c1-6000
) > gpurun_out/r2p_multi.txt 2>&1
cat gpurun_out/r2p_multi.txt | cut -c1-900
