# GPU box with N GPUs: PCIe / NUMA topology and what concurrent SM-issued and copy-engine writes to pinned host memory reach.
cd $GRAFT_REPO_ROOT
out=gpurun_out/topo_8gpu.txt
( nvidia-smi topo -m
  echo "--- nproc $(nproc); affinity $(python -c 'import os; print(sorted(os.sched_getaffinity(0)))')"
  lscpu | grep -i -E "numa|socket|model name|^CPU\(s\)"
  for d in /sys/bus/pci/devices/*; do if [ "$(cat $d/class 2>/dev/null)" = "0x030200" ]; then echo "$d numa_node $(cat $d/numa_node) link $(cat $d/current_link_speed 2>/dev/null) x$(cat $d/current_link_width 2>/dev/null)"; fi; done
  nvidia-smi --query-gpu=index,pci.bus_id --format=csv,noheader
  n=$(nvidia-smi -L | wc -l)
  echo "=== one GPU at a time (GPU 0)"; CUDA_VISIBLE_DEVICES=0 tools/microbench/host_write 4096 | grep "^host"
  echo "=== GPUs 0 and 1 together"
  for i in 0 1; do CUDA_VISIBLE_DEVICES=$i tools/microbench/host_write 16384 > /tmp/hw_$i.txt & done; wait
  for i in 0 1; do echo "GPU $i"; grep "^host" /tmp/hw_$i.txt | grep "mode [0167]"; done
  if [ $n -ge 8 ]; then
  echo "=== GPUs 0 and 4 together"
  for i in 0 4; do CUDA_VISIBLE_DEVICES=$i tools/microbench/host_write 16384 > /tmp/hw_$i.txt & done; wait
  for i in 0 4; do echo "GPU $i"; grep "^host" /tmp/hw_$i.txt | grep "mode [0167]"; done
  echo "=== all $n GPUs together"
  for i in $(seq 0 $((n-1))); do CUDA_VISIBLE_DEVICES=$i tools/microbench/host_write 16384 > /tmp/hw_$i.txt & done; wait
  for i in $(seq 0 $((n-1))); do echo "GPU $i"; grep "^host" /tmp/hw_$i.txt | grep "mode [0167]"; done
  fi
) > $out 2>&1
cat $out
