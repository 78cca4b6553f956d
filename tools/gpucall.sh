#!/bin/bash
# usage: tools/gpucall.sh <name> <timeout_s> <script>   -- runs `bash <script>` on the GPU box, retrying while the pod is busy
name=$1; to=$2; script=$3
for i in 1 2 3 4 5 6 7 8 9 10 11 12; do
  /usr/local/graft/bin/gpurun --timeout $to -- "bash $script" > gpurun_out/${name}_call.log 2>&1
  rc=$?
  if [ $rc -ne 3 ]; then exit $rc; fi
  sleep 90
done
exit 3
