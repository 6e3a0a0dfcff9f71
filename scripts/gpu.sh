#!/bin/bash
# usage: scripts/gpu.sh [--gpus N] <timeout-seconds> <script>   -- runs a script on the GPU box, retrying while the pod answers "busy"
G=""
if [ "$1" = "--gpus" ]; then G="--gpus $2"; shift 2; fi
for i in $(seq 1 20); do
  /usr/local/graft/bin/gpurun $G --timeout $1 -- "bash $2" > /tmp/gpu_last.log 2>&1
  rc=$?
  if grep -q "status=transient" /tmp/gpu_last.log; then sleep 120; continue; fi
  break
done
cat /tmp/gpu_last.log
