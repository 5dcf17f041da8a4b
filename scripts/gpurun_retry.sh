#!/bin/bash
# usage: scripts/gpurun_retry.sh <timeout> <gpus> <script>   — retries while the pod answers "busy"
T=$1; G=$2; S=$3
for i in $(seq 1 30); do
  if [ "$G" = "1" ]; then OUT=$(/usr/local/graft/bin/gpurun --timeout $T -- bash $S 2>&1); else OUT=$(/usr/local/graft/bin/gpurun --gpus $G --timeout $T -- bash $S 2>&1); fi
  if echo "$OUT" | grep -q "status=transient\|retry in a few minutes"; then sleep 90; continue; fi
  echo "$OUT" | tail -40; exit 0
done
echo "gave up"; exit 3
