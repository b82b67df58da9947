#!/bin/bash
# usage: gpuretry.sh <script> <log>   -- run a gpurun job script, retrying while the pod answers busy (status=transient)
for i in $(seq 1 40); do
  bash "$1" > "$2" 2>&1
  if grep -q "status=transient" "$2"; then sleep 45; continue; fi
  break
done
