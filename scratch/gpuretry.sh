#!/bin/bash
# usage: gpuretry.sh <timeout> <command string>; retries while the pod answers transient (rc 3)
for i in $(seq 1 15); do
  /usr/local/graft/bin/gpurun --timeout "$1" -- "$2" > /tmp/w/gpuretry.out 2>&1
  rc=$?
  if ! grep -q "status=transient" /tmp/w/gpuretry.out; then break; fi
  sleep 150
done
tail -25 /tmp/w/gpuretry.out
