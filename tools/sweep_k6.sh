#!/bin/bash
# sweep K6 configs: usage sweep_k6.sh <nl> <m> cfg1 cfg2 ...
nl=$1; m=$2; shift 2
for cfg in "$@"; do
  echo -n "cfg=$cfg  "
  SLAM_K6_CFG=$cfg timeout 120 python tools/probe_update.py $nl $m 20 2>&1 | tail -1 | sed 's/.*ms)  //'
done
