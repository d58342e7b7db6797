#!/bin/bash
# min-of-5 layer times under role-ablation bits (see tools/dbg_sweep.sh for the bit list)
for d in ${DBG_LIST:-0}; do
  echo "--- NST_TC_DBG=$d"
  NST_TC_DBG=$d NST_TC_PROF=0 tools/quick.sh ${1:-johnson} | awk '{printf "%s %s | ", $1, $2} END {print ""}'
done
