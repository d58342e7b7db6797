#!/bin/bash
# A/B of one build with and without a NST_TC_DBG bit (e.g. 65536 = unpipelined plain epilogue): tools/ab_dbg.sh 65536 "johnson nst reconet"
for r in 1 2; do for a in ${2:-johnson}; do for d in 0 $1; do
  echo "--- $a dbg=$d (round $r)"
  NST_TC_DBG=$d NST_TC_PROF=0 tools/quick.sh $a | awk '{printf "%s %s | ", $1, $2} END {print ""}'
done; done; done
