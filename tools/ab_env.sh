#!/bin/bash
# A/B of one build with and without an environment knob: tools/ab_env.sh NST_NO_RES_RING=1 "johnson reconet"
for r in 1 2; do for a in ${2:-johnson}; do for e in "NST_AB_NOOP=1" "$1"; do
  echo "--- $a $e (round $r)"
  env $e NST_TC_PROF=0 tools/quick.sh $a | awk '{printf "%s %s | ", $1, $2} END {print ""}'
done; done; done
