#!/bin/bash
# A/B/... of several builds in ONE gpurun call (box-to-box and clock noise is ~5-10 %): tools/abn.sh "base new x" [arch]
for r in 1 2; do for v in $1; do
  echo "--- $v (round $r)"
  NST_LIB_PATH=$PWD/build/lib_$v.so NST_TC_PROF=0 tools/quick.sh ${2:-johnson} | awk '{printf "%s %s | ", $1, $2} END {print ""}'
done; done
