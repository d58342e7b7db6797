#!/bin/bash
# A/B two builds of libb200nst.so in ONE gpurun call (box-to-box and clock noise is ~5-10 %): build/lib_base.so vs build/lib_new.so
for r in 1 2; do for v in base new; do
  echo "--- $v (round $r)"
  NST_LIB_PATH=$PWD/build/lib_$v.so NST_TC_PROF=0 tools/quick.sh ${1:-johnson} | awk '{printf "%s %s | ", $1, $2} END {print ""}'
done; done
