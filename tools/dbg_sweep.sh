#!/bin/bash
# Ablation of conv_tc_kernel roles. NST_TC_DBG bits: 2 producers idle (no loads/stores), 4 epilogue idle, 8 loader idle (no copies),
# 16 producers absent, 32 epilogue absent, 64 loader absent (the MMA warp then skips the matching waits),
# 128 epilogue: no global stores, 256 no statistics, 512 no tcgen05.ld; producers: 1024 no global loads, 2048 no smem stores, 4096 no side output
for d in ${DBG_LIST:-0 112 48 80 96 16 32 64}; do
  echo "=== NST_TC_DBG=$d"
  NST_TC_DBG=$d NST_TC_PROF=1 python tools/tcprof.py johnson 8 2>&1 | grep "${LAYERS:-res1.conv2\|deconv1\|conv3 }"
done
