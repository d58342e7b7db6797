#!/bin/bash
# ncu --set full capture of chosen conv_tc_kernel launches of the SECOND forward (the first is warm-up).
# usage: SKIPS="4 3" NAME=r02 bash tools/ncu_layer.sh   (launch index inside a forward: 0 = conv1 ... 15 = deconv3)
set -u
mkdir -p gpurun_out
for s in ${SKIPS:-4}; do
  skip=$((16 + s))
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:conv_tc -s $skip -c 1 -f -o gpurun_out/${NAME:-ncu}_l$s \
     python tools/one_forward.py ${ARCH:-johnson} 8 > gpurun_out/${NAME:-ncu}_l$s.log 2>&1
  echo "layer $s rc=$?"; tail -2 gpurun_out/${NAME:-ncu}_l$s.log
done
