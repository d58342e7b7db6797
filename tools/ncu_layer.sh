#!/bin/bash
# ncu --set full capture of chosen conv_tc_kernel launches of the SECOND forward (the first is warm-up), condensed on the box
# (the .ncu-rep files are ~12 MB each; gpurun returns at most 64 MiB): raw-page CSV + top stalled source lines per launch.
# usage: SKIPS="4 3" NAME=r02 [KEEP_REP=4] bash tools/ncu_layer.sh   (launch index inside a forward: 0 = conv1 ... 15 = deconv3)
set -u
mkdir -p gpurun_out
for s in ${SKIPS:-4}; do
  skip=$((16 + s))
  rep=gpurun_out/${NAME:-ncu}_l$s
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:conv_tc -s $skip -c 1 -f -o $rep \
     python tools/one_forward.py ${ARCH:-johnson} 8 > $rep.log 2>&1
  echo "layer $s rc=$?"
  ncu -i $rep.ncu-rep --page raw --csv > $rep.raw.csv 2>/dev/null
  ncu -i $rep.ncu-rep --page source --csv --print-source cuda,sass > /tmp/src_$s.csv 2>/dev/null
  python tools/ncu_lines.py /tmp/src_$s.csv 40 > $rep.top_lines.txt 2>&1
  [ "${KEEP_REP:-}" = "$s" ] || rm -f $rep.ncu-rep
done
