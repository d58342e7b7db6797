#!/bin/bash
# Which part of the error compensation of ReCoNet's encoder matters: 3-term MMA, or just fp32 storage of the early activations?
for env in "NST_AB_NOOP=1" "NST_X3_LAYERS=0" "NST_X3_LAYERS=0 NST_ACT16=0" "NST_X3_LAYERS=1" "NST_X3_LAYERS=3" "NST_X3_LAYERS=1 NST_ACT16=0"; do
  for sz in "272 480" "540 960" "1080 1920"; do
    echo "--- $env $sz"; env $env python tools/parity_margin.py $sz 2>&1 | grep reconet
  done
done
