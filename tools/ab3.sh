#!/bin/bash
# A/B (build/lib_base.so vs build/lib_new.so) of the per-frame sum of layer times for the three architectures
for a in johnson nst reconet; do for v in base new; do
  echo -n "$a $v: "; NST_LIB_PATH=$PWD/build/lib_$v.so NST_TC_PROF=0 REPS=3 tools/quick.sh $a | tail -1
done; done
