#!/bin/bash
# steady-state role profile (no per-layer sync): last of 3 profiled forwards, then the per-layer times of the same build
a=${1:-johnson}
NST_TC_PROF=1 NST_PROF_REPS=3 timeout 300 python tools/tcprof.py $a 8 2> gpurun_out/tcprof_$a.err > gpurun_out/tcprof_$a.out
n=$(grep -c tcprof gpurun_out/tcprof_$a.err); per=$((n / 3))
grep "tcprof\|tcspan" gpurun_out/tcprof_$a.err | tail -$((2*per)) | cut -c1-330
grep -v "tcspan\|tcprof" gpurun_out/tcprof_$a.out | awk '{printf "%s %s | ", $1, $2} END {print ""}'
