#!/bin/bash
# per-layer timings (8 x 1080p) without the role profiler
NST_PROF_REPS=${REPS:-5} python tools/tcprof.py ${1:-johnson} 8 2>&1 | grep -v tcprof | awk '{print; s+=$2} END {print "sum ms/frame", s}'
