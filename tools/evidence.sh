#!/bin/bash
# Round evidence in one call: benches (N=1), steady-state role profile, ncu launch list of the bench command, ncu --set full of the
# dominant kernels.  Every ncu run follows an un-profiled run of the same command that exited 0.
set -u
R=${ROUND:-r02}
mkdir -p gpurun_out
[ -n "${SKIP_BENCH:-}" ] || tools/bench_all.sh 1 ${CFGS:-c2 nst c3_4k c4_reconet300 c5_region}
tools/prof2.sh johnson > gpurun_out/${R}_tcprof_johnson_8x1080p.txt 2>&1
tools/prof2.sh reconet > gpurun_out/${R}_tcprof_reconet_8x1080p.txt 2>&1
python bench.py --steps 1 --warmup 3 --frames 8 --no-cpu-baseline > gpurun_out/${R}_bench8.json 2> gpurun_out/${R}_bench8.err && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${R}_launches_bench8x1080p.csv \
   python bench.py --steps 1 --warmup 3 --frames 8 --no-cpu-baseline > gpurun_out/${R}_ncu_launches.log 2>&1
echo "launch list rc=$?"
python tools/one_forward.py johnson 8 && SKIPS="${SKIPS:-4 5 1 0 14 15}" NAME=${R} bash tools/ncu_layer.sh
