#!/bin/bash
# One GPU-box pass: parity tests, bench line, ncu launch list, ncu full capture of the residual conv kernel.
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
python bench.py --steps 3 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
cat gpurun_out/bench.json | cut -c1-1800
NST_TC_PROF=1 python tools/tcprof.py johnson 8 > gpurun_out/tcprof.log 2>&1; tail -16 gpurun_out/tcprof.log
PROF="python bench.py --frames 8 --steps 1 --warmup 3 --no-cpu-baseline"
$PROF > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $PROF > gpurun_out/ncu1.log 2>&1
echo "ncu launches rc=$?"
$PROF > gpurun_out/plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:conv_tc -s 52 -c 3 -f -o gpurun_out/prof $PROF > gpurun_out/ncu2.log 2>&1
echo "ncu full rc=$?"
