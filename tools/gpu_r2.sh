#!/bin/bash
# Round-2 GPU pass: full parity suite (all failures listed), per-layer timings with role waits, short bench line.
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -q --tb=short -p no:cacheprovider ${PYTEST_ARGS:-} > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -40 gpurun_out/pytest_gpu.log
for a in ${ARCHS:-johnson}; do
  NST_TC_PROF=1 timeout 300 python tools/tcprof.py $a 8 > gpurun_out/tcprof_$a.log 2>&1; tail -40 gpurun_out/tcprof_$a.log
  NST_TC_PROF=0 NST_PROF_REPS=5 timeout 300 python tools/tcprof.py $a 8 2>&1 | grep -v tcprof > gpurun_out/layers_$a.log; cat gpurun_out/layers_$a.log
done
timeout 600 python bench.py --steps 3 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
cut -c1-1500 gpurun_out/bench.json
