#!/bin/bash
# One perf iteration on a fresh box: small parity triage (own process, bounded), selected pytest cases, per-layer timings with role waits.
set -u
mkdir -p gpurun_out
for arch in ${ARCHS:-johnson}; do
  timeout 180 python - "$arch" <<'PY' 2>&1 | tail -8
import sys, os
sys.path.insert(0, os.getcwd())
import torch
import neuralstyletransferv1_b200 as m
from neuralstyletransferv1_b200 import synth
from oracle import nst_oracle as O
arch = sys.argv[1]
cls = {"johnson": m.TransformerNet, "nst": m.TransformerNetNST, "reconet": m.ReCoNet}[arch]
preset = {"johnson": "raw_255", "nst": "raw_01", "reconet": "tanh"}[arch]
sd = synth.make_state_dict(arch, 0)
model = cls().to("cuda:0"); model.load_state_dict(sd); model.eval()
for (n, h, w) in ((2, 64, 96), (3, 120, 200), (1, 272, 480)):
    frames = synth.synthetic_frames_u8(n, h, w, seed=3)
    out = model.stylize_u8(frames.cuda(), preset).cpu()
    torch.cuda.synchronize()
    ref = O.stylize_u8(arch, sd, frames, preset)
    print(arch, (n, h, w), "max|d|, PSNR =", O.u8_metrics(out, ref))
PY
done
if [ -n "${PYTEST_K:-}" ]; then
  timeout 900 python -m pytest tests -m gpu -q -x --tb=short -p no:cacheprovider -k "$PYTEST_K" 2>&1 | tail -15
fi
for a in ${ARCHS:-johnson}; do
  NST_TC_PROF=1 timeout 300 python tools/tcprof.py $a 8 > gpurun_out/tcprof_$a.log 2>&1; grep tcprof gpurun_out/tcprof_$a.log | cut -c1-400
  NST_TC_PROF=0 NST_PROF_REPS=5 timeout 300 python tools/tcprof.py $a 8 2>&1 | grep -v tcprof > gpurun_out/layers_$a.log; cat gpurun_out/layers_$a.log
done
if [ -n "${BENCH:-}" ]; then
  timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; cut -c1-600 gpurun_out/bench.json
fi
