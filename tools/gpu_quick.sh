#!/bin/bash
# Fast triage on a fresh box: the same small stylisation with the operand gather, then TMA-fed, each in its own process
# (a kernel fault poisons the CUDA context), for all three generator families.
mkdir -p gpurun_out
for env in "NST_NO_TMA=1" "NST_NO_TMA="; do
  for arch in johnson nst reconet; do
    echo "=== $env $arch"
    env $env timeout 120 python - "$arch" <<'PY' 2>&1 | tail -6
import sys, os
if not os.environ.get("NST_NO_TMA"): os.environ.pop("NST_NO_TMA", None)
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
for (n, h, w) in ((2, 64, 96), (2, 120, 200), (1, 272, 480)):
    frames = synth.synthetic_frames_u8(n, h, w, seed=3)
    out = model.stylize_u8(frames.cuda(), preset).cpu()
    torch.cuda.synchronize()
    ref = O.stylize_u8(arch, sd, frames, preset)
    print(arch, (n, h, w), "max|d|, PSNR =", O.u8_metrics(out, ref), [d["feed"] for d in model.engine.layer_info()].count("tma"), "tma layers")
PY
  done
done
