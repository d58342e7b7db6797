"""Per-layer role/wait breakdown of conv_tc_kernel (debug): NST_TC_PROF=1 python tools/tcprof.py [arch] [frames] [H] [W]"""
import os, sys
os.environ.setdefault("NST_TC_PROF", "1")
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import neuralstyletransferv1_b200 as m
from neuralstyletransferv1_b200 import synth

arch = sys.argv[1] if len(sys.argv) > 1 else "johnson"
F = int(sys.argv[2]) if len(sys.argv) > 2 else 8
H = int(sys.argv[3]) if len(sys.argv) > 3 else 1080
W = int(sys.argv[4]) if len(sys.argv) > 4 else 1920
cls = {"johnson": m.TransformerNet, "nst": m.TransformerNetNST, "reconet": m.ReCoNet}[arch]
model = cls().to("cuda:0")
model.load_state_dict(synth.make_state_dict(arch, 0))
frames = synth.synthetic_frames_u8(min(F, 2), H, W, seed=1).repeat((F + 1) // 2, 1, 1, 1)[:F].cuda()
preset = {"johnson": "raw_255", "nst": "raw_01", "reconet": "tanh"}[arch]
model.stylize_u8(frames, preset)
torch.cuda.synchronize()
R = int(os.environ.get("NST_PROF_REPS", "1"))
runs = [model.engine.profile_layers(frames) for _ in range(R)]
best = runs[0]
for r in runs[1:]:
    for d, e in zip(best, r):
        d["ms"] = min(d["ms"], e["ms"])
for d in best:
    print(f"{d['name']:28s} {d['ms'] / F:8.4f} ms/frame  {d['flops'] * F / d['ms'] / 1e9:8.1f} TFLOP/s  {d['bytes'] * F / d['ms'] / 1e6:8.1f} GB/s")
