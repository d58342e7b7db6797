"""One warm-up + one stylisation of F frames (ncu target): python tools/one_forward.py [arch] [F] [H] [W]"""
import os, sys
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
for _ in range(int(os.environ.get("NST_FORWARDS", "2"))):
    model.stylize_u8(frames, preset)
torch.cuda.synchronize()
