"""Print max|d| / PSNR of the CUDA path vs the CPU oracle for each architecture (margin against 2 / 45 dB)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import neuralstyletransferv1_b200 as m
from neuralstyletransferv1_b200 import synth
from oracle import nst_oracle as O

H, W = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (272, 480)
for arch, cls, preset in (("johnson", m.TransformerNet, "raw_255"), ("nst", m.TransformerNetNST, "raw_01"), ("reconet", m.ReCoNet, "tanh")):
    for seed in (0, 1):
        sd = synth.make_state_dict(arch, seed)
        model = cls().to("cuda:0"); model.load_state_dict(sd); model.eval()
        frames = synth.synthetic_frames_u8(2, H, W, seed=5 + seed)
        out = model.stylize_u8(frames.cuda(), preset).cpu()
        ref = O.stylize_u8(arch, sd, frames, preset)
        mx, psnr = O.u8_metrics(out, ref)
        print(f"{arch:8s} seed {seed} {H}x{W}: max|d|={mx} PSNR={psnr:.2f} dB")
        d = (out.int() - ref.int()).abs()
        if mx >= 3:
            idx = (d >= 3).nonzero()
            print("   |d|>=3 count", idx.shape[0], "of", d.numel(), "first:", idx[:6].tolist())
