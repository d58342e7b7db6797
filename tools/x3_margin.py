"""Margin of two error-compensation masks of the ReCoNet encoder against the fp32 oracle over several weight seeds and frames:
python tools/x3_margin.py "7 3" H W  -> per mask: worst max|d|, how many values sit at |d| = 2, PSNR range."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import neuralstyletransferv1_b200 as m
from neuralstyletransferv1_b200 import synth
from oracle import nst_oracle as O

masks = sys.argv[1].split()
H, W = int(sys.argv[2]), int(sys.argv[3])
stats = {k: dict(mx=0, at2=0, at3=0, n=0, psnr=[]) for k in masks}
for seed in range(int(os.environ.get("SEEDS", "5"))):
    sd = synth.make_state_dict("reconet", seed)
    frames = synth.synthetic_frames_u8(2, H, W, seed=40 + seed)
    frames[1] = torch.randint(0, 256, frames[1].shape, dtype=torch.uint8, generator=torch.Generator().manual_seed(seed))   # a pure-noise frame
    ref = O.stylize_u8("reconet", sd, frames, "tanh")
    for k in masks:
        os.environ["NST_X3_LAYERS"] = k
        model = m.ReCoNet().to("cuda:0"); model.load_state_dict(sd); model.eval()
        out = model.stylize_u8(frames.cuda(), "tanh").cpu()
        d = (out.int() - ref.int()).abs()
        mx, psnr = O.u8_metrics(out, ref)
        s = stats[k]
        s["mx"] = max(s["mx"], mx); s["at2"] += int((d == 2).sum()); s["at3"] += int((d >= 3).sum()); s["n"] += d.numel(); s["psnr"].append(round(psnr, 2))
        del model
    print("seed", seed, {k: (stats[k]["mx"], stats[k]["at2"], stats[k]["at3"]) for k in masks}, flush=True)
for k in masks:
    s = stats[k]
    print(f"mask {k} {H}x{W}: worst max|d| {s['mx']}, |d|=2 on {s['at2']} of {s['n']} values ({s['at2'] / s['n'] * 1e6:.2f} ppm), |d|>=3 on {s['at3']}, PSNR {min(s['psnr'])}..{max(s['psnr'])}")
