"""Which ReCoNet layers need error-compensated operands for max|d| <= 2?  NST_X3_LAYERS ablation against the fp32 oracle."""
import os, sys, itertools
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import neuralstyletransferv1_b200 as m
from neuralstyletransferv1_b200 import synth
from oracle import nst_oracle as O
sizes = [(272, 480), (540, 960)] + ([(1080, 1920)] if os.environ.get("BIG") else [])
refs = {}
for seed in [int(x) for x in os.environ.get("SEEDS", "0,1").split(",")]:
    sd = synth.make_state_dict("reconet", seed)
    for (h, w) in sizes:
        f = synth.synthetic_frames_u8(2, h, w, seed=100 + seed)
        f[-1] = torch.randint(0, 256, (h, w, 3), dtype=torch.uint8, generator=torch.Generator().manual_seed(seed + 101))
        refs[(seed, h, w)] = (sd, f, O.stylize_u8("reconet", sd, f, "tanh"))
for mask in sys.argv[1:] or ["0x7", "0x3", "0x5", "0x6", "0x1", "0x2", "0x4", "0x0"]:
    os.environ["NST_X3_LAYERS"] = mask
    worst, n3, n2, tot = 0, 0, 0, 0
    for (seed, h, w), (sd, f, ref) in refs.items():
        model = m.ReCoNet().to("cuda:0"); model.load_state_dict(sd); model.eval()
        out = model.stylize_u8(f.cuda(), "tanh").cpu()
        d = (out.int() - ref.int()).abs()
        worst = max(worst, int(d.max())); n3 += int((d >= 3).sum()); n2 += int((d >= 2).sum()); tot += d.numel()
    print(f"NST_X3_LAYERS={mask}: max|d| {worst}, values with |d| >= 3: {n3}, |d| >= 2: {n2} of {tot} ({1e6 * n2 / tot:.1f} ppm)")
