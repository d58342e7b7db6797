"""Empirical tile-shape sweep: for each layer kind force (TH, MB) candidates via NST_TC_FORCE and time that layer (min of 3).
usage: python tools/tile_sweep.py [arch]   -> prints the table; the winners inform the cost model in nst_api.cu"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import neuralstyletransferv1_b200 as m
from neuralstyletransferv1_b200 import synth

arch = sys.argv[1] if len(sys.argv) > 1 else "johnson"
cls = {"johnson": m.TransformerNet, "nst": m.TransformerNetNST, "reconet": m.ReCoNet}[arch]
preset = {"johnson": "raw_255", "nst": "raw_01", "reconet": "tanh"}[arch]
frames = synth.synthetic_frames_u8(2, 1080, 1920, seed=1).repeat(4, 1, 1, 1).cuda()
sd = synth.make_state_dict(arch, 0)
LAYERS = {"johnson": [0, 1, 2, 4, 5, 13, 14, 15]}.get(arch, list(range(14)))

def run(force):
    os.environ["NST_TC_FORCE"] = force
    model = cls().to("cuda:0"); model.load_state_dict(sd); model.eval()
    try:
        model.stylize_u8(frames, preset)
        torch.cuda.synchronize()
        runs = [model.engine.profile_layers(frames) for _ in range(3)]
    except Exception as ex:
        return None
    return [min(r[i]["ms"] for r in runs) / 8 for i in range(len(runs[0]))], [d["name"] for d in runs[0]]

base, names = run("")
for li in LAYERS:
    best = (base[li], "auto")
    out = []
    for mb in (8, 4, 2, 1):
        for th in (4, 6, 8, 10, 12, 15, 16, 18, 20, 24, 30, 32):
            r = run(f"{li}:{th}:{mb}")
            if r is None or r[0][li] <= 0: continue
            t = r[0][li]
            if abs(t - base[li]) < 1e-9 and False: continue
            out.append((t, th, mb))
    out.sort()
    print(f"{names[li]:12s} auto {base[li]:.4f} | best " + "  ".join(f"TH{th}/MB{mb}:{t:.4f}" for t, th, mb in out[:5]), flush=True)
