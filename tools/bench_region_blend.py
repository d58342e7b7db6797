"""BASELINE config 5: four TransformerNet weight sets over masked 1080p frames, composited on the GPU.
Per frame: 4 stylisations (stylize_u8) + one composite kernel (uint8 styled frames, fp32 region masks).
Prints frames/s and the composite kernel's achieved HBM GB/s (algorithmic bytes: 4 u8 frames + original + 4 fp32 masks + output)."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import neuralstyletransferv1_b200 as m
from neuralstyletransferv1_b200 import synth

H, W, F, K = 1080, 1920, 8, 4
dev = torch.device("cuda", 0)
models = []
for k in range(K):
    mod = m.TransformerNet().to(dev); mod.load_state_dict(synth.make_state_dict("johnson", k)); mod.eval(); models.append(mod)
frames = synth.synthetic_frames_u8(2, H, W, seed=3).repeat(F // 2, 1, 1, 1).to(dev)
# soft 4-region masks (vertical bands with feathered edges), normalised to sum 1
xs = torch.linspace(0, 1, W, device=dev)[None, :].expand(H, W)
masks = torch.stack([torch.clamp(1 - (xs - (k + 0.5) / K).abs() * K, 0, 1) for k in range(K)]).contiguous()
masks = masks / masks.sum(0, keepdim=True).clamp_min(1e-6)
out = torch.empty_like(frames)

def step():
    styled = [mod.stylize_u8(frames, "raw_255") for mod in models]
    for i in range(F):
        out[i] = m.composite_regions([s[i] for s in styled], masks, list(range(K)), original=frames[i])

for _ in range(2): step()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(3): step()
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 3
styled = [mod.stylize_u8(frames, "raw_255") for mod in models]
c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
c0.record()
for i in range(F):
    out[i] = m.composite_regions([s[i] for s in styled], masks, list(range(K)), original=frames[i])
c1.record(); torch.cuda.synchronize()
cms = c0.elapsed_time(c1) / F
bytes_per_frame = (K + 2) * H * W * 3 + K * H * W * 4
print(json.dumps({"config": "C5 region_blend 4 x TransformerNet 1080p", "frames_per_s": F / (ms / 1e3), "ms_per_frame": ms / F,
                  "composite_ms_per_frame": cms, "composite_GBs": bytes_per_frame / (cms / 1e3) / 1e9}))
