"""Achieved HBM bandwidth of the post-network / mask kernels at 1080p (CUDA events, 3 warm-ups + 20 timed calls each, inputs larger than
nothing fancy: each call streams its own operands once).  Algorithmic bytes per pixel are stated per line; peak = MEASURED_PEAKS.json
copy bandwidth.  python tools/post_bench.py > profiles/r02_post_kernels.txt"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import neuralstyletransferv1_b200 as m
from neuralstyletransferv1_b200 import engine, synth, temporal

H, W = 1080, 1920
PX = H * W
dev = torch.device("cuda:0")
peak = 6553.6
try:
    peak = float(json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))["hbm_gbs"])
except Exception:
    pass


def timed(fn, reps=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def report(name, ms, bytes_per_call, note):
    gbs = bytes_per_call / ms / 1e6
    print(f"{name:34s} {ms * 1000:9.1f} us  {bytes_per_call / 1e6:8.1f} MB  {gbs:8.1f} GB/s = {gbs / peak:5.3f} of {peak:.0f}   {note}")


frames = synth.synthetic_frames_u8(8, H, W, seed=1).to(dev)
orig = synth.synthetic_frames_u8(8, H, W, seed=2).to(dev)
flow = (torch.randn(H, W, 2, generator=torch.Generator().manual_seed(0)) * 4).to(dev)
mask8 = torch.rand(8, H, W, generator=torch.Generator().manual_seed(1)).to(dev)
masks = m.generate_region_masks(H, W, "voronoi", 4, seed=42, feather=20)
print(f"# 1080p, B200, peak = {peak:.1f} GB/s (copy); algorithmic bytes = every operand once")

lab = m.LabEma(True, True, 0.7, 0.85)
lab(frames)
report("lab_ema (8 frames / launch)", timed(lambda: lab(frames)), PX * (8 * (3 + 3 + 8) + 24), "u8 in + out, two 4-byte table look-ups per pixel, state planes once per chunk")
report("motion_blend (2 launches)", timed(lambda: m.motion_blend(frames[0], orig[0], flow, 0.9)), PX * (8 + 3 + 3 + 3 + 8), "flow, styled, original, out, row-filter plane w + r; 25 taps per pass from L1/L2")
report("blend_pack (8 frames, mask)", timed(lambda: temporal.blend_pack(frames, orig, 0.8, mask8)), PX * 8 * (3 + 3 + 4 + 3), "styled, original, mask, out")
c01 = torch.rand(H, W, 3, device=dev)
p01 = torch.rand(H, W, 3, device=dev)
o01 = torch.empty_like(c01)
report("flow_ema", timed(lambda: engine.flow_ema(c01, p01, flow, 0.8, out=o01)), PX * (12 + 12 + 8 + 12), "curr, prev (gather), flow, out, fp32")
out01 = torch.rand(8, 3, H, W, device=dev)
report("resize_post_pack (identity, 8 fr)", timed(lambda: engine.resize_post_pack(out01, H, W)), PX * 8 * (12 + 3), "fp32 planes in, u8 out")
report("rotate_all_masks (R = 4)", timed(lambda: m.rotate_all_masks(masks, 17.0)), PX * 32, "4 masks in (bilinear gather) + 4 out, fp32")
morph = m.MorphAnimation(True, 1.0, 0.15, 3.0, 3, "blob", 42)
report("warp_all_masks_organic (blob, R = 4)", timed(lambda: m.warp_all_masks_organic(masks, morph, 5), reps=10), PX * (64 + 32 + 16 + 1), "8 noise planes w + r (double sin / cos), masks in + out + normalise, gap map")
report("feather_masks (feather 10, R = 4)", timed(lambda: m.feather_masks(masks, 10)), PX * (16 + 16 + 16 + 16), "two separable passes, 21 taps each from L1/L2")
report("generate_region_masks voronoi", timed(lambda: m.generate_region_masks(H, W, "voronoi", 4, seed=42, feather=20), reps=10), PX * (1 + 1 + 16 + 16 + 16), "labels + two 41-tap passes: tap-bound, not HBM-bound")
report("generate_region_masks spiral", timed(lambda: m.generate_region_masks(H, W, "spiral", 4, seed=42, feather=20), reps=10), PX * (8 + 1 + 1 + 16 + 16 + 16), "field w + r, labels, two 41-tap passes")
report("mask_bboxes (R = 4)", timed(lambda: m.mask_bboxes(masks)), PX * 16, "4 masks once; includes the 16-int read-back to the host")
styled = [frames, orig, frames, orig]
report("composite_regions_batch (4 styles, 8 fr)", timed(lambda: engine.composite_regions_batch(styled, masks, [[0], [1], [2], [3]], original=orig)), PX * (8 * (4 * 3 + 3 + 3) + 16), "4 styled + original + out per frame, fp32 masks once")
