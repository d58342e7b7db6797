"""ncu target: three calls each of the streaming post-network kernels at 1080p, 8 frames."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import neuralstyletransferv1_b200 as m
from neuralstyletransferv1_b200 import engine, synth, temporal
H, W = 1080, 1920
fr = synth.synthetic_frames_u8(8, H, W, seed=1).cuda()
orig = synth.synthetic_frames_u8(8, H, W, seed=2).cuda()
mask8 = torch.rand(8, H, W).cuda()
out01 = torch.rand(8, 3, H, W, device="cuda")
masks = m.generate_region_masks(H, W, "voronoi", 4, seed=42, feather=20)
for _ in range(3):
    engine.resize_post_pack(out01, H, W)
    temporal.blend_pack(fr, orig, 0.8, mask8)
    engine.composite_regions_batch([fr, orig, fr, orig], masks, [[0], [1], [2], [3]], original=orig)
torch.cuda.synchronize()
