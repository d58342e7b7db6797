"""ncu target: a few calls of the mask-animation / temporal kernels at 1080p (launch list: which kernel takes the time)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import neuralstyletransferv1_b200 as m
from neuralstyletransferv1_b200 import synth
H, W = 1080, 1920
masks = m.generate_region_masks(H, W, "voronoi", 4, seed=42, feather=20)
morph = m.MorphAnimation(True, 1.0, 0.15, 3.0, 3, "blob", 42)
fr = synth.synthetic_frames_u8(2, H, W, seed=1).cuda()
flow = (torch.randn(H, W, 2) * 4).cuda()
for i in range(2):
    m.warp_all_masks_organic(masks, morph, 5 + i)
    m.rotate_all_masks(masks, 17.0)
    m.feather_masks(masks, 10)
    m.motion_blend(fr[0], fr[1], flow, 0.9)
torch.cuda.synchronize()
