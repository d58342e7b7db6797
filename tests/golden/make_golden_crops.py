"""Fixture for prepare_region_crops (region_blend.py:1969-2059) from the real reference: masks, tight and padded boxes, an empty mask.
Build container only:  python tests/golden/make_golden_crops.py"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("NST_REFERENCE_DIR", "/root/reference")


def main():
    sys.path.insert(0, REF)
    import region_blend as rb
    sys.path.remove(REF)
    H, W = 96, 128
    masks = rb.generate_region_masks(H, W, "voronoi", 3, seed=5, feather=5)
    masks = [m * (torch.arange(W).view(1, 1, 1, W) < W - 24).float() for m in masks]        # a strip no region covers
    masks.append(torch.zeros(1, 1, H, W))                                                   # an empty region: full-frame box
    small = torch.zeros(1, 1, H, W)
    small[:, :, 40:43, 70:90] = 0.0101                                                      # just above the 0.01 threshold
    small[:, :, 43:50, 60:100] = 0.0099                                                     # just below it
    masks.append(small)
    configs = [rb.RegionConfig([0], [1.0], 1.0) for _ in masks]
    pack = {"masks": torch.cat([m.reshape(1, H, W) for m in masks]).numpy()}
    for pad in (12, 64):
        crops = rb.prepare_region_crops(masks, configs, H, W, padding=pad)
        pack[f"bbox_p{pad}"] = np.array([c.bbox for c in crops])
        pack[f"padded_p{pad}"] = np.array([c.padded_bbox for c in crops])
        print(pad, pack[f"bbox_p{pad}"].tolist(), pack[f"padded_p{pad}"].tolist())
    np.savez_compressed(os.path.join(HERE, "region_crops_96x128.npz"), **pack)


if __name__ == "__main__":
    main()
