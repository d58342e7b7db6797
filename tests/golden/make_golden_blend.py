"""Fixture for region_blend.blend_by_regions / assign_models_to_regions (region_blend.py:983-1046, 1690-1787) from the real reference:
three frames of a static, a rotating and a morphing region blend over seeded uint8-valued "styled" frames, plus a table of model
assignments.  Build container only:  python tests/golden/make_golden_blend.py"""
import itertools
import json
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = os.environ.get("NST_REFERENCE_DIR", "/root/reference")
sys.path.insert(0, ROOT)
from neuralstyletransferv1_b200 import synth  # noqa: E402

CASES = {
    "static": dict(mode="voronoi", region_count=4, assignment="sequential", feather=6, seed=42, original_chance=0.0, rotation_rate=0.0, morph=None),
    "rotating": dict(mode="voronoi", region_count=4, assignment="random", feather=8, seed=7, original_chance=0.3, rotation_rate=9.0, morph=None),
    "morphing": dict(mode="grid", region_count=4, assignment="weighted", feather=8, seed=11, original_chance=0.0, rotation_rate=0.0,
                     morph=dict(enabled=True, speed=1.5, amplitude=0.12, frequency=3.0, octaves=3, mode="blob", seed=42)),
    "both": dict(mode="radial", region_count=3, assignment="sequential", feather=10, seed=3, original_chance=0.5, rotation_rate=-4.0,
                 morph=dict(enabled=True, speed=1.0, amplitude=0.1, frequency=2.0, octaves=3, mode="pulse", seed=1)),
}


def main():
    sys.path.insert(0, REF)
    import region_blend as rb
    sys.path.remove(REF)
    H, W, K, N = 60, 84, 3, 3
    styled_u8 = [synth.synthetic_frames_u8(N, H, W, seed=60 + k).numpy() for k in range(K)]
    orig_u8 = synth.synthetic_frames_u8(N, H, W, seed=70).numpy()
    pack = {"orig": orig_u8, **{f"styled{k}": styled_u8[k] for k in range(K)}}
    for name, c in CASES.items():
        rb.clear_mask_cache()
        morph = rb.MorphAnimation(**c["morph"]) if c["morph"] else None
        outs = []
        for i in range(N):
            styled = [torch.from_numpy(styled_u8[k][i]).permute(2, 0, 1).float() / 255.0 for k in range(K)]
            original = torch.from_numpy(orig_u8[i]).permute(2, 0, 1).float() / 255.0
            out = rb.blend_by_regions(styled, H, W, c["mode"], c["region_count"], c["assignment"], [0.5, 0.3, 0.2], c["feather"], c["seed"], original,
                                      c["original_chance"], frame_idx=i + 4, rotation_rate=c["rotation_rate"], morph=morph)
            outs.append((out.clamp(0, 1) * 255.0).to(torch.uint8).permute(1, 2, 0).numpy())          # ToPILImage pack
        pack["out_" + name] = np.stack(outs)
        print(name, pack["out_" + name].mean())
    np.savez_compressed(os.path.join(HERE, "blend_by_regions_60x84.npz"), **pack)
    table = []
    for nr, nm, mode, seed, oc in itertools.product([3, 4, 7], [2, 4], ["sequential", "random", "weighted"], [0, 5, 42], [0.0, 0.3]):
        w = [0.4, 0.3, 0.2, 0.1][:nm]
        table.append([[nr, nm, mode, w, seed, oc], rb.assign_models_to_regions(nr, nm, mode, w, seed, oc)])
    json.dump(table, open(os.path.join(HERE, "region_assignments.json"), "w"))
    print(len(table), "assignments")


if __name__ == "__main__":
    main()
