"""Generate tests/golden/*.npz by running the REAL reference modules.

Run in the build container only (needs /root/reference, which does not exist on
the GPU box):  python tests/golden/make_golden.py

Each fixture stores the seeded input, the reference module's fp32 output, and a
digest of the synthetic state dict (regenerated from the seed at test time by
``neuralstyletransferv1_b200.synth``) so RNG drift is detected rather than
silently compared.
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = os.environ.get("NST_REFERENCE_DIR", "/root/reference")
sys.path.insert(0, ROOT)

from neuralstyletransferv1_b200 import synth  # noqa: E402


def _ref_modules():
    sys.path.insert(0, REF)
    import transformer_net          # noqa: E402
    import transformer_net_nst      # noqa: E402
    import model as reconet_model   # noqa: E402
    import region_blend             # noqa: E402
    sys.path.remove(REF)
    return transformer_net, transformer_net_nst, reconet_model, region_blend


def _input(arch, n, h, w, seed):
    g = torch.Generator().manual_seed(seed)
    u8 = synth.synthetic_frames_u8(n, h, w, seed=seed)
    x01 = u8.permute(0, 3, 1, 2).float() / 255.0
    if arch == "johnson":
        return x01 * 255.0                       # raw_255  (pipeline.py:1473)
    if arch == "nst":
        return x01                               # raw_01   (pipeline.py:611-614)
    return x01 * 2.0 - 1.0                       # tanh     (utils.py:99-101)


CASES = [
    # name, arch, N, H, W, weight seed, input seed
    ("johnson_64x64", "johnson", 1, 64, 64, 0, 11),
    ("johnson_odd_50x70_n2", "johnson", 2, 50, 70, 1, 12),     # 50 -> 52, 70 -> 72 output
    ("nst_64x48", "nst", 1, 64, 48, 0, 13),
    ("nst_odd_53x47", "nst", 1, 53, 47, 1, 14),
    ("reconet_64x64", "reconet", 1, 64, 64, 0, 15),
    ("reconet_odd_38x54_n2", "reconet", 2, 38, 54, 1, 16),
]


def main():
    tn, tn_nst, rc, rb = _ref_modules()
    ctor = {"johnson": tn.TransformerNet, "nst": tn_nst.TransformerNet, "reconet": rc.ReCoNet}
    torch.set_num_threads(os.cpu_count())
    for name, arch, n, h, w, wseed, iseed in CASES:
        sd = synth.make_state_dict(arch, wseed)
        m = ctor[arch]()
        missing, unexpected = m.load_state_dict(sd, strict=True)
        assert not missing and not unexpected
        m.eval()
        x = _input(arch, n, h, w, iseed)
        with torch.no_grad():
            y = m(x)
        np.savez_compressed(
            os.path.join(HERE, name + ".npz"),
            arch=arch, weight_seed=wseed, x=x.numpy().astype(np.float32), y=y.numpy().astype(np.float32),
            digest=synth.state_dict_digest(sd),
        )
        print(f"{name}: x{tuple(x.shape)} -> y{tuple(y.shape)}  y mean {y.mean():.3f} std {y.std():.3f} "
              f"min {y.min():.3f} max {y.max():.3f}")

    # region compositing (region_blend.py:1049-1108) with the reference's own mask generator
    H, W = 48, 64
    g = torch.Generator().manual_seed(77)
    styled = [torch.rand(3, H, W, generator=g) for _ in range(4)]
    original = torch.rand(3, H, W, generator=g)
    masks = rb.generate_region_masks(H, W, "voronoi", 4, seed=42, feather=4)
    out_a = rb.composite_regions(styled, masks, [0, 1, 2, 3], original)
    out_b = rb.composite_regions(styled, masks, [0, -1, 2, 1], original)
    np.savez_compressed(
        os.path.join(HERE, "composite_voronoi4_48x64.npz"),
        styled=torch.stack(styled).numpy(), original=original.numpy(),
        masks=torch.cat([m.reshape(1, H, W) for m in masks]).numpy(),
        out_seq=out_a.numpy(), out_mixed=out_b.numpy(),
    )
    print("composite:", tuple(out_a.shape), float(out_a.mean()))


if __name__ == "__main__":
    main()
