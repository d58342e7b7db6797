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

    # ---- round 2 fixtures: every post-network path pinned by the reference's own functions -----------------------------------
    import torch.nn.functional as F
    # (1) composite_regions_advanced with per-region model blends AND a half-resolution scale (region_blend.py:1589-1679)
    H, W = 60, 84
    g = torch.Generator().manual_seed(78)
    full = [torch.rand(3, H, W, generator=g) for _ in range(3)]
    half = [torch.rand(3, H // 2, W // 2, generator=g) for _ in range(3)]
    original = torch.rand(3, H, W, generator=g)
    masks = rb.generate_region_masks(H, W, "voronoi", 4, seed=7, feather=6)
    configs = [rb.RegionConfig([0, 1], [0.7, 0.3], 1.0), rb.RegionConfig([1, -1], [0.5, 0.5], 0.5),
               rb.RegionConfig([2], [1.0], 0.5), rb.RegionConfig([2, 0], [0.2, 0.8], 1.0)]
    out_adv = rb.composite_regions_advanced({1.0: full, 0.5: half}, masks, configs, original, H, W)
    np.savez_compressed(os.path.join(HERE, "composite_advanced_60x84.npz"), full=torch.stack(full).numpy(), half=torch.stack(half).numpy(),
                        original=original.numpy(), masks=torch.cat([m.reshape(1, H, W) for m in masks]).numpy(),
                        idx=np.array([[0, 1], [1, -1], [2, -2], [2, 0]]), wts=np.array([[0.7, 0.3], [0.5, 0.5], [1.0, 0.0], [0.2, 0.8]], np.float32),
                        scales=np.array([1.0, 0.5, 0.5, 1.0], np.float32), out=out_adv.numpy())
    print("composite_advanced:", tuple(out_adv.shape), float(out_adv.mean()))

    # (2) prepare_region_crops + composite_from_crops (region_blend.py:2019-2059, 2186-2294): with an original (gap fill by the original)
    #     and without (max-pool dilation fill); one styled crop is delivered at half size (resized inside composite_from_crops)
    #     (canvas large enough that every crop resize has > 12k outputs: below that ATen takes a scalar path whose rounding differs
    #     by 1 ulp from the vectorised kernel every real frame size uses -- tests/test_oracle.py::test_bilinear_formula...)
    H, W = 180, 240
    g = torch.Generator().manual_seed(79)
    q8 = lambda t: torch.round(t * 255) / 255                                               # image-like values (k / 255): compressible fixture
    original = q8(torch.rand(3, H, W, generator=g))
    masks = rb.generate_region_masks(H, W, "voronoi", 3, seed=5, feather=5)
    masks = [m * (torch.arange(W).view(1, 1, 1, W) < W - 40).float() for m in masks]       # a strip no region covers: a gap
    configs = [rb.RegionConfig([0], [1.0], 1.0), rb.RegionConfig([1, -1], [0.6, 0.4], 1.0), rb.RegionConfig([0, 1], [0.5, 0.5], 1.0)]
    crops = rb.prepare_region_crops(masks, configs, H, W, padding=12)
    styled = {0: {}, 1: {}}
    for c in crops:
        x1, y1, x2, y2 = c.padded_bbox
        for mi in c.config.model_indices:
            if mi >= 0:
                hh, ww = (y2 - y1, x2 - x1) if not (mi == 1 and c.region_idx == 2) else ((y2 - y1) // 2, (x2 - x1) // 2)
                styled[mi][c.region_idx] = q8(torch.rand(3, hh, ww, generator=g))
    out_o = rb.composite_from_crops(styled, crops, original, H, W)
    crops_no = [c for c in crops if -1 not in c.config.model_indices]
    out_n = rb.composite_from_crops(styled, crops_no, None, H, W)
    pack = {"original": original.numpy(), "out_with_original": out_o.numpy(), "out_without_original": out_n.numpy(),
            "n_regions": np.array(len(crops)), "no_orig_regions": np.array([c.region_idx for c in crops_no])}
    for c in crops:
        r = c.region_idx
        pack[f"bbox{r}"] = np.array(c.padded_bbox)
        pack[f"mask{r}"] = c.crop_mask.numpy()
        pack[f"idx{r}"] = np.array(c.config.model_indices)
        pack[f"wts{r}"] = np.array(c.config.model_weights, np.float32)
        for mi in c.config.model_indices:
            if mi >= 0:
                pack[f"styled{mi}_{r}"] = styled[mi][r].numpy()
    pack["hw"] = np.array([H, W])
    for c in crops:
        print("  crop", c.region_idx, c.padded_bbox)
    np.savez_compressed(os.path.join(HERE, "composite_crops_180x240.npz"), **pack)
    print("composite_from_crops:", float(out_o.mean()), float(out_n.mean()), "gap px", int((sum(m for m in masks) < 0.1).sum()))

    # (2b) the C5 benchmark masks: generate_region_masks(1080, 1920, "voronoi", 4, seed=42, feather=20) (pipeline.py:1126-1127,1152) --
    #      per-mask sums, 8192 sampled values (positions seeded) and the hard labels' histogram; the full masks would be 33 MB
    big = rb.generate_region_masks(1080, 1920, "voronoi", 4, seed=42, feather=20)
    big = torch.cat([m.reshape(1, 1080, 1920) for m in big])
    pos = torch.randint(0, 1080 * 1920, (8192,), generator=torch.Generator().manual_seed(1))
    np.savez_compressed(os.path.join(HERE, "voronoi4_1080p_samples.npz"), sums=big.double().sum((1, 2)).numpy(), pos=pos.numpy(),
                        values=big.reshape(4, -1)[:, pos].numpy(), argmax_hist=torch.bincount(big.argmax(0).flatten(), minlength=4).numpy())
    print("voronoi 1080p sums", big.double().sum((1, 2)).tolist())

    # (3) flow-guided EMA: the reference's `_warp_with_flow` (pipeline.py:425-439) + the lerp of :1927-1929
    import importlib.util
    spec = importlib.util.spec_from_file_location("ref_pipeline", os.path.join(REF, "pipeline.py"))
    H, W = 54, 70
    rng = np.random.default_rng(3)
    prev = rng.random((H, W, 3), dtype=np.float32)
    curr = rng.random((H, W, 3), dtype=np.float32)
    flow = (rng.standard_normal((H, W, 2)) * 4).astype(np.float32)
    try:
        sys.path.insert(0, REF)
        ref_pipeline = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(ref_pipeline)
        warped = ref_pipeline._warp_with_flow(prev, flow)
    finally:
        sys.path.remove(REF)
    a = 0.65
    fused = (a * curr + (1.0 - a) * warped).clip(0.0, 1.0).astype("float32")
    np.savez_compressed(os.path.join(HERE, "flow_ema_54x70.npz"), prev=prev, curr=curr, flow=flow, alpha=np.array(a), warped=warped, fused=fused)
    print("flow_ema:", float(fused.mean()))

    # (4) the resize-back of odd-sized frames + mask composite + blend + ToPILImage pack with the pipeline's literal arithmetic
    #     (pipeline.py:1473-1476, 1512-1516, 2040-2043, 2088-2094) around the real TransformerNet
    from torchvision import transforms
    sd = synth.make_state_dict("johnson", 1)
    m = tn.TransformerNet(); m.load_state_dict(sd); m.eval()
    u8 = synth.synthetic_frames_u8(1, 50, 70, seed=21)
    x_orig01 = u8.permute(0, 3, 1, 2).float() / 255.0
    alpha = torch.rand(1, 50, 70, generator=torch.Generator().manual_seed(4))
    with torch.no_grad():
        y = m((x_orig01 * 255.0))
    out01 = (y / 255.0).clamp(0, 1).squeeze(0)
    out01 = F.interpolate(out01.unsqueeze(0), size=(50, 70), mode="bilinear", align_corners=False).squeeze(0)
    S = out01.permute(1, 2, 0).numpy().astype(np.float32); O_ = x_orig01.squeeze(0).permute(1, 2, 0).numpy().astype(np.float32)
    al = alpha.permute(1, 2, 0).numpy().astype(np.float32)
    Cc = (al * S + (1.0 - al) * O_).clip(0.0, 1.0)
    sm = torch.from_numpy(Cc).permute(2, 0, 1).unsqueeze(0)
    blend = 0.8
    fin = (blend * sm + (1.0 - blend) * x_orig01).clamp(0, 1).squeeze(0)
    pil = np.array(transforms.ToPILImage()(fin))
    pil_plain = np.array(transforms.ToPILImage()(out01))
    np.savez_compressed(os.path.join(HERE, "fit_johnson_50x70.npz"), frames=u8.numpy(), alpha=alpha.numpy(), blend=np.array(blend),
                        out_u8=pil, out_u8_plain=pil_plain, weight_seed=np.array(1))
    print("fit:", pil.shape, pil.mean())


if __name__ == "__main__":
    main()
