"""Fixtures for the rest of SURVEY.md 8f rows N2 / N3, produced by the REAL reference functions (build container only; needs
/root/reference):  python tests/golden/make_golden_masks.py

  masks_patterns_60x84.npz     generate_region_masks for every pattern (region_blend.py:109-521, 925-981)
  masks_patterns_1080p.npz     the analytic patterns at 1080p: hard-label histograms + sampled feathered values
  masks_rotate_60x84.npz       rotate_all_masks (region_blend.py:25-66)
  masks_organic_60x84.npz      warp_all_masks_organic, four modes + one amplitude large enough to open gaps (region_blend.py:670-810)
  lab_ema_48x64.npz            the LAB lightness / chroma EMA of pipeline.py:1938-1975 (literal arithmetic, Pillow's LAB transform)
  motion_blend_54x70.npz       the motion-adaptive blend of pipeline.py:2072-2086 (cv2.GaussianBlur of the flow magnitude)
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = os.environ.get("NST_REFERENCE_DIR", "/root/reference")
sys.path.insert(0, ROOT)

PATTERNS = ["grid", "diagonal", "fractal", "radial", "waves", "spiral", "concentric"]


def _rb():
    sys.path.insert(0, REF)
    import region_blend
    sys.path.remove(REF)
    return region_blend


def _stack(masks):
    return torch.cat([m.reshape(1, *m.shape[-2:]) for m in masks]).numpy()


def main():
    rb = _rb()
    H, W = 60, 84
    pack = {}
    for mode in PATTERNS:
        for count, seed in ((4, 7), (5, 11)):
            pack[f"{mode}_c{count}_s{seed}_f5"] = _stack(rb.generate_region_masks(H, W, mode, count, seed=seed, feather=5))
            pack[f"{mode}_c{count}_s{seed}_f0"] = _stack(rb.generate_region_masks(H, W, mode, count, seed=seed, feather=0)).astype(np.uint8)
    pack["random_c4_s3_f0"] = _stack(rb.generate_region_masks(H, W, "random", 4, seed=3, feather=0)).astype(np.uint8)
    np.savez_compressed(os.path.join(HERE, "masks_patterns_60x84.npz"), **pack)
    print("patterns:", len(pack), "arrays")

    big = {}
    pos = torch.randint(0, 1080 * 1920, (4096,), generator=torch.Generator().manual_seed(2))
    big["pos"] = pos.numpy()
    for mode in ("radial", "waves", "spiral", "concentric", "diagonal"):
        hard = torch.from_numpy(_stack(rb.generate_region_masks(1080, 1920, mode, 4, seed=42, feather=0)))
        lab = torch.where(hard.sum(0) > 0, hard.argmax(0), torch.full((1080, 1920), 4))
        big[f"{mode}_hist"] = torch.bincount(lab.flatten(), minlength=5).numpy()
        big[f"{mode}_labels_at_pos"] = lab.flatten()[pos].numpy().astype(np.uint8)
        print(mode, big[f"{mode}_hist"])
    soft = torch.from_numpy(_stack(rb.generate_region_masks(1080, 1920, "spiral", 4, seed=42, feather=20)))
    big["spiral_f20_values"] = soft.reshape(4, -1)[:, pos].numpy()
    np.savez_compressed(os.path.join(HERE, "masks_patterns_1080p.npz"), **big)

    base = rb.generate_region_masks(H, W, "voronoi", 4, seed=7, feather=6)
    rot = {"masks": _stack(base)}
    for ang in (33.0, -90.0, 180.5):
        rot[f"rot_{ang}"] = _stack(rb.rotate_all_masks(base, ang))
    np.savez_compressed(os.path.join(HERE, "masks_rotate_60x84.npz"), **rot)
    print("rotate: sum range", rot["rot_33.0"].sum(0).min(), rot["rot_33.0"].sum(0).max())

    org = {"masks": _stack(base)}
    for mode in ("blob", "tentacle", "wave", "pulse"):
        morph = rb.MorphAnimation(enabled=True, speed=1.5, amplitude=0.15, frequency=3.0, octaves=3, mode=mode, seed=42)
        org[f"{mode}_f7"] = _stack(rb.warp_all_masks_organic(base, morph, 7))
    hard = rb.generate_region_masks(H, W, "voronoi", 4, seed=7, feather=0)
    morph = rb.MorphAnimation(enabled=True, speed=1.0, amplitude=0.9, frequency=5.0, octaves=3, mode="blob", seed=5)
    org["hard_masks"] = _stack(hard).astype(np.uint8)
    # count the gap pixels the reference saw before filling (same arithmetic as region_blend.py:769-778)
    warped = [rb.warp_mask_organic(m, morph.amplitude, morph.frequency, 3, morph.speed, morph.seed + i * 100, morph.mode) for i, m in enumerate(hard)]
    org["gaps_before_fill"] = np.array(int((sum(warped) < 0.1).sum()))
    org["gappy_f3"] = _stack(rb.warp_all_masks_organic(hard, morph, 3))
    np.savez_compressed(os.path.join(HERE, "masks_organic_60x84.npz"), **org)
    print("organic: gaps before fill", int(org["gaps_before_fill"]))

    # ---- LAB EMA: pipeline.py:1938-1975, literally, over four frames (lightness + chroma, then lightness only) ----------------
    from PIL import Image
    from neuralstyletransferv1_b200 import synth
    frames = synth.synthetic_frames_u8(4, 48, 64, seed=31).numpy()
    lab_pack = {"frames": frames, "smooth_alpha": np.array(0.7), "chroma_alpha": np.array(0.85)}
    for tag, (sl, sc) in {"both": (True, True), "light": (True, False), "chroma": (False, True)}.items():
        prev_L = prev_aC = prev_bC = None
        smooth_alpha, chroma_alpha = 0.7, 0.85
        outs = []
        for f in frames:
            out_img = Image.fromarray(f, mode="RGB")
            lab = out_img.convert("LAB")
            lab_np = np.array(lab, dtype=np.uint8).astype(np.float32)
            L = lab_np[..., 0]; aC = lab_np[..., 1]; bC = lab_np[..., 2]
            if sl:
                if (prev_L is None) or (prev_L.shape != L.shape):
                    prev_L = L.copy()
                L_sm = smooth_alpha * L + (1.0 - smooth_alpha) * prev_L
                prev_L = L_sm
                lab_np[..., 0] = np.clip(L_sm, 0, 255)
            if sc:
                if (prev_aC is None) or (prev_aC.shape != aC.shape):
                    prev_aC = aC.copy()
                if (prev_bC is None) or (prev_bC.shape != bC.shape):
                    prev_bC = bC.copy()
                a_sm = chroma_alpha * aC + (1.0 - chroma_alpha) * prev_aC
                b_sm = chroma_alpha * bC + (1.0 - chroma_alpha) * prev_bC
                prev_aC, prev_bC = a_sm, b_sm
                lab_np[..., 1] = np.clip(a_sm, 0, 255)
                lab_np[..., 2] = np.clip(b_sm, 0, 255)
            out_img = Image.fromarray(lab_np.astype(np.uint8), mode="LAB").convert("RGB")
            outs.append(np.array(out_img))
        lab_pack[f"out_{tag}"] = np.stack(outs)
    np.savez_compressed(os.path.join(HERE, "lab_ema_48x64.npz"), **lab_pack)
    print("lab_ema: mean |out - in|", float(np.abs(lab_pack["out_both"].astype(int) - frames.astype(int)).mean()))

    # ---- motion-adaptive blend: pipeline.py:2072-2086 (MOTION_NORM 8, MIN_ALPHA 0.4, GAUSS_SIGMA 3: :1075-1077) -----------------
    import cv2
    from torchvision import transforms
    MOTION_NORM, MIN_ALPHA, GAUSS_SIGMA = 8.0, 0.40, 3.0
    mb = {}
    for (H2, W2) in ((54, 70), (40, 64)):                                  # 70: SIMD tails of both filter passes; 64: none
        rng = np.random.default_rng(H2)
        styled = synth.synthetic_frames_u8(1, H2, W2, seed=41)[0].numpy()
        orig = synth.synthetic_frames_u8(1, H2, W2, seed=42)[0].numpy()
        last_flow = (rng.standard_normal((H2, W2, 2)) * 6).astype(np.float32)
        blend = 0.9
        out01_smooth = transforms.ToTensor()(Image.fromarray(styled)).unsqueeze(0)
        x_orig01_rs = transforms.ToTensor()(Image.fromarray(orig)).unsqueeze(0)
        mag = np.sqrt(last_flow[..., 0] ** 2 + last_flow[..., 1] ** 2).astype(np.float32)
        m = np.clip(mag / MOTION_NORM, 0.0, 1.0)
        m = cv2.GaussianBlur(m, (0, 0), GAUSS_SIGMA)
        base_alpha = float(blend); min_alpha = float(MIN_ALPHA); max_alpha = base_alpha
        alpha_map = max_alpha - (max_alpha - min_alpha) * m
        alpha_map3 = np.repeat(alpha_map[..., None], 3, axis=2)
        s_np = out01_smooth.squeeze(0).permute(1, 2, 0).cpu().numpy().astype(np.float32)
        o_np = x_orig01_rs.squeeze(0).permute(1, 2, 0).cpu().numpy().astype(np.float32)
        fused = (alpha_map3 * s_np + (1.0 - alpha_map3) * o_np).clip(0.0, 1.0).astype(np.float32)
        out01_final = torch.from_numpy(fused).permute(2, 0, 1)
        out = np.array(transforms.ToPILImage()(out01_final))
        t = f"{H2}x{W2}"
        mb.update({f"styled_{t}": styled, f"orig_{t}": orig, f"flow_{t}": last_flow, f"alpha_map_{t}": alpha_map.astype(np.float32), f"out_{t}": out})
        print("motion_blend", t, alpha_map.dtype, float(alpha_map.mean()))
    mb["blend"] = np.array(0.9)
    np.savez_compressed(os.path.join(HERE, "motion_blend.npz"), **mb)


if __name__ == "__main__":
    main()
