"""Fixture for the whole per-frame tail of the reference loop (pipeline.py:1884-2094) over a short clip: flow-guided EMA -> ToPILImage ->
LAB lightness / chroma EMA -> ToTensor -> mask composite -> motion-adaptive or uniform blend -> ToPILImage, with the loop's literal
arithmetic and its own library calls (`_warp_with_flow` from the real pipeline.py, Pillow, OpenCV, torchvision).  The loop is inline in
`style_frames`, so it cannot be called; the statements below are its statements, in order, with the file I/O and the network removed
(the network outputs `out01` are inputs of the fixture).  Build container only:  python tests/golden/make_golden_chain.py
"""
import importlib.util
import os
import sys

import cv2
import numpy as np
import torch
from PIL import Image
from torchvision import transforms

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = os.environ.get("NST_REFERENCE_DIR", "/root/reference")
sys.path.insert(0, ROOT)
from neuralstyletransferv1_b200 import synth  # noqa: E402

MOTION_NORM, MIN_ALPHA, GAUSS_SIGMA = 8.0, 0.40, 3.0          # pipeline.py:1075-1077


def _ref_pipeline():
    spec = importlib.util.spec_from_file_location("ref_pipeline", os.path.join(REF, "pipeline.py"))
    sys.path.insert(0, REF)
    try:
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
    finally:
        sys.path.remove(REF)
    return mod


def run_loop(ref, out01_seq, orig_u8_seq, flows, cfg, masks=None):
    """The loop body of pipeline.py:1884-2094 for every frame; returns the uint8 frames it would save."""
    to_tensor, to_pil = transforms.ToTensor(), transforms.ToPILImage()
    prev_styled01 = None
    prev_L = prev_aC = prev_bC = None
    smooth_alpha, chroma_alpha = cfg["smooth_alpha"], cfg["chroma_alpha"]
    outs = []
    for idx in range(len(out01_seq)):
        out01 = out01_seq[idx]                                              # [3,H,W] fp32, clamped 0..1 (pipeline.py:1486)
        x_orig01_rs = to_tensor(Image.fromarray(orig_u8_seq[idx])).unsqueeze(0)
        flow = flows[idx]
        # ---- Flow-guided EMA (pre-LAB/pre-blend) ---- (:1884-1936)
        if cfg["flow_ema"] and (flow is not None) and (prev_styled01 is not None):
            curr_styled01 = out01.detach().permute(1, 2, 0).cpu().numpy().astype("float32")
            prev_warp01 = ref._warp_with_flow(prev_styled01, flow)
            a = float(max(0.0, min(1.0, cfg["flow_alpha"])))
            fused01 = (a * curr_styled01 + (1.0 - a) * prev_warp01).clip(0.0, 1.0).astype("float32")
            out01 = torch.from_numpy(fused01).permute(2, 0, 1)
            last_flow = flow
        else:
            last_flow = None
        prev_styled01 = out01.permute(1, 2, 0).cpu().numpy().astype("float32")
        # ---- LAB smoothing ---- (:1942-1978)
        out_img = transforms.ToPILImage()(out01)
        if cfg["smooth_lightness"] or cfg["smooth_chroma"]:
            lab = out_img.convert("LAB")
            lab_np = np.array(lab, dtype=np.uint8).astype(np.float32)
            L = lab_np[..., 0]
            aC = lab_np[..., 1]
            bC = lab_np[..., 2]
            if cfg["smooth_lightness"]:
                if (prev_L is None) or (prev_L.shape != L.shape):
                    prev_L = L.copy()
                L_sm = smooth_alpha * L + (1.0 - smooth_alpha) * prev_L
                prev_L = L_sm
                lab_np[..., 0] = np.clip(L_sm, 0, 255)
            if cfg["smooth_chroma"]:
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
        # ---- blend / composite ---- (:1981-2094)
        out01_smooth = to_tensor(out_img).unsqueeze(0)
        mask_used = False
        if masks is not None:
            alpha = masks[idx][..., None]                                   # HxWx1 fp32, what _load_mask_fit returns
            S = out01_smooth.squeeze(0).permute(1, 2, 0).cpu().numpy().astype(np.float32)
            O = x_orig01_rs.squeeze(0).permute(1, 2, 0).cpu().numpy().astype(np.float32)
            if cfg["composite_mode"] == "keep":
                C = (alpha * S + (1.0 - alpha) * O).clip(0.0, 1.0)
            else:
                C = ((1.0 - alpha) * S + alpha * O).clip(0.0, 1.0)
            out01_smooth = torch.from_numpy(C).permute(2, 0, 1).unsqueeze(0)
            mask_used = True
        blend = cfg["blend"]
        if cfg["motion_blend"] and (last_flow is not None) and not mask_used:
            mag = np.sqrt(last_flow[..., 0] ** 2 + last_flow[..., 1] ** 2).astype(np.float32)
            m = np.clip(mag / MOTION_NORM, 0.0, 1.0)
            m = cv2.GaussianBlur(m, (0, 0), GAUSS_SIGMA)
            base_alpha = float(blend)
            min_alpha = float(MIN_ALPHA)
            max_alpha = base_alpha
            alpha_map = max_alpha - (max_alpha - min_alpha) * m
            alpha_map3 = np.repeat(alpha_map[..., None], 3, axis=2)
            s_np = out01_smooth.squeeze(0).permute(1, 2, 0).cpu().numpy().astype(np.float32)
            o_np = x_orig01_rs.squeeze(0).permute(1, 2, 0).cpu().numpy().astype(np.float32)
            fused = (alpha_map3 * s_np + (1.0 - alpha_map3) * o_np).clip(0.0, 1.0).astype(np.float32)
            out01_final = torch.from_numpy(fused).permute(2, 0, 1)
        else:
            if 0.0 <= blend < 1.0:
                out01_final = (blend * out01_smooth + (1.0 - blend) * x_orig01_rs).clamp(0, 1).squeeze(0)
            else:
                out01_final = out01_smooth.squeeze(0)
        outs.append(np.array(to_pil(out01_final)))
    return np.stack(outs)


CONFIGS = {
    "ema_lab_blend": dict(flow_ema=True, flow_alpha=0.7, smooth_lightness=True, smooth_chroma=True, smooth_alpha=0.7, chroma_alpha=0.85,
                          blend=0.8, motion_blend=False, composite_mode="keep"),
    "ema_motion": dict(flow_ema=True, flow_alpha=0.85, smooth_lightness=False, smooth_chroma=False, smooth_alpha=0.7, chroma_alpha=0.85,
                       blend=0.9, motion_blend=True, composite_mode="keep"),
    "mask_light": dict(flow_ema=True, flow_alpha=0.6, smooth_lightness=True, smooth_chroma=False, smooth_alpha=0.5, chroma_alpha=0.85,
                       blend=0.75, motion_blend=True, composite_mode="replace"),          # a mask disables the motion blend (:2072)
    "plain": dict(flow_ema=False, flow_alpha=0.7, smooth_lightness=False, smooth_chroma=False, smooth_alpha=0.7, chroma_alpha=0.85,
                  blend=1.0, motion_blend=False, composite_mode="keep"),
}


def make_inputs(n=5, H=48, W=72, seed=51):
    rng = np.random.default_rng(seed)
    base = synth.synthetic_frames_u8(n, H, W, seed=seed).float() / 255.0                  # smooth, image-like
    out01 = (base + torch.from_numpy(rng.normal(0, 0.04, base.shape).astype(np.float32))).clamp(0, 1)   # off the k/255 grid
    out01 = (torch.round(out01 * 4096) / 4096).permute(0, 3, 1, 2).contiguous()         # k / 4096: stored as uint16 (load_inputs)
    orig = synth.synthetic_frames_u8(n, H, W, seed=seed + 1).numpy()
    flows = [None] + [(np.round(rng.standard_normal((H, W, 2)) * 40) / 8).astype(np.float32) for _ in range(n - 1)]   # k / 8: exact in fp16
    flows[3] = None                                                                       # a frame whose flow estimate failed (:1933)
    masks = (np.round(rng.random((n, H, W), dtype=np.float32) * 1024) / 1024).astype(np.float32)
    masks[:, :6] = 0.0
    masks[:, 6:12] = 1.0
    return out01, orig, flows, masks


def main():
    ref = _ref_pipeline()
    out01, orig, flows, masks = make_inputs()
    n, _, H, W = out01.shape
    pack = {"out01_q4096": np.round(out01.numpy() * 4096).astype(np.uint16), "orig": orig, "masks_q1024": np.round(masks * 1024).astype(np.uint16),
            "flows": np.stack([f if f is not None else np.zeros((H, W, 2), np.float32) for f in flows]).astype(np.float16),
            "has_flow": np.array([f is not None for f in flows])}
    for name, cfg in CONFIGS.items():
        res = run_loop(ref, list(out01), orig, flows, cfg, masks if name == "mask_light" else None)
        pack["out_" + name] = res
        print(name, res.shape, float(res.mean()))
    np.savez_compressed(os.path.join(HERE, "post_chain_48x72.npz"), **pack)


if __name__ == "__main__":
    main()
