"""Operation-order check of the region-mask / temporal kernels WITHOUT a GPU: the per-pixel functions the CUDA kernels call
(neuralstyletransferv1_b200/csrc/post_math.cuh) are compiled for the host by g++ (tests/hostcheck/hostcheck.cpp, test infrastructure) and
looped over the fixtures the real reference produced (tests/golden/make_golden_masks.py).  The device differs from this build only in its
libm (atan2f / sinf / sin / cos are within 1-2 ulp of glibc's), which the `-m gpu` tests in tests/test_gpu_masks.py then cover."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN = os.path.join(HERE, "golden")
SRC = os.path.join(HERE, "hostcheck", "hostcheck.cpp")
SO = os.path.join(HERE, "hostcheck", "_hostcheck.so")


@pytest.fixture(scope="module")
def hc():
    hdr = os.path.join(HERE, "..", "neuralstyletransferv1_b200", "csrc", "post_math.cuh")
    if not os.path.exists(SO) or os.path.getmtime(SO) < max(os.path.getmtime(SRC), os.path.getmtime(hdr)):
        subprocess.run(["g++", "-O2", "-std=c++17", "-ffp-contract=off", "-fPIC", "-shared", "-x", "c++", SRC, "-o", SO], check=True)
    return C.CDLL(SO)


def _p(a, t=C.c_void_p):
    return a.ctypes.data_as(t)


def _labels(hc, spec, H, W):
    from neuralstyletransferv1_b200.region_masks import PAT_RECTS
    R = spec.regions
    out = np.zeros((H, W), np.uint8)
    f32 = lambda v: np.asarray(v, np.float32)
    p, lo, hi = f32(spec.p), f32(spec.lo if spec.lo else [0]), f32(spec.hi if spec.hi else [0])
    rects = np.asarray([v for r in spec.rects for v in r] or [0], np.int32)
    hc.hc_pattern_labels(spec.mode, spec.norm, _p(p), _p(lo) if spec.mode != PAT_RECTS else None, _p(hi) if spec.mode != PAT_RECTS else None,
                         _p(rects) if spec.mode == PAT_RECTS else None, R, H, W, _p(out))
    return out


def test_pattern_labels_follow_the_reference(hc):
    import random
    from neuralstyletransferv1_b200.region_masks import pattern_spec
    g = np.load(os.path.join(GOLDEN, "masks_patterns_60x84.npz"))
    H, W = 60, 84
    worst = 0
    for key in g.files:
        mode, c, s, f = key.split("_")
        if f != "f0" or mode == "random":
            continue
        count, seed = int(c[1:]), int(s[1:])
        spec = pattern_spec(H, W, mode, count, random.Random(seed))
        lab = _labels(hc, spec, H, W)
        hard = np.stack([(lab == i) for i in range(spec.regions)]).astype(np.uint8)
        ref = g[key][:spec.regions]
        bad = int((hard != ref).sum())
        worst = max(worst, bad)
        # glibc's atan2f / sinf against torch's SLEEF: a pixel within an ulp of a band edge may differ; exact for the rational patterns
        assert bad <= (4 if mode in ("radial", "waves", "spiral") else 0), (key, bad)
    print("worst label mismatch", worst)


def test_rotation_is_opencvs_bit_for_bit(hc):
    g = np.load(os.path.join(GOLDEN, "masks_rotate_60x84.npz"))
    masks = np.ascontiguousarray(g["masks"])
    for key in g.files:
        if not key.startswith("rot_"):
            continue
        out = np.empty_like(masks)
        hc.hc_rotate_masks(_p(masks), _p(out), masks.shape[0], 60, 84, C.c_double(float(key[4:])))
        assert np.array_equal(out, g[key]), key


@pytest.mark.parametrize("mode", ["blob", "tentacle", "wave", "pulse", "gappy"])
def test_organic_warp_follows_the_reference(hc, mode):
    from neuralstyletransferv1_b200.region_masks import ORGANIC_MODE, organic_offsets
    g = np.load(os.path.join(GOLDEN, "masks_organic_60x84.npz"))
    if mode == "gappy":
        masks, name, amp, freq, speed, seed, frame, want = g["hard_masks"].astype(np.float32), "blob", 0.9, 5.0, 1.0, 5, 3, g["gappy_f3"]
    else:
        masks, name, amp, freq, speed, seed, frame, want = g["masks"], mode, 0.15, 3.0, 1.5, 42, 7, g[f"{mode}_f7"]
    masks = np.ascontiguousarray(masks)
    R = masks.shape[0]
    t = frame * speed * 0.02
    offs = np.asarray([v for i in range(R) for v in organic_offsets(seed + i * 100, t)], np.float64)
    out = np.empty_like(masks)
    hc.hc_warp_masks_organic.restype = C.c_int
    gaps = hc.hc_warp_masks_organic(_p(masks), _p(out), R, 60, 84, ORGANIC_MODE[name], C.c_double(amp), C.c_double(freq), C.c_double(t), _p(offs))
    if mode == "gappy":
        assert gaps == int(g["gaps_before_fill"]) and gaps > 0
    assert np.array_equal(out, want), float(np.abs(out - want).max())


def test_lab_ema_and_motion_blend_follow_the_reference(hc):
    from neuralstyletransferv1_b200.temporal import lab_tables_host
    rgb2lab, lab2rgb = lab_tables_host()
    g = np.load(os.path.join(GOLDEN, "lab_ema_48x64.npz"))
    frames = np.ascontiguousarray(g["frames"])
    for tag, (sl, sc) in {"both": (1, 1), "light": (1, 0), "chroma": (0, 1)}.items():
        out = np.empty_like(frames)
        hc.hc_lab_ema(_p(frames), _p(out), frames.shape[0], 48, 64, _p(rgb2lab), _p(lab2rgb), sl, sc, C.c_double(float(g["smooth_alpha"])),
                      C.c_double(float(g["chroma_alpha"])))
        assert np.array_equal(out, g[f"out_{tag}"]), tag
    m = np.load(os.path.join(GOLDEN, "motion_blend.npz"))
    for t in ("54x70", "40x64"):
        H, W = (int(v) for v in t.split("x"))
        out = np.empty((H, W, 3), np.uint8)
        amap = np.empty((H, W), np.float32)
        hc.hc_motion_blend(_p(np.ascontiguousarray(m[f"styled_{t}"])), _p(np.ascontiguousarray(m[f"orig_{t}"])), _p(np.ascontiguousarray(m[f"flow_{t}"])),
                           _p(out), _p(amap), H, W, C.c_double(float(m["blend"])), C.c_double(0.40), C.c_double(8.0), C.c_double(3.0))
        assert np.array_equal(amap, m[f"alpha_map_{t}"]), t           # cv2.GaussianBlur's fp32 order, SIMD tails included
        assert np.array_equal(out, m[f"out_{t}"]), t


def test_blend_pack_follows_the_reference_arithmetic(hc):
    """The uint8 blend stage (to_tensor -> mask composite -> uniform blend -> ToPILImage, pipeline.py:1981, 2040-2043, 2088-2094) against
    the oracle's torch expression of the same lines, every uint8 pair included."""
    import torch
    from oracle import nst_oracle as O
    rng = np.random.default_rng(5)
    H, W = 64, 256 * 4
    styled = rng.integers(0, 256, (H, W, 3), dtype=np.uint8)
    orig = rng.integers(0, 256, (H, W, 3), dtype=np.uint8)
    styled[0, :, 0] = np.repeat(np.arange(256, dtype=np.uint8), 4)           # a sweep of every level against random originals
    mask = rng.random((H, W), dtype=np.float32)
    mask[:4] = 0.0
    mask[4:8] = 1.0
    s01 = torch.from_numpy(styled).permute(2, 0, 1).float().div(255)[None]
    x01 = torch.from_numpy(orig).permute(2, 0, 1).float().div(255)[None]
    for blend, use_mask, mode in ((0.8, False, 0), (1.0, True, 0), (0.6, True, 1), (0.0, True, 0), (-1.0, False, 0), (0.35, True, 1)):
        out = np.empty_like(styled)
        hc.hc_blend_pack(_p(styled), _p(orig), _p(mask) if use_mask else None, C.c_double(blend), mode, _p(out), C.c_size_t(H * W))
        want = O.post_u8(s01, x01, blend, torch.from_numpy(mask)[None] if use_mask else None, "keep" if mode == 0 else "replace")[0].numpy()
        assert np.array_equal(out, want), (blend, use_mask, mode)
