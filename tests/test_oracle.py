"""The CPU oracle against the fixtures produced by the REAL reference modules (tests/golden)."""
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN, golden_names
from neuralstyletransferv1_b200 import synth
from oracle import nst_oracle as O


@pytest.mark.parametrize("name", golden_names())
def test_oracle_matches_reference_modules(name):
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    arch, seed = str(g["arch"]), int(g["weight_seed"])
    sd = synth.make_state_dict(arch, seed)
    assert synth.state_dict_digest(sd) == str(g["digest"]), "synthetic checkpoint drifted from the fixture's"
    y = O.forward(arch, sd, torch.from_numpy(g["x"]))
    ref = torch.from_numpy(g["y"])
    assert y.shape == ref.shape
    # same torch ops in the same order -> bit-identical on the same build; allow 1e-5 rel across builds
    scale = float(ref.abs().max())
    assert float((y - ref).abs().max()) <= 1e-5 * scale


def test_output_shapes_follow_reference():
    sd = synth.make_state_dict("johnson", 0)
    y = O.forward("johnson", sd, torch.zeros(1, 3, 50, 70))
    assert tuple(y.shape) == (1, 3, 52, 72)          # 4*ceil(ceil(H/2)/2)
    sdn = synth.make_state_dict("nst", 0)
    y = O.forward("nst", sdn, torch.zeros(1, 3, 53, 47))
    assert tuple(y.shape) == (1, 3, 53, 47)


@pytest.mark.parametrize("preset", ["raw_255", "raw_01", "tanh", "imagenet_01", "imagenet_255", "caffe_bgr"])
def test_io_preset_roundtrip_identity_net(preset):
    """pre then post of an identity 'network' returns the input (pipeline.py:1447-1486 pairs).

    imagenet_255 and caffe_bgr are exceptions the reference itself has (post is not the inverse
    of pre) -- pinned here as they are."""
    x = torch.rand(2, 3, 5, 7)
    back = O.io_post(O.io_pre(x, preset), preset)
    if preset == "imagenet_255":
        m = torch.tensor(O.IMAGENET_MEAN).view(1, 3, 1, 1); s = torch.tensor(O.IMAGENET_STD).view(1, 3, 1, 1)
        expect = (((x * 255 - m * 255) / (s * 255)) / 255).clamp(0, 1)
        assert torch.allclose(back, expect, atol=1e-6)
    elif preset == "caffe_bgr":      # post does not add the Caffe mean back (pipeline.py:1462-1470)
        c = torch.tensor(O.CAFFE_MEAN).view(1, 3, 1, 1)
        expect = (((x * 255)[:, [2, 1, 0]] - c)[:, [2, 1, 0]] / 255).clamp(0, 1)
        assert torch.allclose(back, expect, atol=1e-6)
    else:
        assert torch.allclose(back, x, atol=1e-5)


def test_pack_truncates():
    v = torch.tensor([0.0, 0.999 / 255, 1.0 / 255 + 1e-4, 254.99 / 255, 1.0, 1.7, -0.3]).view(1, 1, 1, 7).repeat(1, 3, 1, 1)
    u = O.pack_u8(v)[0, 0, :, 0]
    assert u.tolist() == [0, 0, 1, 254, 255, 255, 0]


def test_composite_matches_reference_fixture():
    g = np.load(os.path.join(GOLDEN, "composite_voronoi4_48x64.npz"))
    styled = [torch.from_numpy(s) for s in g["styled"]]
    masks = [torch.from_numpy(m)[None, None] for m in g["masks"]]
    orig = torch.from_numpy(g["original"])
    a = O.composite_regions(styled, masks, [0, 1, 2, 3], orig)
    b = O.composite_regions(styled, masks, [0, -1, 2, 1], orig)
    assert torch.equal(a, torch.from_numpy(g["out_seq"]))
    assert torch.equal(b, torch.from_numpy(g["out_mixed"]))
    adv = O.composite_regions_advanced(styled, masks, [[0], [1], [2], [3]], [[1.0]] * 4, orig)
    assert torch.allclose(adv, a, atol=1e-7)


def test_reconet_run_u8_equals_stylize_tanh():
    sd = synth.make_state_dict("reconet", 0)
    u8 = synth.synthetic_frames_u8(1, 32, 40, seed=2)
    a = O.reconet_run_u8(sd, u8)
    b = O.stylize_u8("reconet", sd, u8, "tanh")
    mx, _ = O.u8_metrics(a, b)
    assert mx <= 1       # (y+1)/2*255 vs ((y+1)*0.5).clamp*255: same value up to 1 ulp at a truncation edge


# ---------------------------------------------------------------------------------------------------------------------
# round 2: the post-network restatements against fixtures produced by the reference's OWN functions (make_golden.py)
# ---------------------------------------------------------------------------------------------------------------------
def _crops_from_fixture(g, only=None):
    crops, styled = [], {}
    for r in range(int(g["n_regions"])):
        if only is not None and r not in only:
            continue
        idx = [int(i) for i in g[f"idx{r}"]]
        crops.append({"region_idx": r, "padded_bbox": tuple(int(v) for v in g[f"bbox{r}"]), "crop_mask": torch.from_numpy(g[f"mask{r}"]),
                      "model_indices": idx, "model_weights": [float(w) for w in g[f"wts{r}"]]})
        for mi in idx:
            if mi >= 0:
                styled.setdefault(mi, {})[r] = torch.from_numpy(g[f"styled{mi}_{r}"])
    return crops, styled


def test_composite_advanced_multiscale_matches_reference():
    g = np.load(os.path.join(GOLDEN, "composite_advanced_60x84.npz"))
    full, half = [torch.from_numpy(a) for a in g["full"]], [torch.from_numpy(a) for a in g["half"]]
    masks = [torch.from_numpy(m)[None, None] for m in g["masks"]]
    configs = [([int(i) for i in ix if i > -2], [float(w) for i, w in zip(ix, ws) if i > -2], float(sc))
               for ix, ws, sc in zip(g["idx"], g["wts"], g["scales"])]
    out = O.composite_regions_multiscale({1.0: full, 0.5: half}, masks, configs, torch.from_numpy(g["original"]), 60, 84)
    assert torch.equal(out, torch.from_numpy(g["out"]))


def test_composite_from_crops_matches_reference():
    g = np.load(os.path.join(GOLDEN, "composite_crops_180x240.npz"))
    crops, styled = _crops_from_fixture(g)
    out = O.composite_from_crops(styled, crops, torch.from_numpy(g["original"]), 180, 240)
    assert torch.equal(out, torch.from_numpy(g["out_with_original"]))
    crops, styled = _crops_from_fixture(g, only=set(int(r) for r in g["no_orig_regions"]))
    out = O.composite_from_crops(styled, crops, None, 180, 240)
    assert torch.equal(out, torch.from_numpy(g["out_without_original"]))


def test_flow_ema_matches_reference():
    g = np.load(os.path.join(GOLDEN, "flow_ema_54x70.npz"))
    fused = O.flow_ema(g["curr"], g["prev"], g["flow"], float(g["alpha"]))
    assert np.array_equal(fused, g["fused"])


def test_fit_path_matches_pipeline_arithmetic():
    """Odd-sized frame: net -> /255 -> clamp -> bilinear resize back -> mask composite -> blend -> ToPILImage, generated with the
    pipeline's literal numpy / torch arithmetic around the real TransformerNet (make_golden.py (4))."""
    g = np.load(os.path.join(GOLDEN, "fit_johnson_50x70.npz"))
    sd = synth.make_state_dict("johnson", int(g["weight_seed"]))
    frames = torch.from_numpy(g["frames"])
    out = O.stylize_u8_fit("johnson", sd, frames, "raw_255", blend=float(g["blend"]), mask=torch.from_numpy(g["alpha"]))
    assert torch.equal(out[0], torch.from_numpy(g["out_u8"]))
    plain = O.stylize_u8_fit("johnson", sd, frames, "raw_255")
    assert torch.equal(plain[0], torch.from_numpy(g["out_u8_plain"]))


def test_bilinear_formula_used_by_the_kernels_is_torchs():
    """csrc/postproc.cu evaluates F.interpolate(bilinear, align_corners=False) as: source index fma(scale, dst + 0.5, -0.5), value
    fma(wy0, fma(wx0, a, wx1*b), wy1 * fma(wx0, c, wx1*d)).  That is bit-for-bit what ATen's CPU kernel computes (sizes below ~100
    outputs take a scalar path that differs by 1 ulp)."""
    f32 = np.float32

    def fma(a, b, c):
        return (a.astype(np.float64) * b.astype(np.float64) + c.astype(np.float64)).astype(np.float32)

    def coords(i_sz, o_sz):
        scale = f32(i_sz) / f32(o_sz)
        d = np.arange(o_sz, dtype=np.float32)
        real = np.maximum(fma(np.full_like(d, scale), d + f32(0.5), np.full_like(d, -0.5)), f32(0))
        i0 = np.minimum(np.floor(real).astype(np.int64), i_sz - 1)
        lam = np.clip(real - i0.astype(np.float32), f32(0), f32(1)).astype(np.float32)
        return i0, i0 + (i0 < i_sz - 1), (f32(1) - lam).astype(np.float32), lam

    for hi, wi, h, w in ((272, 484, 271, 481), (135, 240, 270, 480), (252, 272, 250, 270)):
        x = torch.rand(1, 3, hi, wi, generator=torch.Generator().manual_seed(hi))
        ref = torch.nn.functional.interpolate(x, size=(h, w), mode="bilinear", align_corners=False).numpy()
        xn = x.numpy()
        y0, y1, wy0, wy1 = coords(hi, h)
        x0, x1, wx0, wx1 = coords(wi, w)
        a, b, c, d = xn[:, :, y0][..., x0], xn[:, :, y0][..., x1], xn[:, :, y1][..., x0], xn[:, :, y1][..., x1]
        bx = lambda v: np.broadcast_to(v[None, None, None, :], a.shape)
        by = lambda v: np.broadcast_to(v[None, None, :, None], a.shape)
        top, bot = fma(bx(wx0), a, bx(wx1) * b), fma(bx(wx0), c, bx(wx1) * d)
        mine = fma(by(wy0), top, by(wy1) * bot)
        assert np.array_equal(mine, ref), (hi, wi, h, w)


def test_u8_to_unit_shortcut_is_exact():
    """csrc/nst_kernels.cuh::u8_to_unit: q = k * fl(1/255); q += fma(-q, 255, k) * fl(1/255) equals the IEEE fp32 division k / 255
    (ToTensor) for all 256 byte values -- the multiply alone is wrong for 126 of them."""
    k = np.arange(256, dtype=np.float32)
    exact = (k / np.float32(255)).astype(np.float32)
    r = np.frombuffer(np.uint32(0x3b808081).tobytes(), dtype=np.float32)[0]
    assert r == np.float32(1) / np.float32(255)
    fma = lambda a, b, c: (a.astype(np.float64) * np.float64(b) + c.astype(np.float64)).astype(np.float32)
    q = (k * r).astype(np.float32)
    assert int((q != exact).sum()) > 100
    q2 = fma(fma(-q, 255.0, k), r, q)
    assert np.array_equal(q2, exact)


# ---- region masks / temporal post-processing: the restatements against the reference's own outputs -------------
def test_mask_patterns_rotation_and_organic_warp_match_reference():
    g = np.load(os.path.join(GOLDEN, "masks_patterns_60x84.npz"))
    for key in g.files:
        mode, c, s, f = key.split("_")
        got = O.pattern_masks(60, 84, mode, int(c[1:]), int(s[1:]), int(f[1:])).numpy()
        assert np.array_equal(got, g[key].astype(np.float32)), key
    r = np.load(os.path.join(GOLDEN, "masks_rotate_60x84.npz"))
    for key in r.files:
        if key.startswith("rot_"):
            assert np.array_equal(O.rotate_all_masks(torch.from_numpy(r["masks"]), float(key[4:])).numpy(), r[key]), key
    o = np.load(os.path.join(GOLDEN, "masks_organic_60x84.npz"))
    for mode in ("blob", "tentacle", "wave", "pulse"):
        assert np.array_equal(O.warp_all_masks_organic(torch.from_numpy(o["masks"]), mode, 0.15, 3.0, 1.5, 42, 7).numpy(), o[f"{mode}_f7"]), mode
    got = O.warp_all_masks_organic(torch.from_numpy(o["hard_masks"]).float(), "blob", 0.9, 5.0, 1.0, 5, 3).numpy()
    assert np.array_equal(got, o["gappy_f3"]) and int(o["gaps_before_fill"]) > 0


def test_lab_ema_and_motion_blend_match_reference():
    g = np.load(os.path.join(GOLDEN, "lab_ema_48x64.npz"))
    for tag, (sl, sc) in {"both": (True, True), "light": (True, False), "chroma": (False, True)}.items():
        assert np.array_equal(O.lab_ema(g["frames"], sl, sc, float(g["smooth_alpha"]), float(g["chroma_alpha"])), g[f"out_{tag}"]), tag
    m = np.load(os.path.join(GOLDEN, "motion_blend.npz"))
    for t in ("54x70", "40x64"):
        assert np.array_equal(O.motion_blend(m[f"styled_{t}"], m[f"orig_{t}"], m[f"flow_{t}"], float(m["blend"])), m[f"out_{t}"]), t


def test_lab_tables_are_pillows_conversion():
    """The 2^24-entry tables the LAB EMA kernel reads (temporal.lab_tables_host) against Pillow on random colours."""
    from PIL import Image
    from neuralstyletransferv1_b200.temporal import lab_tables_host
    rgb2lab, lab2rgb = lab_tables_host()
    px = np.random.default_rng(0).integers(0, 256, (64, 64, 3), dtype=np.uint8)
    idx = (px[..., 0].astype(np.uint32) << 16) | (px[..., 1].astype(np.uint32) << 8) | px[..., 2]
    unpack = lambda v: np.stack([v & 255, (v >> 8) & 255, (v >> 16) & 255], -1).astype(np.uint8)
    assert np.array_equal(unpack(rgb2lab[idx]), np.array(Image.fromarray(px, mode="RGB").convert("LAB")))
    assert np.array_equal(unpack(lab2rgb[idx]), np.array(Image.fromarray(px, mode="LAB").convert("RGB")))


CHAIN_CONFIGS = {      # tests/golden/make_golden_chain.py: CONFIGS
    "ema_lab_blend": dict(flow_ema_on=True, flow_alpha=0.7, smooth_lightness=True, smooth_chroma=True, smooth_alpha=0.7, chroma_alpha=0.85,
                          blend=0.8, motion_blend_on=False, composite_mode="keep"),
    "ema_motion": dict(flow_ema_on=True, flow_alpha=0.85, smooth_lightness=False, smooth_chroma=False, smooth_alpha=0.7, chroma_alpha=0.85,
                       blend=0.9, motion_blend_on=True, composite_mode="keep"),
    "mask_light": dict(flow_ema_on=True, flow_alpha=0.6, smooth_lightness=True, smooth_chroma=False, smooth_alpha=0.5, chroma_alpha=0.85,
                       blend=0.75, motion_blend_on=True, composite_mode="replace"),
    "plain": dict(flow_ema_on=False, flow_alpha=0.7, smooth_lightness=False, smooth_chroma=False, smooth_alpha=0.7, chroma_alpha=0.85,
                  blend=1.0, motion_blend_on=False, composite_mode="keep"),
}


def load_chain_fixture():
    g = np.load(os.path.join(GOLDEN, "post_chain_48x72.npz"))
    out01 = torch.from_numpy(g["out01_q4096"].astype(np.float32) / 4096.0)
    flows = [g["flows"][i].astype(np.float32) if g["has_flow"][i] else None for i in range(len(out01))]
    masks = g["masks_q1024"].astype(np.float32) / 1024.0
    return g, out01, g["orig"], flows, masks


@pytest.mark.parametrize("name", sorted(CHAIN_CONFIGS))
def test_post_chain_matches_the_reference_loop(name):
    """The loop tail restated from its parts (flow EMA, pack, LAB EMA, composite / blends) against the literal loop of
    pipeline.py:1884-2094 run over a 5-frame clip (one frame without a flow estimate)."""
    g, out01, orig, flows, masks = load_chain_fixture()
    got = O.post_chain(out01, orig, flows, masks if name == "mask_light" else None, **CHAIN_CONFIGS[name])
    assert np.array_equal(got, g["out_" + name]), name


def test_prepare_region_crops_matches_reference():
    g = np.load(os.path.join(GOLDEN, "region_crops_96x128.npz"))
    masks = torch.from_numpy(g["masks"])
    cfgs = [dict(model_indices=[0], model_weights=[1.0], scale=1.0)] * masks.shape[0]
    for pad in (12, 64):
        crops = O.prepare_region_crops(masks, cfgs, 96, 128, padding=pad)
        assert [list(c["bbox"]) for c in crops] == g[f"bbox_p{pad}"].tolist()
        assert [list(c["padded_bbox"]) for c in crops] == g[f"padded_p{pad}"].tolist()
