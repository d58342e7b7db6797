"""GPU parity of the region-mask generators / animation (csrc/masks.cu) and of the LAB EMA / motion-adaptive blend (csrc/temporal.cu)
against the fixtures the reference's own functions produced (tests/golden/make_golden_masks.py) and, at 1080p, the CPU oracle.

Tolerances.  Index / byte work (rectangle patterns, warpAffine / remap fixed-point sampling, LAB tables, uint8 packs): bit-exact.
Analytic patterns: the band label of a pixel is decided by atan2f / sinf, whose device versions are within 2 ulp of torch's; a pixel whose
field value lies within that of a band edge may change band -- at most MAX_FLIPS pixels per frame, everything else identical.  The
separable feather differs from the reference's 2-D kernel by fp32 summation order: 2e-5."""
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN
from neuralstyletransferv1_b200 import synth
from oracle import nst_oracle as O

pytestmark = pytest.mark.gpu
MAX_FLIPS = 8
TRANSCENDENTAL = ("radial", "waves", "spiral")


def _hard(m, count):
    return (m.cpu().numpy() > 0.5).astype(np.uint8)


def test_every_pattern_matches_the_reference_generator():
    import neuralstyletransferv1_b200 as m
    g = np.load(os.path.join(GOLDEN, "masks_patterns_60x84.npz"))
    H, W = 60, 84
    for key in g.files:
        mode, c, s, f = key.split("_")
        count, seed, feather = int(c[1:]), int(s[1:]), int(f[1:])
        got = m.generate_region_masks(H, W, mode, count, seed=seed, feather=feather).cpu().numpy()
        assert got.shape == (count, H, W)
        ref = g[key].astype(np.float32)
        if feather == 0:
            flips = int((got != ref).sum())
            assert flips <= (2 * MAX_FLIPS if mode in TRANSCENDENTAL or mode == "random" else 0), (key, flips)
            assert set(np.unique(got).tolist()) <= {0.0, 1.0}
        else:
            hard = m.generate_region_masks(H, W, mode, count, seed=seed, feather=0).cpu().numpy()
            if np.array_equal(hard, g[f"{mode}_{c}_{s}_f0"].astype(np.float32)):       # same cells: only the summation order differs
                assert float(np.abs(got - ref).max()) <= 2e-5, key
            else:
                assert float(np.abs(got - ref).max()) <= 0.1, key


def test_analytic_patterns_at_1080p():
    import neuralstyletransferv1_b200 as m
    g = np.load(os.path.join(GOLDEN, "masks_patterns_1080p.npz"))
    pos = torch.from_numpy(g["pos"]).cuda()
    for mode in ("radial", "waves", "spiral", "concentric", "diagonal"):
        hard = m.generate_region_masks(1080, 1920, mode, 4, seed=42, feather=0)
        lab = torch.where(hard.sum(0) > 0, hard.argmax(0), torch.full((1080, 1920), 4, device="cuda"))
        hist = torch.bincount(lab.flatten(), minlength=5).cpu().numpy()
        tol = MAX_FLIPS if mode in TRANSCENDENTAL else 0
        assert int(np.abs(hist - g[f"{mode}_hist"]).sum()) <= 2 * tol, (mode, hist, g[f"{mode}_hist"])
        assert int((lab.flatten()[pos].cpu().numpy() != g[f"{mode}_labels_at_pos"]).sum()) <= (1 if tol else 0), mode
    soft = m.generate_region_masks(1080, 1920, "spiral", 4, seed=42, feather=20)
    d = np.abs(soft.reshape(4, -1)[:, pos].cpu().numpy() - g["spiral_f20_values"])
    assert int((d > 2e-5).sum()) <= 16 and float(d.max()) <= 5e-3
    assert float((soft.sum(0) - 1).abs().max()) <= 1e-4


def test_rotation_is_opencvs_bit_for_bit():
    import neuralstyletransferv1_b200 as m
    g = np.load(os.path.join(GOLDEN, "masks_rotate_60x84.npz"))
    masks = torch.from_numpy(g["masks"]).cuda()
    for key in g.files:
        if key.startswith("rot_"):
            assert np.array_equal(m.rotate_all_masks(masks, float(key[4:])).cpu().numpy(), g[key]), key
    assert m.rotate_all_masks(masks, 0) is masks
    big = m.generate_region_masks(1080, 1920, "voronoi", 4, seed=42, feather=20)
    want = O.rotate_all_masks(big.cpu(), 17.25)
    assert torch.equal(m.rotate_all_masks(big, 17.25).cpu(), want)


@pytest.mark.parametrize("mode", ["blob", "tentacle", "wave", "pulse", "gappy"])
def test_organic_warp_matches_the_reference(mode):
    import neuralstyletransferv1_b200 as m
    g = np.load(os.path.join(GOLDEN, "masks_organic_60x84.npz"))
    if mode == "gappy":                                   # amplitude 0.9 on hard masks: 1766 gap pixels, the dilation fill runs
        masks, morph, frame, want = g["hard_masks"].astype(np.float32), m.MorphAnimation(True, 1.0, 0.9, 5.0, 3, "blob", 5), 3, g["gappy_f3"]
    else:
        masks, morph, frame, want = g["masks"], m.MorphAnimation(True, 1.5, 0.15, 3.0, 3, mode, 42), 7, g[f"{mode}_f7"]
    got = m.warp_all_masks_organic(torch.from_numpy(masks).cuda(), morph, frame).cpu().numpy()
    # double sin / cos of the device vs numpy's: identical after the fp32 narrowing and the 1/32-pixel map quantisation except, at most,
    # for isolated pixels
    bad = np.abs(got - want) > 0
    assert int(bad.any(0).sum()) <= MAX_FLIPS, int(bad.any(0).sum())
    assert float(np.abs(got.sum(0) - 1).max()) <= 1e-5
    off = m.MorphAnimation(enabled=False)
    t = torch.from_numpy(masks).cuda()
    assert m.warp_all_masks_organic(t, off, 3) is t


def test_organic_warp_at_1080p_against_the_oracle():
    import neuralstyletransferv1_b200 as m
    big = m.generate_region_masks(1080, 1920, "voronoi", 4, seed=42, feather=20)
    for mode, amp in (("blob", 0.15), ("pulse", 0.1)):
        morph = m.MorphAnimation(True, 1.0, amp, 3.0, 3, mode, 42)
        got = m.warp_all_masks_organic(big, morph, 12).cpu()
        want = O.warp_all_masks_organic(big.cpu(), mode, amp, 3.0, 1.0, 42, 12)
        bad = (got != want).any(0)
        assert int(bad.sum()) <= 64, int(bad.sum())             # of 2 073 600 pixels
        assert float((got - want).abs().max()) <= 0.05


def test_lab_ema_is_pillows_bit_for_bit():
    import neuralstyletransferv1_b200 as m
    g = np.load(os.path.join(GOLDEN, "lab_ema_48x64.npz"))
    frames = torch.from_numpy(g["frames"]).cuda()
    for tag, (sl, sc) in {"both": (True, True), "light": (True, False), "chroma": (False, True)}.items():
        ema = m.LabEma(sl, sc, float(g["smooth_alpha"]), float(g["chroma_alpha"]))
        assert np.array_equal(ema(frames).cpu().numpy(), g[f"out_{tag}"]), tag
        ema.reset()                                                   # the state carries across calls: 1 + 3 frames == 4 frames
        parts = torch.cat([ema(frames[0])[None], ema(frames[1:])])
        assert np.array_equal(parts.cpu().numpy(), g[f"out_{tag}"]), tag
    assert m.LabEma(False, False)(frames) is frames
    clip = synth.synthetic_frames_u8(3, 1080, 1920, seed=5)
    got = m.LabEma(True, True, 0.6, 0.8)(clip.cuda()).cpu().numpy()
    assert np.array_equal(got, O.lab_ema(clip.numpy(), True, True, 0.6, 0.8))


def test_motion_blend_is_opencvs_bit_for_bit():
    import neuralstyletransferv1_b200 as m
    g = np.load(os.path.join(GOLDEN, "motion_blend.npz"))
    for t in ("54x70", "40x64"):
        got = m.motion_blend(torch.from_numpy(g[f"styled_{t}"]).cuda(), torch.from_numpy(g[f"orig_{t}"]).cuda(), torch.from_numpy(g[f"flow_{t}"]).cuda(),
                             float(g["blend"]))
        assert np.array_equal(got.cpu().numpy(), g[f"out_{t}"]), t
    s = synth.synthetic_frames_u8(2, 1080, 1920, seed=6)
    flow = torch.randn(1080, 1920, 2, generator=torch.Generator().manual_seed(1)) * 5
    got = m.motion_blend(s[0].cuda(), s[1].cuda(), flow.cuda(), 0.85)
    assert np.array_equal(got.cpu().numpy(), O.motion_blend(s[0].numpy(), s[1].numpy(), flow.numpy(), 0.85))


# ---- the whole per-frame tail (post_chain.PostChain) against the literal reference loop ------------------------------------------------
@pytest.mark.parametrize("name", ["ema_lab_blend", "ema_motion", "mask_light", "plain"])
def test_post_chain_matches_the_reference_loop_bit_for_bit(name):
    import neuralstyletransferv1_b200 as m
    from test_oracle import CHAIN_CONFIGS, load_chain_fixture
    g, out01, orig, flows, masks = load_chain_fixture()
    c = CHAIN_CONFIGS[name]
    mk = lambda: m.PostChain(flow_ema=c["flow_ema_on"], flow_alpha=c["flow_alpha"], smooth_lightness=c["smooth_lightness"],
                             smooth_chroma=c["smooth_chroma"], smooth_alpha=c["smooth_alpha"], chroma_alpha=c["chroma_alpha"], blend=c["blend"],
                             motion_blend=c["motion_blend_on"], composite_mode=c["composite_mode"])
    dflows = [torch.from_numpy(f).cuda() if f is not None else None for f in flows]
    dmasks = torch.from_numpy(masks).cuda() if name == "mask_light" else None
    chain = mk()
    got = chain(out01.cuda(), torch.from_numpy(orig).cuda(), dflows, dmasks).cpu().numpy()
    assert np.array_equal(got, g["out_" + name]), name
    chain = mk()                                           # the state carries across calls: 2 + 3 frames == 5 frames
    parts = [chain(out01[a:b].cuda(), torch.from_numpy(orig[a:b]).cuda(), dflows[a:b], dmasks[a:b] if dmasks is not None else None)
             for a, b in ((0, 2), (2, 5))]
    assert np.array_equal(torch.cat(parts).cpu().numpy(), g["out_" + name]), name


def test_post_chain_at_1080p_behind_the_generator():
    """Generator -> fp32 out01 -> PostChain for a 3-frame 1080p clip: bit-identical to the oracle's chain applied to the SAME device out01
    (the tail is byte work), and the frames stay within the network tolerance of the all-CPU path."""
    import neuralstyletransferv1_b200 as m
    model = m.TransformerNet().to("cuda:0")
    sd = synth.make_state_dict("johnson", 0)
    model.load_state_dict(sd)
    model.eval()
    frames = synth.synthetic_frames_u8(3, 1080, 1920, seed=9)
    x = frames.permute(0, 3, 1, 2).float().cuda()                       # raw_255 input (pipeline.py:1473)
    with torch.no_grad():
        out01 = (model(x) / 255.0).clamp(0, 1)
    gen = torch.Generator().manual_seed(3)
    flows = [None] + [torch.randn(1080, 1920, 2, generator=gen) * 3 for _ in range(2)]
    cfg = dict(flow_alpha=0.8, smooth_lightness=True, smooth_chroma=True, smooth_alpha=0.7, chroma_alpha=0.85, blend=0.9)
    chain = m.PostChain(flow_ema=True, motion_blend=True, **cfg)
    got = chain(out01, frames.cuda(), [f.cuda() if f is not None else None for f in flows]).cpu().numpy()
    want = O.post_chain(out01.cpu(), frames.numpy(), [f.numpy() if f is not None else None for f in flows], None, flow_ema_on=True,
                        motion_blend_on=True, **cfg)
    assert np.array_equal(got, want)


BLEND_CASES = {      # tests/golden/make_golden_blend.py: CASES
    "static": dict(mode="voronoi", region_count=4, assignment="sequential", feather=6, seed=42, original_chance=0.0, rotation_rate=0.0, morph=None),
    "rotating": dict(mode="voronoi", region_count=4, assignment="random", feather=8, seed=7, original_chance=0.3, rotation_rate=9.0, morph=None),
    "morphing": dict(mode="grid", region_count=4, assignment="weighted", feather=8, seed=11, original_chance=0.0, rotation_rate=0.0,
                     morph=dict(enabled=True, speed=1.5, amplitude=0.12, frequency=3.0, octaves=3, mode="blob", seed=42)),
    "both": dict(mode="radial", region_count=3, assignment="sequential", feather=10, seed=3, original_chance=0.5, rotation_rate=-4.0,
                 morph=dict(enabled=True, speed=1.0, amplitude=0.1, frequency=2.0, octaves=3, mode="pulse", seed=1)),
}


@pytest.mark.parametrize("name", sorted(BLEND_CASES))
def test_region_blender_matches_blend_by_regions(name):
    """RegionBlender (masks kept for the clip, rotation + re-feather, organic morph + re-feather, assignment, one batched composite) against
    three frames of the reference's blend_by_regions.  The feathers are separable here and 2-D in the reference (mask values agree to ~1e-6).
    The sources are uint8 frames, so inside a region the composite (m * k/255) / m sits exactly ON a uint8 edge and the last bit of m
    decides whether it truncates to k or k - 1: at most 1 level, on a few per cent of the values (with identical masks the composite is
    bit-exact: tests/test_gpu_post.py)."""
    import neuralstyletransferv1_b200 as m
    g = np.load(os.path.join(GOLDEN, "blend_by_regions_60x84.npz"))
    c = BLEND_CASES[name]
    styled = [torch.from_numpy(g[f"styled{k}"]).cuda() for k in range(3)]
    blender = m.RegionBlender(3, 60, 84, mode=c["mode"], region_count=c["region_count"], assignment=c["assignment"], weights=[0.5, 0.3, 0.2],
                              feather=c["feather"], seed=c["seed"], original_chance=c["original_chance"], rotation_rate=c["rotation_rate"],
                              morph=m.MorphAnimation(**c["morph"]) if c["morph"] else None)
    got = blender(styled, torch.from_numpy(g["orig"]).cuda(), frame_idx=4).cpu().numpy().astype(int)
    want = g["out_" + name].astype(int)
    d = np.abs(got - want)
    assert d.max() <= 1 and (d > 0).mean() < 0.05, (int(d.max()), float((d > 0).mean()))


def test_region_blender_with_per_region_model_blends():
    """blend_by_regions_advanced's composite (region_blend.py:1832-1957 -> composite_regions_advanced): the blender with RegionConfigs is
    the batched composite with those indices / weights over the blender's own masks."""
    import neuralstyletransferv1_b200 as m
    from neuralstyletransferv1_b200 import engine
    g = np.load(os.path.join(GOLDEN, "blend_by_regions_60x84.npz"))
    styled = [torch.from_numpy(g[f"styled{k}"]).cuda() for k in range(3)]
    orig = torch.from_numpy(g["orig"]).cuda()
    cfgs = [m.RegionConfig([0, 1], [0.7, 0.3]), m.RegionConfig([1, -1], [0.5, 0.5]), m.RegionConfig([2], [1.0]), m.RegionConfig([2, 0], [0.2, 0.8])]
    blender = m.RegionBlender(3, 60, 84, mode="voronoi", region_count=4, feather=6, seed=42, configs=cfgs)
    got = blender(styled, orig, frame_idx=0)
    masks = m.generate_region_masks(60, 84, "voronoi", 4, seed=42, feather=6)
    want = engine.composite_regions_batch(styled, masks, [c.model_indices for c in cfgs], [c.model_weights for c in cfgs], original=orig)
    assert torch.equal(got, want)
    ref = O.composite_regions_advanced([s[0].cpu().permute(2, 0, 1).float() / 255 for s in styled], [mk[None, None] for mk in masks.cpu()],
                                       [c.model_indices for c in cfgs], [c.model_weights for c in cfgs], orig[0].cpu().permute(2, 0, 1).float() / 255)
    assert torch.equal(got[0].cpu(), O.pack_u8(ref[None])[0])
