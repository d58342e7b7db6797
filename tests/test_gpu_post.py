"""GPU parity of everything between the network output and the encoder (csrc/postproc.cu, frame_pipeline.py) against the
fixtures the reference's own functions produced (tests/golden/make_golden.py) and the CPU oracle.  Byte / fp32 elementwise
work: bit-exact unless stated."""
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN
from neuralstyletransferv1_b200 import synth

pytestmark = pytest.mark.gpu


def _johnson(seed=0):
    import neuralstyletransferv1_b200 as m
    model = m.TransformerNet().to("cuda:0")
    sd = synth.make_state_dict("johnson", seed)
    model.load_state_dict(sd)
    model.eval()
    return model, sd


@pytest.mark.parametrize("hi,wi,h,w", [(272, 484, 271, 481), (252, 272, 250, 270), (1080, 1920, 1079, 1918), (120, 160, 120, 160)])
def test_resize_post_pack_is_bit_exact(hi, wi, h, w):
    """pipeline.py:1512-1516 + 2040-2094 on a given fp32 out01 frame: bilinear resize (ATen's arithmetic), mask composite, blend, pack."""
    from neuralstyletransferv1_b200.engine import resize_post_pack
    from oracle import nst_oracle as O
    g = torch.Generator().manual_seed(hi + w)
    out01 = torch.rand(2, 3, hi, wi, generator=g)
    frames = torch.randint(0, 256, (2, h, w, 3), dtype=torch.uint8, generator=g)
    mask = torch.rand(2, h, w, generator=g)
    x01 = frames.permute(0, 3, 1, 2).float() / 255.0
    rs = torch.nn.functional.interpolate(out01, size=(h, w), mode="bilinear", align_corners=False) if (hi, wi) != (h, w) else out01
    for blend, use_mask, mode in ((1.0, False, "keep"), (0.7, False, "keep"), (1.0, True, "keep"), (0.4, True, "replace")):
        want = O.post_u8(rs, x01, blend, mask if use_mask else None, mode)
        got = resize_post_pack(out01.cuda(), h, w, frames.cuda(), mask.cuda() if use_mask else None, blend, mode).cpu()
        assert torch.equal(got, want), (blend, use_mask, mode, int((got.int() - want.int()).abs().max()))


def test_fit_surface_vs_reference_fixture_and_oracle():
    """Odd-sized frames through the whole fused path: fixture made with the pipeline's literal arithmetic around the real module."""
    from oracle import nst_oracle as O
    g = np.load(os.path.join(GOLDEN, "fit_johnson_50x70.npz"))
    model, sd = _johnson(int(g["weight_seed"]))
    frames = torch.from_numpy(g["frames"])
    alpha = torch.from_numpy(g["alpha"])
    got = model.stylize_u8(frames.cuda(), "raw_255", blend=float(g["blend"]), mask=alpha.cuda(), fit=True).cpu()
    assert tuple(got.shape) == (1, 50, 70, 3)
    mx, psnr = O.u8_metrics(got[0], torch.from_numpy(g["out_u8"]))
    assert mx <= 2 and psnr >= 45.0, (mx, psnr)
    mx, psnr = O.u8_metrics(model.stylize_u8(frames.cuda(), "raw_255", fit=True).cpu()[0], torch.from_numpy(g["out_u8_plain"]))
    assert mx <= 2 and psnr >= 45.0, (mx, psnr)
    # a larger odd size, two frames (micro-batch loop), ReCoNet (tanh preset)
    import neuralstyletransferv1_b200 as m
    rc = m.ReCoNet().to("cuda:0")
    sdr = synth.make_state_dict("reconet", 1)
    rc.load_state_dict(sdr)
    fr = synth.synthetic_frames_u8(2, 133, 201, seed=8)
    got = rc.stylize_u8(fr.cuda(), "tanh", blend=0.75, fit=True).cpu()
    want = O.stylize_u8_fit("reconet", sdr, fr, "tanh", blend=0.75)
    mx, psnr = O.u8_metrics(got, want)
    assert tuple(got.shape) == (2, 133, 201, 3) and mx <= 2 and psnr >= 45.0, (mx, psnr)
    # multiples of 4: fit is the plain path
    f4 = synth.synthetic_frames_u8(1, 64, 80, seed=2).cuda()
    assert torch.equal(model.stylize_u8(f4, "raw_255", fit=True), model.stylize_u8(f4, "raw_255"))


def test_composite_batch_multiscale_masks_and_frames():
    """composite_regions_advanced with per-region blends and a half-resolution scale (fixture by the reference), as ONE batched launch
    over several frames; fp32 / fp16 / uint8 masks; per-frame mask sets."""
    import neuralstyletransferv1_b200 as m
    from oracle import nst_oracle as O
    g = np.load(os.path.join(GOLDEN, "composite_advanced_60x84.npz"))
    H, W = 60, 84
    to_u8 = lambda t: (t * 255).to(torch.uint8)
    full8 = [to_u8(torch.from_numpy(a)) for a in g["full"]]            # [3,H,W] uint8
    half8 = [to_u8(torch.from_numpy(a)) for a in g["half"]]
    orig8 = to_u8(torch.from_numpy(g["original"]))
    masks = torch.from_numpy(g["masks"])
    idx = [[int(i) for i in ix if i > -2] for ix in g["idx"]]
    wts = [[float(w) for i, w in zip(ix, ws) if i > -2] for ix, ws in zip(g["idx"], g["wts"])]
    scales = [float(s) for s in g["scales"]]
    # the library takes one source list: sources 0-2 full resolution, 3-5 the half-resolution renditions of the same models
    lib_idx = [[(i if (i < 0 or sc == 1.0) else i + 3) for i in ix] for ix, sc in zip(idx, scales)]
    N = 3
    rot = lambda t, n: torch.roll(t, shifts=5 * n, dims=-1)            # frame n = the fixture's frame shifted by 5n columns
    srcs = [torch.stack([rot(s, n) for n in range(N)]).permute(0, 2, 3, 1).contiguous().cuda() for s in full8 + half8]
    orig = torch.stack([rot(orig8, n) for n in range(N)]).permute(0, 2, 3, 1).contiguous().cuda()
    got = m.composite_regions_batch(srcs, masks.cuda(), lib_idx, wts, original=orig).cpu()
    for n in range(N):
        f = lambda t: rot(t, n).float() / 255.0
        want = O.composite_regions_multiscale({1.0: [f(s) for s in full8], 0.5: [f(s) for s in half8]}, [mk[None, None] for mk in masks],
                                              list(zip(idx, wts, scales)), f(orig8), H, W)
        assert torch.equal(got[n], O.pack_u8(want[None])[0]), n
    # unquantised check against the reference's own output: u8 sources cost at most 1 level + the truncation
    ref = O.pack_u8(torch.from_numpy(g["out"])[None])[0]
    assert int((got[0].int() - ref.int()).abs().max()) <= 2
    # fp16 masks are exact when the mask values are fp16-representable; uint8 masks mean value / 255
    m16 = masks.half()
    got16 = m.composite_regions_batch(srcs, m16.cuda(), lib_idx, wts, original=orig).cpu()
    f = lambda t: t.float() / 255.0
    want16 = O.composite_regions_multiscale({1.0: [f(s) for s in full8], 0.5: [f(s) for s in half8]}, [mk.float()[None, None] for mk in m16],
                                            list(zip(idx, wts, scales)), f(orig8), H, W)
    assert torch.equal(got16[0], O.pack_u8(want16[None])[0])
    m8 = (masks * 255).to(torch.uint8)
    got8 = m.composite_regions_batch(srcs, m8.cuda(), lib_idx, wts, original=orig).cpu()
    want8 = O.composite_regions_multiscale({1.0: [f(s) for s in full8], 0.5: [f(s) for s in half8]}, [(mk.float() / 255.0)[None, None] for mk in m8],
                                           list(zip(idx, wts, scales)), f(orig8), H, W)
    assert torch.equal(got8[0], O.pack_u8(want8[None])[0])
    # per-frame mask sets [N,R,H,W]
    per = torch.stack([torch.roll(masks, shifts=3 * n, dims=-2) for n in range(N)])
    gotp = m.composite_regions_batch(srcs, per.cuda(), lib_idx, wts, original=orig).cpu()
    f2 = lambda t: rot(t, 2).float() / 255.0
    wantp = O.composite_regions_multiscale({1.0: [f2(s) for s in full8], 0.5: [f2(s) for s in half8]}, [mk[None, None] for mk in per[2]],
                                           list(zip(idx, wts, scales)), f2(orig8), H, W)
    assert torch.equal(gotp[2], O.pack_u8(wantp[None])[0])
    # the batched call equals the single-frame entry point on full-resolution sources
    single = m.composite_regions([s[0] for s in srcs[:3]], masks.cuda(), [[0, 1], [1, -1], [2], [2, 0]], wts, original=orig[0]).cpu()
    batch = m.composite_regions_batch(srcs[:3], masks.cuda(), [[0, 1], [1, -1], [2], [2, 0]], wts, original=orig).cpu()
    assert int((single.int() - batch[0].int()).abs().max()) <= 1       # the older kernel multiplies by mask before summing models


def test_composite_from_crops_matches_reference_bit_for_bit():
    import neuralstyletransferv1_b200 as m
    from test_oracle import _crops_from_fixture
    g = np.load(os.path.join(GOLDEN, "composite_crops_180x240.npz"))
    cuda = lambda d: {k: ({kk: vv.cuda() for kk, vv in v.items()} if isinstance(v, dict) else v) for k, v in d.items()}
    crops, styled = _crops_from_fixture(g)
    for c in crops:
        c["crop_mask"] = c["crop_mask"].cuda()
    out, out8 = m.composite_from_crops(cuda(styled), crops, torch.from_numpy(g["original"]).cuda(), 180, 240, return_u8=True)
    ref = torch.from_numpy(g["out_with_original"])
    assert torch.equal(out.cpu(), ref)
    assert torch.equal(out8.cpu(), (ref * 255).to(torch.uint8).permute(1, 2, 0))
    crops, styled = _crops_from_fixture(g, only=set(int(r) for r in g["no_orig_regions"]))
    for c in crops:
        c["crop_mask"] = c["crop_mask"].cuda()
    out = m.composite_from_crops(cuda(styled), crops, None, 180, 240)
    assert torch.equal(out.cpu(), torch.from_numpy(g["out_without_original"]))       # max-pool 5 / 11 / 21 gap fill
    with pytest.raises(ValueError, match="no original"):
        crops2, styled2 = _crops_from_fixture(g)
        for c in crops2:
            c["crop_mask"] = c["crop_mask"].cuda()
        m.composite_from_crops(cuda(styled2), crops2, None, 180, 240)


def test_flow_ema_matches_opencv_bit_for_bit():
    import neuralstyletransferv1_b200 as m
    from oracle import nst_oracle as O
    g = np.load(os.path.join(GOLDEN, "flow_ema_54x70.npz"))
    t = lambda a: torch.from_numpy(a).cuda()
    fused = m.flow_ema(t(g["curr"]), t(g["prev"]), t(g["flow"]), float(g["alpha"])).cpu().numpy()
    assert np.array_equal(fused, g["fused"])
    # 1080p, flow that leaves the frame (replicate border), alpha outside [0,1] is clamped, a chain of three frames
    rng = np.random.default_rng(9)
    H, W = 1080, 1920
    prev = rng.random((H, W, 3), dtype=np.float32)
    for step, alpha in enumerate((0.5, 1.3, 0.0)):
        curr = rng.random((H, W, 3), dtype=np.float32)
        flow = (rng.standard_normal((H, W, 2)) * (8 + 40 * step)).astype(np.float32)
        want = O.flow_ema(curr, prev, flow, alpha)
        got = m.flow_ema(t(curr), t(prev), t(flow), alpha).cpu().numpy()
        assert np.array_equal(got, want), step
        prev = want


def test_frame_pipeline_preserves_order_and_matches_per_frame_calls():
    import neuralstyletransferv1_b200 as m
    model, _ = _johnson(0)
    frames = synth.synthetic_frames_u8(21, 48, 64, seed=5)             # 21 frames: chunks of 8, 8, 5
    want = torch.cat([model.stylize_u8(frames[i:i + 1].cuda(), "raw_255") for i in range(21)]).cpu()
    pipe = m.FramePipeline(model, "raw_255", chunk=8, slots=2)
    got = pipe(frames)                                                 # pageable in, pinned out
    assert got.is_pinned() and torch.equal(got, want)
    out = torch.empty_like(frames)                                     # pageable out: staged
    assert torch.equal(pipe(frames.pin_memory(), out=out), want)
    assert torch.equal(pipe(frames.numpy()), want)                     # numpy in, second call reuses the staging
    # >= 4 chunks: the optional ramped schedule (a quarter-size first and last chunk) -- here 1, 4, 4, 4, 4, 3, 1 -- with 2 and 3 slots, staged and direct
    for slots in (2, 3):
        p4 = m.FramePipeline(model, "raw_255", chunk=4, slots=slots, ramp=True)
        assert torch.equal(p4(frames), want)
        assert torch.equal(p4(frames.pin_memory(), out=torch.empty_like(frames)), want)
    assert torch.equal(m.FramePipeline(model, "raw_255", chunk=4)(frames), want)
    # odd size + mask + blend go through the resize-back path, in order
    odd = synth.synthetic_frames_u8(5, 50, 70, seed=6)
    mask = torch.rand(5, 50, 70, generator=torch.Generator().manual_seed(2))
    p2 = m.FramePipeline(model, "raw_255", chunk=2, blend=0.6)
    got = p2(odd, masks=mask)
    want = torch.cat([model.stylize_u8(odd[i:i + 1].cuda(), "raw_255", blend=0.6, mask=mask[i:i + 1].cuda(), fit=True) for i in range(5)]).cpu()
    assert tuple(got.shape) == (5, 50, 70, 3) and torch.equal(got, want)


def test_multi_gpu_pipeline_one_process():
    """One worker thread + replica + stream set per device.  With one GPU the two workers share it; with >= 2 GPUs
    (gpurun --gpus 2) they run on different devices.  Output order and values equal per-frame calls."""
    import neuralstyletransferv1_b200 as m
    sd = synth.make_state_dict("johnson", 0)

    def factory(d):
        model = m.TransformerNet().to(f"cuda:{d}")
        model.load_state_dict(sd)
        model.eval()
        return model

    ndev = torch.cuda.device_count()
    devices = [0, 1] if ndev >= 2 else [0, 0]
    frames = synth.synthetic_frames_u8(13, 40, 56, seed=7)
    ref_model = factory(0)
    want = torch.cat([ref_model.stylize_u8(frames[i:i + 1].cuda(), "raw_255") for i in range(13)]).cpu()
    mg = m.MultiGpuFramePipeline(factory, "raw_255", devices=devices, chunk=4)
    assert torch.equal(mg(frames), want)
    assert torch.equal(mg(frames[:1]), want[:1])                       # fewer frames than workers


def test_voronoi_masks_match_the_reference_generator():
    """generate_region_masks(..., "voronoi", 4, seed=42, feather=...) (region_blend.py:925, 174-236): same host RNG sequence for the seed
    points, labelling + reflect-padded Gaussian feather on the device.  The reference convolves with the 2-D outer-product kernel
    (1681 taps at feather 20), the device with the separable form: fp32 summation order differs, tolerance 2e-5."""
    import neuralstyletransferv1_b200 as m
    g = np.load(os.path.join(GOLDEN, "composite_voronoi4_48x64.npz"))
    got = m.voronoi_masks(48, 64, 4, seed=42, feather=4).cpu()
    assert float((got - torch.from_numpy(g["masks"])).abs().max()) <= 2e-5
    s = np.load(os.path.join(GOLDEN, "voronoi4_1080p_samples.npz"))
    big = m.voronoi_masks(1080, 1920, 4, seed=42, feather=20)
    vals = big.reshape(4, -1)[:, torch.from_numpy(s["pos"]).cuda()].cpu().numpy()
    assert np.abs(vals - s["values"]).max() <= 2e-5
    assert np.allclose(big.double().sum((1, 2)).cpu().numpy(), s["sums"], rtol=1e-6)
    assert np.array_equal(torch.bincount(big.argmax(0).flatten(), minlength=4).cpu().numpy(), s["argmax_hist"])
    assert float((big.sum(0) - 1).abs().max()) <= 1e-4          # the cells partition the frame
    hard = m.voronoi_masks(48, 64, 4, seed=42, feather=0).cpu()
    assert set(hard.unique().tolist()) == {0.0, 1.0} and torch.equal(hard.sum(0), torch.ones(48, 64))


@pytest.mark.parametrize("mask_dtype", [torch.float32, torch.float16, torch.uint8])
def test_composite_batch_vector_path_is_bit_exact(mask_dtype):
    """All sources at full resolution and H*W % 4 == 0: the 4-pixels-per-thread kernel; same arithmetic as the reference's tensor ops."""
    import neuralstyletransferv1_b200 as m
    from oracle import nst_oracle as O
    g = torch.Generator().manual_seed(31)
    N, K, R, H, W = 3, 4, 4, 36, 52
    srcs = [torch.randint(0, 256, (N, H, W, 3), dtype=torch.uint8, generator=g) for _ in range(K)]
    orig = torch.randint(0, 256, (N, H, W, 3), dtype=torch.uint8, generator=g)
    masks = torch.rand(R, H, W, generator=g)
    masks = (masks / masks.sum(0, keepdim=True))
    masks = (masks * 255).to(torch.uint8) if mask_dtype == torch.uint8 else masks.to(mask_dtype)
    mf = masks.float() / 255.0 if mask_dtype == torch.uint8 else masks.float()
    idx, wts = [[0, 1], [1, -1], [2], [3, 0]], [[0.7, 0.3], [0.5, 0.5], [1.0], [0.2, 0.8]]
    got = m.composite_regions_batch([s.cuda() for s in srcs], masks.cuda(), idx, wts, original=orig.cuda()).cpu()
    f = lambda t: t.permute(2, 0, 1).float() / 255.0
    for n in range(N):
        want = O.composite_regions_advanced([f(s[n]) for s in srcs], [mk[None, None] for mk in mf], idx, wts, f(orig[n]))
        assert torch.equal(got[n], O.pack_u8(want[None])[0]), n


def test_region_crops_bboxes_and_crop_styling():
    """SURVEY 8f N4.  (1) prepare_region_crops on the device against the reference's boxes (tight box of mask > 0.01, padding, the
    full-frame box of an empty mask, values just either side of the threshold).  (2) RegionCropStyler -- crop, optional down-scale,
    preset, generator, post, composite_from_crops -- against the all-CPU oracle of pipeline.py:1293-1400: the frame goes through two
    generators, so the bar is the network tolerance (max |d| <= 2 on the uint8 pack, PSNR >= 45 dB)."""
    import neuralstyletransferv1_b200 as m
    from oracle import nst_oracle as O
    g = np.load(os.path.join(GOLDEN, "region_crops_96x128.npz"))
    masks = torch.from_numpy(g["masks"]).cuda()
    cfgs = [m.RegionConfig([0], [1.0], 1.0) for _ in range(masks.shape[0])]
    for pad in (12, 64):
        crops = m.prepare_region_crops(masks, cfgs, 96, 128, padding=pad)
        assert [list(c.bbox) for c in crops] == g[f"bbox_p{pad}"].tolist()
        assert [list(c.padded_bbox) for c in crops] == g[f"padded_p{pad}"].tolist()
        for c in crops:
            x1, y1, x2, y2 = c.padded_bbox
            assert torch.equal(c.crop_mask, masks[c.region_idx, y1:y2, x1:x2])

    H, W = 180, 240
    rmasks = m.generate_region_masks(H, W, "voronoi", 3, seed=5, feather=5)
    rmasks = rmasks * (torch.arange(W, device="cuda").view(1, 1, W) < W - 40)            # a strip no region covers: gap fill by the original
    cfgs = [m.RegionConfig([0], [1.0], 1.0), m.RegionConfig([1, -1], [0.6, 0.4], 1.0), m.RegionConfig([0, 1], [0.5, 0.5], 0.5)]
    sds = [synth.make_state_dict("johnson", 0), synth.make_state_dict("johnson", 1)]
    models = []
    for sd in sds:
        mod = m.TransformerNet().to("cuda:0")
        mod.load_state_dict(sd)
        mod.eval()
        models.append(mod)
    presets = ["raw_255", "imagenet_255"]
    frame = synth.synthetic_frames_u8(1, H, W, seed=17)[0]
    styler = m.RegionCropStyler(models, presets, rmasks, cfgs, padding=12, use_original=True)
    got01 = styler(frame.cuda()).cpu()
    ocfgs = [dict(model_indices=c.model_indices, model_weights=c.model_weights, scale=c.scale) for c in cfgs]
    want01 = O.region_crop_style(frame, rmasks.cpu(), ocfgs, ["johnson", "johnson"], sds, presets, padding=12, use_original=True)
    assert got01.shape == want01.shape == (3, H, W)
    mx, psnr = O.u8_metrics(O.pack_u8(got01[None]), O.pack_u8(want01[None]))
    assert mx <= 2 and psnr >= 45.0, (mx, psnr)
    strip = (rmasks.sum(0) < 0.1).cpu()                                               # the uncovered strip is the original, exactly
    x01 = frame.permute(2, 0, 1).float() / 255.0
    assert strip.any() and torch.equal(got01[:, strip], x01[:, strip])


def test_region_crop_styler_animations_are_the_static_path_with_animated_scalars():
    """blend / scale animations only change the per-frame weights and scales handed to the same kernels: frame 7 of an animated styler is
    bit-identical to a static styler configured with animated_weights / animated_scale of frame 7."""
    import neuralstyletransferv1_b200 as m
    H, W = 120, 160
    masks = m.generate_region_masks(H, W, "voronoi", 2, seed=3, feather=4)
    mod = m.TransformerNet().to("cuda:0")
    mod.load_state_dict(synth.make_state_dict("johnson", 0))
    mod.eval()
    cfgs = [m.RegionConfig([0, -1], [0.7, 0.3], 1.0), m.RegionConfig([0], [1.0], 1.0)]
    ba = [m.BlendAnimation(enabled=True, period=20.0, min_opacity=0.1, max_opacity=1.0), m.BlendAnimation(enabled=False)]
    sa = [m.ScaleAnimation(enabled=False), m.ScaleAnimation(enabled=True, period=30.0, min_scale=0.5, max_scale=0.9, waveform="triangle")]
    frame = synth.synthetic_frames_u8(1, H, W, seed=4)[0].cuda()
    animated = m.RegionCropStyler([mod], ["raw_255"], masks, cfgs, padding=16, blend_animations=ba, scale_animations=sa)(frame, frame_idx=7)
    w7 = m.animated_weights([0.7, 0.3], 7, ba[0])
    s7 = m.animated_scale(1.0, 7, sa[1])
    assert w7 != [0.7, 0.3] and s7 < 1.0
    static = m.RegionCropStyler([mod], ["raw_255"], masks, [m.RegionConfig([0, -1], w7, 1.0), m.RegionConfig([0], [1.0], s7)], padding=16)(frame)
    assert torch.equal(animated, static)
