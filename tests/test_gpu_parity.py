"""Parity of the CUDA path (through the C ABI) against the CPU oracle and the golden fixtures.

Tolerance (BASELINE.json north_star): on the uint8 output after the reference's truncating pack,
max|delta| <= 2 and PSNR >= 45 dB.  The fp32 direct mode is additionally held to 2e-3 of the
output range on the raw fp32 tensor.
"""
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN, golden_names
from neuralstyletransferv1_b200 import synth

pytestmark = pytest.mark.gpu

MAX_D, MIN_PSNR = 2, 45.0
CLS = {"johnson": "TransformerNet", "nst": "TransformerNetNST", "reconet": "ReCoNet"}
PRESET = {"johnson": "raw_255", "nst": "raw_01", "reconet": "tanh"}


def _model(arch, seed=0, precision="f16"):
    import neuralstyletransferv1_b200 as m
    model = getattr(m, CLS[arch])(precision=precision).to("cuda:0")
    sd = synth.make_state_dict(arch, seed)
    model.load_state_dict(sd)
    model.eval()
    return model, sd


def _u8_from_net_out(y, arch):
    from oracle import nst_oracle as O
    return O.pack_u8(O.io_post(y.cpu(), PRESET[arch]))


@pytest.mark.parametrize("name", golden_names())
@pytest.mark.parametrize("precision", ["fp32_direct", "f16"])
def test_golden_fixtures(name, precision):
    """fp32 NCHW module call vs the REAL reference modules' outputs (tests/golden)."""
    from oracle import nst_oracle as O
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    arch, seed = str(g["arch"]), int(g["weight_seed"])
    model, _ = _model(arch, seed, precision)
    x = torch.from_numpy(g["x"]).cuda()
    y = model(x)
    ref = torch.from_numpy(g["y"])
    assert tuple(y.shape) == tuple(ref.shape)
    assert y.dtype == torch.float32 and y.is_cuda
    mx, psnr = O.u8_metrics(_u8_from_net_out(y, arch), _u8_from_net_out(ref, arch))
    assert mx <= MAX_D and psnr >= MIN_PSNR, (mx, psnr)
    if precision == "fp32_direct":
        rng = float(ref.max() - ref.min())
        assert float((y.cpu() - ref).abs().max()) <= 2e-3 * rng


@pytest.mark.parametrize("arch", ["johnson", "nst", "reconet"])
def test_all_layers_run_on_tcgen05(arch):
    model, _ = _model(arch)
    model.stylize_u8(synth.synthetic_frames_u8(1, 64, 64).cuda(), PRESET[arch])
    kinds = [d["kernel"] for d in model.engine.layer_info()]
    assert all(k == "tcgen05" for k in kinds), kinds
    assert len(kinds) <= model.engine.launch_count <= len(kinds) + 1      # ReCoNet decoder.layers.1 runs as two channel halves


@pytest.mark.parametrize("arch,n,h,w", [("johnson", 1, 64, 64), ("johnson", 3, 120, 200), ("johnson", 2, 272, 480),
                                         ("reconet", 2, 96, 128), ("nst", 2, 100, 84)])
def test_stylize_u8_vs_oracle(arch, n, h, w):
    from oracle import nst_oracle as O
    model, sd = _model(arch, seed=1)
    frames = synth.synthetic_frames_u8(n, h, w, seed=h + w)
    out = model.stylize_u8(frames.cuda(), PRESET[arch]).cpu()
    ref = O.stylize_u8(arch, sd, frames, PRESET[arch])
    mx, psnr = O.u8_metrics(out, ref)
    assert mx <= MAX_D and psnr >= MIN_PSNR, (mx, psnr)


@pytest.mark.parametrize("arch,preset", [("johnson", "raw_255"), ("johnson", "imagenet_255"), ("johnson", "caffe_bgr"),
                                         ("nst", "raw_01"), ("nst", "imagenet_01"), ("reconet", "tanh")])
def test_io_presets(arch, preset):
    """Every io_preset branch of pipeline.py:1447-1486, each on a family whose (synthetic) output
    range matches the preset's post-scale (0..255 nets for the /255 presets, 0..1 / tanh nets for
    the others) so the uint8 tolerance is meaningful."""
    from oracle import nst_oracle as O
    model, sd = _model(arch, seed=0)
    frames = synth.synthetic_frames_u8(1, 48, 64, seed=4)
    out = model.stylize_u8(frames.cuda(), preset).cpu()
    ref = O.stylize_u8(arch, sd, frames, preset)
    mx, psnr = O.u8_metrics(out, ref)
    assert mx <= MAX_D and psnr >= MIN_PSNR, (arch, preset, mx, psnr)


def test_blend_and_mask_composite():
    from oracle import nst_oracle as O
    model, sd = _model("johnson", seed=0)
    frames = synth.synthetic_frames_u8(2, 64, 80, seed=6)
    ref = O.stylize_u8("johnson", sd, frames, "raw_255", blend=0.6)
    out = model.stylize_u8(frames.cuda(), "raw_255", blend=0.6).cpu()
    mx, psnr = O.u8_metrics(out, ref)
    assert mx <= MAX_D and psnr >= MIN_PSNR
    mask = torch.rand(2, 64, 80, generator=torch.Generator().manual_seed(1))
    for mode in ("keep", "replace"):
        ref = O.stylize_u8("johnson", sd, frames, "raw_255", mask=mask, composite_mode=mode)
        out = model.stylize_u8(frames.cuda(), "raw_255", mask=mask.cuda(), composite_mode=mode).cpu()
        mx, psnr = O.u8_metrics(out, ref)
        assert mx <= MAX_D and psnr >= MIN_PSNR, (mode, mx, psnr)


def test_high_precision_mode_is_tighter():
    from oracle import nst_oracle as O
    frames = synth.synthetic_frames_u8(1, 96, 96, seed=8)
    sd = synth.make_state_dict("johnson", 0)
    ref = O.stylize_u8("johnson", sd, frames, "raw_255")
    res = {}
    for prec in ("f16", "f16x3"):
        model, _ = _model("johnson", 0, prec)
        res[prec] = O.u8_metrics(model.stylize_u8(frames.cuda(), "raw_255").cpu(), ref)
    assert res["f16x3"][0] <= 1 and res["f16x3"][1] >= res["f16"][1] + 6.0, res


def test_batch_equals_single_frames_and_microbatching():
    """Frames are independent: a batch larger than the internal micro-batch equals per-frame calls."""
    from oracle import nst_oracle as O
    model, _ = _model("johnson", seed=0)
    frames = synth.synthetic_frames_u8(11, 40, 56, seed=10).cuda()      # 11 > micro-batch of 8
    whole = model.stylize_u8(frames, "raw_255")
    single = torch.cat([model.stylize_u8(frames[i:i + 1], "raw_255") for i in range(11)])
    mx, psnr = O.u8_metrics(whole.cpu(), single.cpu())
    assert mx <= 1 and psnr >= 60.0, (mx, psnr)      # only atomics ordering in the statistics differs


def test_module_surface_on_device():
    model, sd = _model("johnson", seed=0)
    x = torch.rand(1, 3, 50, 70, device="cuda") * 255
    y = model(x)
    assert tuple(y.shape) == (1, 3, 52, 72)          # odd sizes grow like the reference's
    assert float(y.min()) < float(y.max())           # .min()/.max()/.cpu() as pipeline.py:1475 uses them
    with pytest.raises(RuntimeError):
        model(torch.rand(1, 3, 50, 70))              # CPU tensor: no fallback
    # in-place weight edits are picked up
    with torch.no_grad():
        model.deconv3.conv2d.bias.add_(10.0)
    y2 = model(x)
    assert abs(float((y2 - y).mean()) - 10.0) < 1e-2


def test_errors_are_exceptions():
    from neuralstyletransferv1_b200._lib import NstError
    model, _ = _model("johnson")
    with pytest.raises((NstError, ValueError)):
        model.stylize_u8(torch.zeros(1, 50, 70, 3, dtype=torch.uint8, device="cuda"), "raw_255")   # not /4
    nst, _ = _model("nst")
    with pytest.raises(NstError):
        nst(torch.zeros(1, 3, 32, 32, device="cuda"))    # <= 40: ReflectionPad2d(40) impossible


def test_composite_regions_gpu():
    import neuralstyletransferv1_b200 as m
    from oracle import nst_oracle as O
    g = np.load(os.path.join(GOLDEN, "composite_voronoi4_48x64.npz"))
    styled = [torch.from_numpy(s).cuda() for s in g["styled"]]
    masks = torch.from_numpy(g["masks"]).cuda()
    orig = torch.from_numpy(g["original"]).cuda()
    a = m.composite_regions(styled, masks, [0, 1, 2, 3], original=orig).cpu()
    b = m.composite_regions(styled, masks, [0, -1, 2, 1], original=orig).cpu()
    assert float((a - torch.from_numpy(g["out_seq"])).abs().max()) <= 2e-6
    assert float((b - torch.from_numpy(g["out_mixed"])).abs().max()) <= 2e-6
    # uint8 form + per-region model blends (composite_regions_advanced, region_blend.py:1589-1679)
    s8 = [(s * 255).to(torch.uint8).permute(1, 2, 0).contiguous() for s in styled]
    o8 = (orig * 255).to(torch.uint8).permute(1, 2, 0).contiguous()
    idx, wts = [[0, 1], [1, -1], [2, 3], [3, 0]], [[0.7, 0.3], [0.5, 0.5], [1.0, 0.0], [0.2, 0.8]]
    got = m.composite_regions(s8, masks, idx, wts, original=o8).cpu()
    sf = [s.permute(2, 0, 1).float().cpu() / 255 for s in s8]
    want = O.composite_regions_advanced(sf, [mk[None, None].cpu() for mk in masks], idx, wts, o8.permute(2, 0, 1).float().cpu() / 255)
    mx, psnr = O.u8_metrics(got, O.pack_u8(want[None])[0])
    assert mx <= 1 and psnr >= 55.0, (mx, psnr)


def test_full_size_1080p_properties():
    """BASELINE config sizes: 1080p Johnson.  Oracle parity on one frame (4 s of CPU) plus a
    size-independent property: transposing the frame and every filter transposes the output
    (conv, stride, reflection and nearest-upsample are symmetric in H and W), which drives the
    same data through a completely different tiling (1920x1080 instead of 1080x1920)."""
    from oracle import nst_oracle as O
    model, sd = _model("johnson", seed=0)
    frames = synth.synthetic_frames_u8(2, 1080, 1920, seed=21)
    out = model.stylize_u8(frames.cuda(), "raw_255").cpu()
    ref = O.stylize_u8("johnson", sd, frames[:1], "raw_255")
    mx, psnr = O.u8_metrics(out[:1], ref)
    assert mx <= MAX_D and psnr >= MIN_PSNR, (mx, psnr)
    import neuralstyletransferv1_b200 as m
    sd_t = {k: (v.transpose(-1, -2).contiguous() if v.dim() == 4 else v) for k, v in sd.items()}
    tr = m.TransformerNet().to("cuda:0")
    tr.load_state_dict(sd_t)
    out_t = tr.stylize_u8(frames[1:].transpose(1, 2).contiguous().cuda(), "raw_255").cpu().transpose(1, 2)
    mx, psnr = O.u8_metrics(out_t, out[1:])
    assert mx <= MAX_D and psnr >= 50.0, (mx, psnr)


@pytest.mark.parametrize("arch", ["johnson", "nst", "reconet"])
@pytest.mark.parametrize("env", [{"NST_NO_PHASE": "1"}, {"NST_NO_FOLD": "1"}, {"NST_NO_QUAD": "1"}, {"NST_NO_SPLIT": "1"},
                                 {"NST_PROD_WARPS": "8"}, {"NST_PROD_WARPS": "10"}, {"NST_PROD_WARPS": "5"},
                                 {"NST_TC_FORCE": ";".join(f"{i}:3:1" for i in range(16))}])
def test_kernel_forms_agree(arch, env, monkeypatch):
    """The sub-pixel (phase) form of upsample+conv / ConvTranspose, the folded image layer and any tile shape
    (and the number of operand-producer warps) are re-formulations of the same layer (only the fp32 accumulation order changes): switching them off (or
    forcing 3-row tiles) must leave the uint8 output within the tolerance of itself at PSNR >= 50 dB, and both
    must meet the oracle tolerance."""
    from oracle import nst_oracle as O
    frames = synth.synthetic_frames_u8(2, 88, 136, seed=9)
    model, sd = _model(arch, seed=2)
    out = model.stylize_u8(frames.cuda(), PRESET[arch]).cpu()
    for k, v in env.items():
        monkeypatch.setenv(k, v)
    alt_model, _ = _model(arch, seed=2)
    alt = alt_model.stylize_u8(frames.cuda(), PRESET[arch]).cpu()
    mx, psnr = O.u8_metrics(alt, out)
    assert mx <= MAX_D and psnr >= 50.0, (mx, psnr)
    ref = O.stylize_u8(arch, sd, frames, PRESET[arch])
    for y in (out, alt):
        mx, psnr = O.u8_metrics(y, ref)
        assert mx <= MAX_D and psnr >= MIN_PSNR, (mx, psnr)


def test_composite_u8_odd_size_scalar_path():
    """H*W not a multiple of 4 takes the one-pixel-per-thread kernel; it must agree with the oracle like the vector path."""
    import neuralstyletransferv1_b200 as m
    from oracle import nst_oracle as O
    g = torch.Generator().manual_seed(11)
    H, W, K = 7, 9, 3
    s8 = [torch.randint(0, 256, (H, W, 3), dtype=torch.uint8, generator=g) for _ in range(K)]
    o8 = torch.randint(0, 256, (H, W, 3), dtype=torch.uint8, generator=g)
    masks = torch.rand(K, H, W, generator=g)
    idx = [[0], [1], [-1]]
    got = m.composite_regions([s.cuda() for s in s8], masks.cuda(), idx, original=o8.cuda()).cpu()
    sf = [s.permute(2, 0, 1).float() / 255 for s in s8]
    want = O.composite_regions_advanced(sf, [mk[None, None] for mk in masks], idx, [[1.0]] * 3, o8.permute(2, 0, 1).float() / 255)
    mx, _ = O.u8_metrics(got, O.pack_u8(want[None])[0])
    assert mx <= 1, mx
