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
