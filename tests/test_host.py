"""Host logic and the C-ABI boundary, without a GPU."""
import os
import re
import sys

import pytest
import torch

from conftest import ROOT
from neuralstyletransferv1_b200 import synth


def test_cabi_exports_every_declared_symbol(built_lib):
    hdr = open(os.path.join(ROOT, "include", "b200nst.h")).read()
    declared = set(re.findall(r"\b(nst_[a-z0-9_]+)\s*\(", hdr))
    assert len(declared) >= 15
    from neuralstyletransferv1_b200 import _lib
    assert declared == set(_lib.PROTOTYPES), "ctypes prototypes out of sync with include/b200nst.h"
    for name in declared:
        assert hasattr(built_lib, name), f"libb200nst.so does not export {name}"
    assert b"sm_100a" in built_lib.nst_version()


def test_create_fails_loudly_without_gpu(built_lib):
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    import ctypes as C
    h = C.c_void_p()
    rc = built_lib.nst_create(0, 0, 0, C.byref(h))
    assert rc != 0 and not h.value
    assert built_lib.nst_last_error()


@pytest.mark.parametrize("arch,cls", [("johnson", "TransformerNet"), ("nst", "TransformerNetNST"), ("reconet", "ReCoNet")])
def test_dropin_state_dict_surface(arch, cls):
    import neuralstyletransferv1_b200 as m
    model = getattr(m, cls)()
    sd = synth.make_state_dict(arch, 0)
    assert list(model.state_dict().keys()) == list(sd.keys())
    for k, v in model.state_dict().items():
        assert tuple(v.shape) == tuple(sd[k].shape), k
    res = model.load_state_dict(sd, strict=True)
    assert not res.missing_keys and not res.unexpected_keys
    # legacy InstanceNorm buffers are dropped like pipeline.py:565-566 does
    legacy = dict(sd)
    first_norm = [k for k in sd if k.endswith(".weight") and sd[k].dim() == 1][0][:-len("weight")]
    legacy[first_norm + "running_mean"] = torch.zeros(4)
    legacy[first_norm + "running_var"] = torch.ones(4)
    res = model.load_state_dict(legacy, strict=True)
    assert not res.unexpected_keys
    # strict=False returns (missing, unexpected) whose len() pipeline.py:569 prints
    partial = {k: v for k, v in sd.items() if not k.startswith("res1") and "layers.3" not in k}
    partial["bogus.weight"] = torch.zeros(1)
    missing, unexpected = model.load_state_dict(partial, strict=False)
    assert len(missing) > 0 and unexpected == ["bogus.weight"]
    model.eval()
    assert torch.equal(model.state_dict()[list(sd)[0]], sd[list(sd)[0]])


def test_dropin_has_no_cpu_fallback():
    import neuralstyletransferv1_b200 as m
    model = m.TransformerNet()
    with pytest.raises(RuntimeError, match="no CPU"):
        model(torch.zeros(1, 3, 16, 16))
    with pytest.raises(NotImplementedError):
        m.ReCoNet(frn=True)


def test_shims_resolve_like_pipeline_imports():
    shim = os.path.join(ROOT, "neuralstyletransferv1_b200", "shims")
    sys.path.insert(0, shim)
    try:
        for name in ("transformer_net", "transformer_net_nst", "lib"):
            sys.modules.pop(name, None)
        from transformer_net import TransformerNet as A            # pipeline.py:64
        from transformer_net_nst import TransformerNet as B        # pipeline.py:69
        from lib import ReCoNet as Cc                               # pipeline.py:50
        assert A.ARCH == "johnson" and B.ARCH == "nst" and Cc.ARCH == "reconet"
    finally:
        sys.path.remove(shim)
        for name in ("transformer_net", "transformer_net_nst", "lib"):
            sys.modules.pop(name, None)


@pytest.mark.parametrize("preset", ["raw_255", "raw_01", "tanh", "imagenet_01", "imagenet_255", "caffe_bgr"])
def test_io_preset_affine_equals_oracle(preset):
    from neuralstyletransferv1_b200 import io_preset_affine
    from oracle import nst_oracle as O
    s_in, b_in, d_out, s_out, b_out, swap = io_preset_affine(preset)
    u8 = torch.arange(0, 256, dtype=torch.float32).view(1, 1, 16, 16).repeat(1, 3, 1, 1)
    u8[:, 1] = u8[:, 1].flip(-1)
    u8[:, 2] = (u8[:, 2] * 7) % 256
    want_in = O.io_pre(u8 / 255.0, preset)
    src = u8[:, [2, 1, 0]] if swap else u8
    got_in = src * torch.tensor(s_in).view(1, 3, 1, 1) + torch.tensor(b_in).view(1, 3, 1, 1)
    assert torch.allclose(got_in, want_in, rtol=1e-5, atol=1e-4)
    y = torch.randn(1, 3, 16, 16) * 100 + 100
    want_out = O.io_post(y, preset)
    ysrc = y[:, [2, 1, 0]] if swap else y
    got_out = ((ysrc / torch.tensor(d_out).view(1, 3, 1, 1)) * torch.tensor(s_out).view(1, 3, 1, 1) + torch.tensor(b_out).view(1, 3, 1, 1)).clamp(0, 1)
    assert torch.equal(got_out, want_out)         # the same fp32 operation sequence as the reference's: bit-identical


def test_shard_ranges_partition_frames():
    from neuralstyletransferv1_b200 import shard_range, shard_sizes
    for n in (0, 1, 7, 64, 300):
        for w in (1, 2, 3, 4, 8):
            sizes = shard_sizes(n, w)
            assert sum(sizes) == n and max(sizes) - min(sizes) <= 1
            cover = []
            for r in range(w):
                a, b = shard_range(n, w, r)
                cover += list(range(a, b))
            assert cover == list(range(n))
    with pytest.raises(ValueError):
        shard_range(4, 2, 2)


def test_synthetic_frames_are_deterministic():
    a = synth.synthetic_frames_u8(3, 32, 48, seed=5)
    b = synth.synthetic_frames_u8(3, 32, 48, seed=5)
    assert a.dtype == torch.uint8 and tuple(a.shape) == (3, 32, 48, 3) and torch.equal(a, b)
    assert int(a.max()) > 200 and int(a.min()) < 50


# ---- planner (host only): production shapes stay on the tensor cores, TMA-fed where the source is an fp16 activation tensor ----
@pytest.mark.parametrize("arch,n,h,w", [("johnson", 8, 1080, 1920), ("johnson", 8, 2160, 3840), ("johnson", 1, 3024, 4032),
                                         ("johnson", 2, 1081, 1923), ("nst", 8, 1080, 1920), ("nst", 1, 2160, 3840),
                                         ("reconet", 8, 1080, 1920), ("reconet", 2, 2160, 3840), ("johnson", 3, 50, 70),
                                         ("nst", 1, 53, 47), ("reconet", 2, 38, 54)])
def test_planner_keeps_production_shapes_on_tcgen05(arch, n, h, w):
    from neuralstyletransferv1_b200 import _lib
    text = _lib.describe_plan(arch, "f16", n, h, w)
    lines = text.strip().splitlines()[1:]
    assert len(lines) == {"johnson": 16, "nst": 16, "reconet": 14}[arch]
    assert all(" tc " in ln and " direct " not in ln for ln in lines), text
    gather = {"johnson": 1, "nst": 1, "reconet": 3}[arch]        # image layer; + ReCoNet's error-compensated encoder layers
    assert all(" tc gather" in ln for ln in lines[:gather]), text
    assert all(" tc tma" in ln for ln in lines[gather:]), text
    for ln in lines:
        smem = int(ln.split("smem=")[1].split("B")[0])
        assert smem <= 227 * 1024, ln


def test_planner_refuses_instead_of_degrading():
    """A shape with no tcgen05 tiling is an error, never a silent fall to the fp32 FFMA kernel."""
    from neuralstyletransferv1_b200 import _lib
    assert " direct " in _lib.describe_plan("johnson", "fp32_direct", 1, 64, 64)
    assert " direct " not in _lib.describe_plan("johnson", "f16", 1, 64, 64)


def test_param_shape_table_matches_synth():
    from neuralstyletransferv1_b200 import modules
    for arch in synth.ARCHS:
        sd = synth.make_state_dict(arch, 0)
        assert [(k, tuple(v.shape)) for k, v in sd.items()] == [(k, tuple(sh)) for k, sh in modules.PARAM_SHAPES[arch]]


def test_region_animation_scalars_match_the_reference_functions():
    """Host arithmetic of the blend / scale animations (region_blend.py:1127-1247, 1334-1370) against a table the reference's own
    functions produced (tests/golden/region_anim.json, written by tests/golden/make_golden_anim.py): exact double equality."""
    import json
    from neuralstyletransferv1_b200 import region_anim as ra
    table = json.load(open(os.path.join(ROOT, "tests", "golden", "region_anim.json")))
    assert len(table["harmonic"]) >= 190 and len(table["weights"]) >= 60 and len(table["scale"]) >= 20
    for args, want in table["harmonic"]:
        assert ra.harmonic_value(*args) == want, args
    for base, idx, anim, want in table["weights"]:
        assert ra.animated_weights(base, idx, ra.BlendAnimation(**anim)) == want, (base, idx, anim)
    for base, idx, anim, want in table["scale"]:
        assert ra.animated_scale(base, idx, ra.ScaleAnimation(**anim)) == want, (base, idx, anim)


def test_model_assignments_match_the_reference_function():
    """assign_models_to_regions (region_blend.py:983-1046): same `random.Random` draws, same lists."""
    import json
    from neuralstyletransferv1_b200.region_masks import assign_models_to_regions
    table = json.load(open(os.path.join(ROOT, "tests", "golden", "region_assignments.json")))
    assert len(table) >= 100
    for (nr, nm, mode, w, seed, oc), want in table:
        assert assign_models_to_regions(nr, nm, mode, w, seed, oc) == want, (nr, nm, mode, seed, oc)


def test_crop_bookkeeping_helpers():
    """region_blend.py:2112-2183 (models needed, regions of a model, merged box, coverage) on hand-checked values -- host list arithmetic."""
    from neuralstyletransferv1_b200 import region_crops as rc
    mk = lambda i, box, idx: rc.RegionCrop(i, box, box, torch.zeros(1, 1), rc.RegionConfig(idx, [1.0 / len(idx)] * len(idx)))
    crops = [mk(0, (0, 0, 40, 30), [2]), mk(1, (30, 10, 100, 50), [0, -1]), mk(2, (5, 5, 25, 45), [2, 0])]
    assert rc.models_needed(crops) == [0, 2]
    assert [c.region_idx for c in rc.regions_for_model(crops, 0)] == [1, 2] and rc.regions_for_model(crops, 1) == []
    assert [c.region_idx for c in rc.regions_for_model(crops, -1)] == [1]
    assert rc.merge_bboxes([c.padded_bbox for c in crops]) == (0, 0, 100, 50) and rc.merge_bboxes([]) == (0, 0, 0, 0)
    assert rc.crop_coverage(crops, 50, 100) == (40 * 30 + 70 * 40 + 20 * 40) / 5000
    assert rc.pad_bbox((5, 5, 25, 45), 12, 100, 50) == (0, 0, 37, 50)
