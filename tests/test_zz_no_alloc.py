"""The allocation-free entry point nst_motion_blend_u8_noalloc (static device scratch inside the library; no reference counterpart)
against the existing nst_motion_blend_u8, through the C ABI.  Byte work: bit-exact, the bar of
tests/test_gpu_masks.py.  Collected last on purpose (file name)."""
import ctypes as C

import pytest
import torch

from neuralstyletransferv1_b200 import synth

pytestmark = pytest.mark.gpu


def _call_pair(lib, H, W, seed, blend):
    dev = torch.device("cuda:0")
    fr = synth.synthetic_frames_u8(2, H, W, seed=seed).to(dev)
    styled, orig = fr[0].contiguous(), fr[1].contiguous()
    flow = (torch.randn(H, W, 2, generator=torch.Generator().manual_seed(seed)) * 5).to(dev)
    want = torch.empty_like(styled)
    got = torch.full_like(styled, 7)
    scratch = torch.empty(H, W, dtype=torch.float32, device=dev)
    st = torch.cuda.current_stream(dev).cuda_stream
    with torch.cuda.device(dev):
        rc = lib.nst_motion_blend_u8(styled.data_ptr(), orig.data_ptr(), flow.data_ptr(), want.data_ptr(), scratch.data_ptr(), H, W,
                                     blend, 0.40, 8.0, 3.0, st)
        assert rc == 0, lib.nst_last_error()
        before = torch.cuda.memory_allocated(dev)
        rc = lib.nst_motion_blend_u8_noalloc(styled.data_ptr(), orig.data_ptr(), flow.data_ptr(), got.data_ptr(), H, W, blend, 0.40, 8.0, 3.0, st)
        assert rc == 0, lib.nst_last_error()
        assert torch.cuda.memory_allocated(dev) == before
    torch.cuda.synchronize(dev)
    return got, want


def test_noalloc_motion_blend_equals_the_scratch_version_twice_in_a_row():
    from neuralstyletransferv1_b200 import _lib
    lib = _lib.load()
    # two calls in a row share the static scratch; the second frame is smaller, so stale rows of the first stay in the array behind it
    for (H, W, seed, blend) in ((270, 484, 3, 0.9), (135, 250, 4, 0.75)):
        got, want = _call_pair(lib, H, W, seed, blend)
        assert torch.equal(got, want), (H, W)
    got, want = _call_pair(lib, 1080, 1920, 5, 0.85)          # the documented maximum class of frame (1080p fits 2^21 pixels)
    assert torch.equal(got, want)


def test_noalloc_motion_blend_rejects_frames_beyond_the_static_scratch():
    from neuralstyletransferv1_b200 import _lib
    from neuralstyletransferv1_b200.temporal import MOTION_STATIC_MAX_PIXELS
    lib = _lib.load()
    dev = torch.device("cuda:0")
    H, W = 1100, 1920
    assert H * W > MOTION_STATIC_MAX_PIXELS
    s = torch.zeros(H, W, 3, dtype=torch.uint8, device=dev)
    f = torch.zeros(H, W, 2, device=dev)
    o = torch.empty_like(s)
    with torch.cuda.device(dev):
        rc = lib.nst_motion_blend_u8_noalloc(s.data_ptr(), s.data_ptr(), f.data_ptr(), o.data_ptr(), H, W, 0.9, 0.4, 8.0, 3.0,
                                             torch.cuda.current_stream(dev).cuda_stream)
    assert rc != 0 and b"exceed" in lib.nst_last_error()


def test_noalloc_python_wrapper():
    import neuralstyletransferv1_b200.temporal as T
    dev = torch.device("cuda:0")
    fr = synth.synthetic_frames_u8(2, 64, 96, seed=8).to(dev)
    flow = (torch.randn(64, 96, 2, generator=torch.Generator().manual_seed(2)) * 4).to(dev)
    out = torch.empty_like(fr[0])
    assert torch.equal(T.motion_blend_noalloc(fr[0], fr[1], flow, out, 0.8), T.motion_blend(fr[0], fr[1], flow, 0.8))
