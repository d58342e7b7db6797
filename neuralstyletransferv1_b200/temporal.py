"""Temporal post-processing on the device (second part): the LAB lightness / chroma EMA and the
motion-adaptive blend of the reference's per-frame loop (pipeline.py:1938-1975, 2072-2086).  The flow-guided EMA that precedes them is
``engine.flow_ema``.

``LabEma`` keeps the running (L, a, b) planes on the device and smooths a whole chunk of frames per launch.  Pillow's RGB <-> LAB
conversion (littleCMS, what ``Image.convert("LAB")`` runs in the reference) is a pure function of the uint8 triple: the two 2^24-entry
tables the kernel reads are filled once per process by running Pillow's own transform over every colour, so the device result is
Pillow's by construction.
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, Optional, Tuple

import numpy as np
import torch

from . import _lib

_LUT_HOST: Optional[Tuple[np.ndarray, np.ndarray]] = None
_LUT_DEV: Dict[str, Tuple[torch.Tensor, torch.Tensor]] = {}


def _pack(img: np.ndarray) -> np.ndarray:
    a = img.reshape(-1, 3).astype(np.uint32)
    return a[:, 0] | (a[:, 1] << 8) | (a[:, 2] << 16)


def lab_tables_host() -> Tuple[np.ndarray, np.ndarray]:
    """(rgb2lab, lab2rgb): uint32 [2^24]; index c0 << 16 | c1 << 8 | c2 -> c0' | c1' << 8 | c2' << 16, from Pillow's own conversion."""
    global _LUT_HOST
    if _LUT_HOST is None:
        from PIL import Image
        v = np.arange(1 << 24, dtype=np.uint32)
        cube = np.stack([(v >> 16) & 255, (v >> 8) & 255, v & 255], -1).astype(np.uint8).reshape(4096, 4096, 3)
        rgb2lab = _pack(np.array(Image.fromarray(cube, mode="RGB").convert("LAB"), dtype=np.uint8))
        lab2rgb = _pack(np.array(Image.fromarray(cube, mode="LAB").convert("RGB"), dtype=np.uint8))
        _LUT_HOST = (rgb2lab, lab2rgb)
    return _LUT_HOST


def lab_tables(device) -> Tuple[torch.Tensor, torch.Tensor]:
    dev = torch.device(device)
    key = str(dev)
    if key not in _LUT_DEV:
        a, b = lab_tables_host()
        _LUT_DEV[key] = (torch.from_numpy(a.view(np.int32)).to(dev), torch.from_numpy(b.view(np.int32)).to(dev))
    return _LUT_DEV[key]


class LabEma:
    """The LAB smoothing block of pipeline.py:1938-1975 for a stream of uint8 RGB frames.

    ``ema(frames_u8)`` takes ``[N, H, W, 3]`` (or ``[H, W, 3]``) CUDA uint8 frames in display order and returns the smoothed frames;
    the running state carries over between calls and resets when the frame size changes (pipeline.py:1948, 1959-1962).
    """

    def __init__(self, smooth_lightness: bool = True, smooth_chroma: bool = False, smooth_alpha: float = 0.7, chroma_alpha: float = 0.85):
        self.smooth_lightness, self.smooth_chroma = bool(smooth_lightness), bool(smooth_chroma)
        self.smooth_alpha, self.chroma_alpha = float(smooth_alpha), float(chroma_alpha)
        self._state: Optional[torch.Tensor] = None       # [3, H, W] fp32: prev_L, prev_a, prev_b

    def reset(self) -> None:
        self._state = None

    def __call__(self, frames: torch.Tensor) -> torch.Tensor:
        if not frames.is_cuda:
            raise RuntimeError("LabEma needs CUDA tensors; there is no CPU path")
        single = frames.dim() == 3
        x = (frames[None] if single else frames).contiguous()
        if x.dtype != torch.uint8 or x.dim() != 4 or x.shape[-1] != 3:
            raise ValueError("frames must be uint8 [N, H, W, 3]")
        if not (self.smooth_lightness or self.smooth_chroma):
            return frames                                 # the reference skips the LAB round trip entirely (pipeline.py:1940)
        lib = _lib.load()
        N, H, W, _ = x.shape
        init = self._state is None or tuple(self._state.shape[1:]) != (H, W) or self._state.device != x.device
        if init:
            self._state = torch.empty(3, H, W, dtype=torch.float32, device=x.device)
        rgb2lab, lab2rgb = lab_tables(x.device)
        out = torch.empty_like(x)
        st = self._state
        with torch.cuda.device(x.device):
            rc = lib.nst_lab_ema_u8(x.data_ptr(), out.data_ptr(), N, H, W, rgb2lab.data_ptr(), lab2rgb.data_ptr(), st[0].data_ptr(), st[1].data_ptr(),
                                    st[2].data_ptr(), int(init), int(self.smooth_lightness), int(self.smooth_chroma), self.smooth_alpha,
                                    self.chroma_alpha, torch.cuda.current_stream(x.device).cuda_stream)
        _lib.check(rc, "nst_lab_ema_u8")
        return out[0] if single else out


MOTION_NORM, MIN_ALPHA, GAUSS_SIGMA = 8.0, 0.40, 3.0        # pipeline.py:1075-1077


def motion_blend(styled_u8: torch.Tensor, original_u8: torch.Tensor, flow: torch.Tensor, blend: float, min_alpha: float = MIN_ALPHA,
                 motion_norm: float = MOTION_NORM, sigma: float = GAUSS_SIGMA) -> torch.Tensor:
    """Motion-adaptive blend of pipeline.py:2072-2086: less style where the optical flow is large.

    styled_u8 / original_u8: uint8 [H, W, 3] CUDA frames; flow: fp32 [H, W, 2] = (dx, dy) of the last flow estimate.
    alpha = blend - (blend - min_alpha) * GaussianBlur(clip(|flow| / motion_norm, 0, 1), sigma); out = pack(alpha * S + (1 - alpha) * O).
    """
    if not styled_u8.is_cuda:
        raise RuntimeError("motion_blend needs CUDA tensors; there is no CPU path")
    H, W, _ = styled_u8.shape
    for t, shp, dt, name in ((styled_u8, (H, W, 3), torch.uint8, "styled_u8"), (original_u8, (H, W, 3), torch.uint8, "original_u8"),
                             (flow, (H, W, 2), torch.float32, "flow")):
        if tuple(t.shape) != shp or t.dtype != dt or t.device != styled_u8.device:
            raise ValueError(f"{name} must be {dt} {shp} on {styled_u8.device}")
    lib = _lib.load()
    styled_u8, original_u8, flow = styled_u8.contiguous(), original_u8.contiguous(), flow.contiguous()
    out = torch.empty_like(styled_u8)
    scratch = torch.empty(H, W, dtype=torch.float32, device=styled_u8.device)
    with torch.cuda.device(styled_u8.device):
        rc = lib.nst_motion_blend_u8(styled_u8.data_ptr(), original_u8.data_ptr(), flow.data_ptr(), out.data_ptr(), scratch.data_ptr(), H, W,
                                     float(blend), float(min_alpha), float(motion_norm), float(sigma), torch.cuda.current_stream(styled_u8.device).cuda_stream)
    _lib.check(rc, "nst_motion_blend_u8")
    return out


def blend_pack(styled_u8: torch.Tensor, original_u8: torch.Tensor, blend: float = 1.0, mask: Optional[torch.Tensor] = None,
               composite_mode: str = "keep") -> torch.Tensor:
    """Mask composite, then uniform blend, then pack, of frames that are already uint8 (pipeline.py:1981, 2040-2043, 2088-2094).
    styled_u8 / original_u8: uint8 [N, H, W, 3] CUDA; mask: fp32 [N, H, W] or None."""
    if not styled_u8.is_cuda:
        raise RuntimeError("blend_pack needs CUDA tensors; there is no CPU path")
    N, H, W, _ = styled_u8.shape
    if tuple(original_u8.shape) != (N, H, W, 3) or styled_u8.dtype != torch.uint8 or original_u8.dtype != torch.uint8:
        raise ValueError("styled_u8 and original_u8 must be uint8 [N, H, W, 3]")
    if mask is not None and (tuple(mask.shape) != (N, H, W) or mask.dtype != torch.float32):
        raise ValueError("mask must be float32 [N, H, W]")
    lib = _lib.load()
    styled_u8, original_u8 = styled_u8.contiguous(), original_u8.contiguous()
    m = mask.contiguous() if mask is not None else None
    out = torch.empty_like(styled_u8)
    with torch.cuda.device(styled_u8.device):
        rc = lib.nst_blend_pack_u8(styled_u8.data_ptr(), original_u8.data_ptr(), m.data_ptr() if m is not None else None, float(blend),
                                   _lib.COMPOSITE_MODE[composite_mode], out.data_ptr(), N, H, W, torch.cuda.current_stream(styled_u8.device).cuda_stream)
    _lib.check(rc, "nst_blend_pack_u8")
    return out


MOTION_STATIC_MAX_PIXELS = 1 << 21          # include/b200nst.h: NST_MOTION_STATIC_MAX_PIXELS


def motion_blend_noalloc(styled_u8: torch.Tensor, original_u8: torch.Tensor, flow: torch.Tensor, out: torch.Tensor, blend: float,
                         min_alpha: float = MIN_ALPHA, motion_norm: float = MOTION_NORM, sigma: float = GAUSS_SIGMA) -> torch.Tensor:
    """``motion_blend`` without any allocation (no reference counterpart; added for callers under CUDA graph capture or in loops that may
    not allocate): the scratch plane is a static array inside the library, ``out`` is the caller's buffer, inputs must already be
    contiguous.  H * W <= MOTION_STATIC_MAX_PIXELS; calls that overlap on one GPU share that scratch and race."""
    if not styled_u8.is_cuda:
        raise RuntimeError("motion_blend_noalloc needs CUDA tensors; there is no CPU path")
    H, W, _ = styled_u8.shape
    for t, shp, dt, name in ((styled_u8, (H, W, 3), torch.uint8, "styled_u8"), (original_u8, (H, W, 3), torch.uint8, "original_u8"),
                             (flow, (H, W, 2), torch.float32, "flow"), (out, (H, W, 3), torch.uint8, "out")):
        if tuple(t.shape) != shp or t.dtype != dt or t.device != styled_u8.device or not t.is_contiguous():
            raise ValueError(f"{name} must be a contiguous {dt} {shp} tensor on {styled_u8.device}")
    lib = _lib.load()
    with torch.cuda.device(styled_u8.device):
        rc = lib.nst_motion_blend_u8_noalloc(styled_u8.data_ptr(), original_u8.data_ptr(), flow.data_ptr(), out.data_ptr(), H, W, float(blend),
                                             float(min_alpha), float(motion_norm), float(sigma), torch.cuda.current_stream(styled_u8.device).cuda_stream)
    _lib.check(rc, "nst_motion_blend_u8_noalloc")
    return out
