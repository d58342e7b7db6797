"""Host-side engine: one ``nst_handle`` (one GPU, one model) driven through the C ABI.

PyTorch is used for device memory, streams and tensors only; every arithmetic
operation of the stylisation path happens inside ``libb200nst.so``.
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, Mapping, Optional, Sequence, Tuple

import torch

from . import _lib

IMAGENET_MEAN = (0.485, 0.456, 0.406)     # pipeline.py:272
IMAGENET_STD = (0.229, 0.224, 0.225)      # pipeline.py:273
CAFFE_MEAN = (103.939, 116.779, 123.68)   # pipeline.py:1465


def io_preset_affine(preset: str):
    """uint8 -> network input and network output -> out01, per channel.

    Folds ToTensor (/255, pipeline.py:1099) and the io_preset branches of pipeline.py:1447-1486 into
    ``in = s_in*u8 + b_in`` and ``out01 = clamp((y / d_out) * s_out + b_out, 0, 1)`` (+ R/B swap for
    caffe_bgr).  The output side keeps the reference's own operation per preset -- a true division for
    the ``/255`` presets, multiply-then-add for ``imagenet_01`` / ``tanh`` -- because the uint8 pack that
    follows truncates and must be bit-identical.  Returns
    (in_scale, in_shift, out_div, out_scale, out_shift, swap_rb).  An unknown preset takes the
    reference's final ``else`` branch (pipeline.py:1482-1486), i.e. raw_255.
    """
    one, zero = (1.0, 1.0, 1.0), (0.0, 0.0, 0.0)
    d255 = (255.0,) * 3
    inv255 = (1.0 / 255.0,) * 3
    if preset == "raw_01":
        return inv255, zero, one, one, zero, 0
    if preset == "tanh":
        return (2.0 / 255.0,) * 3, (-1.0,) * 3, one, (0.5,) * 3, (0.5,) * 3, 0      # (y + 1) * 0.5 == y * 0.5 + 0.5 exactly in fp32
    if preset in ("imagenet_01", "imagenet_255"):
        s_in = tuple(1.0 / (255.0 * s) for s in IMAGENET_STD)
        b_in = tuple(-m / s for m, s in zip(IMAGENET_MEAN, IMAGENET_STD))
        if preset == "imagenet_01":
            return s_in, b_in, one, IMAGENET_STD, IMAGENET_MEAN, 0
        return s_in, b_in, d255, one, zero, 0
    if preset == "caffe_bgr":
        return one, tuple(-m for m in CAFFE_MEAN), d255, one, zero, 1
    return one, zero, d255, one, zero, 0          # raw_255 and the reference's default branch


def _f3(v: Sequence[float]):
    return (C.c_float * 3)(*[float(x) for x in v])


class Engine:
    """Owns one C handle.  Not re-entrant; use one Engine per GPU / worker thread."""

    def __init__(self, arch: str, device: int = 0, precision: str = "f16"):
        self._lib = _lib.load()
        if arch not in _lib.ARCH:
            raise ValueError(f"unknown arch {arch!r}")
        if precision not in _lib.PRECISION:
            raise ValueError(f"unknown precision {precision!r}; one of {sorted(_lib.PRECISION)}")
        if not torch.cuda.is_available():
            raise RuntimeError("neuralstyletransferv1_b200 needs a CUDA (sm_100a) device; there is no CPU fallback")
        self.arch, self.device, self.precision = arch, int(device), precision
        h = C.c_void_p()
        _lib.check(self._lib.nst_create(self.device, _lib.ARCH[arch], _lib.PRECISION[precision], C.byref(h)), "nst_create")
        self._h = h
        self._ws: Optional[torch.Tensor] = None
        self._ws_key = None
        self.loaded = False

    def close(self):
        if getattr(self, "_h", None):
            self._lib.nst_destroy(self._h)
            self._h = None

    def __del__(self):  # pragma: no cover
        try:
            self.close()
        except Exception:
            pass

    # ---- weights -------------------------------------------------------------------------
    def load_state_dict(self, sd: Mapping[str, torch.Tensor]) -> Tuple[int, int]:
        """Replaces model.load_state_dict (pipeline.py:568): returns (n_missing, n_unexpected)."""
        names, keep = [], []
        for k, v in sd.items():
            if not torch.is_tensor(v) or v.dim() > 4 or not v.is_floating_point():
                continue
            names.append(k)
            keep.append(v.detach().to("cpu", torch.float32).contiguous())
        n = len(names)
        c_names = (C.c_char_p * n)(*[s.encode() for s in names])
        c_ptrs = (C.c_void_p * n)(*[t.data_ptr() for t in keep])
        shapes = []
        for t in keep:
            s = list(t.shape) + [1] * (4 - t.dim())
            shapes.extend(s)
        c_shapes = (C.c_int64 * (4 * n))(*shapes)
        miss, unexp = C.c_int(0), C.c_int(0)
        rc = self._lib.nst_load_weights(self._h, n, c_names, C.cast(c_ptrs, C.POINTER(C.c_void_p)), c_shapes,
                                        C.byref(miss), C.byref(unexp))
        _lib.check(rc, "nst_load_weights")
        self.loaded = True
        return miss.value, unexp.value

    # ---- helpers -------------------------------------------------------------------------
    def output_shape(self, H: int, W: int) -> Tuple[int, int]:
        ho, wo = C.c_int(0), C.c_int(0)
        _lib.check(self._lib.nst_output_shape(self._h, H, W, C.byref(ho), C.byref(wo)), "nst_output_shape")
        return ho.value, wo.value

    def _workspace(self, N: int, H: int, W: int) -> torch.Tensor:
        key = (N, H, W)
        if self._ws_key != key:
            nbytes = C.c_size_t(0)
            _lib.check(self._lib.nst_workspace_bytes(self._h, N, H, W, C.byref(nbytes)), "nst_workspace_bytes")
            if self._ws is None or self._ws.numel() < nbytes.value:
                self._ws = None
                self._ws = torch.empty(nbytes.value, dtype=torch.uint8, device=f"cuda:{self.device}")
            self._ws_key = key
        return self._ws

    def _stream(self) -> int:
        return torch.cuda.current_stream(self.device).cuda_stream

    def _check_dev(self, t: torch.Tensor, name: str):
        if not t.is_cuda or t.device.index != self.device:
            raise RuntimeError(f"{name} must live on cuda:{self.device} (got {t.device}); there is no CPU path")

    # ---- forward -------------------------------------------------------------------------
    def forward_f32(self, x: torch.Tensor) -> torch.Tensor:
        """``y = model(x)`` (pipeline.py:1474): fp32 NCHW in, fp32 NCHW out, same device."""
        if x.dim() != 4 or x.shape[1] != 3:
            raise ValueError(f"expected [N,3,H,W], got {tuple(x.shape)}")
        self._check_dev(x, "input")
        x = x.to(torch.float32).contiguous()
        N, _, H, W = x.shape
        Ho, Wo = self.output_shape(H, W)
        y = torch.empty(N, 3, Ho, Wo, dtype=torch.float32, device=x.device)
        ws = self._workspace(N, H, W)
        with torch.cuda.device(self.device):
            rc = self._lib.nst_forward_f32_nchw(self._h, x.data_ptr(), y.data_ptr(), N, H, W,
                                                ws.data_ptr(), ws.numel(), self._stream())
        _lib.check(rc, "nst_forward_f32_nchw")
        return y

    def stylize_u8(self, frames: torch.Tensor, preset: str, blend: float = 1.0,
                   mask: Optional[torch.Tensor] = None, composite_mode: str = "keep",
                   out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """uint8 NHWC frames -> stylised uint8 NHWC (whole per-frame path of pipeline.py, fused).

        The result is [N, Ho, Wo, 3] with (Ho, Wo) = ``output_shape(H, W)``: the input size when H, W are
        multiples of 4 (always for the NST variant), else the reference module's larger output
        (pipeline.py:1512-1516 resizes it back; ``FramePipeline`` does that on the device).
        """
        if frames.dim() == 3:
            return self.stylize_u8(frames[None], preset, blend, None if mask is None else mask[None], composite_mode,
                                   None if out is None else out[None])[0]
        if frames.dtype != torch.uint8 or frames.dim() != 4 or frames.shape[-1] != 3:
            raise ValueError(f"expected uint8 [N,H,W,3], got {frames.dtype} {tuple(frames.shape)}")
        self._check_dev(frames, "frames")
        frames = frames.contiguous()
        N, H, W, _ = frames.shape
        Ho, Wo = self.output_shape(H, W)
        if out is None:
            out = torch.empty(N, Ho, Wo, 3, dtype=torch.uint8, device=frames.device)
        else:
            self._check_dev(out, "out")
            if tuple(out.shape) != (N, Ho, Wo, 3) or out.dtype != torch.uint8 or not out.is_contiguous():
                raise ValueError(f"out must be a contiguous uint8 tensor of shape {(N, Ho, Wo, 3)}")
        mptr = None
        if mask is not None:
            self._check_dev(mask, "mask")
            mask = mask.to(torch.float32).contiguous()
            if tuple(mask.shape) != (N, H, W):
                raise ValueError(f"mask must be [N,H,W]={N, H, W}, got {tuple(mask.shape)}")
            mptr = mask.data_ptr()
        s_in, b_in, d_out, s_out, b_out, swap = io_preset_affine(preset)
        ws = self._workspace(N, H, W)
        with torch.cuda.device(self.device):
            rc = self._lib.nst_forward_u8_nhwc(self._h, frames.data_ptr(), out.data_ptr(), N, H, W,
                                               _f3(s_in), _f3(b_in), _f3(d_out), _f3(s_out), _f3(b_out), swap, float(blend),
                                               mptr, _lib.COMPOSITE_MODE[composite_mode],
                                               ws.data_ptr(), ws.numel(), self._stream())
        _lib.check(rc, "nst_forward_u8_nhwc")
        return out

    # ---- introspection -------------------------------------------------------------------
    def layer_info(self):
        n = self._lib.nst_num_layers(self._h)
        out = []
        for i in range(n):
            name = C.create_string_buffer(96)
            fl, by, kind = C.c_double(0), C.c_double(0), C.c_int(0)
            _lib.check(self._lib.nst_layer_info(self._h, i, name, 96, C.byref(fl), C.byref(by), C.byref(kind)), "nst_layer_info")
            out.append({"name": name.value.decode(), "flops": fl.value, "bytes": by.value,
                        "kernel": "tcgen05" if kind.value >= 1 else "direct",
                        "feed": {0: "direct", 1: "gather", 2: "tma"}[kind.value]})
        return out

    def profile_layers(self, frames: torch.Tensor):
        """Per-layer CUDA-event milliseconds for one u8 forward (synchronises)."""
        self._check_dev(frames, "frames")
        frames = frames.contiguous()
        N, H, W, _ = frames.shape
        Ho, Wo = self.output_shape(H, W)
        out = torch.empty(N, Ho, Wo, 3, dtype=torch.uint8, device=frames.device)
        ws = self._workspace(N, H, W)
        n = self._lib.nst_num_layers(self._h)
        ms = (C.c_float * n)()
        with torch.cuda.device(self.device):
            rc = self._lib.nst_profile_layers_u8(self._h, frames.data_ptr(), out.data_ptr(), N, H, W,
                                                 ws.data_ptr(), ws.numel(), self._stream(), ms)
        _lib.check(rc, "nst_profile_layers_u8")
        info = self.layer_info()
        for i, d in enumerate(info):
            d["ms"] = float(ms[i])
        return info

    @property
    def launch_count(self) -> int:
        return int(self._lib.nst_launch_count(self._h))


# ---- region compositing (region_blend.py:1049-1108, 1589-1679) ------------------------------
def composite_regions(styled: Sequence[torch.Tensor], masks: torch.Tensor,
                      model_indices: Sequence[Sequence[int]], model_weights: Optional[Sequence[Sequence[float]]] = None,
                      original: Optional[torch.Tensor] = None) -> torch.Tensor:
    """On-GPU multi-style composite.

    styled: K tensors, all uint8 [H,W,3] or all fp32 [3,H,W] (0..1); masks: fp32 [R,H,W];
    model_indices[r] lists the source(s) of region r (-1 = original), model_weights[r] their
    blend weights (default 1.0 each).  Returns the composited frame in the input's format.
    """
    lib = _lib.load()
    K, R = len(styled), masks.shape[0]
    if len(model_indices) != R or (model_weights is not None and len(model_weights) != R):
        raise ValueError(f"model_indices / model_weights must have one entry per region ({R})")
    u8 = styled[0].dtype == torch.uint8
    dev = styled[0].device
    if not styled[0].is_cuda:
        raise RuntimeError("composite_regions needs CUDA tensors; there is no CPU path")
    H, W = (styled[0].shape[0], styled[0].shape[1]) if u8 else (styled[0].shape[1], styled[0].shape[2])
    idx = [list(ix) if isinstance(ix, (list, tuple)) else [int(ix)] for ix in model_indices]
    J = max(len(ix) for ix in idx)
    wts = [list(w) for w in model_weights] if model_weights is not None else [[1.0] * len(ix) for ix in idx]
    flat_i, flat_w = [], []
    for ix, w in zip(idx, wts):
        flat_i += ix + [0] * (J - len(ix))
        flat_w += list(w) + [0.0] * (J - len(w))
    want = (H, W, 3) if u8 else (3, H, W)
    for t in list(styled) + ([original] if original is not None else []):
        if tuple(t.shape) != want or t.dtype != styled[0].dtype or t.device != dev:
            raise ValueError(f"every frame must be {styled[0].dtype} {want} on {dev}; got {t.dtype} {tuple(t.shape)} on {t.device}")
    if tuple(masks.shape) != (R, H, W) or masks.device != dev:
        raise ValueError(f"masks must be [R,H,W] = {(R, H, W)} on {dev}, got {tuple(masks.shape)} on {masks.device}")
    srcs = [s.contiguous() for s in styled]
    masks = masks.to(torch.float32).contiguous()
    orig = original.contiguous() if original is not None else None
    out = torch.empty_like(srcs[0])
    c_srcs = (C.c_void_p * K)(*[s.data_ptr() for s in srcs])
    fn = lib.nst_composite_regions_u8 if u8 else lib.nst_composite_regions_f32
    with torch.cuda.device(dev):
        rc = fn(C.cast(c_srcs, C.POINTER(C.c_void_p)), K, orig.data_ptr() if orig is not None else None,
                masks.data_ptr(), R, J, (C.c_int * (R * J))(*flat_i), (C.c_float * (R * J))(*flat_w),
                out.data_ptr(), H, W, torch.cuda.current_stream(dev).cuda_stream)
    _lib.check(rc, "nst_composite_regions")
    return out
