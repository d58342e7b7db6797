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

    def stylize_u8_fit(self, frames: torch.Tensor, preset: str, blend: float = 1.0,
                       mask: Optional[torch.Tensor] = None, composite_mode: str = "keep",
                       out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """``stylize_u8`` whose result always has the input's H x W: sizes that are not multiples of 4 are resized back on the
        device with the reference's bilinear resize before blending and packing (pipeline.py:1512-1516)."""
        return self.stylize_u8(frames, preset, blend, mask, composite_mode, out, fit=True)

    def stylize_u8(self, frames: torch.Tensor, preset: str, blend: float = 1.0,
                   mask: Optional[torch.Tensor] = None, composite_mode: str = "keep",
                   out: Optional[torch.Tensor] = None, fit: bool = False) -> torch.Tensor:
        """uint8 NHWC frames -> stylised uint8 NHWC (whole per-frame path of pipeline.py, fused).

        The result is [N, Ho, Wo, 3] with (Ho, Wo) = ``output_shape(H, W)``: the input size when H, W are
        multiples of 4 (always for the NST variant), else the reference module's larger output
        (pipeline.py:1512-1516 resizes it back; ``FramePipeline`` does that on the device).
        """
        if frames.dim() == 3:
            return self.stylize_u8(frames[None], preset, blend, None if mask is None else mask[None], composite_mode,
                                   None if out is None else out[None], fit)[0]
        if frames.dtype != torch.uint8 or frames.dim() != 4 or frames.shape[-1] != 3:
            raise ValueError(f"expected uint8 [N,H,W,3], got {frames.dtype} {tuple(frames.shape)}")
        self._check_dev(frames, "frames")
        frames = frames.contiguous()
        N, H, W, _ = frames.shape
        Ho, Wo = (H, W) if fit else self.output_shape(H, W)
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
            fn = self._lib.nst_forward_u8_nhwc_fit if fit else self._lib.nst_forward_u8_nhwc
            rc = fn(self._h, frames.data_ptr(), out.data_ptr(), N, H, W,
                    _f3(s_in), _f3(b_in), _f3(d_out), _f3(s_out), _f3(b_out), swap, float(blend),
                    mptr, _lib.COMPOSITE_MODE[composite_mode], ws.data_ptr(), ws.numel(), self._stream())
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


def composite_regions_batch(styled: Sequence[torch.Tensor], masks: torch.Tensor,
                            model_indices: Sequence[Sequence[int]], model_weights: Optional[Sequence[Sequence[float]]] = None,
                            original: Optional[torch.Tensor] = None, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``composite_regions`` / ``composite_regions_advanced`` over a BATCH of frames in one kernel launch.

    styled: K uint8 tensors [N, Hk, Wk, 3]; a source smaller than the frame (a region styled at scale < 1,
    region_blend.py:1662-1668) is upscaled bilinearly per pixel.  masks: [R,H,W] (shared by the batch) or [N,R,H,W];
    float32, float16 or uint8 (value / 255).  original: uint8 [N,H,W,3] (index -1).  Returns uint8 [N,H,W,3].
    """
    lib = _lib.load()
    K = len(styled)
    if K < 1 or not styled[0].is_cuda:
        raise RuntimeError("composite_regions_batch needs CUDA tensors; there is no CPU path")
    dev = styled[0].device
    per_frame = masks.dim() == 4
    R = masks.shape[1] if per_frame else masks.shape[0]
    H, W = masks.shape[-2], masks.shape[-1]
    N = styled[0].shape[0]
    if len(model_indices) != R or (model_weights is not None and len(model_weights) != R):
        raise ValueError(f"model_indices / model_weights must have one entry per region ({R})")
    if str(masks.dtype).replace("torch.", "") not in _lib.MASK_TYPE:
        raise ValueError(f"masks must be float32, float16 or uint8, got {masks.dtype}")
    if masks.device != dev or (per_frame and masks.shape[0] != N):
        raise ValueError(f"masks must be [R,H,W] or [N,R,H,W] on {dev}")
    for t in styled:
        if t.dtype != torch.uint8 or t.dim() != 4 or t.shape[0] != N or t.shape[-1] != 3 or t.device != dev:
            raise ValueError(f"every styled batch must be uint8 [N={N},h,w,3] on {dev}; got {t.dtype} {tuple(t.shape)}")
    if original is not None and (tuple(original.shape) != (N, H, W, 3) or original.dtype != torch.uint8 or original.device != dev):
        raise ValueError(f"original must be uint8 {(N, H, W, 3)} on {dev}")
    idx = [list(ix) if isinstance(ix, (list, tuple)) else [int(ix)] for ix in model_indices]
    J = max(len(ix) for ix in idx)
    wts = [list(w) for w in model_weights] if model_weights is not None else [[1.0] * len(ix) for ix in idx]
    flat_i, flat_w = [], []
    for ix, w in zip(idx, wts):
        flat_i += ix + [-2] * (J - len(ix))
        flat_w += list(w) + [0.0] * (J - len(w))
    srcs = [t.contiguous() for t in styled]
    masks = masks.contiguous()
    orig = original.contiguous() if original is not None else None
    if out is None:
        out = torch.empty(N, H, W, 3, dtype=torch.uint8, device=dev)
    elif tuple(out.shape) != (N, H, W, 3) or out.dtype != torch.uint8 or not out.is_contiguous() or out.device != dev:
        raise ValueError(f"out must be a contiguous uint8 tensor {(N, H, W, 3)} on {dev}")
    c_srcs = (C.c_void_p * K)(*[t.data_ptr() for t in srcs])
    hw = []
    for t in srcs:
        hw += [t.shape[1], t.shape[2]]
    with torch.cuda.device(dev):
        rc = lib.nst_composite_regions_batch_u8(C.cast(c_srcs, C.POINTER(C.c_void_p)), (C.c_int * (2 * K))(*hw), K,
                                                orig.data_ptr() if orig is not None else None, masks.data_ptr(),
                                                _lib.MASK_TYPE[str(masks.dtype).replace("torch.", "")], 1 if per_frame else 0, R, J,
                                                (C.c_int * (R * J))(*flat_i), (C.c_float * (R * J))(*flat_w),
                                                out.data_ptr(), N, H, W, torch.cuda.current_stream(dev).cuda_stream)
    _lib.check(rc, "nst_composite_regions_batch_u8")
    return out


def composite_from_crops(styled_crops, crops, original: Optional[torch.Tensor], H: int, W: int,
                         return_u8: bool = False):
    """region_blend.composite_from_crops (region_blend.py:2186-2294) on the device.

    styled_crops: ``{model_idx: {region_idx: fp32 [3,h,w] CUDA tensor}}`` exactly like the reference's nested dict.
    crops: a sequence of objects / dicts with ``region_idx``, ``padded_bbox`` (x1, y1, x2, y2), ``crop_mask``
    (fp32 [1,1,h,w] or [h,w]), ``model_indices`` and ``model_weights`` (the reference's ``RegionCrop`` with its
    ``config`` flattened is accepted as is).  original: fp32 [3,H,W] CUDA tensor or None.  Returns the composited frame
    fp32 [3,H,W] (and its uint8 [H,W,3] pack when ``return_u8``).
    """
    lib = _lib.load()
    MAXJ = 4
    R = len(crops)
    dev = None
    bbox, mask_hw, nterms, term_hw, term_w = [], [], [], [0] * (R * MAXJ * 2), [0.0] * (R * MAXJ)
    mask_ptrs, term_ptrs, keep = [], [0] * (R * MAXJ), []

    def field(c, name):
        if isinstance(c, dict):
            return c[name]
        if hasattr(c, name):
            return getattr(c, name)
        return getattr(c.config, name)          # RegionCrop.config.model_indices / model_weights

    for r, c in enumerate(crops):
        x1, y1, x2, y2 = [int(v) for v in field(c, "padded_bbox")]
        bbox += [x1, y1, x2, y2]
        m = field(c, "crop_mask")
        m = m.reshape(m.shape[-2], m.shape[-1]).to(torch.float32).contiguous()
        if not m.is_cuda:
            raise RuntimeError("composite_from_crops needs CUDA tensors; there is no CPU path")
        dev = m.device
        keep.append(m)
        mask_ptrs.append(m.data_ptr())
        mask_hw += [m.shape[0], m.shape[1]]
        idxs, wts = list(field(c, "model_indices")), list(field(c, "model_weights"))
        if not 1 <= len(idxs) <= MAXJ:
            raise ValueError(f"region {r}: 1..{MAXJ} blended models per region are supported")
        nterms.append(len(idxs))
        for j, (mi, wt) in enumerate(zip(idxs, wts)):
            term_w[r * MAXJ + j] = float(wt)
            if mi == -1:
                if original is None:
                    raise ValueError("Region uses original but no original provided")           # region_blend.py:2231
                continue
            if mi not in styled_crops:
                raise ValueError(f"Model {mi} not in styled_crops")                             # region_blend.py:2236
            ridx = int(field(c, "region_idx"))
            if ridx not in styled_crops[mi]:
                raise ValueError(f"Region {ridx} not styled by model {mi}")                     # region_blend.py:2238
            t = styled_crops[mi][ridx].to(torch.float32).contiguous()
            keep.append(t)
            term_ptrs[r * MAXJ + j] = t.data_ptr()
            term_hw[(r * MAXJ + j) * 2], term_hw[(r * MAXJ + j) * 2 + 1] = t.shape[-2], t.shape[-1]
    orig = None
    if original is not None:
        orig = original.reshape(3, H, W).to(torch.float32).contiguous()
        dev = orig.device
    out = torch.empty(3, H, W, dtype=torch.float32, device=dev)
    out_u8 = torch.empty(H, W, 3, dtype=torch.uint8, device=dev) if return_u8 else None
    scratch = None if orig is not None else torch.empty(8 * H * W, dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        rc = lib.nst_composite_from_crops_f32(
            R, (C.c_int * (4 * R))(*bbox), C.cast((C.c_void_p * R)(*mask_ptrs), C.POINTER(C.c_void_p)), (C.c_int * (2 * R))(*mask_hw),
            (C.c_int * R)(*nterms), C.cast((C.c_void_p * (R * MAXJ))(*[p or None for p in term_ptrs]), C.POINTER(C.c_void_p)),
            (C.c_int * (R * MAXJ * 2))(*term_hw), (C.c_float * (R * MAXJ))(*term_w),
            orig.data_ptr() if orig is not None else None, out.data_ptr(), out_u8.data_ptr() if out_u8 is not None else None,
            scratch.data_ptr() if scratch is not None else None, H, W, torch.cuda.current_stream(dev).cuda_stream)
    _lib.check(rc, "nst_composite_from_crops_f32")
    return (out, out_u8) if return_u8 else out


def flow_ema(curr01: torch.Tensor, prev01: torch.Tensor, flow: torch.Tensor, alpha: float,
             out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Flow-guided EMA of pipeline.py:1918-1931: ``clip(a * curr + (1 - a) * remap(prev, grid + flow), 0, 1)``.

    curr01 / prev01: fp32 [H,W,3] CUDA tensors in 0..1 (the reference's numpy layout); flow: fp32 [H,W,2] = (dx, dy).
    """
    lib = _lib.load()
    if not curr01.is_cuda:
        raise RuntimeError("flow_ema needs CUDA tensors; there is no CPU path")
    H, W, _ = curr01.shape
    for t, shp, name in ((curr01, (H, W, 3), "curr01"), (prev01, (H, W, 3), "prev01"), (flow, (H, W, 2), "flow")):
        if tuple(t.shape) != shp or t.dtype != torch.float32 or t.device != curr01.device:
            raise ValueError(f"{name} must be float32 {shp} on {curr01.device}")
    curr01, prev01, flow = curr01.contiguous(), prev01.contiguous(), flow.contiguous()
    if out is None:
        out = torch.empty_like(curr01)
    with torch.cuda.device(curr01.device):
        rc = lib.nst_flow_ema_f32(curr01.data_ptr(), prev01.data_ptr(), flow.data_ptr(), out.data_ptr(), H, W, float(alpha),
                                  torch.cuda.current_stream(curr01.device).cuda_stream)
    _lib.check(rc, "nst_flow_ema_f32")
    return out


def resize_post_pack(out01: torch.Tensor, H: int, W: int, original: Optional[torch.Tensor] = None, mask: Optional[torch.Tensor] = None,
                     blend: float = 1.0, composite_mode: str = "keep") -> torch.Tensor:
    """fp32 out01 frames [N,3,h,w] -> uint8 [N,H,W,3]: bilinear resize to the content size (pipeline.py:1512-1516), mask composite,
    blend, truncating pack (pipeline.py:2040-2094).  original: uint8 [N,H,W,3]; mask: fp32 [N,H,W]."""
    lib = _lib.load()
    if not out01.is_cuda:
        raise RuntimeError("resize_post_pack needs CUDA tensors; there is no CPU path")
    out01 = out01.to(torch.float32).contiguous()
    N, _, hi, wi = out01.shape
    out = torch.empty(N, H, W, 3, dtype=torch.uint8, device=out01.device)
    orig = original.contiguous() if original is not None else None
    m = mask.to(torch.float32).contiguous() if mask is not None else None
    with torch.cuda.device(out01.device):
        rc = lib.nst_resize_post_pack_u8(out01.data_ptr(), N, hi, wi, orig.data_ptr() if orig is not None else None,
                                         m.data_ptr() if m is not None else None, float(blend), _lib.COMPOSITE_MODE[composite_mode],
                                         out.data_ptr(), H, W, torch.cuda.current_stream(out01.device).cuda_stream)
    _lib.check(rc, "nst_resize_post_pack_u8")
    return out
