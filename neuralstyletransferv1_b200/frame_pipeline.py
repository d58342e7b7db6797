"""Batched frame I/O around the stylisation call.

The reference walks a video one frame at a time: PIL decode -> ``ToTensor`` -> blocking H2D ->
``model(x)`` -> ``.cpu()`` -> ``float(y.min())`` syncs -> ``ToPILImage`` -> JPEG
(pipeline.py:1080-1100, 1473-1476, 1493-1497, 2094-2119).  With the forward at ~1 ms per 1080p
frame that host path is the whole cost, so this module is the batched caller a maintainer would
put in its place: uint8 NHWC frames in (host memory), uint8 NHWC frames out (host memory), in
order.

``FramePipeline`` (one GPU): frames are cut into chunks; chunk i+1 is uploaded (copy stream) and
chunk i-1 downloaded (second copy stream: PCIe is full duplex) while chunk i is stylised
(compute stream).  Device staging is a fixed ring of slots, host staging is pinned; nothing is
allocated per chunk.  Frame sizes that are not a multiple of 4 come back from the generator 4
pixels larger (stride-2 convs round up) and are resized back to the input size on the device
with the reference's own bilinear resize *before* blending / packing (pipeline.py:1512-1516),
through ``Engine.stylize_u8_fit``.

``MultiGpuFramePipeline`` (one process, G GPUs): one worker thread + one model replica + one
``FramePipeline`` per GPU, contiguous frame ranges (``sharding.shard_range``), no collective.
ctypes releases the GIL inside every C-ABI call, so the workers overlap.
"""
from __future__ import annotations

import threading
from typing import List, Optional, Sequence

import numpy as np
import torch

from .sharding import shard_range


def _as_u8_frames(frames) -> torch.Tensor:
    if isinstance(frames, np.ndarray):
        frames = torch.from_numpy(np.ascontiguousarray(frames))
    if not torch.is_tensor(frames) or frames.dtype != torch.uint8 or frames.dim() != 4 or frames.shape[-1] != 3:
        raise ValueError("frames must be uint8 [F, H, W, 3] (torch tensor or numpy array)")
    if frames.is_cuda:
        raise ValueError("FramePipeline takes HOST frames; call model.stylize_u8 directly for device tensors")
    return frames.contiguous()


class FramePipeline:
    """frames in -> stylised frames out, in order, on one GPU.

    model        a drop-in generator (``TransformerNet`` / ``TransformerNetNST`` / ``ReCoNet``) already on its CUDA device
    io_preset    pipeline.py:1447-1486 preset name
    chunk        frames per ``stylize_u8`` call (the library's micro-batch is 8)
    slots        depth of the device staging ring (>= 2)
    ramp         clips of >= 4 chunks start and end with a quarter-size chunk (shorter un-overlapped first upload / last download); off by
                 default: measured on 64 x 1080p it does not pay (e2e 1061-1077 against 1109 frames/s -- the small chunks run the kernels
                 at a less efficient micro-batch)
    """

    def __init__(self, model, io_preset: str, *, chunk: int = 8, slots: int = 3, blend: float = 1.0,
                 composite_mode: str = "keep", ramp: bool = False):
        p = next(model.parameters())
        if p.device.type != "cuda":
            raise RuntimeError("FramePipeline needs the model on a CUDA device; there is no CPU path")
        if chunk < 1 or slots < 2:
            raise ValueError("chunk >= 1 and slots >= 2")
        self.model, self.preset, self.chunk, self.slots = model, io_preset, int(chunk), int(slots)
        self.blend, self.composite_mode = float(blend), composite_mode
        self.ramp = bool(ramp)
        self.device = p.device
        self._s_in = torch.cuda.Stream(self.device)
        self._s_run = torch.cuda.Stream(self.device)
        self._s_out = torch.cuda.Stream(self.device)
        self._shape = None
        self._d_in: List[torch.Tensor] = []
        self._d_out: List[torch.Tensor] = []
        self._d_mask: List[Optional[torch.Tensor]] = []
        self._h_in: List[torch.Tensor] = []
        self._h_out: List[torch.Tensor] = []

    # -- staging ---------------------------------------------------------------------------
    def _prepare(self, H: int, W: int, with_mask: bool, pin_in: bool, pin_out: bool):
        key = (H, W, with_mask, pin_in, pin_out)
        if self._shape == key:
            return
        c, dev = self.chunk, self.device
        self._d_in = [torch.empty(c, H, W, 3, dtype=torch.uint8, device=dev) for _ in range(self.slots)]
        self._d_out = [torch.empty(c, H, W, 3, dtype=torch.uint8, device=dev) for _ in range(self.slots)]
        self._d_mask = [torch.empty(c, H, W, dtype=torch.float32, device=dev) if with_mask else None for _ in range(self.slots)]
        # pinned host staging only where the caller's buffers are pageable
        self._h_in = [torch.empty(c, H, W, 3, dtype=torch.uint8).pin_memory() for _ in range(self.slots)] if pin_in else []
        self._h_out = [torch.empty(c, H, W, 3, dtype=torch.uint8).pin_memory() for _ in range(self.slots)] if pin_out else []
        self._shape = key

    # -- the batched call ------------------------------------------------------------------
    def __call__(self, frames, out: Optional[torch.Tensor] = None, masks: Optional[torch.Tensor] = None) -> torch.Tensor:
        """Stylise ``frames`` (host uint8 [F,H,W,3]); returns host uint8 [F,H,W,3] (``out`` if given).

        ``masks``: optional host fp32 [F,H,W] alpha for the mask composite (pipeline.py:2040-2043).
        Pinned ``frames`` / ``out`` are copied directly; pageable ones go through pinned staging.
        """
        frames = _as_u8_frames(frames)
        F, H, W, _ = frames.shape
        if out is None:
            out = torch.empty(F, H, W, 3, dtype=torch.uint8).pin_memory()
        if tuple(out.shape) != (F, H, W, 3) or out.dtype != torch.uint8 or out.is_cuda or not out.is_contiguous():
            raise ValueError(f"out must be a contiguous host uint8 tensor of shape {(F, H, W, 3)}")
        if masks is not None and (tuple(masks.shape) != (F, H, W) or masks.is_cuda):
            raise ValueError(f"masks must be host fp32 [F,H,W] = {(F, H, W)}")
        if masks is not None:
            masks = masks.to(torch.float32).contiguous()
        pin_in, pin_out = not frames.is_pinned(), not out.is_pinned()
        self._prepare(H, W, masks is not None, pin_in, pin_out)
        S, c = self.slots, self.chunk
        # chunk schedule: full chunks, but a short first and last one when the clip is long enough -- the first upload and the last download
        # are the only copies nothing overlaps, so they are kept small (c / 4 frames); everything in between moves c frames at a time
        r = max(1, c // 4)
        if self.ramp and F >= 4 * c and r < c:
            k, rem = divmod(F - 2 * r, c)
            sizes = [r] + [c] * k + ([rem] if rem else []) + [r]
        else:
            sizes = [c] * (F // c) + ([F % c] if F % c else [])
        bounds, a0 = [], 0
        for n_ in sizes:
            bounds.append((a0, a0 + n_))
            a0 += n_
        n_chunks = len(bounds)
        ev_in = [None] * n_chunks      # upload of chunk i finished
        ev_run = [None] * n_chunks     # stylisation of chunk i finished
        ev_out = [None] * n_chunks     # download of chunk i finished
        with torch.cuda.device(self.device):
            for i in range(n_chunks):
                a, b = bounds[i]
                n, s = b - a, i % S
                # ---- upload (copy stream 1); slot s was last read by the stylisation of chunk i - S
                if i >= S:
                    self._s_in.wait_event(ev_run[i - S])
                    if pin_in:
                        ev_in[i - S].synchronize()           # the pinned staging slot is free once its upload has finished
                src = frames[a:b]
                if pin_in:
                    self._h_in[s][:n].copy_(src)
                    src = self._h_in[s][:n]
                with torch.cuda.stream(self._s_in):
                    self._d_in[s][:n].copy_(src, non_blocking=True)
                    if masks is not None:
                        self._d_mask[s][:n].copy_(masks[a:b], non_blocking=True)
                    ev_in[i] = torch.cuda.Event()
                    ev_in[i].record(self._s_in)
                # ---- stylise (compute stream); the output slot was last read by the download of chunk i - S
                with torch.cuda.stream(self._s_run):
                    self._s_run.wait_event(ev_in[i])
                    if i >= S:
                        self._s_run.wait_event(ev_out[i - S])
                    self.model.engine.stylize_u8_fit(self._d_in[s][:n], self.preset, self.blend,
                                                     None if masks is None else self._d_mask[s][:n],
                                                     self.composite_mode, out=self._d_out[s][:n])
                    ev_run[i] = torch.cuda.Event()
                    ev_run[i].record(self._s_run)
                # ---- download (copy stream 2)
                if pin_out and i >= S:
                    ev_out[i - S].synchronize()
                    pa, pb = bounds[i - S]
                    out[pa:pb].copy_(self._h_out[s][:pb - pa])
                with torch.cuda.stream(self._s_out):
                    self._s_out.wait_event(ev_run[i])
                    dst = self._h_out[s][:n] if pin_out else out[a:b]
                    dst.copy_(self._d_out[s][:n], non_blocking=True)
                    ev_out[i] = torch.cuda.Event()
                    ev_out[i].record(self._s_out)
            self._s_out.synchronize()
            if pin_out:                                   # drain the last staged chunks
                for i in range(max(0, n_chunks - S), n_chunks):
                    a, b = bounds[i]
                    out[a:b].copy_(self._h_out[i % S][:b - a])
        return out


class MultiGpuFramePipeline:
    """One process, one worker thread + model replica + ``FramePipeline`` per GPU.

    ``model_factory(device_index) -> model on cuda:device_index`` builds each replica (weights are
    replicated; frames are independent, so there is no collective).  ``devices`` may repeat an index
    (several workers on one GPU) -- useful for testing on a single-GPU box.
    """

    def __init__(self, model_factory, io_preset: str, devices: Optional[Sequence[int]] = None, **pipeline_kwargs):
        if devices is None:
            devices = list(range(torch.cuda.device_count()))
        if not devices:
            raise RuntimeError("MultiGpuFramePipeline needs at least one CUDA device; there is no CPU path")
        self.devices = [int(d) for d in devices]
        self.pipes = [FramePipeline(model_factory(d), io_preset, **pipeline_kwargs) for d in self.devices]

    def __call__(self, frames, out: Optional[torch.Tensor] = None, masks: Optional[torch.Tensor] = None) -> torch.Tensor:
        frames = _as_u8_frames(frames)
        F, H, W, _ = frames.shape
        if out is None:
            out = torch.empty(F, H, W, 3, dtype=torch.uint8).pin_memory()
        G = len(self.pipes)
        errors: List[BaseException] = []

        def work(rank: int):
            a, b = shard_range(F, G, rank)
            if a == b:
                return
            try:
                self.pipes[rank](frames[a:b], out=out[a:b], masks=None if masks is None else masks[a:b])
            except BaseException as e:  # surfaced in the caller's thread
                errors.append(e)

        threads = [threading.Thread(target=work, args=(r,), daemon=True) for r in range(G)]
        for t in threads:
            t.start()
        for t in threads:
            t.join()
        if errors:
            raise errors[0]
        return out
