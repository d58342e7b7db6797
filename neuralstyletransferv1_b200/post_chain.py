"""The per-frame tail of the reference loop on the device: what `style_frames` does to a stylised
frame between the network and the encoder (pipeline.py:1884-2094), for a stream of frames.

    out01 (network output, clamped 0..1, content size)
      -> flow-guided EMA with the previous fused frame            pipeline.py:1884-1936   engine.flow_ema           (needs the optical flow)
      -> ToPILImage (truncating uint8 pack)                       :1942                   engine.resize_post_pack
      -> LAB lightness / chroma EMA                               :1943-1978              temporal.LabEma           (a chunk per launch)
      -> ToTensor, mask composite, motion-adaptive / uniform blend, ToPILImage   :1981-2094   temporal.motion_blend / temporal.blend_pack

Every step is one of the library's kernels; this class only carries the state (previous fused frame, LAB planes) and the order.  The
optical flow itself stays where the reference computes it (OpenCV DIS / Farneback on the host): the caller passes one
`[H, W, 2]` flow per frame, or None for a frame without one (first frame, failed estimate -- pipeline.py:1933-1936).
"""
from __future__ import annotations

from typing import Optional, Sequence

import torch

from . import engine
from .temporal import GAUSS_SIGMA, MIN_ALPHA, MOTION_NORM, LabEma, blend_pack, motion_blend


class PostChain:
    def __init__(self, flow_ema: bool = False, flow_alpha: float = 0.85, smooth_lightness: bool = False, smooth_chroma: bool = False,
                 smooth_alpha: float = 0.7, chroma_alpha: float = 0.85, blend: float = 1.0, motion_blend: bool = False,
                 composite_mode: str = "keep"):
        self.flow_ema, self.flow_alpha = bool(flow_ema), float(flow_alpha)
        self.lab = LabEma(smooth_lightness, smooth_chroma, smooth_alpha, chroma_alpha)
        self.blend, self.motion_blend, self.composite_mode = float(blend), bool(motion_blend), composite_mode
        self._prev: Optional[torch.Tensor] = None        # previous fused frame, fp32 [H, W, 3] (pipeline.py: prev_styled01)

    def reset(self) -> None:
        self._prev = None
        self.lab.reset()

    def __call__(self, out01: torch.Tensor, originals_u8: torch.Tensor, flows: Optional[Sequence[Optional[torch.Tensor]]] = None,
                 masks: Optional[torch.Tensor] = None) -> torch.Tensor:
        """out01: fp32 [N, 3, H, W] CUDA, the network output after the preset's post-affine and clamp, at the content size;
        originals_u8: uint8 [N, H, W, 3]; flows: N entries, fp32 [H, W, 2] = (dx, dy) or None; masks: fp32 [N, H, W] or None.
        Frames are consecutive frames of one clip; state carries over to the next call.  Returns uint8 [N, H, W, 3]."""
        if not out01.is_cuda:
            raise RuntimeError("PostChain needs CUDA tensors; there is no CPU path")
        N, _, H, W = out01.shape
        if flows is None:
            flows = [None] * N
        if len(flows) != N or tuple(originals_u8.shape) != (N, H, W, 3):
            raise ValueError("one flow entry and one original frame per frame")
        out01 = out01.to(torch.float32)
        fused = out01
        last_flow = [None] * N
        if self.flow_ema:
            # sequential by construction: frame i blends with the FUSED frame i - 1
            hwc = out01.permute(0, 2, 3, 1).contiguous()                     # the reference's numpy layout
            if self._prev is not None and tuple(self._prev.shape) != (H, W, 3):
                self._prev = None
            for i in range(N):
                if flows[i] is not None and self._prev is not None:
                    engine.flow_ema(hwc[i], self._prev, flows[i], self.flow_alpha, out=hwc[i])
                    last_flow[i] = flows[i]
                self._prev = hwc[i]
            self._prev = self._prev.clone()                                   # detach the state from this call's buffer
            fused = hwc.permute(0, 3, 1, 2).contiguous()
        # ToPILImage of the fused frame, then the LAB EMA over the whole chunk
        smooth = self.lab(engine.resize_post_pack(fused, H, W))
        moving = [self.motion_blend and last_flow[i] is not None and masks is None for i in range(N)]
        blend_on = 0.0 <= self.blend < 1.0
        if masks is None and not blend_on and not any(moving):
            return smooth                                                     # to_pil(to_tensor(u8)) is the identity
        out = torch.empty_like(smooth)
        rest = [i for i in range(N) if not moving[i]]
        if rest:
            idx = torch.tensor(rest, device=smooth.device)
            out[idx] = blend_pack(smooth[idx], originals_u8[idx], self.blend, masks[idx] if masks is not None else None, self.composite_mode)
        for i in range(N):
            if moving[i]:
                out[i] = motion_blend(smooth[i], originals_u8[i], last_flow[i], self.blend, MIN_ALPHA, MOTION_NORM, GAUSS_SIGMA)
        return out
