"""Drop-in ``nn.Module`` classes with the reference's construction / state_dict / call surface.

``TransformerNet()`` (transformer_net.py:4), ``TransformerNetNST()``
(transformer_net_nst.py:62) and ``ReCoNet(frn=False)`` (model.py:107) keep real
``nn.Parameter`` objects under the reference's exact key names, so
``.to(device)``, ``.state_dict()``, ``.load_state_dict(state, strict=...)`` and
``.eval()`` behave as the callers in pipeline.py:554-619 expect.  ``forward``
runs the sm_100a kernels through the C ABI; it never falls back to PyTorch ops.
"""
from __future__ import annotations

import math
from typing import Optional

import torch
from torch import nn

from . import synth
from .engine import Engine

_DROP_TOKENS = ("running_mean", "running_var", "num_batches_tracked")   # pipeline.py:565


class _Holder(nn.Module):
    """Parameter container; exists only so dotted state_dict keys match the reference."""

    def forward(self, *a, **k):  # pragma: no cover
        raise RuntimeError("parameter holder is not callable; call the top-level model")


class _B200Generator(nn.Module):
    ARCH = ""

    def __init__(self, precision: str = "f16"):
        super().__init__()
        self.precision = precision
        shapes = {k: tuple(v.shape) for k, v in synth.make_state_dict(self.ARCH, 0).items()}
        for key, shape in shapes.items():
            *path, leaf = key.split(".")
            mod = self
            for part in path:
                if part not in mod._modules:
                    mod.add_module(part, _Holder())
                mod = mod._modules[part]
            t = torch.empty(shape)
            if len(shape) == 4:                      # conv weight: PyTorch default init
                nn.init.kaiming_uniform_(t, a=math.sqrt(5))
            elif leaf == "weight":                   # InstanceNorm gamma
                t.fill_(1.0)
            else:                                    # conv bias / InstanceNorm beta
                t.zero_()
            mod.register_parameter(leaf, nn.Parameter(t))
        self._engine: Optional[Engine] = None
        self._synced_version = None

    # -- weight synchronisation ------------------------------------------------------------
    def _weights_version(self):
        return tuple((p.data_ptr(), p._version) for p in self.parameters())

    def load_state_dict(self, state_dict, strict: bool = True, assign: bool = False):
        # legacy InstanceNorm buffers of pytorch/examples checkpoints carry no information
        # (track_running_stats=False); the reference drops them too (pipeline.py:565-566).
        state_dict = {k: v for k, v in state_dict.items() if not any(t in k for t in _DROP_TOKENS)}
        self._synced_version = None
        return super().load_state_dict(state_dict, strict=strict, assign=assign)

    def _ensure_engine(self, device: torch.device) -> Engine:
        if device.type != "cuda":
            raise RuntimeError(
                f"{type(self).__name__} (B200 drop-in) only runs on CUDA sm_100a devices, got {device}; "
                "there is no CPU/MPS fallback -- use the reference module for those devices")
        idx = device.index if device.index is not None else torch.cuda.current_device()
        if self._engine is None or self._engine.device != idx:
            if self._engine is not None:
                self._engine.close()
            self._engine = Engine(self.ARCH, idx, self.precision)
            self._synced_version = None
        ver = self._weights_version()
        if self._synced_version != ver:
            self._engine.load_state_dict({k: v for k, v in self.state_dict().items()})
            self._synced_version = ver
        return self._engine

    @property
    def engine(self) -> Engine:
        p = next(self.parameters())
        return self._ensure_engine(p.device)

    # -- reference call surface ------------------------------------------------------------
    def forward(self, x: torch.Tensor) -> torch.Tensor:
        """``model(x_in)``: fp32 NCHW [N,3,H,W] on the model's device -> fp32 NCHW (pipeline.py:1474)."""
        if x.dim() != 4 or x.shape[1] != 3:
            raise RuntimeError(f"expected input [N,3,H,W], got {tuple(x.shape)}")
        eng = self._ensure_engine(x.device)
        with torch.no_grad():
            return eng.forward_f32(x)

    def stylize_u8(self, frames_u8_nhwc: torch.Tensor, io_preset: str, blend: float = 1.0,
                   mask: Optional[torch.Tensor] = None, composite_mode: str = "keep",
                   out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """Fused uint8 path (mirrors lib.ReCoNetModel.run, lib.py:27-51, for every family)."""
        eng = self._ensure_engine(frames_u8_nhwc.device)
        return eng.stylize_u8(frames_u8_nhwc, io_preset, blend, mask, composite_mode, out)


class TransformerNet(_B200Generator):
    """Johnson-style generator; key names of transformer_net.py:8-25."""
    ARCH = "johnson"


class TransformerNetNST(_B200Generator):
    """NST_Train variant; key names of transformer_net_nst.py:73-93 (selected by `down1.` keys, pipeline.py:72-79)."""
    ARCH = "nst"


class ReCoNet(_B200Generator):
    """ReCoNet(frn=False); key names of model.py:75-116.  ``frn=True`` is unreachable from pipeline.py:602."""
    ARCH = "reconet"

    def __init__(self, frn: bool = False, precision: str = "f16"):
        if frn:
            raise NotImplementedError("ReCoNet(frn=True) is not part of the pipeline hot path (pipeline.py:602 builds frn=False)")
        super().__init__(precision)
