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

from .engine import Engine

_DROP_TOKENS = ("running_mean", "running_var", "num_batches_tracked")   # pipeline.py:565


def _conv(name: str, cout: int, cin: int, k: int, w: str = ".weight", b: str = ".bias"):
    return [(name + w, (cout, cin, k, k)), (name + b, (cout,))]


def _norm(name: str, c: int):
    return [(name + ".weight", (c,)), (name + ".bias", (c,))]


def _johnson_shapes():
    """state_dict keys / shapes of transformer_net.TransformerNet (transformer_net.py:8-25)."""
    out = []
    for name, cout, cin, k, nrm in (("conv1", 32, 3, 9, "in1"), ("conv2", 64, 32, 3, "in2"), ("conv3", 128, 64, 3, "in3")):
        out += _conv(name + ".conv2d", cout, cin, k) + _norm(nrm, cout)
    for i in range(1, 6):
        for j in (1, 2):
            out += _conv(f"res{i}.conv{j}.conv2d", 128, 128, 3) + _norm(f"res{i}.in{j}", 128)
    out += _conv("deconv1.conv2d", 64, 128, 3) + _norm("in4", 64)
    out += _conv("deconv2.conv2d", 32, 64, 3) + _norm("in5", 32)
    out += _conv("deconv3.conv2d", 3, 32, 9)
    return out


def _nst_shapes():
    """transformer_net_nst.TransformerNet (transformer_net_nst.py:73-93); ConvTranspose2d weights are (in, out, k, k)."""
    out = []
    for name, cout, cin, k in (("down1", 32, 3, 9), ("down2", 64, 32, 3), ("down3", 128, 64, 3)):
        out += _conv(name + ".conv", cout, cin, k) + _norm(name + ".norm", cout)
    for i in range(1, 6):
        for j in (1, 2):
            out += _conv(f"res{i}.conv{j}", 128, 128, 3) + _norm(f"res{i}.norm{j}", 128)
    for name, cin, cout in (("up1", 128, 64), ("up2", 64, 32)):
        out += [(name + ".conv.weight", (cin, cout, 3, 3)), (name + ".conv.bias", (cout,))] + _norm(name + ".norm", cout)
    out += _conv("final", 3, 32, 9)
    return out


def _reconet_shapes():
    """model.ReCoNet(frn=False) (model.py:75-116): ConvNormLayer = Sequential(ConvLayer(Sequential(pad, conv)), IN[, ReLU])."""
    out = []

    def cnl(prefix, cout, cin, k):
        return _conv(prefix + ".layers.0.layers.1", cout, cin, k) + _norm(prefix + ".layers.1", cout)

    out += cnl("encoder.layers.0", 48, 3, 9) + cnl("encoder.layers.1", 96, 48, 3) + cnl("encoder.layers.2", 192, 96, 3)
    for i in range(3, 7):
        out += cnl(f"encoder.layers.{i}.branch.0", 192, 192, 3) + cnl(f"encoder.layers.{i}.branch.1", 192, 192, 3)
    out += cnl("decoder.layers.1", 96, 192, 3) + cnl("decoder.layers.3", 48, 96, 3)
    out += _conv("decoder.layers.4.layers.0.layers.1", 3, 48, 9)
    return out


PARAM_SHAPES = {"johnson": _johnson_shapes(), "nst": _nst_shapes(), "reconet": _reconet_shapes()}


class _Holder(nn.Module):
    """Parameter container; exists only so dotted state_dict keys match the reference."""

    def forward(self, *a, **k):  # pragma: no cover
        raise RuntimeError("parameter holder is not callable; call the top-level model")


class _B200Generator(nn.Module):
    ARCH = ""

    def __init__(self, precision: str = "f16"):
        super().__init__()
        self.precision = precision
        for key, shape in PARAM_SHAPES[self.ARCH]:
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
                   out: Optional[torch.Tensor] = None, fit: bool = False) -> torch.Tensor:
        """Fused uint8 path (mirrors lib.ReCoNetModel.run, lib.py:27-51, for every family).  ``fit=True`` returns the input's
        H x W even when it is not a multiple of 4 (the reference's resize-back of pipeline.py:1512-1516)."""
        eng = self._ensure_engine(frames_u8_nhwc.device)
        return eng.stylize_u8(frames_u8_nhwc, io_preset, blend, mask, composite_mode, out, fit)


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
