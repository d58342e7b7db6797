"""Seeded synthetic checkpoints for the three generator families.

The reference mount ships no ``.pth`` files (``.MISSING_LARGE_BLOBS:1-12``), so
parity tests, golden fixtures and ``bench.py`` all use state dicts generated
here.  Key names / shapes follow the reference modules exactly
(``transformer_net.py:4-99``, ``transformer_net_nst.py:62-93``,
``model.py:75-116``) so the same dict loads into the reference modules and into
the drop-in modules of this package.

Recipe: PyTorch default conv init, InstanceNorm
gamma ~ 1 + 0.3 N(0,1), beta ~ 0.3 N(0,1), and a *calibrated* last conv so the
fp32 output spans and overflows the displayable range like a trained net
(otherwise the uint8 tolerance is vacuous).
"""
from __future__ import annotations

import hashlib
import math
from collections import OrderedDict

import torch

ARCHS = ("johnson", "nst", "reconet")


def _conv(g, cout, cin, k):
    bound = 1.0 / math.sqrt(cin * k * k)
    w = (torch.rand(cout, cin, k, k, generator=g) * 2 - 1) * bound
    b = (torch.rand(cout, generator=g) * 2 - 1) * bound
    return w, b


def _norm(g, c):
    gamma = 1.0 + 0.3 * torch.randn(c, generator=g)
    beta = 0.3 * torch.randn(c, generator=g)
    return gamma, beta


def _last(g, cout, cin, k, sigma_out, bias):
    # input of the last conv is relu(IN(.)) with E[a^2] ~ 0.5..0.6 per channel
    fan = cin * k * k
    std = sigma_out / math.sqrt(0.55 * fan)
    w = torch.randn(cout, cin, k, k, generator=g) * std
    b = torch.full((cout,), float(bias)) + 0.05 * sigma_out * torch.randn(cout, generator=g)
    return w, b


def johnson_state_dict(seed: int = 0) -> "OrderedDict[str, torch.Tensor]":
    """Keys of ``transformer_net.TransformerNet`` (transformer_net.py:8-25)."""
    g = torch.Generator().manual_seed(1000 + seed)
    sd = OrderedDict()

    def conv(name, cout, cin, k):
        sd[name + ".conv2d.weight"], sd[name + ".conv2d.bias"] = _conv(g, cout, cin, k)

    def norm(name, c):
        sd[name + ".weight"], sd[name + ".bias"] = _norm(g, c)

    conv("conv1", 32, 3, 9); norm("in1", 32)
    conv("conv2", 64, 32, 3); norm("in2", 64)
    conv("conv3", 128, 64, 3); norm("in3", 128)
    for i in range(1, 6):
        conv(f"res{i}.conv1", 128, 128, 3); norm(f"res{i}.in1", 128)
        conv(f"res{i}.conv2", 128, 128, 3); norm(f"res{i}.in2", 128)
    conv("deconv1", 64, 128, 3); norm("in4", 64)
    conv("deconv2", 32, 64, 3); norm("in5", 32)
    w, b = _last(g, 3, 32, 9, sigma_out=60.0, bias=127.5)
    sd["deconv3.conv2d.weight"], sd["deconv3.conv2d.bias"] = w, b
    return sd


def nst_state_dict(seed: int = 0) -> "OrderedDict[str, torch.Tensor]":
    """Keys of ``transformer_net_nst.TransformerNet`` (transformer_net_nst.py:73-93).

    ConvTranspose2d weights are (in, out, 3, 3) (transformer_net_nst.py:51-54).
    The pipeline forces io_preset raw_01 for this family (pipeline.py:611-614),
    so the last conv is calibrated to mean 0.5, sigma 0.25.
    """
    g = torch.Generator().manual_seed(2000 + seed)
    sd = OrderedDict()

    def block(name, cout, cin, k):
        sd[name + ".conv.weight"], sd[name + ".conv.bias"] = _conv(g, cout, cin, k)
        sd[name + ".norm.weight"], sd[name + ".norm.bias"] = _norm(g, cout)

    block("down1", 32, 3, 9)
    block("down2", 64, 32, 3)
    block("down3", 128, 64, 3)
    for i in range(1, 6):
        for j in (1, 2):
            sd[f"res{i}.conv{j}.weight"], sd[f"res{i}.conv{j}.bias"] = _conv(g, 128, 128, 3)
            sd[f"res{i}.norm{j}.weight"], sd[f"res{i}.norm{j}.bias"] = _norm(g, 128)
    for name, cin, cout in (("up1", 128, 64), ("up2", 64, 32)):
        bound = 1.0 / math.sqrt(cin * 9)
        # torch ConvTranspose2d layout: (in, out, kh, kw)
        sd[name + ".conv.weight"] = (torch.rand(cin, cout, 3, 3, generator=g) * 2 - 1) * bound
        sd[name + ".conv.bias"] = (torch.rand(cout, generator=g) * 2 - 1) * bound
        sd[name + ".norm.weight"], sd[name + ".norm.bias"] = _norm(g, cout)
    w, b = _last(g, 3, 32, 9, sigma_out=0.25, bias=0.5)
    sd["final.weight"], sd["final.bias"] = w, b
    return sd


def reconet_state_dict(seed: int = 0) -> "OrderedDict[str, torch.Tensor]":
    """Keys of ``model.ReCoNet(frn=False)`` (model.py:75-116); pre-tanh sigma ~ 1."""
    g = torch.Generator().manual_seed(3000 + seed)
    sd = OrderedDict()

    def cnl(prefix, cout, cin, k):
        # ConvNormLayer.layers = Sequential(ConvLayer(.layers = Sequential(pad, conv)), IN[, ReLU])
        w, b = _conv(g, cout, cin, k)
        sd[prefix + ".layers.0.layers.1.weight"], sd[prefix + ".layers.0.layers.1.bias"] = w, b
        sd[prefix + ".layers.1.weight"], sd[prefix + ".layers.1.bias"] = _norm(g, cout)

    cnl("encoder.layers.0", 48, 3, 9)
    cnl("encoder.layers.1", 96, 48, 3)
    cnl("encoder.layers.2", 192, 96, 3)
    for i in range(3, 7):
        cnl(f"encoder.layers.{i}.branch.0", 192, 192, 3)
        cnl(f"encoder.layers.{i}.branch.1", 192, 192, 3)
    cnl("decoder.layers.1", 96, 192, 3)
    cnl("decoder.layers.3", 48, 96, 3)
    w, b = _last(g, 3, 48, 9, sigma_out=1.0, bias=0.0)
    sd["decoder.layers.4.layers.0.layers.1.weight"] = w
    sd["decoder.layers.4.layers.0.layers.1.bias"] = b
    return sd


def make_state_dict(arch: str, seed: int = 0):
    if arch == "johnson":
        return johnson_state_dict(seed)
    if arch == "nst":
        return nst_state_dict(seed)
    if arch == "reconet":
        return reconet_state_dict(seed)
    raise ValueError(f"unknown arch {arch!r}; expected one of {ARCHS}")


def state_dict_digest(sd) -> str:
    """sha256 over the fp32 bytes of every tensor, in key order (RNG drift detector)."""
    h = hashlib.sha256()
    for k, v in sd.items():
        h.update(k.encode())
        h.update(v.detach().to(torch.float32).contiguous().numpy().tobytes())
    return h.hexdigest()


def synthetic_frames_u8(n: int, h: int, w: int, seed: int = 1234, device="cpu") -> torch.Tensor:
    """Image-like uint8 NHWC frames: low-pass noise, every 10th frame pure noise."""
    g = torch.Generator(device="cpu").manual_seed(seed)
    out = torch.empty(n, h, w, 3, dtype=torch.uint8)
    for i in range(n):
        u = torch.rand(1, 3, h // 8 + 2, w // 8 + 2, generator=g)
        if i % 10 == 9:
            f = torch.rand(1, 3, h, w, generator=g)
        else:
            f = torch.nn.functional.interpolate(u, size=(h, w), mode="bilinear", align_corners=False)
            f = f + 0.08 * torch.rand(1, 3, h, w, generator=g)
            f = (f - f.amin()) / (f.amax() - f.amin() + 1e-6)
        out[i] = (f[0].permute(1, 2, 0) * 255.0).to(torch.uint8)
    return out.to(device)
