"""CPU oracle for the per-frame stylisation hot path.  TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` /
``--impl reference`` legs of ``bench.py`` may import this module; the product
path (``neuralstyletransferv1_b200``) never does.

It is a plain ``torch.nn.functional`` fp32 restatement of the reference
modules (which are themselves pure PyTorch; the arithmetic lives in the
un-vendored, un-pinned ``torch`` dependency, ``requirements.txt:1`` -- here
torch 2.11.0 CPU/oneDNN).  Each function cites the reference lines it follows.

Parity pin: ``tests/golden/*.npz`` were produced by importing the REAL reference
modules from ``/root/reference`` (``tests/golden/make_golden.py``); the oracle is
checked against them in ``tests/test_oracle.py``.  The reference has no tests or
golden vectors of its own for this path, so the pin is
"reference executed in the build container", not "reference's own fixtures".
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence

import torch
import torch.nn.functional as F

Tensor = torch.Tensor
SD = Dict[str, Tensor]

IMAGENET_MEAN = (0.485, 0.456, 0.406)   # pipeline.py:272
IMAGENET_STD = (0.229, 0.224, 0.225)    # pipeline.py:273
CAFFE_MEAN = (103.939, 116.779, 123.68)  # pipeline.py:1465


# --------------------------------------------------------------------------- #
# building blocks
# --------------------------------------------------------------------------- #
def _rconv(x: Tensor, w: Tensor, b: Tensor, stride: int = 1) -> Tensor:
    """ReflectionPad2d(k//2) + Conv2d  (transformer_net.py:44-54, model.py:5-15)."""
    p = w.shape[-1] // 2
    return F.conv2d(F.pad(x, (p, p, p, p), mode="reflect"), w, b, stride=stride)


def _inorm(x: Tensor, g: Tensor, b: Tensor) -> Tensor:
    """InstanceNorm2d(affine=True), instance statistics in eval, eps 1e-5, biased var."""
    return F.instance_norm(x, None, None, g, b, use_input_stats=True, eps=1e-5)


def _up2(x: Tensor) -> Tensor:
    """F.interpolate(mode='nearest', scale_factor=2)  (transformer_net.py:96)."""
    return F.interpolate(x, scale_factor=2, mode="nearest")


# --------------------------------------------------------------------------- #
# transformer_net.TransformerNet  (transformer_net.py:4-41)
# --------------------------------------------------------------------------- #
def johnson_forward(sd: SD, x: Tensor) -> Tensor:
    def conv(name, t, stride=1):
        return _rconv(t, sd[name + ".conv2d.weight"], sd[name + ".conv2d.bias"], stride)

    def norm(name, t):
        return _inorm(t, sd[name + ".weight"], sd[name + ".bias"])

    y = F.relu(norm("in1", conv("conv1", x)))                 # :30
    y = F.relu(norm("in2", conv("conv2", y, 2)))              # :31
    y = F.relu(norm("in3", conv("conv3", y, 2)))              # :32
    for i in range(1, 6):                                     # :33-37, block :71-76
        r = y
        o = F.relu(norm(f"res{i}.in1", conv(f"res{i}.conv1", y)))
        o = norm(f"res{i}.in2", conv(f"res{i}.conv2", o))
        y = o + r                                             # no ReLU after the add
    y = F.relu(norm("in4", conv("deconv1", _up2(y))))         # :38, :93-99
    y = F.relu(norm("in5", conv("deconv2", _up2(y))))         # :39
    return conv("deconv3", y)                                 # :40 raw output


# --------------------------------------------------------------------------- #
# transformer_net_nst.TransformerNet  (transformer_net_nst.py:62-127)
# --------------------------------------------------------------------------- #
def nst_forward(sd: SD, x: Tensor) -> Tensor:
    h, w = x.shape[-2:]
    y = F.pad(x, (40, 40, 40, 40), mode="reflect")            # :74,:100

    def block(name, t, stride, pad):
        t = F.conv2d(t, sd[name + ".conv.weight"], sd[name + ".conv.bias"], stride=stride, padding=pad)
        return F.relu(_inorm(t, sd[name + ".norm.weight"], sd[name + ".norm.bias"]))

    y = block("down1", y, 1, 4)                               # :77  (zero padding)
    y = block("down2", y, 2, 1)                               # :78
    y = block("down3", y, 2, 1)                               # :79
    for i in range(1, 6):                                     # :82-86, block :28-43
        r = y
        o = F.conv2d(y, sd[f"res{i}.conv1.weight"], sd[f"res{i}.conv1.bias"], padding=1)
        o = F.relu(_inorm(o, sd[f"res{i}.norm1.weight"], sd[f"res{i}.norm1.bias"]))
        o = F.conv2d(o, sd[f"res{i}.conv2.weight"], sd[f"res{i}.conv2.bias"], padding=1)
        o = _inorm(o, sd[f"res{i}.norm2.weight"], sd[f"res{i}.norm2.bias"])
        y = o + r
    for name in ("up1", "up2"):                               # :89-90, block :46-59
        y = F.conv_transpose2d(y, sd[name + ".conv.weight"], sd[name + ".conv.bias"],
                               stride=2, padding=1, output_padding=1)
        y = F.relu(_inorm(y, sd[name + ".norm.weight"], sd[name + ".norm.bias"]))
    y = F.conv2d(y, sd["final.weight"], sd["final.bias"], padding=4)   # :93
    oh, ow = y.shape[-2:]
    ch, cw = (oh - h) // 2, (ow - w) // 2                     # :122-125
    return y[:, :, ch:ch + h, cw:cw + w]


# --------------------------------------------------------------------------- #
# model.ReCoNet(frn=False)  (model.py:75-116)
# --------------------------------------------------------------------------- #
def reconet_forward(sd: SD, x: Tensor) -> Tensor:
    def cnl(prefix, t, stride=1, act=True):                   # ConvNormLayer model.py:18-40
        t = _rconv(t, sd[prefix + ".layers.0.layers.1.weight"], sd[prefix + ".layers.0.layers.1.bias"], stride)
        t = _inorm(t, sd[prefix + ".layers.1.weight"], sd[prefix + ".layers.1.bias"])
        return F.relu(t) if act else t

    y = cnl("encoder.layers.0", x)
    y = cnl("encoder.layers.1", y, 2)
    y = cnl("encoder.layers.2", y, 2)
    for i in range(3, 7):                                     # ResLayer model.py:43-60
        b = cnl(f"encoder.layers.{i}.branch.0", y)
        b = cnl(f"encoder.layers.{i}.branch.1", b, act=False)
        y = F.relu(y + b)                                     # ReLU AFTER the add
    y = cnl("decoder.layers.1", _up2(y))
    y = cnl("decoder.layers.3", _up2(y))
    p = "decoder.layers.4.layers.0.layers.1"                  # ConvTanhLayer model.py:63-72
    return torch.tanh(_rconv(y, sd[p + ".weight"], sd[p + ".bias"]))


FORWARD = {"johnson": johnson_forward, "nst": nst_forward, "reconet": reconet_forward}


def forward(arch: str, sd: SD, x: Tensor) -> Tensor:
    with torch.no_grad():
        return FORWARD[arch](sd, x.to(torch.float32))


# --------------------------------------------------------------------------- #
# io presets (pipeline.py:1447-1486), pack (pipeline.py:1943,2094; lib.py:46)
# --------------------------------------------------------------------------- #
def io_pre(x01: Tensor, preset: str) -> Tensor:
    """x01: [N,3,H,W] fp32 in 0..1 -> network input."""
    if preset == "tanh":
        return x01 * 2.0 - 1.0
    if preset == "imagenet_01":
        m = torch.tensor(IMAGENET_MEAN).view(1, 3, 1, 1); s = torch.tensor(IMAGENET_STD).view(1, 3, 1, 1)
        return (x01 - m) / s
    if preset == "imagenet_255":
        m = torch.tensor(IMAGENET_MEAN).view(1, 3, 1, 1); s = torch.tensor(IMAGENET_STD).view(1, 3, 1, 1)
        return (x01 * 255.0 - m * 255.0) / (s * 255.0)
    if preset == "caffe_bgr":
        c = torch.tensor(CAFFE_MEAN).view(1, 3, 1, 1)
        return (x01 * 255.0)[:, [2, 1, 0]] - c
    if preset == "raw_01":
        return x01
    return x01 * 255.0  # raw_255 and the default branch


def io_post(y: Tensor, preset: str) -> Tensor:
    """network output -> out01 (clamped), pipeline.py:1451-1486."""
    if preset == "tanh":
        o = (y + 1.0) * 0.5
    elif preset == "imagenet_01":
        m = torch.tensor(IMAGENET_MEAN).view(1, 3, 1, 1); s = torch.tensor(IMAGENET_STD).view(1, 3, 1, 1)
        o = y * s + m
    elif preset == "caffe_bgr":
        o = y[:, [2, 1, 0]] / 255.0
    elif preset == "raw_01":
        o = y
    else:  # imagenet_255, raw_255, default
        o = y / 255.0
    return o.clamp(0, 1)


def pack_u8(out01: Tensor) -> Tensor:
    """ToPILImage semantics: (x*255).astype(uint8) == truncation; NCHW -> NHWC."""
    return (out01.clamp(0, 1) * 255.0).to(torch.uint8).permute(0, 2, 3, 1).contiguous()


def reconet_run_u8(sd: SD, images_u8_nhwc: Tensor) -> Tensor:
    """lib.ReCoNetModel.run (lib.py:27-51) with utils.py:99-106."""
    x = images_u8_nhwc.permute(0, 3, 1, 2).to(torch.float32) / 255
    y = forward("reconet", sd, x * 2 - 1)
    y = (y + 1) / 2
    return torch.clamp(y * 255, 0, 255).to(torch.uint8).permute(0, 2, 3, 1).contiguous()


def stylize_u8(arch: str, sd: SD, frames_u8_nhwc: Tensor, preset: str,
               blend: float = 1.0, mask: Optional[Tensor] = None,
               composite_mode: str = "keep") -> Tensor:
    """Whole per-frame path of pipeline.py with uint8 in/out.

    ToTensor (/255) -> io_pre -> model -> io_post -> optional mask composite
    (pipeline.py:2040-2043), THEN the uniform blend when ``0 <= blend < 1`` ("still applies
    after region composite", pipeline.py:2088-2092) -> pack.
    ``mask``: [N,H,W] fp32 alpha in 0..1 or None.
    """
    x01 = frames_u8_nhwc.permute(0, 3, 1, 2).to(torch.float32) / 255.0
    out01 = io_post(forward(arch, sd, io_pre(x01, preset)), preset)
    return post_u8(out01, x01, blend, mask, composite_mode)


def post_u8(out01: Tensor, x01: Tensor, blend: float = 1.0, mask: Optional[Tensor] = None,
            composite_mode: str = "keep") -> Tensor:
    """Everything after io_post: mask composite (pipeline.py:2040-2043), uniform blend
    (pipeline.py:2088-2092), truncating pack (pipeline.py:2094).  fp32, one rounding per op."""
    if out01.shape[-2:] != x01.shape[-2:]:
        if mask is not None or 0.0 <= blend < 1.0:
            raise ValueError("blend / mask need H,W multiples of 4 (the reference resizes first, pipeline.py:1512-1516)")
        return pack_u8(out01)
    if mask is not None:
        a = mask.unsqueeze(1).to(torch.float32)
        if composite_mode == "keep":
            out01 = (a * out01 + (1.0 - a) * x01).clamp(0, 1)
        else:
            out01 = ((1.0 - a) * out01 + a * x01).clamp(0, 1)
    if 0.0 <= blend < 1.0:
        out01 = (blend * out01 + (1.0 - blend) * x01).clamp(0, 1)
    return pack_u8(out01)


# --------------------------------------------------------------------------- #
# region compositing (region_blend.py:1049-1108, 1589-1679)
# --------------------------------------------------------------------------- #
def composite_regions(styled: Sequence[Tensor], masks: Sequence[Tensor], assignments: Sequence[int],
                      original: Optional[Tensor] = None) -> Tensor:
    """styled: list of [3,H,W] fp32 0..1; masks: list of [1,1,H,W]; -> [3,H,W]."""
    C, H, W = styled[0].shape[-3:]
    result = torch.zeros(1, C, H, W)
    wsum = torch.zeros(1, 1, H, W)
    for m, a in zip(masks, assignments):
        src = original if a == -1 else styled[a]
        src = src.reshape(1, C, H, W)
        result += src * m.expand(1, C, H, W)
        wsum += m
    result = result / wsum.expand(1, C, H, W).clamp(min=1e-6)
    return result.squeeze(0).clamp(0, 1)


def composite_regions_advanced(styled: Sequence[Tensor], masks: Sequence[Tensor],
                               model_indices: Sequence[Sequence[int]],
                               model_weights: Sequence[Sequence[float]],
                               original: Optional[Tensor] = None) -> Tensor:
    """Single-scale form of region_blend.py:1589-1679 (per-region model blends)."""
    C, H, W = styled[0].shape[-3:]
    result = torch.zeros(1, C, H, W)
    wsum = torch.zeros(1, 1, H, W)
    for m, idxs, ws in zip(masks, model_indices, model_weights):
        rb = torch.zeros(1, C, H, W)
        for i, wt in zip(idxs, ws):
            src = original if i == -1 else styled[i]
            rb += wt * src.reshape(1, C, H, W)
        result += rb * m.expand(1, C, H, W)
        wsum += m
    result = result / wsum.expand(1, C, H, W).clamp(min=1e-6)
    return result.squeeze(0).clamp(0, 1)


# --------------------------------------------------------------------------- #
# metrics
# --------------------------------------------------------------------------- #
def u8_metrics(a_u8: Tensor, b_u8: Tensor):
    """(max|delta|, PSNR dB) between two uint8 tensors -- the north_star tolerance."""
    d = a_u8.to(torch.int16) - b_u8.to(torch.int16)
    mx = int(d.abs().max())
    mse = float((d.to(torch.float64) ** 2).mean())
    psnr = float("inf") if mse == 0 else 10.0 * torch.log10(torch.tensor(255.0 ** 2 / mse)).item()
    return mx, psnr


# --------------------------------------------------------------------------- #
# round 2: resize-back, multi-scale / batched composite, crop composite, flow EMA
# (test infrastructure: pinned against the reference's own functions by tests/golden/make_golden.py)
# --------------------------------------------------------------------------- #
def stylize_u8_fit(arch: str, sd: SD, frames_u8_nhwc: Tensor, preset: str, blend: float = 1.0,
                   mask: Optional[Tensor] = None, composite_mode: str = "keep") -> Tensor:
    """Per-frame path INCLUDING pipeline.py:1512-1516: when the generator returns a larger frame (H or W not a
    multiple of 4) out01 is resized back with ``F.interpolate(bilinear, align_corners=False)`` before the mask
    composite, the blend and the pack."""
    x01 = frames_u8_nhwc.permute(0, 3, 1, 2).to(torch.float32) / 255.0
    out01 = io_post(forward(arch, sd, io_pre(x01, preset)), preset)
    if out01.shape[-2:] != x01.shape[-2:]:
        out01 = F.interpolate(out01, size=tuple(x01.shape[-2:]), mode="bilinear", align_corners=False)
    return post_u8(out01, x01, blend, mask, composite_mode)


def composite_regions_multiscale(styled_by_scale, masks: Sequence[Tensor], configs, original: Optional[Tensor],
                                 H: int, W: int) -> Tensor:
    """region_blend.composite_regions_advanced (region_blend.py:1589-1679) with its per-scale sources:
    ``styled_by_scale`` = {scale: [frame per model, fp32 [3,h,w]]}, ``configs`` = [(model_indices, model_weights, scale)]."""
    C = 3
    if original is not None:
        original = original.reshape(1, C, *original.shape[-2:])
        if original.shape[-2:] != (H, W):
            original = F.interpolate(original, size=(H, W), mode="bilinear", align_corners=False)
    result = torch.zeros(1, C, H, W)
    wsum = torch.zeros(1, 1, H, W)
    for m, (idxs, ws, scale) in zip(masks, configs):
        m = m.reshape(1, 1, *m.shape[-2:])
        if m.shape[-2:] != (H, W):
            m = F.interpolate(m, size=(H, W), mode="bilinear", align_corners=False)
        if scale not in styled_by_scale:
            scale = min(styled_by_scale.keys(), key=lambda s: abs(s - scale))
        outs = styled_by_scale[scale]
        rb = torch.zeros(1, C, H, W)
        for i, wt in zip(idxs, ws):
            if i == -1:
                src = original
            else:
                src = outs[i].reshape(1, C, *outs[i].shape[-2:])
                if src.shape[-2:] != (H, W):
                    src = F.interpolate(src, size=(H, W), mode="bilinear", align_corners=False)
            rb += wt * src
        result += rb * m.expand(1, C, H, W)
        wsum += m
    result = result / wsum.expand(1, C, H, W).clamp(min=1e-6)
    return result.squeeze(0).clamp(0, 1)


def composite_from_crops(styled_crops, crops, original: Optional[Tensor], H: int, W: int) -> Tensor:
    """region_blend.composite_from_crops + place_crop (region_blend.py:2186-2294, 2080-2109), static blend weights.
    ``crops``: dicts with region_idx, padded_bbox, crop_mask [1,1,h,w], model_indices, model_weights."""
    C = 3
    canvas = torch.zeros(C, H, W)
    wsum = torch.zeros(1, H, W)
    for c in crops:
        x1, y1, x2, y2 = c["padded_bbox"]
        ch, cw = y2 - y1, x2 - x1
        rb = torch.zeros(C, ch, cw)
        for mi, wt in zip(c["model_indices"], c["model_weights"]):
            src = original[:, y1:y2, x1:x2].clone() if mi == -1 else styled_crops[mi][c["region_idx"]]
            if src.shape[1] != ch or src.shape[2] != cw:
                src = F.interpolate(src.unsqueeze(0), size=(ch, cw), mode="bilinear", align_corners=False).squeeze(0)
            rb += wt * src
        m = c["crop_mask"]
        if m.shape[2] != ch or m.shape[3] != cw:
            m = F.interpolate(m, size=(ch, cw), mode="bilinear", align_corners=False)
        canvas[:, y1:y2, x1:x2] += rb * m.squeeze(0).expand(C, -1, -1)
        wsum[:, y1:y2, x1:x2] += m.squeeze(0)
    gap = (wsum < 0.1).float()
    if gap.sum() > 0:
        gap3 = gap.expand(C, -1, -1)
        if original is not None:
            canvas = canvas + original * gap3
            wsum = wsum + gap
        else:
            for k in (5, 11, 21):
                cd = F.max_pool2d(canvas.unsqueeze(0), kernel_size=k, stride=1, padding=k // 2).squeeze(0)
                wd = F.max_pool2d(wsum.unsqueeze(0), kernel_size=k, stride=1, padding=k // 2).squeeze(0)
                canvas = canvas * (1 - gap3) + cd * gap3
                wsum = wsum * (1 - gap) + wd * gap
                gap = (wsum < 0.1).float()
                gap3 = gap.expand(C, -1, -1)
                if gap.sum() == 0:
                    break
    return (canvas / wsum.expand(C, -1, -1).clamp(min=1e-6)).clamp(0, 1)


def flow_ema(curr01_hwc, prev01_hwc, flow_hw2, alpha: float):
    """pipeline.py:425-439 `_warp_with_flow` (cv2.remap, the reference's own OpenCV call) + :1927-1929 EMA lerp; numpy fp32 HWC."""
    import cv2
    import numpy as np
    H, W = flow_hw2.shape[:2]
    gx, gy = np.meshgrid(np.arange(W, dtype=np.float32), np.arange(H, dtype=np.float32))
    map_x = (gx + flow_hw2[..., 0]).astype(np.float32)
    map_y = (gy + flow_hw2[..., 1]).astype(np.float32)
    warped = cv2.remap(prev01_hwc, map_x, map_y, interpolation=cv2.INTER_LINEAR, borderMode=cv2.BORDER_REPLICATE)
    a = float(max(0.0, min(1.0, alpha)))
    return (a * curr01_hwc + (1.0 - a) * warped).clip(0.0, 1.0).astype("float32")


# --------------------------------------------------------------------------- #
# region masks: patterns, rotation, organic warp
# --------------------------------------------------------------------------- #
def gaussian_blur_mask(mask: Tensor, sigma: float) -> Tensor:
    """region_blend.py:69-94: 2-D outer-product Gaussian, reflect padding."""
    if sigma <= 0:
        return mask
    ks = int(6 * sigma + 1)
    ks += ks % 2 == 0
    ks = max(3, ks)
    x = torch.arange(ks, dtype=torch.float32) - ks // 2
    k1 = torch.exp(-x ** 2 / (2 * sigma ** 2))
    k1 = k1 / k1.sum()
    k2 = (k1.view(-1, 1) @ k1.view(1, -1)).view(1, 1, ks, ks)
    return F.conv2d(F.pad(mask, (ks // 2,) * 4, mode="reflect"), k2)


def feather_mask(mask: Tensor, feather_px: int) -> Tensor:
    """region_blend.py:97-102."""
    return mask if feather_px <= 0 else gaussian_blur_mask(mask, feather_px / 3.0)


def pattern_masks(H: int, W: int, mode: str, count: int, seed: int, feather: int) -> Tensor:
    """generate_region_masks for grid / diagonal / fractal / radial / waves / spiral / concentric (region_blend.py:109-172, 307-521,
    925-981) restated with the reference's torch expressions and its `random.Random` draw order (own restatement: nothing is shared with
    the product's host helper; both are pinned by the fixtures of tests/golden/make_golden_masks.py).  -> [count,H,W]."""
    import math
    import random
    rng = random.Random(seed)
    if mode == "random":
        mode = rng.choice(["grid", "diagonal", "voronoi", "fractal", "radial", "waves", "spiral", "concentric"])
    yy = torch.arange(H, dtype=torch.float32).view(H, 1).expand(H, W)
    xx = torch.arange(W, dtype=torch.float32).view(1, W).expand(H, W)

    def bands(f):
        return [((f >= i / count) & (f < (i + 1) / count)).float() for i in range(count)]

    def boxes(rects):
        out = []
        for (y1, y2, x1, x2) in rects:
            m = torch.zeros(H, W)
            m[y1:y2, x1:x2] = 1.0
            out.append(m)
        return out

    if mode == "grid":                                                    # :109-135
        gs = int(math.ceil(math.sqrt(count)))
        ch, cw = H / gs, W / gs
        hard = boxes([(int((i // gs) * ch), min(int((i // gs + 1) * ch), H), int((i % gs) * cw), min(int((i % gs + 1) * cw), W)) for i in range(count)])
    elif mode == "fractal":                                               # :307-364
        regions = []

        def subdivide(y1, y2, x1, x2, depth):
            if len(regions) >= count:
                return
            if depth >= 4 or (y2 - y1) < 20 or (x2 - x1) < 20:
                regions.append((y1, y2, x1, x2))
                return
            if rng.random() > 0.4 and depth > 0:
                regions.append((y1, y2, x1, x2))
                return
            my = (y1 + y2) // 2 + rng.randint(-10, 10)
            mx = (x1 + x2) // 2 + rng.randint(-10, 10)
            my = max(y1 + 10, min(y2 - 10, my))
            mx = max(x1 + 10, min(x2 - 10, mx))
            quads = [(y1, my, x1, mx), (y1, my, mx, x2), (my, y2, x1, mx), (my, y2, mx, x2)]
            rng.shuffle(quads)
            for q in quads:
                if len(regions) >= count:
                    break
                subdivide(*q, depth + 1)

        subdivide(0, H, 0, W, 0)
        hard = boxes(regions[:count])
    elif mode == "diagonal":                                              # :138-171
        f = (xx + yy) if rng.random() > 0.5 else ((W - 1 - xx) + yy)
        hard = bands(f / f.max())
    elif mode == "radial":                                                # :367-401
        cx = W // 2 + rng.randint(-W // 4, W // 4)
        cy = H // 2 + rng.randint(-H // 4, H // 4)
        rotation = rng.random() * 2 * math.pi
        ang = (torch.atan2(yy - cy, xx - cx) + math.pi + rotation) % (2 * math.pi)
        wedge = 2 * math.pi / count
        hard = [((ang >= i * wedge) & (ang < (i + 1) * wedge)).float() for i in range(count)]
    elif mode == "waves":                                                 # :404-447
        frequency = rng.uniform(1.5, 4.0)
        amplitude = rng.uniform(0.05, 0.15)
        direction = rng.choice(["horizontal", "vertical", "diagonal"])
        phase = rng.random() * 2 * math.pi
        yn, xn = yy / H, xx / W
        if direction == "horizontal":
            pos = yn + torch.sin(xn * frequency * 2 * math.pi + phase) * amplitude
        elif direction == "vertical":
            pos = xn + torch.sin(yn * frequency * 2 * math.pi + phase) * amplitude
        else:
            d = (xn + yn) / 2
            pos = d + torch.sin(d * frequency * 2 * math.pi + phase) * amplitude
        hard = bands((pos - pos.min()) / (pos.max() - pos.min() + 1e-6))
    elif mode == "spiral":                                                # :450-485
        tightness = rng.uniform(2.0, 5.0)
        rotation = rng.random() * 2 * math.pi
        yc, xc = yy - H // 2, xx - W // 2
        r = torch.sqrt(xc ** 2 + yc ** 2)
        theta = torch.atan2(yc, xc) + math.pi + rotation
        hard = bands(((theta + r / max(H, W) * tightness * 2 * math.pi) % (2 * math.pi)) / (2 * math.pi))
    elif mode == "concentric":                                            # :488-521
        cx = W // 2 + rng.randint(-W // 6, W // 6)
        cy = H // 2 + rng.randint(-H // 6, H // 6)
        r = torch.sqrt((xx - cx) ** 2 + (yy - cy) ** 2)
        hard = bands(r / r.max())
    else:
        raise ValueError(mode)
    masks = [feather_mask(m.view(1, 1, H, W), feather) for m in hard]
    while len(masks) < count:
        masks.append(masks[-1].clone())
    return torch.cat([m.view(1, H, W) for m in masks[:count]])


def rotate_all_masks(masks: Tensor, angle: float) -> Tensor:
    """region_blend.py:25-66 with the reference's own OpenCV calls.  masks [R,H,W] -> [R,H,W]."""
    import cv2
    if angle == 0:
        return masks
    R, H, W = masks.shape
    M = cv2.getRotationMatrix2D((W / 2, H / 2), angle, 1.0)
    rot = [torch.from_numpy(cv2.warpAffine(m.numpy(), M, (W, H), flags=cv2.INTER_LINEAR, borderMode=cv2.BORDER_REPLICATE)) for m in masks]
    s = torch.zeros_like(rot[0])
    for m in rot:
        s += m
    s = s.clamp(min=1e-6)
    return torch.stack([m / s for m in rot])


def _simplex_noise_2d(H, W, frequency, octaves, seed, time_offset):
    """region_blend.py:604-652."""
    import numpy as np
    rng = np.random.default_rng(seed)
    xx, yy = np.meshgrid(np.linspace(0, frequency, W), np.linspace(0, frequency, H))
    result = np.zeros((H, W), dtype=np.float32)
    amplitude, total, fm = 1.0, 0.0, 1.0
    for octave in range(octaves):
        ox = time_offset * (0.5 + 0.3 * octave) + rng.random() * 1000
        oy = time_offset * (0.3 + 0.2 * octave) + rng.random() * 1000
        noise = np.sin(xx * fm + ox) * np.cos(yy * fm + oy)
        noise += np.sin((xx + yy) * fm * 0.7 + ox * 0.8) * 0.5
        noise += np.cos((xx - yy) * fm * 0.5 + oy * 0.6) * 0.3
        result += noise * amplitude
        total += amplitude
        amplitude *= 0.5
        fm *= 2.0
    result = result / total
    return (result - result.min()) / (result.max() - result.min() + 1e-6)


def warp_all_masks_organic(masks: Tensor, mode: str, amplitude: float, frequency: float, speed: float, seed: int, frame_idx: int) -> Tensor:
    """region_blend.py:655-810 (flow field, cv2.remap with BORDER_REFLECT, dilation gap fill, renormalisation).  masks [R,H,W]."""
    import cv2
    import numpy as np
    R, H, W = masks.shape
    t = frame_idx * speed * 0.02
    warped = []
    for i in range(R):
        s = seed + i * 100
        if mode == "tentacle":
            fx = _simplex_noise_2d(H, W, frequency * 2, 2, s, t) * 2 - 1
            fy = _simplex_noise_2d(H, W, frequency * 2, 2, s + 1000, t * 1.3) * 2 - 1
            fy += np.sin(np.linspace(0, 1, H)[:, None] * np.pi * 3 + t) * 0.5
        elif mode == "wave":
            fx = np.sin(np.linspace(0, np.pi * frequency, H)[:, None] + t * 2) * np.ones((H, W))
            fy = np.cos(np.linspace(0, np.pi * frequency, W)[None, :] + t * 1.5) * np.ones((H, W))
        elif mode == "pulse":
            yc, xc = np.arange(H)[:, None] - H // 2, np.arange(W)[None, :] - W // 2
            r = np.sqrt(xc ** 2 + yc ** 2) + 1e-6
            theta = np.arctan2(yc, xc)
            pulse = np.sin(r * 0.05 - t * 3) * 0.5 + 0.5
            fx, fy = np.cos(theta) * pulse, np.sin(theta) * pulse
        else:
            fx = _simplex_noise_2d(H, W, frequency, 2, s, t) * 2 - 1
            fy = _simplex_noise_2d(H, W, frequency, 2, s + 1000, t * 1.3) * 2 - 1
        md = max(H, W) * amplitude
        fx, fy = fx * md, fy * md
        yc = np.arange(H, dtype=np.float32)[:, None].repeat(W, axis=1)
        xc = np.arange(W, dtype=np.float32)[None, :].repeat(H, axis=0)
        w = cv2.remap(masks[i].numpy(), (xc + fx).astype(np.float32), (yc + fy).astype(np.float32), cv2.INTER_LINEAR, borderMode=cv2.BORDER_REFLECT)
        warped.append(torch.from_numpy(w).view(1, 1, H, W))
    msum = torch.zeros_like(warped[0])
    for m in warped:
        msum += m
    gap = (msum < 0.1).float()
    if gap.sum() > 0:
        filled = list(warped)
        for k in (5, 11, 21, 41):
            filled = [m * (1 - gap) + F.max_pool2d(m, kernel_size=k, stride=1, padding=k // 2) * gap for m in filled]
            msum = torch.zeros_like(filled[0])
            for m in filled:
                msum += m
            gap = (msum < 0.1).float()
            if gap.sum() == 0:
                break
        warped = filled
    msum = msum.clamp(min=1e-6)
    return torch.cat([(m / msum).view(1, H, W) for m in warped])


# --------------------------------------------------------------------------- #
# temporal: LAB EMA, motion-adaptive blend
# --------------------------------------------------------------------------- #
def lab_ema(frames_u8, smooth_lightness: bool, smooth_chroma: bool, smooth_alpha: float, chroma_alpha: float):
    """pipeline.py:1938-1975 over a clip; numpy uint8 [N,H,W,3] -> same (Pillow's LAB transform, like the reference)."""
    import numpy as np
    from PIL import Image
    pL = pa = pb = None
    outs = []
    for f in frames_u8:
        lab_np = np.array(Image.fromarray(f, mode="RGB").convert("LAB"), dtype=np.uint8).astype(np.float32)
        L, aC, bC = lab_np[..., 0], lab_np[..., 1], lab_np[..., 2]
        if smooth_lightness:
            if pL is None:
                pL = L.copy()
            pL = smooth_alpha * L + (1.0 - smooth_alpha) * pL
            lab_np[..., 0] = np.clip(pL, 0, 255)
        if smooth_chroma:
            if pa is None:
                pa, pb = aC.copy(), bC.copy()
            pa = chroma_alpha * aC + (1.0 - chroma_alpha) * pa
            pb = chroma_alpha * bC + (1.0 - chroma_alpha) * pb
            lab_np[..., 1] = np.clip(pa, 0, 255)
            lab_np[..., 2] = np.clip(pb, 0, 255)
        outs.append(np.array(Image.fromarray(lab_np.astype(np.uint8), mode="LAB").convert("RGB")) if (smooth_lightness or smooth_chroma) else f)
    return np.stack(outs)


def motion_blend(styled_u8, original_u8, flow, blend: float, min_alpha: float = 0.40, motion_norm: float = 8.0, sigma: float = 3.0):
    """pipeline.py:2072-2086 + the ToPILImage pack of :2094; numpy uint8 [H,W,3], flow fp32 [H,W,2] -> uint8 [H,W,3]."""
    import cv2
    import numpy as np
    mag = np.sqrt(flow[..., 0] ** 2 + flow[..., 1] ** 2).astype(np.float32)
    m = cv2.GaussianBlur(np.clip(mag / motion_norm, 0.0, 1.0), (0, 0), sigma)
    alpha = float(blend) - (float(blend) - float(min_alpha)) * m
    a3 = np.repeat(alpha[..., None], 3, axis=2)
    s = torch.from_numpy(styled_u8).float().div(255).numpy()                     # ToTensor
    o = torch.from_numpy(original_u8).float().div(255).numpy()
    fused = (a3 * s + (1.0 - a3) * o).clip(0.0, 1.0).astype(np.float32)
    return torch.from_numpy(fused).mul(255).byte().numpy()                        # ToPILImage


def post_chain(out01_seq, orig_u8_seq, flows, masks=None, flow_ema_on=False, flow_alpha=0.85, smooth_lightness=False, smooth_chroma=False,
               smooth_alpha=0.7, chroma_alpha=0.85, blend=1.0, motion_blend_on=False, composite_mode="keep"):
    """The tail of the reference's per-frame loop over a clip (pipeline.py:1884-2094), composed from the restatements above:
    flow EMA -> pack -> LAB EMA -> mask composite / motion-adaptive or uniform blend -> pack.  out01_seq: torch fp32 [N,3,H,W] (clamped);
    orig_u8_seq: numpy uint8 [N,H,W,3]; flows: list of numpy [H,W,2] or None; masks: numpy [N,H,W] or None.  -> numpy uint8 [N,H,W,3]."""
    import numpy as np
    N = len(out01_seq)
    prev = None
    fused, last_flow = [], []
    for i in range(N):
        cur = out01_seq[i].permute(1, 2, 0).numpy().astype(np.float32)
        lf = None
        if flow_ema_on and flows[i] is not None and prev is not None:
            cur = flow_ema(cur, prev, flows[i], flow_alpha)
            lf = flows[i]
        prev = cur
        fused.append(pack_u8(torch.from_numpy(cur).permute(2, 0, 1)[None])[0].numpy())
        last_flow.append(lf)
    smooth = lab_ema(np.stack(fused), smooth_lightness, smooth_chroma, smooth_alpha, chroma_alpha)
    outs = []
    for i in range(N):
        if motion_blend_on and last_flow[i] is not None and masks is None:
            outs.append(motion_blend(smooth[i], orig_u8_seq[i], last_flow[i], blend))
            continue
        s01 = torch.from_numpy(smooth[i]).permute(2, 0, 1).float().div(255)[None]
        x01 = torch.from_numpy(orig_u8_seq[i]).permute(2, 0, 1).float().div(255)[None]
        m = torch.from_numpy(masks[i])[None] if masks is not None else None
        outs.append(post_u8(s01, x01, blend, m, composite_mode)[0].numpy())
    return np.stack(outs)


# --------------------------------------------------------------------------- #
# crop-based region styling
# --------------------------------------------------------------------------- #
def prepare_region_crops(masks: Tensor, configs, H: int, W: int, padding: int = 64):
    """region_blend.compute_mask_bbox / pad_bbox / prepare_region_crops (region_blend.py:1969-2059).  masks [R,H,W]; configs: dicts with
    model_indices, model_weights, scale.  -> the crop dicts composite_from_crops above takes (+ bbox, scale)."""
    import numpy as np
    crops = []
    for idx, cfg in enumerate(configs):
        m2 = masks[idx].numpy()
        rows, cols = np.any(m2 > 0.01, axis=1), np.any(m2 > 0.01, axis=0)
        if not np.any(rows) or not np.any(cols):
            bbox = (0, 0, W, H)
        else:
            y1, y2 = np.where(rows)[0][[0, -1]]
            x1, x2 = np.where(cols)[0][[0, -1]]
            bbox = (int(x1), int(y1), int(x2) + 1, int(y2) + 1)
        pb = (max(0, bbox[0] - padding), max(0, bbox[1] - padding), min(W, bbox[2] + padding), min(H, bbox[3] + padding))
        crops.append(dict(region_idx=idx, bbox=bbox, padded_bbox=pb, crop_mask=masks[idx][None, None, pb[1]:pb[3], pb[0]:pb[2]].clone(),
                          model_indices=list(cfg["model_indices"]), model_weights=list(cfg["model_weights"]), scale=float(cfg["scale"])))
    return crops


def region_crop_style(frame_u8: Tensor, masks: Tensor, configs, archs, sds, presets, padding: int = 64, use_original: bool = True) -> Tensor:
    """The `--region_optimize` branch of the loop (pipeline.py:1293-1400) for Transformer / ReCoNet models: style the padded box of every
    region with the models it uses, composite.  frame_u8 [H,W,3]; -> out01 fp32 [3,H,W]."""
    H, W, _ = frame_u8.shape
    x01 = frame_u8.permute(2, 0, 1).float() / 255.0
    crops = prepare_region_crops(masks, configs, H, W, padding)
    mean = torch.tensor(IMAGENET_MEAN).view(1, 3, 1, 1)
    std = torch.tensor(IMAGENET_STD).view(1, 3, 1, 1)
    styled = {}
    for mi in sorted({m for c in crops for m in c["model_indices"] if m >= 0}):
        styled[mi] = {}
        for c in crops:
            if mi not in c["model_indices"]:
                continue
            x1, y1, x2, y2 = c["padded_bbox"]
            ch, cw = y2 - y1, x2 - x1
            crop = x01[:, y1:y2, x1:x2].clone().unsqueeze(0)
            if c["scale"] < 1.0:
                crop = F.interpolate(crop, size=(max(1, int(ch * c["scale"])), max(1, int(cw * c["scale"]))), mode="bilinear", align_corners=False)
            if presets[mi] == "imagenet_255":
                out = (forward(archs[mi], sds[mi], (crop * 255.0 - mean * 255.0) / (std * 255.0)) / 255.0).clamp(0, 1)
            elif presets[mi] == "imagenet_01":
                out = (forward(archs[mi], sds[mi], (crop - mean) / std) * std + mean).clamp(0, 1)
            else:
                out = (forward(archs[mi], sds[mi], crop * 255.0) / 255.0).clamp(0, 1)
            if c["scale"] < 1.0:
                out = F.interpolate(out, size=(ch, cw), mode="bilinear", align_corners=False)
            styled[mi][c["region_idx"]] = out.squeeze(0)
    return composite_from_crops(styled, crops, x01 if use_original else None, H, W)
