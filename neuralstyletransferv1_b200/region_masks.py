"""Region masks on the device (first part): Voronoi cells with feathered edges.

``voronoi_masks(H, W, count, seed, feather)`` equals ``region_blend.generate_region_masks(H, W, "voronoi", count, seed=seed,
feather=feather)`` (region_blend.py:925-981 -> 174-236): the seed points are drawn on the host with the same ``random.Random``
sequence as ``_generate_balanced_points`` (region_blend.py:239-304), the nearest-point labelling and the reflect-padded Gaussian
feather run as CUDA kernels (csrc/postproc.cu).
"""
from __future__ import annotations

import ctypes as C
import math
import random
from typing import List, Optional, Sequence, Tuple

import torch

from . import _lib


def balanced_points(W: int, H: int, count: int, rng: random.Random, jitter_factor: float = 0.3) -> List[Tuple[float, float]]:
    """One jittered point per cell of a cols x rows grid with cols * rows >= count, shuffled (region_blend.py:239-304)."""
    aspect = W / H
    cols = max(1, int(math.sqrt(count * aspect) + 0.5))
    rows = max(1, int(math.sqrt(count / aspect) + 0.5))
    while cols * rows < count:
        if cols / rows < aspect:
            cols += 1
        else:
            rows += 1
    cell_w, cell_h = W / cols, H / rows
    points: List[Tuple[float, float]] = []
    for row in range(rows):
        for col in range(cols):
            if len(points) >= count:
                break
            cx, cy = (col + 0.5) * cell_w, (row + 0.5) * cell_h
            jx = (rng.random() - 0.5) * cell_w * jitter_factor
            jy = (rng.random() - 0.5) * cell_h * jitter_factor
            points.append((max(0, min(W - 1, cx + jx)), max(0, min(H - 1, cy + jy))))
    while len(points) < count:
        points.append((rng.randint(0, W - 1), rng.randint(0, H - 1)))
    rng.shuffle(points)
    return points[:count]


def voronoi_masks(H: int, W: int, count: int, seed: Optional[int] = None, feather: int = 20,
                  region_sizes: Optional[Sequence[float]] = None, device="cuda:0") -> torch.Tensor:
    """fp32 [count, H, W] on ``device``; ``masks[i][None, None]`` is the reference's i-th ``[1,1,H,W]`` mask."""
    lib = _lib.load()
    dev = torch.device(device)
    if dev.type != "cuda":
        raise RuntimeError("voronoi_masks runs on a CUDA device; there is no CPU path")
    rng = random.Random() if seed is None else random.Random(seed)
    pts = balanced_points(W, H, count, rng, jitter_factor=0.3)
    flat = [float(v) for p in pts for v in p]
    masks = torch.empty(count, H, W, dtype=torch.float32, device=dev)
    scratch = torch.empty((H * W + 3) // 4 + count * H * W, dtype=torch.float32, device=dev)
    wts = None
    if region_sizes:
        wts = (C.c_float * count)(*[float(region_sizes[i]) if i < len(region_sizes) else 1.0 for i in range(count)])
    with torch.cuda.device(dev):
        rc = lib.nst_voronoi_masks_f32((C.c_float * (2 * count))(*flat), wts, count, int(feather), masks.data_ptr(), scratch.data_ptr(), H, W,
                                       torch.cuda.current_stream(dev).cuda_stream)
    _lib.check(rc, "nst_voronoi_masks_f32")
    return masks
