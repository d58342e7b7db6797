"""Region masks on the device: every pattern of the reference's generator, and the per-frame mask
animation (rotation, organic warp with gap fill, renormalisation).

``generate_region_masks(H, W, mode, count, seed, feather, region_sizes)`` equals ``region_blend.generate_region_masks``
(region_blend.py:925-981): the handful of random scalars of a pattern (seed points, centre, frequency, ...) are drawn on the host with the
same ``random.Random`` sequence as the reference, the per-pixel labelling and the reflect-padded Gaussian feather run as CUDA kernels
(csrc/postproc.cu, csrc/masks.cu).  ``rotate_all_masks`` / ``warp_all_masks_organic`` mirror region_blend.py:49-66 / 737-810
(cv2.warpAffine / cv2.remap arithmetic in 1/32-pixel fixed point, csrc/post_math.cuh).  Masks are fp32 ``[count, H, W]`` CUDA tensors;
``masks[i][None, None]`` is the reference's i-th ``[1, 1, H, W]`` mask.
"""
from __future__ import annotations

import ctypes as C
import math
import random
from dataclasses import dataclass, field
from typing import List, Optional, Sequence, Tuple

import numpy as np
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
    return _voronoi_from_rng(H, W, count, rng, feather, region_sizes, device)


def _voronoi_from_rng(H: int, W: int, count: int, rng: random.Random, feather: int, region_sizes, device) -> torch.Tensor:
    lib = _lib.load()
    dev = torch.device(device)
    if dev.type != "cuda":
        raise RuntimeError("voronoi_masks runs on a CUDA device; there is no CPU path")
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


# ------------------------------------------------------------------------------------------------------------------------------
# the other patterns (region_blend.py:109-172, 307-521)
# ------------------------------------------------------------------------------------------------------------------------------
AVAILABLE_MODES = ["grid", "diagonal", "voronoi", "fractal", "radial", "waves", "spiral", "concentric", "random"]   # region_blend.py:879
PAT_RECTS, PAT_DIAGONAL, PAT_RADIAL, PAT_WAVES, PAT_SPIRAL, PAT_CONCENTRIC = range(6)
NORM_NONE, NORM_DIV_MAX, NORM_MINMAX = range(3)
ORGANIC_MODE = {"blob": 0, "tentacle": 1, "wave": 2, "pulse": 3}


@dataclass
class PatternSpec:
    """What the device needs to label one pattern: the scalars the reference draws from its ``random.Random``."""
    mode: int
    norm: int = NORM_NONE
    p: List[float] = field(default_factory=lambda: [0.0] * 6)
    lo: List[float] = field(default_factory=list)       # band i = lo[i] <= field < hi[i]
    hi: List[float] = field(default_factory=list)
    rects: List[Tuple[int, int, int, int]] = field(default_factory=list)    # (y1, y2, x1, x2)

    @property
    def regions(self) -> int:
        return len(self.rects) if self.mode == PAT_RECTS else len(self.lo)


def _bands(count: int):
    return [i / count for i in range(count)], [(i + 1) / count for i in range(count)]


def pattern_spec(H: int, W: int, mode: str, count: int, rng: random.Random) -> PatternSpec:
    """The host half of grid / diagonal / fractal / radial / waves / spiral / concentric: same draws, same order, as the reference."""
    if mode == "grid":                                                   # region_blend.py:109-135
        gs = int(math.ceil(math.sqrt(count)))
        ch, cw = H / gs, W / gs
        rects = []
        for i in range(count):
            row, col = i // gs, i % gs
            rects.append((int(row * ch), min(int((row + 1) * ch), H), int(col * cw), min(int((col + 1) * cw), W)))
        return PatternSpec(PAT_RECTS, rects=rects)
    if mode == "fractal":                                                # region_blend.py:307-364
        regions: List[Tuple[int, int, int, int]] = []
        max_depth = 4

        def subdivide(y1, y2, x1, x2, depth):
            if len(regions) >= count:
                return
            if depth >= max_depth or (y2 - y1) < 20 or (x2 - x1) < 20:
                regions.append((y1, y2, x1, x2))
                return
            if rng.random() > 0.4 and depth > 0:
                regions.append((y1, y2, x1, x2))
                return
            mid_y = (y1 + y2) // 2 + rng.randint(-10, 10)
            mid_x = (x1 + x2) // 2 + rng.randint(-10, 10)
            mid_y = max(y1 + 10, min(y2 - 10, mid_y))
            mid_x = max(x1 + 10, min(x2 - 10, mid_x))
            quads = [(y1, mid_y, x1, mid_x), (y1, mid_y, mid_x, x2), (mid_y, y2, x1, mid_x), (mid_y, y2, mid_x, x2)]
            rng.shuffle(quads)
            for q in quads:
                if len(regions) >= count:
                    break
                subdivide(q[0], q[1], q[2], q[3], depth + 1)

        subdivide(0, H, 0, W, 0)
        return PatternSpec(PAT_RECTS, rects=regions[:count])
    if mode == "diagonal":                                               # :138-171
        direction = 1.0 if rng.random() > 0.5 else 0.0
        lo, hi = _bands(count)
        return PatternSpec(PAT_DIAGONAL, NORM_DIV_MAX, [direction, 0, 0, 0, 0, 0], lo, hi)
    if mode == "radial":                                                 # :367-401
        cx = W // 2 + rng.randint(-W // 4, W // 4)
        cy = H // 2 + rng.randint(-H // 4, H // 4)
        rotation = rng.random() * 2 * math.pi
        wedge = 2 * math.pi / count
        return PatternSpec(PAT_RADIAL, NORM_NONE, [cx, cy, rotation, 0, 0, 0], [i * wedge for i in range(count)], [(i + 1) * wedge for i in range(count)])
    if mode == "waves":                                                  # :404-447
        frequency = rng.uniform(1.5, 4.0)
        amplitude = rng.uniform(0.05, 0.15)
        direction = rng.choice(["horizontal", "vertical", "diagonal"])
        phase = rng.random() * 2 * math.pi
        lo, hi = _bands(count)
        return PatternSpec(PAT_WAVES, NORM_MINMAX, [frequency, amplitude, {"horizontal": 0, "vertical": 1, "diagonal": 2}[direction], phase, 0, 0], lo, hi)
    if mode == "spiral":                                                 # :450-485
        tightness = rng.uniform(2.0, 5.0)
        rotation = rng.random() * 2 * math.pi
        lo, hi = _bands(count)
        return PatternSpec(PAT_SPIRAL, NORM_NONE, [W // 2, H // 2, tightness, rotation, max(H, W), 0], lo, hi)
    if mode == "concentric":                                             # :488-521
        cx = W // 2 + rng.randint(-W // 6, W // 6)
        cy = H // 2 + rng.randint(-H // 6, H // 6)
        lo, hi = _bands(count)
        return PatternSpec(PAT_CONCENTRIC, NORM_DIV_MAX, [cx, cy, 0, 0, 0, 0], lo, hi)
    raise ValueError(f"Unknown region mode: {mode}. Available: {AVAILABLE_MODES}")


def _f32(values: Sequence[float]):
    return (C.c_float * len(values))(*[float(v) for v in values])


def pattern_masks(spec: PatternSpec, H: int, W: int, feather: int = 20, device="cuda:0") -> torch.Tensor:
    """fp32 [spec.regions, H, W]: labelling + feather of one analytic / rectangle pattern on the device."""
    lib = _lib.load()
    dev = torch.device(device)
    if dev.type != "cuda":
        raise RuntimeError("pattern_masks runs on a CUDA device; there is no CPU path")
    R = spec.regions
    masks = torch.empty(R, H, W, dtype=torch.float32, device=dev)
    scratch = torch.empty(lib.nst_mask_scratch_floats(R, H, W), dtype=torch.float32, device=dev)
    rects = (C.c_int * (4 * R))(*[int(v) for r in spec.rects for v in r]) if spec.mode == PAT_RECTS else None
    lo = _f32(spec.lo) if spec.mode != PAT_RECTS else None
    hi = _f32(spec.hi) if spec.mode != PAT_RECTS else None
    with torch.cuda.device(dev):
        rc = lib.nst_pattern_masks_f32(spec.mode, spec.norm, _f32(spec.p), lo, hi, rects, R, int(feather), masks.data_ptr(), scratch.data_ptr(),
                                       H, W, torch.cuda.current_stream(dev).cuda_stream)
    _lib.check(rc, "nst_pattern_masks_f32")
    return masks


def generate_region_masks(H: int, W: int, mode: str, count: int, seed: Optional[int] = None, feather: int = 20,
                          region_sizes: Optional[Sequence[float]] = None, device="cuda:0") -> torch.Tensor:
    """region_blend.generate_region_masks (region_blend.py:925-981) on the device: fp32 [count, H, W]."""
    rng = random.Random() if seed is None else random.Random(seed)
    if mode == "random":
        mode = rng.choice([m for m in AVAILABLE_MODES if m != "random"])
    if mode == "voronoi":
        masks = _voronoi_from_rng(H, W, count, rng, feather, region_sizes, device)
    else:
        masks = pattern_masks(pattern_spec(H, W, mode, count, rng), H, W, feather, device)
    if masks.shape[0] < count:                                           # "some generators may produce fewer" (:976-978): repeat the last
        masks = torch.cat([masks, masks[-1:].expand(count - masks.shape[0], -1, -1)])
    return masks[:count].contiguous()


# ------------------------------------------------------------------------------------------------------------------------------
# animation: rotation (region_blend.py:25-66) and organic warp (:524-533, 604-810)
# ------------------------------------------------------------------------------------------------------------------------------
def _check_masks(masks: torch.Tensor, who: str) -> torch.Tensor:
    if not masks.is_cuda:
        raise RuntimeError(f"{who} needs CUDA tensors; there is no CPU path")
    if masks.dim() != 3 or masks.dtype != torch.float32:
        raise ValueError(f"{who}: masks must be float32 [R, H, W]")
    return masks.contiguous()


def rotate_all_masks(masks: torch.Tensor, angle_degrees: float) -> torch.Tensor:
    """Rotate every mask about the frame centre (cv2.warpAffine, bilinear, replicate border) and renormalise to sum 1."""
    if angle_degrees == 0:
        return masks
    masks = _check_masks(masks, "rotate_all_masks")
    lib = _lib.load()
    R, H, W = masks.shape
    out = torch.empty_like(masks)
    with torch.cuda.device(masks.device):
        rc = lib.nst_rotate_masks_f32(masks.data_ptr(), out.data_ptr(), R, H, W, float(angle_degrees), torch.cuda.current_stream(masks.device).cuda_stream)
    _lib.check(rc, "nst_rotate_masks_f32")
    return out


@dataclass
class MorphAnimation:
    """region_blend.MorphAnimation (region_blend.py:524-533)."""
    enabled: bool = False
    speed: float = 1.0
    amplitude: float = 0.15
    frequency: float = 3.0
    octaves: int = 3
    mode: str = "blob"
    seed: int = 42


def organic_offsets(seed: int, time_offset: float) -> List[float]:
    """The 8 phase offsets of one region's flow field: `_generate_flow_field` -> two `_simplex_noise_2d(.., 2 octaves, ..)` calls
    (region_blend.py:604-667), drawn from numpy's default_rng exactly like the reference."""
    out: List[float] = []
    for s, t in ((seed, time_offset), (seed + 1000, time_offset * 1.3)):
        rng = np.random.default_rng(s)
        for octave in range(2):
            out.append(t * (0.5 + 0.3 * octave) + rng.random() * 1000)
            out.append(t * (0.3 + 0.2 * octave) + rng.random() * 1000)
    return out


def warp_all_masks_organic(masks: torch.Tensor, morph: MorphAnimation, frame_idx: int) -> torch.Tensor:
    """region_blend.warp_all_masks_organic (region_blend.py:737-810): per-region noise flow, cv2.remap, gap fill, renormalise."""
    if not morph.enabled:
        return masks
    masks = _check_masks(masks, "warp_all_masks_organic")
    lib = _lib.load()
    R, H, W = masks.shape
    mode = ORGANIC_MODE.get(morph.mode, 0)                               # any other name is the "blob" default (:715-716)
    time_offset = frame_idx * morph.speed * 0.02
    offs: List[float] = []
    for i in range(R):
        offs += organic_offsets(morph.seed + i * 100, time_offset) if mode in (0, 1) else [0.0] * 8
    out = torch.empty_like(masks)
    scratch = torch.empty(lib.nst_mask_scratch_floats(R, H, W), dtype=torch.float32, device=masks.device)
    with torch.cuda.device(masks.device):
        rc = lib.nst_warp_masks_organic_f32(masks.data_ptr(), out.data_ptr(), R, H, W, mode, float(morph.amplitude), float(morph.frequency),
                                            float(time_offset), (C.c_double * len(offs))(*offs), scratch.data_ptr(),
                                            torch.cuda.current_stream(masks.device).cuda_stream)
    _lib.check(rc, "nst_warp_masks_organic_f32")
    return out


def feather_masks(masks: torch.Tensor, feather_px: int) -> torch.Tensor:
    """feather_mask (region_blend.py:97-102) of soft masks, in place on a fresh copy: reflect-padded Gaussian, sigma = feather / 3."""
    masks = _check_masks(masks, "feather_masks")
    if feather_px <= 0:
        return masks
    lib = _lib.load()
    R, H, W = masks.shape
    out = masks.clone()
    scratch = torch.empty(R * H * W, dtype=torch.float32, device=masks.device)
    with torch.cuda.device(masks.device):
        rc = lib.nst_feather_masks_f32(out.data_ptr(), R, int(feather_px), scratch.data_ptr(), H, W, torch.cuda.current_stream(masks.device).cuda_stream)
    _lib.check(rc, "nst_feather_masks_f32")
    return out


def assign_models_to_regions(num_regions: int, num_models: int, assignment: str = "random", weights: Optional[Sequence[float]] = None,
                             seed: Optional[int] = None, original_chance: float = 0.0) -> List[int]:
    """Which model styles which region; -1 = the original frame (region_blend.py:983-1046, same `random.Random` draws)."""
    rng = random.Random() if seed is None else random.Random(seed)
    pool = list(range(num_models)) + ([-1] if original_chance > 0 else [])
    if assignment == "sequential":
        return [pool[i % len(pool)] for i in range(num_regions)]
    if assignment == "random":
        out = []
        for _ in range(num_regions):
            out.append(-1 if (original_chance > 0 and rng.random() < original_chance) else rng.randint(0, num_models - 1))
        return out
    if assignment == "weighted":
        w = list(weights[:num_models]) if weights is not None else [1.0 / num_models] * num_models
        total = sum(w)
        if original_chance > 0:
            return rng.choices(pool, weights=[(x / total) * (1.0 - original_chance) for x in w] + [original_chance], k=num_regions)
        return rng.choices(range(num_models), weights=[x / total for x in w], k=num_regions)
    raise ValueError(f"Unknown assignment mode: {assignment}")


class RegionBlender:
    """region_blend.blend_by_regions (region_blend.py:1690-1787) for batches of frames on the device: base masks from the pattern generator
    (kept for the clip), per-frame rotation (+ re-feather by feather / 2) and organic morph (+ re-feather by max(5, feather / 4)), model
    assignment, one batched composite launch (engine.composite_regions_batch)."""

    def __init__(self, num_models: int, H: int, W: int, mode: str = "voronoi", region_count: Optional[int] = None, assignment: str = "random",
                 weights: Optional[Sequence[float]] = None, feather: int = 20, seed: Optional[int] = None, original_chance: float = 0.0,
                 rotation_rate: float = 0.0, morph: Optional[MorphAnimation] = None, device="cuda:0", configs=None):
        self.num_models, self.H, self.W, self.mode = int(num_models), H, W, mode
        self.region_count = region_count if region_count is not None else self.num_models
        self.assignment, self.weights, self.feather, self.seed = assignment, weights, int(feather), seed
        self.original_chance, self.rotation_rate, self.morph, self.device = float(original_chance), float(rotation_rate), morph, device
        self.animated = self.rotation_rate != 0 or bool(morph and morph.enabled)
        # blend_by_regions_advanced (region_blend.py:1832-1957): per-region model blends instead of one model per region.  configs: one
        # object per region with `model_indices` / `model_weights` (region_crops.RegionConfig, or the reference's own RegionConfig, which
        # `parse_region_configs` builds from the CLI's blend / scale strings -- that parsing stays in the reference's CLI layer)
        self.configs = list(configs) if configs is not None else None
        self._base = None
        self._assignments: Optional[List[int]] = None

    def _fresh(self):
        masks = generate_region_masks(self.H, self.W, self.mode, self.region_count, self.seed, self.feather, device=self.device)
        return masks, assign_models_to_regions(masks.shape[0], self.num_models, self.assignment, self.weights, self.seed, self.original_chance)

    def masks_at(self, frame_idx: int):
        """(masks fp32 [R, H, W], assignments) of one frame.  Animated clips and seeded static ones keep their base masks (the reference's
        mask cache, and its regeneration with the same seed); an unseeded static clip draws new masks every frame, like the reference."""
        if self._base is None or (not self.animated and self.seed is None):
            self._base, self._assignments = self._fresh()
        masks = self._base
        if self.rotation_rate != 0:
            masks = feather_masks(rotate_all_masks(masks, frame_idx * self.rotation_rate), self.feather // 2)
        if self.morph and self.morph.enabled:
            masks = feather_masks(warp_all_masks_organic(masks, self.morph, frame_idx), max(5, self.feather // 4))
        return masks, self._assignments

    def __call__(self, styled: Sequence[torch.Tensor], original: Optional[torch.Tensor] = None, frame_idx: int = 0,
                 out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """styled: one uint8 [N, H, W, 3] CUDA batch per model (consecutive frames frame_idx .. frame_idx + N - 1); original: the same
        for the unstyled frames.  Returns the composited uint8 [N, H, W, 3] batch."""
        from . import engine
        N = styled[0].shape[0]
        if not self.animated and self.seed is not None:
            masks, assign = self.masks_at(frame_idx)
        else:
            per = [self.masks_at(frame_idx + i) for i in range(N)]
            masks, assign = torch.stack([p[0] for p in per]), per[-1][1]
            if any(p[1] != assign for p in per):          # unseeded random assignment differs per frame: composite frame by frame
                res = torch.cat([engine.composite_regions_batch([s[i:i + 1] for s in styled], per[i][0], [[a] for a in per[i][1]],
                                                                original=original[i:i + 1] if original is not None else None) for i in range(N)])
                return res if out is None else out.copy_(res)
        if self.configs is not None:
            return engine.composite_regions_batch(list(styled), masks, [list(c.model_indices) for c in self.configs],
                                                  [list(c.model_weights) for c in self.configs], original=original, out=out)
        return engine.composite_regions_batch(list(styled), masks, [[a] for a in assign], original=original, out=out)
