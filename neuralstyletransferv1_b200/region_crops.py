"""Crop-based region styling on the device: `--region_optimize` of the reference styles only the padded
bounding box of every region with the models that region uses and composites the crops (pipeline.py:1293-1400,
region_blend.py:1959-2294).

    prepare_region_crops   region_blend.py:2019-2059   tight bbox of mask > 0.01 (device reduction, `nst_mask_bboxes`) + padding + mask crop
    RegionCropStyler       pipeline.py:1293-1400       per (model, region): extract crop -> optional bilinear down-scale -> preset -> generator
                                                       (any crop size plans; fp32 surface) -> post -> clamp -> up-scale back ->
                                                       composite_from_crops (engine.composite_from_crops, one kernel over the canvas)

Crops are views of the frame (no copies before the generator reads them); bounding boxes are read back to the host once per mask set
(masks are static for a clip unless they are animated, then once per frame -- 4 ints per region).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, List, Optional, Sequence, Tuple

import torch
import torch.nn.functional as F

from . import _lib, engine
from .region_anim import BlendAnimation, ScaleAnimation, animated_scale, animated_weights


@dataclass
class RegionConfig:
    """region_blend.RegionConfig (region_blend.py:1116-1121)."""
    model_indices: List[int]          # -1 = the original frame
    model_weights: List[float]
    scale: float = 1.0


@dataclass
class RegionCrop:
    """region_blend.RegionCrop (region_blend.py:1959-1966) without the full-frame mask copy."""
    region_idx: int
    bbox: Tuple[int, int, int, int]           # (x1, y1, x2, y2) tight
    padded_bbox: Tuple[int, int, int, int]
    crop_mask: torch.Tensor                   # fp32 [h, w] view of the region mask
    config: RegionConfig


def mask_bboxes(masks: torch.Tensor, threshold: float = 0.01) -> List[Tuple[int, int, int, int]]:
    """compute_mask_bbox (region_blend.py:1969-1994) for every mask of fp32 [R, H, W]: (x1, y1, x2 + 1, y2 + 1), full frame when empty."""
    if not masks.is_cuda:
        raise RuntimeError("mask_bboxes needs CUDA tensors; there is no CPU path")
    lib = _lib.load()
    masks = masks.to(torch.float32).contiguous()
    R, H, W = masks.shape
    box = torch.empty(R, 4, dtype=torch.int32, device=masks.device)
    with torch.cuda.device(masks.device):
        rc = lib.nst_mask_bboxes(masks.data_ptr(), R, H, W, float(threshold), box.data_ptr(), torch.cuda.current_stream(masks.device).cuda_stream)
    _lib.check(rc, "nst_mask_bboxes")
    out = []
    for x1, y1, x2, y2 in box.cpu().tolist():
        out.append((0, 0, W, H) if x2 < 0 else (x1, y1, x2 + 1, y2 + 1))
    return out


def pad_bbox(bbox, padding: int, max_w: int, max_h: int):
    """region_blend.pad_bbox (region_blend.py:1997-2016)."""
    x1, y1, x2, y2 = bbox
    return (max(0, x1 - padding), max(0, y1 - padding), min(max_w, x2 + padding), min(max_h, y2 + padding))


def prepare_region_crops(masks: torch.Tensor, configs: Sequence[RegionConfig], H: int, W: int, padding: int = 64) -> List[RegionCrop]:
    """region_blend.prepare_region_crops (region_blend.py:2019-2059); masks: fp32 [R, H, W] CUDA."""
    crops = []
    for idx, (bbox, config) in enumerate(zip(mask_bboxes(masks), configs)):
        px1, py1, px2, py2 = pb = pad_bbox(bbox, padding, W, H)
        crops.append(RegionCrop(idx, bbox, pb, masks[idx, py1:py2, px1:px2], config))
    return crops


def models_needed(crops: Sequence[RegionCrop]) -> List[int]:
    """Sorted model indices any region uses, the original (-1) excluded (region_blend.get_models_needed_for_regions, :2112-2127)."""
    return sorted({mi for c in crops for mi in c.config.model_indices if mi >= 0})


def regions_for_model(crops: Sequence[RegionCrop], model_idx: int) -> List[RegionCrop]:
    """region_blend.get_regions_for_model (:2130-2141)."""
    return [c for c in crops if model_idx in c.config.model_indices]


def merge_bboxes(bboxes: Sequence[Tuple[int, int, int, int]]) -> Tuple[int, int, int, int]:
    """The box containing all boxes; (0, 0, 0, 0) for none (region_blend.merge_bboxes, :2144-2162)."""
    if not bboxes:
        return (0, 0, 0, 0)
    xs1, ys1, xs2, ys2 = zip(*bboxes)
    return (min(xs1), min(ys1), max(xs2), max(ys2))


def crop_coverage(crops: Sequence[RegionCrop], H: int, W: int) -> float:
    """Styled pixels / frame pixels (> 1 with overlap): what `--region_optimize` saves (region_blend.compute_crop_coverage, :2165-2183)."""
    return sum((c.padded_bbox[2] - c.padded_bbox[0]) * (c.padded_bbox[3] - c.padded_bbox[1]) for c in crops) / (H * W)


def _pre(crop01: torch.Tensor, preset: str) -> torch.Tensor:
    """the three branches of pipeline.py:1357-1372"""
    if preset in ("imagenet_255", "imagenet_01"):
        mean = torch.tensor(engine.IMAGENET_MEAN, device=crop01.device).view(1, 3, 1, 1)
        std = torch.tensor(engine.IMAGENET_STD, device=crop01.device).view(1, 3, 1, 1)
        if preset == "imagenet_255":
            return (crop01 * 255.0 - mean * 255.0) / (std * 255.0)
        return (crop01 - mean) / std
    return crop01 * 255.0


def _post(y: torch.Tensor, preset: str) -> torch.Tensor:
    if preset == "imagenet_01":
        mean = torch.tensor(engine.IMAGENET_MEAN, device=y.device).view(1, 3, 1, 1)
        std = torch.tensor(engine.IMAGENET_STD, device=y.device).view(1, 3, 1, 1)
        return (y * std + mean).clamp(0, 1)
    return (y / torch.full((), 255.0, device=y.device)).clamp(0, 1)         # a tensor divisor: torch's CUDA *scalar* division is a reciprocal multiply


class RegionCropStyler:
    """The `--region_optimize` branch of the per-frame loop (pipeline.py:1293-1400) for Transformer / ReCoNet generators.

    models: the drop-in modules (already on the device, weights loaded); presets: one io preset per model (the branch knows
    ``imagenet_255``, ``imagenet_01`` and, for everything else, raw_255 -- pipeline.py:1357-1372).  ``use_original``: pass the frame as
    `original` to the composite (gap fill and the -1 index), like ``region_original > 0`` in the reference.
    """

    def __init__(self, models: Sequence[torch.nn.Module], presets: Sequence[str], masks: torch.Tensor, configs: Sequence[RegionConfig],
                 padding: int = 64, use_original: bool = True, blend_animations: Optional[Sequence[BlendAnimation]] = None,
                 scale_animations: Optional[Sequence[ScaleAnimation]] = None):
        self.models, self.presets = list(models), list(presets)
        self.use_original = bool(use_original)
        self.padding = int(padding)
        self.configs = list(configs)
        self.blend_animations = list(blend_animations) if blend_animations else None      # per region (pipeline.py:1391-1398)
        self.scale_animations = list(scale_animations) if scale_animations else None      # per region (pipeline.py:1313-1319)
        self.set_masks(masks)

    def set_masks(self, masks: torch.Tensor) -> None:
        """New masks (e.g. the next frame of an animated mask set): bounding boxes are recomputed."""
        R, H, W = masks.shape
        self.H, self.W = H, W
        self.crops = prepare_region_crops(masks, self.configs, H, W, self.padding)

    def __call__(self, frame_u8: torch.Tensor, frame_idx: int = 0) -> torch.Tensor:
        """frame_u8: uint8 [H, W, 3] CUDA; frame_idx drives the blend / scale animations.  Returns out01 fp32 [3, H, W], what
        composite_from_crops returns in the reference."""
        if tuple(frame_u8.shape) != (self.H, self.W, 3) or frame_u8.dtype != torch.uint8 or not frame_u8.is_cuda:
            raise ValueError(f"frame_u8 must be a CUDA uint8 [{self.H}, {self.W}, 3] tensor")
        x01 = (frame_u8.permute(2, 0, 1).to(torch.float32) / torch.full((), 255.0, device=frame_u8.device)).contiguous()   # ToTensor, exact
        styled: Dict[int, Dict[int, torch.Tensor]] = {}
        for mi in models_needed(self.crops):
            styled[mi] = {}
            for c in regions_for_model(self.crops, mi):
                x1, y1, x2, y2 = c.padded_bbox
                ch, cw = y2 - y1, x2 - x1
                crop = x01[:, y1:y2, x1:x2].unsqueeze(0)
                scale = c.config.scale
                if self.scale_animations and c.region_idx < len(self.scale_animations):
                    scale = animated_scale(scale, frame_idx, self.scale_animations[c.region_idx])
                if scale < 1.0:
                    sh, sw = max(1, int(ch * scale)), max(1, int(cw * scale))
                    crop = F.interpolate(crop, size=(sh, sw), mode="bilinear", align_corners=False)
                with torch.no_grad():
                    y = self.models[mi](_pre(crop, self.presets[mi]).contiguous())
                # (the reference up-scales a down-scaled region back to the box here, pipeline.py:1375-1379; composite_from_crops does exactly
                # that resize, with the reference's bilinear arithmetic, for any crop whose size differs from its box)
                styled[mi][c.region_idx] = _post(y, self.presets[mi]).squeeze(0).contiguous()
        crops = self.crops
        if self.blend_animations:                  # composite_from_crops(..., frame_idx, blend_animations) (region_blend.py:2217-2222)
            crops = [RegionCrop(c.region_idx, c.bbox, c.padded_bbox, c.crop_mask,
                                RegionConfig(c.config.model_indices, animated_weights(c.config.model_weights, frame_idx, self.blend_animations[c.region_idx])
                                             if c.region_idx < len(self.blend_animations) else c.config.model_weights, c.config.scale))
                     for c in self.crops]
        return engine.composite_from_crops(styled, crops, x01 if self.use_original else None, self.H, self.W)
