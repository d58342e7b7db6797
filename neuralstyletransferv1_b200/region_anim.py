"""Per-frame animation scalars of the region blend (host side of SURVEY.md 8 rows a-R / N4): the harmonic oscillators that turn a
region's static model weights / scale into the weights / scale of frame `idx` before they reach the composite kernels
(region_blend.py:1127-1247 `compute_harmonic_value`, `BlendAnimation`, `compute_animated_weights`; :1334-1370 `ScaleAnimation`,
`compute_animated_scale`).  A handful of doubles per region and frame: host arithmetic, Python doubles like the reference.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import List, Sequence

_WAVES = {
    "sine": lambda t: (math.sin(2 * math.pi * t) + 1) / 2,
    "triangle": lambda t: t * 2 if t < 0.5 else 2 - (t * 2),
    "sawtooth": lambda t: t,
    "sawtooth_down": lambda t: 1 - t,
    "square": lambda t: 1.0 if t < 0.5 else 0.0,
}


def harmonic_value(frame_idx: int, period: float, min_val: float = 0.0, max_val: float = 1.0, phase_offset: float = 0.0,
                   waveform: str = "sine") -> float:
    """Position in the cycle = frame / period + phase / 360 (mod 1) -> waveform in 0..1 -> [min_val, max_val]; an unknown waveform is a
    sine, a non-positive period the mid value (region_blend.py:1127-1181)."""
    if period <= 0:
        return (min_val + max_val) / 2
    t = ((frame_idx / period) + (phase_offset / 360.0)) % 1.0
    return min_val + _WAVES.get(waveform, _WAVES["sine"])(t) * (max_val - min_val)


@dataclass
class BlendAnimation:
    """region_blend.BlendAnimation (region_blend.py:1184-1192)."""
    enabled: bool = False
    period: float = 120.0
    min_opacity: float = 0.0
    max_opacity: float = 1.0
    phase_offset: float = 0.0
    waveform: str = "sine"
    per_model_phase: bool = True


def animated_weights(base_weights: Sequence[float], frame_idx: int, animation: BlendAnimation) -> List[float]:
    """Model i of n oscillates 360 / n degrees behind model i - 1 (or in phase), scaled by its static weight; normalised to sum 1, equal
    weights when everything is (almost) off (region_blend.py:1195-1247)."""
    n = len(base_weights)
    if not animation.enabled or n <= 1:
        return list(base_weights)
    raw = [harmonic_value(frame_idx, animation.period, animation.min_opacity, animation.max_opacity,
                          animation.phase_offset + (i * 360.0 / n if animation.per_model_phase else 0.0), animation.waveform) * w
           for i, w in enumerate(base_weights)]
    total = sum(raw)
    return [1.0 / n] * n if total < 1e-6 else [w / total for w in raw]


@dataclass
class ScaleAnimation:
    """region_blend.ScaleAnimation (region_blend.py:1334-1341)."""
    enabled: bool = False
    period: float = 60.0
    min_scale: float = 0.5
    max_scale: float = 1.0
    phase_offset: float = 0.0
    waveform: str = "sine"


def animated_scale(base_scale: float, frame_idx: int, animation: ScaleAnimation) -> float:
    """region_blend.compute_animated_scale (region_blend.py:1344-1370)."""
    if not animation.enabled:
        return base_scale
    return harmonic_value(frame_idx, animation.period, animation.min_scale, animation.max_scale, animation.phase_offset, animation.waveform)
