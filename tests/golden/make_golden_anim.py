"""tests/golden/region_anim.json: values of the reference's own compute_harmonic_value / compute_animated_weights /
compute_animated_scale (region_blend.py:1127-1247, 1334-1370) on a grid of arguments.  Build container only."""
import dataclasses
import itertools
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("NST_REFERENCE_DIR", "/root/reference")


def main():
    sys.path.insert(0, REF)
    import region_blend as rb
    sys.path.remove(REF)
    waves = ["sine", "triangle", "sawtooth", "sawtooth_down", "square", "bogus"]
    harm = []
    for wf, (period, lo, hi, ph), idx in itertools.product(waves, [(120.0, 0.0, 1.0, 0.0), (37.5, 0.2, 0.9, 45.0), (0.0, 0.3, 0.7, 10.0), (60.0, 1.0, 0.25, 270.0)],
                                                           [0, 1, 17, 30, 59, 60, 119, 1234]):
        args = [idx, period, lo, hi, ph, wf]
        harm.append([args, rb.compute_harmonic_value(*args)])
    weights = []
    for base, anim, idx in itertools.product([[1.0], [0.7, 0.3], [0.5, 0.25, 0.25], [0.0, 0.0]],
                                             [dict(enabled=True, period=48.0, min_opacity=0.0, max_opacity=1.0, phase_offset=0.0, waveform="sine", per_model_phase=True),
                                              dict(enabled=True, period=30.0, min_opacity=0.1, max_opacity=0.8, phase_offset=90.0, waveform="triangle", per_model_phase=False),
                                              dict(enabled=True, period=20.0, min_opacity=0.0, max_opacity=1.0, phase_offset=0.0, waveform="square", per_model_phase=False),
                                              dict(enabled=False)],
                                             [0, 5, 12, 24, 47]):
        a = rb.BlendAnimation(**anim)
        weights.append([base, idx, dataclasses.asdict(a), rb.compute_animated_weights(base, idx, a)])
    scale = []
    for base, anim, idx in itertools.product([1.0, 0.5], [dict(enabled=True, period=60.0, min_scale=0.5, max_scale=1.0, phase_offset=0.0, waveform="sine"),
                                                          dict(enabled=True, period=24.0, min_scale=0.25, max_scale=0.75, phase_offset=180.0, waveform="sawtooth"),
                                                          dict(enabled=False)], [0, 7, 15, 30, 59]):
        a = rb.ScaleAnimation(**anim)
        scale.append([base, idx, dataclasses.asdict(a), rb.compute_animated_scale(base, idx, a)])
    json.dump({"harmonic": harm, "weights": weights, "scale": scale}, open(os.path.join(HERE, "region_anim.json"), "w"))
    print(len(harm), len(weights), len(scale))


if __name__ == "__main__":
    main()
