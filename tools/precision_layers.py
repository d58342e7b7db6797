"""Per-layer precision ablation on the CPU oracle (design study, not a test).

Emulates the default NST_PREC_F16 scheme (fp16 operands, fp16 weights, fp16 storage of raw conv
outputs) and switches a chosen set of conv layers (by call index) to a higher-precision scheme to
see which layers the uint8 tail error comes from.  Usage: python tools/precision_layers.py arch H W
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F

from oracle import nst_oracle as O
from oracle.precision_study import rnd
from neuralstyletransferv1_b200 import synth

_conv, _convT = F.conv2d, F.conv_transpose2d
STATE = {"i": 0, "hi": {}, "base": ("f16", "f16", "f16")}


def _wrap(real):
    def f(x, w, b=None, **kw):
        i = STATE["i"]
        STATE["i"] += 1
        a, ww, s = STATE["hi"].get(i, STATE["base"])
        y = real(rnd(x, a), rnd(w, ww), b, **kw)
        return y if (y.shape[1] == 3 or s is None) else rnd(y, s)
    return f


def run(arch, sd, u8, preset, hi):
    STATE["i"], STATE["hi"] = 0, hi
    F.conv2d, F.conv_transpose2d = _wrap(_conv), _wrap(_convT)
    try:
        return O.stylize_u8(arch, sd, u8, preset)
    finally:
        F.conv2d, F.conv_transpose2d = _conv, _convT


def main():
    arch = sys.argv[1] if len(sys.argv) > 1 else "reconet"
    H, W = (int(sys.argv[2]), int(sys.argv[3])) if len(sys.argv) > 3 else (272, 480)
    preset = {"johnson": "raw_255", "nst": "raw_01", "reconet": "tanh"}[arch]
    nl = 16
    X3 = ("f16x2", "f16x2", None)
    cases = {
        "baseline f16/f16/f16": {},
        "fp32 storage everywhere": {i: ("f16", "f16", None) for i in range(nl)},
        "last conv x3": {nl - 1: X3},
        "last conv x3 + its input stored fp32": {nl - 1: X3, nl - 2: ("f16", "f16", None)},
        "last 2 convs x3": {nl - 1: X3, nl - 2: X3},
        "last 3 convs x3": {nl - 1: X3, nl - 2: X3, nl - 3: X3},
        "first 3 convs x3": {0: X3, 1: X3, 2: X3},
        "trunk x3": {i: X3 for i in range(3, nl - 3)},
        "all x3": {i: X3 for i in range(nl)},
        "A operand f16x2 everywhere (weights f16)": {i: ("f16x2", "f16", "f16") for i in range(nl)},
        "weights f16x2 everywhere": {i: ("f16", "f16x2", "f16") for i in range(nl)},
    }
    torch.set_num_threads(os.cpu_count())
    for seed in (0, 1, 2):
        sd = synth.make_state_dict(arch, seed)
        u8 = synth.synthetic_frames_u8(2, H, W, seed=5 + seed)
        ref = O.stylize_u8(arch, sd, u8, preset)
        for name, hi in cases.items():
            out = run(arch, sd, u8, preset, hi)
            d = (out.int() - ref.int()).abs()
            mx, psnr = O.u8_metrics(out, ref)
            print(f"{arch} seed{seed} {H}x{W} {name:45s} max|d|={mx} n(|d|>=2)={int((d >= 2).sum()):6d} n(>=3)={int((d >= 3).sum()):4d} PSNR={psnr:.2f}", flush=True)


if __name__ == "__main__":
    main()
