"""Emulate operand-rounding schemes of the tensor-core conv on the CPU oracle.

Not a test: a design study (DESIGN.md section "precision") that picks the MMA
operand scheme able to meet max|d| <= 2 / PSNR >= 45 dB against fp32.
Usage: python oracle/precision_study.py [H W]
"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F
from oracle import nst_oracle as O
from neuralstyletransferv1_b200 import synth

_real_conv2d = F.conv2d
_real_convT = F.conv_transpose2d
SCHEME = {"a": None, "w": None, "s": None, "s_only_c": None}   # s: storage rounding of conv outputs that feed an InstanceNorm


def rnd(t, kind):
    if kind is None:
        return t
    if kind == "bf16":
        return t.to(torch.bfloat16).float()
    if kind == "f16":
        return t.to(torch.float16).float()
    if kind == "f16x2":
        hi = t.to(torch.float16).float()
        lo = (t - hi).to(torch.float16).float()
        return hi + lo
    if kind == "bf16x2":
        hi = t.to(torch.bfloat16).float()
        lo = (t - hi).to(torch.bfloat16).float()
        return hi + lo
    raise ValueError(kind)


def _store(y):
    # the kernels keep raw conv outputs (pre-norm) in HBM; the image-producing conv (3 channels) is never stored
    if SCHEME["s"] is None or y.shape[1] == 3:
        return y
    if SCHEME["s_only_c"] is not None and y.shape[1] not in SCHEME["s_only_c"]:
        return y
    return rnd(y, SCHEME["s"])


def conv2d(x, w, b=None, **kw):
    return _store(_real_conv2d(rnd(x, SCHEME["a"]), rnd(w, SCHEME["w"]), b, **kw))


def convT(x, w, b=None, **kw):
    return _store(_real_convT(rnd(x, SCHEME["a"]), rnd(w, SCHEME["w"]), b, **kw))


def main():
    H, W = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (256, 256)
    torch.set_num_threads(os.cpu_count())
    if os.environ.get("STORAGE"):
        # activation-storage study: f16 operands, raw conv outputs rounded to fp16 in HBM (all layers / trunk only)
        for arch, preset in (("johnson", "raw_255"), ("reconet", "tanh"), ("nst", "raw_01")):
            trunk = {"johnson": (128,), "nst": (128,), "reconet": (192,)}[arch]
            for seed in (0, 1, 2):
                sd = synth.make_state_dict(arch, seed)
                u8 = synth.synthetic_frames_u8(2, H, W, seed=5 + seed)
                ref = O.stylize_u8(arch, sd, u8, preset)
                for st, only in ((None, None), ("f16", None), ("f16", trunk), ("bf16", trunk)):
                    SCHEME.update(a="f16", w="f16", s=st, s_only_c=only)
                    F.conv2d, F.conv_transpose2d = conv2d, convT
                    out = O.stylize_u8(arch, sd, u8, preset)
                    F.conv2d, F.conv_transpose2d = _real_conv2d, _real_convT
                    mx, psnr = O.u8_metrics(out, ref)
                    print(f"{arch:8s} seed{seed} operands f16, storage={st} only_c={only}  max|d|={mx:3d}  PSNR={psnr:6.2f} dB", flush=True)
        return
    schemes = [("f16", "f16"), ("f16x2", "f16")] if os.environ.get("QUICK") else [("bf16", "bf16"), ("f16", "f16"), ("f16x2", "f16"), ("f16x2", "bf16"), ("bf16x2", "bf16x2"), ("f16x2", "f16x2")]
    for arch, preset in (("johnson", "raw_255"), ("reconet", "tanh"), ("nst", "raw_01")):
        for seed in (0, 1):
            sd = synth.make_state_dict(arch, seed)
            u8 = synth.synthetic_frames_u8(1 if os.environ.get("QUICK") else 2, H, W, seed=5 + seed)
            F.conv2d, F.conv_transpose2d = _real_conv2d, _real_convT
            ref = O.stylize_u8(arch, sd, u8, preset)
            for a, w in schemes:
                SCHEME["a"], SCHEME["w"] = a, w
                F.conv2d, F.conv_transpose2d = conv2d, convT
                out = O.stylize_u8(arch, sd, u8, preset)
                F.conv2d, F.conv_transpose2d = _real_conv2d, _real_convT
                mx, psnr = O.u8_metrics(out, ref)
                print(f"{arch:8s} seed{seed} A={a:7s} W={w:7s}  max|d|={mx:3d}  PSNR={psnr:6.2f} dB")


if __name__ == "__main__":
    main()
