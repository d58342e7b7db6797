"""GPU bring-up diagnostics: which layers of which family match the oracle on which kernel.

Usage (on the GPU box): python tools/diag_gpu.py [arch ...]
Each configuration runs in a fresh subprocess so a trapped kernel cannot poison the next one.
"""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

CHILD = r"""
import sys, json, torch
sys.path.insert(0, %(root)r)
import neuralstyletransferv1_b200 as m
from neuralstyletransferv1_b200 import synth
from oracle import nst_oracle as O
arch, prec, n, h, w = %(arch)r, %(prec)r, %(n)d, %(h)d, %(w)d
cls = {"johnson": m.TransformerNet, "nst": m.TransformerNetNST, "reconet": m.ReCoNet}[arch]
preset = {"johnson": "raw_255", "nst": "raw_01", "reconet": "tanh"}[arch]
sd = synth.make_state_dict(arch, 0)
model = cls(precision=prec).to("cuda:0"); model.load_state_dict(sd); model.eval()
frames = synth.synthetic_frames_u8(n, h, w, seed=3)
out = model.stylize_u8(frames.cuda(), preset)
torch.cuda.synchronize()
ref = O.stylize_u8(arch, sd, frames, preset)
mx, psnr = O.u8_metrics(out.cpu(), ref)
kinds = [d["kernel"][0] for d in model.engine.layer_info()]
print("RESULT " + json.dumps({"max": mx, "psnr": round(psnr, 2), "kernels": "".join(kinds)}))
"""


def run(arch, prec, env=None, n=1, h=64, w=64):
    e = dict(os.environ)
    e.update(env or {})
    code = CHILD % {"root": ROOT, "arch": arch, "prec": prec, "n": n, "h": h, "w": w}
    try:
        r = subprocess.run([sys.executable, "-c", code], env=e, capture_output=True, text=True, timeout=300)
    except subprocess.TimeoutExpired:
        return "TIMEOUT"
    for line in r.stdout.splitlines():
        if line.startswith("RESULT "):
            return line[7:]
    return "FAIL rc=%d %s" % (r.returncode, (r.stderr.strip().splitlines() or ["?"])[-1][:300])


def main():
    archs = sys.argv[1:] or ["johnson", "reconet", "nst"]
    nlayers = {"johnson": 16, "nst": 16, "reconet": 14}
    for arch in archs:
        print(f"== {arch}", flush=True)
        print("  direct fp32      :", run(arch, "fp32_direct"), flush=True)
        full = run(arch, "f16")
        print("  tcgen05 f16      :", full, flush=True)
        ok = full.startswith("{") and json.loads(full)["max"] <= 2
        if not ok:
            print("  tcgen05 f16 swap :", run(arch, "f16", {"NST_TC_DBG": "1"}), flush=True)
            for li in range(nlayers[arch]):
                print(f"  only layer {li:2d} tc :", run(arch, "f16", {"NST_TC_LAYER_MASK": hex(1 << li)}), flush=True)
        else:
            print("  tcgen05 f16x3    :", run(arch, "f16x3"), flush=True)
            print("  tcgen05 bf16     :", run(arch, "bf16"), flush=True)
            print("  tcgen05 f16 odd  :", run(arch, "f16", n=3, h=132, w=172), flush=True)


if __name__ == "__main__":
    main()
