"""Debug: a re-scheduling of the same arithmetic must give bit-identical frames.  Default: the pipelined plain epilogue against the
unpipelined one (NST_TC_DBG=65536: same per-lane summation order); `python tools/pipe_check.py NST_NO_RES_RING=1` compares any other
knob (here: residual operand staged inside the operand stage instead of its own ring).  One process per setting."""
import os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CHILD = r'''
import sys, os, torch, hashlib
sys.path.insert(0, %r)
import neuralstyletransferv1_b200 as m
from neuralstyletransferv1_b200 import synth
for arch, cls, preset in (("johnson", m.TransformerNet, "raw_255"), ("reconet", m.ReCoNet, "tanh")):
    model = cls().to("cuda:0"); model.load_state_dict(synth.make_state_dict(arch, 0)); model.eval()
    for (n, h, w) in ((1, 16, 16), (2, 24, 40), (1, 64, 96), (3, 120, 200), (1, 272, 480)):
        fr = synth.synthetic_frames_u8(n, h, w, seed=3).cuda()
        out = model.stylize_u8(fr, preset)
        out2 = model.stylize_u8(fr, preset)
        torch.cuda.synchronize()
        print(arch, n, h, w, hashlib.md5(out.cpu().numpy().tobytes()).hexdigest(), bool(torch.equal(out, out2)), int(out.float().mean()))
''' % ROOT
res = {}
knob = sys.argv[1] if len(sys.argv) > 1 else "NST_TC_DBG=65536"
kname, kval = knob.split("=")
for dbg in ("0", "65536"):
    env = dict(os.environ)
    if dbg != "0":
        env[kname] = kval
    res[dbg] = subprocess.run([sys.executable, "-c", CHILD], env=env, capture_output=True, text=True)
    print("dbg", dbg); print(res[dbg].stdout); print(res[dbg].stderr[-600:])
print("IDENTICAL" if res["0"].stdout == res["65536"].stdout and res["0"].stdout else "DIFFERENT")
