#!/usr/bin/env python
"""Headline benchmark: 1080p frames/sec of the TransformerNet stylisation hot path on N B200s.

  python bench.py --gpus N --steps K --warmup W            (N>1: launched under torchrun)
  python bench.py --impl reference ...                     (the reference's CPU path, host cores)

One step = one pass of the hot path over one batch of `--frames` synthetic uint8 1080p frames
per GPU (BASELINE.json configs[1]: TransformerNet, batch of 64 x 1080p on one B200).  Frames are
independent, so ranks take disjoint frame batches and never communicate on the data path
("scaling": "weak"); torch.distributed (NCCL) is used only for the barrier and the max-over-ranks
time.  `value` is timed with inputs resident in HBM; `e2e` goes through the public module API with
pinned HOST buffers (H2D + stylize_u8 + D2H inside the timed region).  Prints ONE JSON line.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "1080p frames/sec"
UNIT = "frames/s"
GFLOP_PER_1080P = {"johnson": 637.81, "nst": 585.31, "reconet": 1214.70}
PRESET = {"johnson": "raw_255", "nst": "raw_01", "reconet": "tanh"}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--arch", default="johnson", choices=["johnson", "nst", "reconet"])
    ap.add_argument("--precision", default="f16")
    ap.add_argument("--frames", type=int, default=64, help="frames per step per GPU")
    ap.add_argument("--height", type=int, default=1080)
    ap.add_argument("--width", type=int, default=1920)
    ap.add_argument("--cpu-frames", type=int, default=3, help="frames of the bounded CPU-baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    return ap.parse_args()


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        d = json.load(open(path))
        return {"tflops": float(d.get("bf16_tflops_sustained", 1393.3)), "hbm": float(d.get("hbm_gbs", 6524.6)), "src": "measured"}
    return {"tflops": 1400.0, "hbm": 6650.0, "src": "fallback"}


def traffic_from_profile():
    """dram bytes (read + write) per launch of the dominant kernel from the committed `ncu --set full` summary, or None."""
    path = os.path.join(ROOT, "profiles", "traffic.json")
    try:
        return json.load(open(path)).get("res_conv_dram_bytes_per_launch")
    except Exception:
        return None


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.samples, self._stop = index, [], threading.Event()
        self._t = threading.Thread(target=self._run, daemon=True)

    def _run(self):
        while not self._stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.samples.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            self._stop.wait(0.2)

    def __enter__(self):
        self._t.start(); return self

    def __exit__(self, *a):
        self._stop.set(); self._t.join(2)

    def summary(self):
        sm = sorted(int(s[0]) for s in self.samples if s and s[0].isdigit())
        reasons = []
        for name, col in (("hw_slowdown", 2), ("hw_thermal_slowdown", 3), ("sw_thermal_slowdown", 4), ("sw_power_cap", 5)):
            if any(len(s) > col and s[col].lower().startswith("active") for s in self.samples):
                reasons.append(name)
        mx = [int(s[1]) for s in self.samples if len(s) > 1 and s[1].isdigit()]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx[0] if mx else None,
                "reasons": reasons, "samples": len(self.samples)}


def cpu_reference_fps(arch, n_frames, h, w, warmup=1):
    """The reference's CPU fp32 path (oracle port of transformer_net.py / model.py) on the host cores."""
    import torch
    from neuralstyletransferv1_b200 import synth
    from oracle import nst_oracle as O
    torch.set_num_threads(os.cpu_count() or 1)
    sd = synth.make_state_dict(arch, 0)
    frames = synth.synthetic_frames_u8(1, h, w, seed=77)
    for _ in range(warmup):
        O.stylize_u8(arch, sd, frames, PRESET[arch])
    t0 = time.perf_counter()
    for _ in range(n_frames):
        O.stylize_u8(arch, sd, frames, PRESET[arch])          # batch 1, like pipeline.py:1080
    dt = time.perf_counter() - t0
    return n_frames / dt, torch.get_num_threads()


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    per_step = 1
    fps_all = []
    fps, cores = cpu_reference_fps(args.arch, per_step, args.height, args.width, warmup=max(1, min(args.warmup, 1)))
    t0 = time.perf_counter()
    import torch
    from neuralstyletransferv1_b200 import synth
    from oracle import nst_oracle as O
    sd = synth.make_state_dict(args.arch, 0)
    frames = synth.synthetic_frames_u8(1, args.height, args.width, seed=77)
    for _ in range(args.steps):
        O.stylize_u8(args.arch, sd, frames, PRESET[args.arch])
    dt = time.perf_counter() - t0
    value = args.steps * per_step / dt
    sample = f"{args.steps} steps x {per_step} frame {args.height}x{args.width}, batch 1, fp32, torch {torch.__version__} CPU"
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1000.0 * dt / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"TransformerNet({args.arch}) {args.height}x{args.width} frames, CPU reference path (oracle port of "
                               "transformer_net.py; /root/reference is Python and its weights are absent, synthetic seeded weights)"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def run_b200(args):
    import torch
    import torch.distributed as dist
    from neuralstyletransferv1_b200 import TransformerNet, TransformerNetNST, ReCoNet, synth

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", local))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    H, W, F = args.height, args.width, args.frames
    cls = {"johnson": TransformerNet, "nst": TransformerNetNST, "reconet": ReCoNet}[args.arch]
    model = cls(precision=args.precision).to(dev)
    model.load_state_dict(synth.make_state_dict(args.arch, 0))
    model.eval()
    preset = PRESET[args.arch]

    # synthetic frames: 8 distinct image-like frames per rank, tiled to the batch; pinned on the host
    uniq = synth.synthetic_frames_u8(min(8, F), H, W, seed=1234 + rank)
    host_in = torch.empty(F, H, W, 3, dtype=torch.uint8).pin_memory()
    for i in range(F):
        host_in[i].copy_(uniq[i % uniq.shape[0]])
    host_out = torch.empty_like(host_in).pin_memory()
    d_in = host_in.to(dev, non_blocking=True)
    d_out = torch.empty_like(d_in)
    torch.cuda.synchronize()

    def step():
        model.stylize_u8(d_in, preset, out=d_out)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        step()
    barrier()
    launches0 = model.engine.launch_count
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local) as clk:
        e0.record()
        for _ in range(args.steps):
            step()
        e1.record()
        barrier()
    ms = e0.elapsed_time(e1)
    launches = model.engine.launch_count - launches0
    t = torch.tensor([ms], device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total = float(t.item())
    value = world * F * args.steps / (ms_total / 1000.0)

    # ---- end to end through the public API with host buffers --------------------------------
    h2d_stream, d2h_stream = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
    chunk = min(F, 8)

    def e2e_step():
        # chunked so the H2D of chunk i+1 (one stream) and the D2H of chunk i-1 (another: PCIe is full duplex)
        # overlap the compute of chunk i
        main = torch.cuda.current_stream()
        keep = []
        for c0 in range(0, F, chunk):
            with torch.cuda.stream(h2d_stream):
                x = host_in[c0:c0 + chunk].to(dev, non_blocking=True)
                ready = torch.cuda.Event(); ready.record(h2d_stream)
            main.wait_event(ready)
            y = model.stylize_u8(x, preset)
            fin = torch.cuda.Event(); fin.record(main)
            with torch.cuda.stream(d2h_stream):
                d2h_stream.wait_event(fin)
                host_out[c0:c0 + chunk].copy_(y, non_blocking=True)
            x.record_stream(main); y.record_stream(d2h_stream)
            keep.append((x, y))
        h2d_stream.synchronize(); d2h_stream.synchronize(); main.synchronize()

    e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        e2e_step()
    barrier()
    e2e_s = time.perf_counter() - t0
    t = torch.tensor([e2e_s], device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_value = world * F * args.steps / float(t.item())

    if rank != 0:
        if world > 1:
            dist.barrier(); dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel: conv_tc_kernel on the 128->128 residual convs -------
    pk = peaks()
    prof_frames = d_in[:min(F, 8)]
    model.engine.profile_layers(prof_frames)
    prof = model.engine.profile_layers(prof_frames)
    nprof = prof_frames.shape[0]
    scale = (H * W) / (1080.0 * 1920.0)
    dom = [d for d in prof if d["kernel"] == "tcgen05" and (".conv" in d["name"] or ".branch." in d["name"])]
    dom = dom or prof
    dom_flops = sum(d["flops"] for d in dom) * nprof
    dom_ms = sum(d["ms"] for d in dom)
    achieved = dom_flops / (dom_ms / 1000.0) / 1e12
    total_ms = sum(d["ms"] for d in prof)
    net_tflops = GFLOP_PER_1080P[args.arch] * scale * 1e9 * world * F * args.steps / (ms_total / 1000.0) / 1e12
    layers = [{"name": d["name"], "kernel": d["kernel"], "ms_per_frame": round(d["ms"] / nprof, 4),
               "tflops": round(d["flops"] * nprof / (d["ms"] / 1000.0) / 1e12, 1),
               "gbs": round(d["bytes"] * nprof / (d["ms"] / 1000.0) / 1e9, 1)} for d in prof]
    roofline = {"bound": "tensor", "kernel": "conv_tc_kernel (tcgen05 implicit GEMM), residual 128->128 3x3 launches",
                "achieved": round(achieved, 2), "peak": pk["tflops"], "peak_source": pk["src"] + " bf16 dense sustained (cuBLAS)",
                "unit": "TFLOP/s", "frac": round(achieved / pk["tflops"], 4), "traffic": traffic_from_profile(),
                "share_of_step": round(dom_ms / total_ms, 3), "launch_ms_avg": round(dom_ms / len(dom), 4),
                "whole_net_tflops": round(net_tflops, 2), "whole_net_frac": round(net_tflops / pk["tflops"], 4)}

    cpu_baseline = None
    if not args.no_cpu_baseline and world == 1:
        fps, cores = cpu_reference_fps(args.arch, args.cpu_frames, H, W)
        cpu_baseline = {"value": fps, "unit": UNIT, "cores": cores, "kind": "port",
                        "sample": f"{args.cpu_frames} frames {H}x{W} batch 1 fp32 (oracle port of the reference torch CPU path), 1 warm-up"}

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f16" if args.precision.startswith("f16") else args.precision, "data": "synthetic",
        "config": {"workload": f"TransformerNet({args.arch}) batch {F} x {H}x{W} uint8 frames per GPU, synthetic seeded weights "
                               f"(BASELINE configs[1]); precision={args.precision}: fp16 tcgen05 operands and fp16 NHWC activation storage, fp32 accumulate and statistics",
                   "frames_per_step_per_gpu": F, "l2": "inputs larger than L2 (batch = %.0f MB uint8, activations ~660 MB/frame)" % (F * H * W * 3 / 1e6),
                   "parallelism": f"frame-batch sharding over {world} GPU(s), no collective"},
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": F * H * W * 3, "d2h_bytes_per_step": F * H * W * 3},
        "gpu_launches": launches,
        "clocks": clk.summary(),
        "roofline": roofline,
        "cpu_baseline": cpu_baseline,
        "layers": layers,
    }
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier(); dist.destroy_process_group()


def main():
    args = parse()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
