#!/usr/bin/env python
"""Headline benchmark: frames/sec of the stylisation hot path on N B200s.

  python bench.py --gpus N --steps K --warmup W [--config c2|c3_4k|c4_reconet300|c5_region|nst]
  python bench.py --impl reference ...            (the reference's CPU path on the host cores)
  (N > 1: launched under torchrun, one rank per GPU)

Configs (BASELINE.json `configs`):
  c2 (default)    TransformerNet, 64 x 1080p uint8 frames per GPU per step            -- the config the metric is quoted on; weak scaling
  nst             the transformer_net_nst variant on the same workload
  c3_4k           TransformerNet, 3840x2160 frames, 16 per GPU per step (micro-batches of 8), frame-batch sharded; weak scaling
  c4_reconet300   ReCoNet, a 300-frame 1080p clip split contiguously over the GPUs (sharding.shard_range); STRONG scaling
  c5_region       4 TransformerNet weight sets over 1080p frames + voronoi region masks (the reference generator's seed 42 /
                  feather 20 set), composited by ONE batched kernel call; weak scaling

One step = one pass of the hot path over one batch of synthetic frames.  Frames are independent, so ranks take disjoint frame
batches and never communicate on the data path; torch.distributed (NCCL) is used only for the barrier and the max-over-ranks
time.  `value` is timed with inputs resident in HBM (CUDA events); `e2e` goes through the package's public batched caller
(`FramePipeline`: pinned HOST buffers, H2D + stylize + D2H inside the timed region).  Prints ONE JSON line.
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

UNIT = "frames/s"
GFLOP_PER_1080P = {"johnson": 637.81, "nst": 585.31, "reconet": 1214.70}     # SURVEY.md 8d (reference formulation)
PRESET = {"johnson": "raw_255", "nst": "raw_01", "reconet": "tanh"}
CONFIGS = {
    #  name            arch       H     W     frames/GPU/step  scaling   metric
    "c2":            ("johnson", 1080, 1920, 64, "weak", "1080p frames/sec"),
    "nst":           ("nst", 1080, 1920, 64, "weak", "1080p frames/sec"),
    "c3_4k":         ("johnson", 2160, 3840, 16, "weak", "4K (3840x2160) frames/sec"),
    "c4_reconet300": ("reconet", 1080, 1920, 300, "strong", "1080p frames/sec (ReCoNet, 300-frame clip)"),
    "c5_region":     ("johnson", 1080, 1920, 16, "weak", "composited 1080p frames/sec (4 styles + region masks)"),
}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", default="c2", choices=sorted(CONFIGS))
    ap.add_argument("--arch", default=None, choices=["johnson", "nst", "reconet"], help="override the config's generator")
    ap.add_argument("--precision", default="f16")
    ap.add_argument("--frames", type=int, default=None, help="frames per step per GPU (strong-scaling configs: frames per step in total)")
    ap.add_argument("--height", type=int, default=None)
    ap.add_argument("--width", type=int, default=None)
    ap.add_argument("--cpu-frames", type=int, default=3, help="frames of the bounded CPU-baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    a = ap.parse_args()
    arch, h, w, f, scaling, metric = CONFIGS[a.config]
    a.arch = a.arch or arch
    a.height, a.width, a.frames = a.height or h, a.width or w, a.frames or f
    a.scaling, a.metric = scaling, metric
    return a


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        d = json.load(open(path))
        return {"tflops": float(d.get("bf16_tflops_sustained", 1393.3)), "hbm": float(d.get("hbm_gbs", 6524.6)), "src": "MEASURED_PEAKS.json"}
    return {"tflops": 1400.0, "hbm": 6650.0, "src": "fallback (B200_PROFILING.md)"}


def traffic_from_profile():
    """dram bytes (read + write) per launch of the dominant kernel, FROM THE COMMITTED `ncu --set full` SUMMARY (not measured by this run)."""
    try:
        d = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        return d.get("res_conv_dram_bytes_per_launch"), d.get("source")
    except Exception:
        return None, None


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
            self._stop.wait(0.1)

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


# ---------------------------------------------------------------------------------------------------------------------------------
# CPU arm: the reference's own modules when the checkout is present (NST_REFERENCE_DIR, default /root/reference), else the oracle port
# ---------------------------------------------------------------------------------------------------------------------------------
def _reference_stylizer(arch):
    """(callable frames_u8 -> u8, kind).  kind "reference": the unmodified reference modules behind the pipeline's io_preset arithmetic;
    kind "port": oracle/nst_oracle.py (the same torch ops restated), used where the reference checkout does not exist (the GPU box)."""
    import torch
    from neuralstyletransferv1_b200 import synth
    from oracle import nst_oracle as O
    sd = synth.make_state_dict(arch, 0)
    ref_dir = os.environ.get("NST_REFERENCE_DIR", "/root/reference")
    try:
        if not os.path.isdir(ref_dir):
            raise ImportError("no reference checkout")
        sys.path.insert(0, ref_dir)
        try:
            if arch == "johnson":
                import transformer_net as mod
                net = mod.TransformerNet()
            elif arch == "nst":
                import transformer_net_nst as mod
                net = mod.TransformerNet()
            else:
                import model as mod
                net = mod.ReCoNet()
        finally:
            sys.path.remove(ref_dir)
        net.load_state_dict(sd)
        net.eval()

        def run(frames):
            x01 = frames.permute(0, 3, 1, 2).to(torch.float32) / 255.0            # ToTensor (pipeline.py:1099)
            with torch.no_grad():
                y = net(O.io_pre(x01, PRESET[arch]))                               # pipeline.py:1447-1486
            return O.pack_u8(O.io_post(y, PRESET[arch]))                           # clamp + ToPILImage (pipeline.py:2094)
        return run, "reference"
    except Exception:
        return (lambda frames: O.stylize_u8(arch, sd, frames, PRESET[arch])), "port"


def cpu_reference_fps(args, n_frames, warmup=1):
    """frames/s of the CPU arm on a bounded sample: n_frames single-frame forwards (batch 1, like pipeline.py:1080) on all host cores.
    c5_region: 4 stylisations + the reference-formulation composite per frame."""
    import torch
    from neuralstyletransferv1_b200 import synth
    torch.set_num_threads(os.cpu_count() or 1)
    run, kind = _reference_stylizer(args.arch)
    frames = synth.synthetic_frames_u8(1, args.height, args.width, seed=77)
    k_styles = 4 if args.config == "c5_region" else 1

    def one():
        outs = [run(frames) for _ in range(k_styles)]
        if k_styles > 1:
            from oracle import nst_oracle as O
            m = torch.full((4, 1, 1, args.height, args.width), 0.25)
            O.composite_regions([o[0].permute(2, 0, 1).float() / 255 for o in outs], list(m), [0, 1, 2, 3])

    for _ in range(warmup):
        one()
    t0 = time.perf_counter()
    for _ in range(n_frames):
        one()
    dt = time.perf_counter() - t0
    return n_frames / dt, torch.get_num_threads(), kind


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import torch
    t0 = time.perf_counter()
    fps, cores, kind = cpu_reference_fps(args, args.steps, warmup=1)
    dt = time.perf_counter() - t0
    sample = (f"{args.steps} steps x 1 frame {args.height}x{args.width} (a bounded sample of the {args.frames}-frame step), batch 1, fp32, "
              f"torch {torch.__version__} CPU, one process")
    line = {
        "impl": "reference", "metric": args.metric, "value": fps, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": 1, "ms_per_step": 1000.0 / fps, "higher_is_better": True, "scaling": args.scaling,
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"{args.config}: {args.arch} {args.height}x{args.width} frames, the reference's CPU torch path "
                               f"({'unmodified reference modules' if kind == 'reference' else 'oracle port of the reference modules: no reference checkout on this box'}), "
                               "synthetic seeded weights; ONE process on all host cores whatever --gpus says"},
        "cpu_baseline": {"value": fps, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": fps, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------------------------------------------
# B200 arm
# ---------------------------------------------------------------------------------------------------------------------------------
def run_b200(args):
    import torch
    import torch.distributed as dist
    import neuralstyletransferv1_b200 as nst
    from neuralstyletransferv1_b200 import synth

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", local))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    H, W = args.height, args.width
    strong = args.scaling == "strong"
    if strong:                                   # a fixed clip split contiguously over the ranks (sharding.shard_range)
        a, b = nst.shard_range(args.frames, world, rank)
        F, F_total = b - a, args.frames
    else:
        F, F_total = args.frames, args.frames * world
    cls = {"johnson": nst.TransformerNet, "nst": nst.TransformerNetNST, "reconet": nst.ReCoNet}[args.arch]
    region = args.config == "c5_region"
    n_models = 4 if region else 1
    models = []
    for k in range(n_models):
        m = cls(precision=args.precision).to(dev)
        m.load_state_dict(synth.make_state_dict(args.arch, k))
        m.eval()
        models.append(m)
    model = models[0]
    preset = PRESET[args.arch]

    # synthetic frames: >= 10 distinct image-like frames per rank (the 10th is pure noise, synth.synthetic_frames_u8), tiled to the batch
    uniq = synth.synthetic_frames_u8(min(max(F, 1), 10), H, W, seed=1234 + rank)
    host_in = torch.empty(max(F, 1), H, W, 3, dtype=torch.uint8).pin_memory()
    for i in range(F):
        host_in[i].copy_(uniq[i % uniq.shape[0]])
    host_in = host_in[:F]
    host_out = torch.empty_like(host_in).pin_memory()
    d_in = host_in.to(dev, non_blocking=True)
    d_out = torch.empty_like(d_in)
    # region_blend.blend_by_regions with the reference's defaults for this config: voronoi, 4 regions, seed 42, feather 20, models assigned
    # A, B, C, D in sequence (region_blend.py:925,1690; pipeline.py:1126-1127,1152) -- masks made on the device, kept for the clip
    blender = nst.RegionBlender(4, H, W, mode="voronoi", region_count=4, assignment="sequential", feather=20, seed=42, device=dev) if region else None
    styled = [torch.empty_like(d_in) for _ in range(n_models)] if region else None
    torch.cuda.synchronize()
    comp_ev = []

    def step(time_composite=False):
        if not region:
            if F:
                model.stylize_u8(d_in, preset, out=d_out)
            return
        for k in range(n_models):
            models[k].stylize_u8(d_in, preset, out=styled[k])
        if time_composite:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
        blender(styled, original=d_in, out=d_out)
        if time_composite:
            e1.record()
            comp_ev.append((e0, e1))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        step()
    barrier()
    launches0 = sum(m.engine.launch_count for m in models)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local) as clk:
        e0.record()
        for _ in range(args.steps):
            step(time_composite=region)
        e1.record()
        barrier()
    ms = e0.elapsed_time(e1)
    launches = sum(m.engine.launch_count for m in models) - launches0 + (args.steps if region else 0)
    t = torch.tensor([ms], device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total = float(t.item())
    value = F_total * args.steps / (ms_total / 1000.0)

    # ---- end to end through the package's batched caller with HOST buffers (frame_pipeline.FramePipeline) ----------------------
    pipes = [nst.FramePipeline(m, preset, chunk=8, slots=3) for m in models]
    host_styled = [torch.empty_like(host_in).pin_memory() for _ in range(n_models)] if region else None

    def e2e_step():
        if not F:
            return
        if not region:
            pipes[0](host_in, out=host_out)
            return
        # 4 styles through the pipelines, composite on the device, one download (the styled frames never visit the host)
        x = host_in.to(dev, non_blocking=True)
        for k in range(n_models):
            models[k].stylize_u8(x, preset, out=styled[k])
        blender(styled, original=x, out=d_out)
        host_out.copy_(d_out, non_blocking=True)
        torch.cuda.synchronize()

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
    e2e_value = F_total * args.steps / float(t.item())

    if rank != 0:
        if world > 1:
            dist.barrier(); dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel ------------------------------------------------------------------------------------------
    pk = peaks()
    prof_frames = d_in[:min(F, 8)]
    model.engine.profile_layers(prof_frames)
    prof = model.engine.profile_layers(prof_frames)
    nprof = prof_frames.shape[0]
    scale = (H * W) / (1080.0 * 1920.0)
    total_ms = sum(d["ms"] for d in prof)
    layers = [{"name": d["name"], "kernel": d["kernel"], "feed": d["feed"], "ms_per_frame": round(d["ms"] / nprof, 4),
               "tflops": round(d["flops"] * nprof / (d["ms"] / 1000.0) / 1e12, 1),
               "gbs": round(d["bytes"] * nprof / (d["ms"] / 1000.0) / 1e9, 1)} for d in prof]
    # whole-job conv throughput per GPU: algorithmic FLOPs of every stylisation / time / number of GPUs
    net_tflops_per_gpu = GFLOP_PER_1080P[args.arch] * scale * 1e9 * n_models * F_total * args.steps / (ms_total / 1000.0) / 1e12 / world
    if region:
        # C5: the composite is the kernel this config adds; it is HBM-bound: 4 styled u8 frames + original + 4 fp32 masks in, u8 out
        comp_ms = sorted(a.elapsed_time(b) for a, b in comp_ev)
        comp_ms = comp_ms[len(comp_ms) // 2]
        comp_bytes = F * ((n_models + 2) * H * W * 3 + 4 * H * W * 4)
        achieved = comp_bytes / (comp_ms / 1000.0) / 1e9
        roofline = {"bound": "hbm", "kernel": "composite_batch_kernel (one launch per step: 4 uint8 styled batches + originals + 4 fp32 region masks -> uint8)",
                    "achieved": round(achieved, 1), "peak": pk["hbm"], "peak_source": pk["src"] + " copy bandwidth", "unit": "GB/s",
                    "frac": round(achieved / pk["hbm"], 4), "traffic": None, "launch_ms": round(comp_ms, 4),
                    "algorithmic_bytes_per_launch": comp_bytes,
                    "share_of_step": round(comp_ms / (ms_total / args.steps), 4),
                    "conv_tflops_per_gpu": round(net_tflops_per_gpu, 2), "conv_frac_of_tensor_peak": round(net_tflops_per_gpu / pk["tflops"], 4)}
    else:
        dom = [d for d in prof if d["kernel"] == "tcgen05" and (".conv" in d["name"] or ".branch." in d["name"])] or prof
        dom_flops = sum(d["flops"] for d in dom) * nprof
        dom_ms = sum(d["ms"] for d in dom)
        achieved = dom_flops / (dom_ms / 1000.0) / 1e12
        traffic, traffic_src = traffic_from_profile()
        roofline = {"bound": "tensor", "kernel": "conv_tc_kernel (tcgen05 implicit GEMM), the residual-trunk 3x3 launches",
                    "achieved": round(achieved, 2), "peak": pk["tflops"], "peak_source": pk["src"] + " bf16 dense sustained (cuBLAS)",
                    "unit": "TFLOP/s", "frac": round(achieved / pk["tflops"], 4),
                    "traffic": traffic if (args.config == "c2" and args.arch == "johnson") else None,
                    "traffic_source": (traffic_src + " -- read from the committed profile, not measured by this run") if traffic else None,
                    "share_of_step": round(dom_ms / total_ms, 3), "launch_ms_avg": round(dom_ms / len(dom), 4),
                    "whole_net_tflops_per_gpu": round(net_tflops_per_gpu, 2), "whole_net_frac": round(net_tflops_per_gpu / pk["tflops"], 4)}

    cpu_baseline = None
    if not args.no_cpu_baseline and world == 1:
        fps, cores, kind = cpu_reference_fps(args, args.cpu_frames)
        cpu_baseline = {"value": fps, "unit": UNIT, "cores": cores, "kind": kind,
                        "sample": f"{args.cpu_frames} frames {H}x{W} batch 1 fp32 ({'the reference modules' if kind == 'reference' else 'oracle port of the reference torch CPU path'}"
                                  f"{', 4 styles + composite per frame' if region else ''}), 1 warm-up"}

    what = {"c2": "BASELINE configs[1]", "nst": "configs[1] workload, transformer_net_nst variant", "c3_4k": "BASELINE configs[2]",
            "c4_reconet300": "BASELINE configs[3]", "c5_region": "BASELINE configs[4]"}[args.config]
    line = {
        "metric": args.metric, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None,
        "dtype": "f16" if args.precision.startswith("f16") else args.precision, "data": "synthetic",
        "config": {"workload": f"{args.config} ({what}): {args.arch} generator{' x 4 weight sets + voronoi(4, seed 42, feather 20) region composite' if region else ''}, "
                               f"{F_total} x {H}x{W} uint8 frames per step over {world} GPU(s) ({F} on rank 0), synthetic seeded weights; precision={args.precision}: "
                               "fp16 tcgen05 operands and fp16 planar-8 activation storage, fp32 accumulate, double statistics",
                   "frames_per_step": F_total, "frames_per_step_rank0": F, "distinct_frames": int(uniq.shape[0]),
                   "l2": "inputs larger than L2 (rank-0 batch = %.0f MB uint8; activations ~320 MB per 1080p frame)" % (F * H * W * 3 / 1e6),
                   "parallelism": f"frame-batch sharding over {world} GPU(s), no collective ({'contiguous split of a fixed clip' if strong else 'fixed batch per GPU'})"},
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": F_total * H * W * 3, "d2h_bytes_per_step": F_total * H * W * 3,
                "api": "neuralstyletransferv1_b200.FramePipeline (pinned host in -> chunked H2D | stylize_u8 | D2H on three streams -> pinned host out)"
                       if not region else "host frames -> H2D -> 4 x stylize_u8 -> RegionBlender (composite_regions_batch) -> D2H"},
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
