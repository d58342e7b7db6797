"""N>1 host path on CPU: two gloo ranks shard a frame list, 'stylise' their range with a stand-in
per-frame function and the assembled result equals the single-process one (no data-path collective:
the only communication is the gather used by the test itself and bench.py's timing reduction)."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _fake_stylize(frames):            # per-frame, order-sensitive stand-in
    return 255 - frames


def _worker(rank, world, port, n_frames, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from neuralstyletransferv1_b200 import shard_range, synth
    frames = synth.synthetic_frames_u8(n_frames, 8, 12, seed=9)
    a, b = shard_range(n_frames, world, rank)
    mine = _fake_stylize(frames[a:b])
    # bench.py's timing reduction: max over ranks
    t = torch.tensor([float(rank + 1)])
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    sizes = [torch.zeros(1, dtype=torch.int64) for _ in range(world)]
    dist.all_gather(sizes, torch.tensor([b - a]))
    parts = [torch.zeros(int(s), 8, 12, 3, dtype=torch.uint8) for s in sizes]
    dist.all_gather(parts, mine) if len({int(s) for s in sizes}) == 1 else None
    if rank == 0:
        q.put((float(t), [int(s) for s in sizes], mine.sum().item()))
    else:
        q.put((float(t), [int(s) for s in sizes], mine.sum().item()))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_frame_sharding():
    world, n_frames = 2, 7
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_frames, q)) for r in range(world)]
    for p in procs: p.start()
    got = [q.get(timeout=120) for _ in range(world)]
    for p in procs: p.join(60)
    assert all(p.exitcode == 0 for p in procs)
    assert all(g[0] == 2.0 for g in got)                 # max over ranks
    assert sorted(got[0][1], reverse=True) == [4, 3]     # contiguous split of 7 frames
    from neuralstyletransferv1_b200 import synth
    total = _fake_stylize(synth.synthetic_frames_u8(n_frames, 8, 12, seed=9)).sum().item()
    assert sum(g[2] for g in got) == total               # every frame stylised exactly once
