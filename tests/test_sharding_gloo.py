"""World-size-2 gloo test (CPU) of the multi-GPU host logic: GOP-8 waves, frame assignment and the
reference-plane publish step (kvazaar_b200/shard.py)."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from kvazaar_b200 import shard


def test_gop8_waves_respect_dependencies():
    waves = shard.gop8_waves(2)
    seen = {0}
    coded = [p for w in waves for p in w]
    assert sorted(coded) == list(range(1, 17))
    for w in waves:
        for p in w:
            e = shard.GOP8[[g["poc"] for g in shard.GOP8].index((p - 1) % 8 + 1)]
            for r in e["refs"]:
                ref = p + r
                if 0 <= ref <= 16:
                    assert ref in seen, (p, ref)
        seen |= set(w)
    # the four layer-4 frames of a GOP are concurrent
    assert any({1, 3}.issubset(w) for w in waves) and any({5, 7}.issubset(w) for w in waves)
    for w in waves:
        parts = shard.assign_wave(w, 2)
        assert sorted(p for part in parts for p in part) == w


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        ok = True
        for step in range(4):
            src = step % world
            plane = torch.full((72, 128), 10 * (rank + 1) + step, dtype=torch.uint8)
            if rank == src:
                plane[3, 5] = 200
            shard.publish_reference(plane, src)
            want = 10 * (src + 1) + step
            ok &= bool((plane[0] == want).all()) and int(plane[3, 5]) == 200
        t = torch.tensor([int(ok)])
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
        if rank == 0:
            out.put(int(t.item()))
    finally:
        dist.destroy_process_group()


def test_publish_reference_world2():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert q.get(timeout=10) == 1
