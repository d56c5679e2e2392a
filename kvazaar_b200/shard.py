"""Multi-GPU sharding of the motion-estimation stage: which frames of the `--gop 8`
pyramid can be searched at the same time, which rank takes which, and the one exchange step --
publishing a reconstructed reference plane to every rank.  Host logic only; runs on gloo (CPU tests)
and NCCL (bench) alike."""
import torch
import torch.distributed as dist

# cfg->gop[] of `--gop 8` (cfg.c:908-942): coding order, poc offset inside the GOP, temporal layer,
# reference offsets (negative = past, positive = future), is_ref
GOP8 = [
    {"poc": 8, "layer": 1, "refs": [-8, -12, -16], "is_ref": True},
    {"poc": 4, "layer": 2, "refs": [-4, -8, 4], "is_ref": True},
    {"poc": 2, "layer": 3, "refs": [-2, -6, 2, 6], "is_ref": True},
    {"poc": 1, "layer": 4, "refs": [-1, 1, 3, 7], "is_ref": False},
    {"poc": 3, "layer": 4, "refs": [-1, -3, 1, 5], "is_ref": False},
    {"poc": 6, "layer": 3, "refs": [-2, -6, 2], "is_ref": True},
    {"poc": 5, "layer": 4, "refs": [-1, -5, 1, 3], "is_ref": False},
    {"poc": 7, "layer": 4, "refs": [-1, -3, -7, 1], "is_ref": False},
]


def gop8_waves(n_gops=1):
    """Frames grouped into waves that may be searched concurrently: a frame is ready when every
    reference it uses (that exists) has been coded.  Returns a list of lists of absolute POCs."""
    frames = {}
    for g in range(n_gops):
        for e in GOP8:
            poc = g * 8 + e["poc"]
            frames[poc] = [poc + r for r in e["refs"]]
    done, waves = {0}, []
    todo = set(frames)
    while todo:
        ready = sorted(p for p in todo if all((r in done) or (r not in frames and r != 0 and r < 0) or (r > max(frames))
                                              for r in frames[p]))
        if not ready:
            raise RuntimeError("dependency cycle in GOP table")
        waves.append(ready)
        done |= set(ready)
        todo -= set(ready)
    return waves


def assign_wave(wave, world_size):
    """rank -> frames of this wave (round robin); every frame goes to exactly one rank."""
    return [wave[r::world_size] for r in range(world_size)]


def publish_reference(plane, src):
    """The exchange step: the rank that reconstructed a reference picture broadcasts its plane(s)
    (C2: 2.07 MB luma) to every rank that will search against it.  NCCL over NVLink on GPUs."""
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.broadcast(plane, src=src)
    return plane
