"""Multi-GPU plumbing of the codec path: one process per GPU, frames of a sequence or tiles of one very
large frame are independent, so the only exchange is collecting byte streams on the rank that writes the file.

`gather_bytes` is that step: an all-gather of the lengths followed by an all-gather of the padded byte tensors over
torch.distributed (NCCL on the GPU box: the tensors live on the rank's device; gloo in the CPU tests).  Nothing here
touches pixels: the hot path has no collective."""
import numpy as np

from .frames import shard_frames, shard_tiles  # noqa: F401  (re-exported: the sharding rules live next to the frame generator)


def gather_bytes(blob, device="cpu", group=None):
    """Every rank passes its own byte string; every rank gets the list of all ranks' strings, in rank order."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group)
    n = torch.tensor([len(blob)], dtype=torch.int64, device=device)
    sizes = [torch.zeros(1, dtype=torch.int64, device=device) for _ in range(world)]
    dist.all_gather(sizes, n, group=group)
    sizes = [int(s.item()) for s in sizes]
    cap = max(max(sizes), 1)
    mine = torch.zeros(cap, dtype=torch.uint8, device=device)
    if blob:
        mine[:len(blob)] = torch.from_numpy(np.frombuffer(blob, dtype=np.uint8).copy()).to(device)
    parts = [torch.zeros(cap, dtype=torch.uint8, device=device) for _ in range(world)]
    dist.all_gather(parts, mine, group=group)
    return [bytes(p[:s].cpu().numpy().tobytes()) for p, s in zip(parts, sizes)]


def join_tile_streams(params, fragments):
    """fragments: tile-part byte strings of all ranks (any order) -> the frame's codestream (j2kcuda_join_tile_streams)"""
    import ctypes as C
    from . import _lib as L
    frags = [f for f in fragments if f]
    bufs = [C.create_string_buffer(f, len(f)) for f in frags]
    ptrs = (C.POINTER(C.c_uint8) * max(len(frags), 1))(*[C.cast(b, C.POINTER(C.c_uint8)) for b in bufs])
    lens = (C.c_size_t * max(len(frags), 1))(*[len(f) for f in frags])
    out = C.POINTER(C.c_uint8)()
    n = C.c_size_t(0)
    rc = L.lib().j2kcuda_join_tile_streams(C.byref(params), ptrs, lens, len(frags), C.byref(out), C.byref(n))
    if rc:
        raise RuntimeError("j2kcuda_join_tile_streams: " + L.last_error())
    data = C.string_at(out, n.value)
    L.lib().j2kcuda_free(out)
    return data
