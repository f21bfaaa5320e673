"""Multi-GPU plumbing of the codec path: one process per GPU, frames of a sequence or tiles of one very
large frame are independent, so the only exchange is collecting byte streams on the rank that writes the file.

`gather_bytes` is that step: an all-gather of the lengths followed by an all-gather of the padded byte tensors over
torch.distributed (NCCL on the GPU box: the tensors live on the rank's device; gloo in the CPU tests).  Nothing here
touches pixels: the hot path has no collective."""
import numpy as np

from .frames import shard_frames, shard_tiles  # noqa: F401  (re-exported: the sharding rules live next to the frame generator)


_staging = {}   # (device, capacity) -> (pinned send buffer, pinned receive buffer): page-locked once, reused per frame


def gather_bytes(blob, device="cpu", group=None, dst=None, as_arrays=False):
    """Every rank passes its own byte string; the ranks get the list of all ranks' strings, in rank order (only rank
    `dst` when given; the others get None).  Two collectives: the lengths, then the payload padded to the longest.
    On a CUDA device the payload goes through page-locked staging buffers that are kept between calls, and
    `as_arrays=True` hands out views of the receive buffer instead of copies (valid until the next call)."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    blob = memoryview(blob).cast("B") if not isinstance(blob, (bytes, bytearray)) else blob
    n = len(blob)
    sizes_t = torch.zeros(world, dtype=torch.int64, device=device)
    dist.all_gather_into_tensor(sizes_t, torch.tensor([n], dtype=torch.int64, device=device), group=group)
    sizes = [int(v) for v in sizes_t.tolist()]
    cap = (max(max(sizes), 1) + 4095) // 4096 * 4096
    cuda = str(device).startswith("cuda")
    if cuda:
        key = (str(device), cap)
        if key not in _staging:
            for k in [k for k in _staging if k[0] == key[0]]:
                del _staging[k]
            _staging[key] = (torch.empty(cap, dtype=torch.uint8).pin_memory(), torch.empty(world * cap, dtype=torch.uint8).pin_memory())
        send, recv = _staging[key]
        if n:
            send.numpy()[:n] = np.frombuffer(blob, dtype=np.uint8)
        mine = send.to(device, non_blocking=True)
    else:
        mine = torch.zeros(cap, dtype=torch.uint8)
        if n:
            mine[:n] = torch.from_numpy(np.frombuffer(blob, dtype=np.uint8).copy())
    allparts = torch.empty(world * cap, dtype=torch.uint8, device=device)
    dist.all_gather_into_tensor(allparts, mine, group=group)
    if dst is not None and rank != dst:
        return None
    if cuda:
        recv.copy_(allparts, non_blocking=True)
        torch.cuda.current_stream().synchronize()
        host = recv.numpy()
    else:
        host = allparts.numpy()
    parts = [host[i * cap:i * cap + sizes[i]] for i in range(world)]
    return parts if as_arrays else [p.tobytes() for p in parts]


def join_tile_streams(params, fragments, length_only=False):
    """fragments: tile-part byte strings of all ranks (any order) -> the frame's codestream (j2kcuda_join_tile_streams);
    length_only: join, then give the buffer straight back to the library (a writer would hand it to the file instead)"""
    import ctypes as C
    from . import _lib as L
    frags = [f for f in fragments if len(f)]
    # bytes are passed as they are, numpy arrays (gather_bytes(as_arrays=True)) by address: no copies on the way in
    keep = [f if isinstance(f, np.ndarray) else np.frombuffer(f, dtype=np.uint8) for f in frags]
    ptrs = (C.POINTER(C.c_uint8) * max(len(frags), 1))(*[C.cast(C.c_void_p(k.ctypes.data), C.POINTER(C.c_uint8)) for k in keep])
    lens = (C.c_size_t * max(len(frags), 1))(*[len(f) for f in frags])
    out = C.POINTER(C.c_uint8)()
    n = C.c_size_t(0)
    rc = L.lib().j2kcuda_join_tile_streams(C.byref(params), ptrs, lens, len(frags), C.byref(out), C.byref(n))
    if rc:
        raise RuntimeError("j2kcuda_join_tile_streams: " + L.last_error())
    if length_only:
        L.lib().j2kcuda_free(out)
        return n.value
    data = C.string_at(out, n.value)
    L.lib().j2kcuda_free(out)
    return data
