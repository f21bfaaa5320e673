"""Deterministic synthetic frames: noise, smooth, ramp.  Used by bench.py and the tests."""
import numpy as np


def make_frame(h, w, ncomp, prec, kind="smooth", seed=0):
    """-> (h, w, ncomp) uint8/uint16 array with values < 2**prec"""
    dt = np.uint8 if prec <= 8 else np.uint16
    top = (1 << prec) - 1
    rng = np.random.default_rng(seed)
    if kind == "noise":
        return rng.integers(0, top + 1, size=(h, w, ncomp), dtype=np.uint32).astype(dt)
    if kind == "ramp":
        y, x = np.mgrid[0:h, 0:w]
        return np.stack([((3 * x + 5 * y + 17 * c) % (top + 1)) for c in range(ncomp)], axis=-1).astype(dt)
    if kind == "smooth":
        y = np.arange(h, dtype=np.float32)[:, None] / max(h, 1)
        x = np.arange(w, dtype=np.float32)[None, :] / max(w, 1)
        out = np.empty((h, w, ncomp), dtype=dt)
        for c in range(ncomp):
            f = np.zeros((h, w), dtype=np.float32)
            for k in range(4):
                fx, fy = rng.uniform(0.5, 6.0, size=2)
                ph = rng.uniform(0, 2 * np.pi)
                f += np.cos(2 * np.pi * (fx * x + fy * y) + ph).astype(np.float32) / (k + 1)
            f = (f - f.min()) / max(float(f.max() - f.min()), 1e-6)
            v = f * (top * 0.9) + rng.integers(0, 4, size=(h, w))
            out[..., c] = np.clip(v, 0, top).astype(dt)
        return out
    raise ValueError(kind)


def shard_frames(nframes, rank, world):
    """Frames of a sequence are independent: frame f goes to rank f % world.  -> this rank's frame indices"""
    if world < 1 or not (0 <= rank < world):
        raise ValueError("bad rank/world")
    return list(range(rank, nframes, world))


def shard_tiles(ntiles, rank, world):
    """Tiles of one large frame never read across a tile edge (tcd.cpp:508-526): contiguous runs of tiles per rank,
    so each rank's tile-parts concatenate in codestream order.  -> (first, last+1)"""
    if world < 1 or not (0 <= rank < world):
        raise ValueError("bad rank/world")
    base, extra = divmod(ntiles, world)
    first = rank * base + min(rank, extra)
    return first, first + base + (1 if rank < extra else 0)
