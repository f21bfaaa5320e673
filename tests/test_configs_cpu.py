"""CPU-only checks of the BASELINE configurations' host side, with the oracle's stage functions standing in for the
kernels (tests/hostpipe.py) and the compiled reference as the arbiter:
  C3  9/7 + ICT, PCRD rate caps, cinema 2K/4K profiles (TLM, POC, tile-parts by component): byte-identical streams
  C4  tiles of one frame sharded over ranks: fragments + j2kcuda_join_tile_streams == the one-context stream == reference,
      with the gather done over torch.distributed (gloo, world_size 2)
  C5  frames of a sequence sharded over ranks: every frame encoded exactly once, gathered in order"""
import os
import sys
import numpy as np
import pytest
import hostpipe
import refgrok
from j2k_b200 import _lib as L
from j2k_b200.frames import make_frame, shard_frames, shard_tiles
from j2k_b200.sharding import join_tile_streams

needs_ref = pytest.mark.skipif(not refgrok.available(), reason="oracle/_ref not built")


def _planes(img):
    return [img[..., c].astype(np.int32) for c in range(img.shape[2])]


@needs_ref
@pytest.mark.parametrize("h,w,prec,kind,ref_kw,our_kw", [
    (96, 128, 8, "smooth", dict(rates=[10.0]), dict(rates=[10.0])),
    (130, 200, 12, "noise", dict(rates=[8.0]), dict(rates=[8.0])),
    (100, 160, 12, "smooth", dict(rates=[40.0, 12.0, 4.0], numlayers=3), dict(rates=[40.0, 12.0, 4.0], numlayers=3)),
    (135, 256, 12, "smooth", dict(rsiz=4), dict(profile=L.PROFILE_CINEMA_4K)),
    (135, 256, 12, "noise", dict(rsiz=4, max_cs_size=20000, max_comp_size=16000),
     dict(profile=L.PROFILE_CINEMA_4K, max_cs_size=20000, max_comp_size=16000)),
    (120, 256, 12, "smooth", dict(rsiz=3), dict(profile=L.PROFILE_CINEMA_2K)),
])
def test_c3_irreversible_rate_capped_streams_identical_to_reference(h, w, prec, kind, ref_kw, our_kw):
    img = make_frame(h, w, 3, prec, kind, seed=3)
    planes = _planes(img)
    ref = refgrok.encode(planes, prec, irreversible=1, mct=1, **ref_kw)
    p = L.default_params(width=w, height=h, ncomp=3, prec=prec, irreversible=1, mct=1, **our_kw)
    ours = hostpipe.encode(planes, prec, p)
    assert len(ours) == len(ref)
    assert ours == ref


def test_sharding_rules():
    for n, world in [(240, 8), (7, 2), (3, 4), (40, 8), (0, 2)]:
        got = sorted(f for r in range(world) for f in shard_frames(n, r, world))
        assert got == list(range(n))
        runs = [shard_tiles(n, r, world) for r in range(world)]
        assert runs[0][0] == 0 and runs[-1][1] == n
        for a, b in zip(runs, runs[1:]):
            assert a[1] == b[0]
        assert max(b - a for a, b in runs) - min(b - a for a, b in runs) <= 1
    with pytest.raises(ValueError):
        shard_frames(4, 2, 2)


@needs_ref
@pytest.mark.parametrize("world", [2, 3])
def test_c4_tile_fragments_join_to_the_reference_stream(world):
    h, w, prec, tile = 150, 200, 16, 64            # 4 x 3 tiles, ragged right and bottom edges
    img = make_frame(h, w, 3, prec, "smooth", seed=4)
    planes = _planes(img)
    whole = L.default_params(width=w, height=h, ncomp=3, prec=prec, mct=1, numresolution=4, tile_w=tile, tile_h=tile)
    ref = refgrok.encode(planes, prec, mct=1, numresolution=4, tile=(tile, tile))
    assert hostpipe.encode(planes, prec, whole) == ref
    frags = []
    for r in range(world):
        first, last = shard_tiles(12, r, world)
        p = L.default_params(width=w, height=h, ncomp=3, prec=prec, mct=1, numresolution=4, tile_w=tile, tile_h=tile,
                             tile_first=first, tile_count=last - first)
        frags.append(hostpipe.encode(planes, prec, p))
        assert frags[-1][:2] == b"\xff\x90"           # a fragment starts at its first SOT
    assert join_tile_streams(whole, frags[::-1]) == ref
    with pytest.raises(RuntimeError):
        join_tile_streams(whole, frags[:-1])           # a missing tile is an error, not a short file
    with pytest.raises(RuntimeError):
        join_tile_streams(whole, frags + frags[:1])


def _rank_main(rank, world, port, q):
    import torch.distributed as dist
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    from j2k_b200.sharding import gather_bytes
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    try:
        # C4: this rank's tiles of one frame
        h, w, prec, tile = 100, 150, 8, 64
        img = make_frame(h, w, 3, prec, "smooth", seed=9)
        planes = _planes(img)
        first, last = shard_tiles(6, rank, world)
        p = L.default_params(width=w, height=h, ncomp=3, prec=prec, mct=1, numresolution=3, tile_w=tile, tile_h=tile,
                             tile_first=first, tile_count=last - first)
        frags = gather_bytes(hostpipe.encode(planes, prec, p))
        # C5: this rank's frames of a sequence
        nframes = 5
        mine = {}
        for f in shard_frames(nframes, rank, world):
            fr = make_frame(48, 64, 1, 8, "ramp", seed=f)
            mine[f] = hostpipe.encode(_planes(fr), 8, L.default_params(width=64, height=48, ncomp=1, prec=8, numresolution=3))
        blob = b"".join(f.to_bytes(4, "little") + len(s).to_bytes(4, "little") + s for f, s in sorted(mine.items()))
        seqs = gather_bytes(blob)
        if rank == 0:
            whole = L.default_params(width=w, height=h, ncomp=3, prec=prec, mct=1, numresolution=3, tile_w=tile, tile_h=tile)
            joined = join_tile_streams(whole, frags)
            frames = {}
            for b in seqs:
                pos = 0
                while pos < len(b):
                    f = int.from_bytes(b[pos:pos + 4], "little"); n = int.from_bytes(b[pos + 4:pos + 8], "little")
                    frames[f] = b[pos + 8:pos + 8 + n]; pos += 8 + n
            q.put((joined, frames))
        dist.barrier()
    finally:
        dist.destroy_process_group()


@needs_ref
def test_world_size_2_gloo_gathers_tile_fragments_and_sequence_frames():
    import torch.multiprocessing as mp
    import socket
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_rank_main, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    joined, frames = q.get(timeout=240)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    h, w, prec, tile = 100, 150, 8, 64
    img = make_frame(h, w, 3, prec, "smooth", seed=9)
    assert joined == refgrok.encode(_planes(img), prec, mct=1, numresolution=3, tile=(tile, tile))
    assert sorted(frames) == list(range(5))
    for f, sbytes in frames.items():
        fr = make_frame(48, 64, 1, 8, "ramp", seed=f)
        assert sbytes == refgrok.encode(_planes(fr), 8, numresolution=3)
