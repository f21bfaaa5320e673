"""Throughput of every BASELINE.json configuration (C1-C5) through the C ABI with pinned HOST buffers (end to end:
H2D, device stages, D2H, host Tier-2), one process per GPU.  Prints one JSON line per configuration.

  python tools/bench_configs.py [--reps 3] [--kind smooth]
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/bench_configs.py

C1-C3, C5: frames of a sequence, frame f -> rank f % N (no collective).  C4: the 40 tiles of each 8K frame are split in
contiguous ranges over the ranks (j2kcuda_params.tile_first / tile_count); the fragments are gathered to rank 0 with
torch.distributed and joined there (j2kcuda_join_tile_streams) -- the one collective of the design, outside the kernels.
Every decoded frame is compared with its source (lossless configs) before the numbers are printed."""
import argparse
import ctypes as C
import json
import os
import sys
import time
import numpy as np

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--kind", default="smooth")
    ap.add_argument("--workers", type=int, default=8)
    ap.add_argument("--only", default="")
    ap.add_argument("--frames8k", type=int, default=4)
    ap.add_argument("--devices", type=int, default=0, help="C4api / C5api: GPUs driven from ONE process through the C ABI (0 = all visible)")
    args = ap.parse_args()
    import torch
    import torch.distributed as dist
    import j2k
    from j2k_b200 import _lib as L
    from j2k_b200.frames import make_frame, shard_frames, shard_tiles
    from j2k_b200.sharding import gather_bytes, join_tile_streams

    rank = int(os.environ.get("RANK", 0)); world = int(os.environ.get("WORLD_SIZE", 1)); local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    lib = L.lib()
    lib.j2kcuda_sequence_create.restype = C.c_void_p

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    def maxtime(dt):
        if world == 1:
            return dt
        t = torch.tensor([dt], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def pinned(frames):
        out = []
        for f in frames:
            t = torch.empty(f.shape, dtype=torch.uint8 if f.dtype == np.uint8 else torch.uint16).pin_memory()
            t.numpy()[...] = f
            out.append(t)
        return out

    def table(tensors):
        keep = [j2k.planes_of(t.numpy()) for t in tensors]
        return (C.POINTER(L.Plane) * len(keep))(*[C.cast(k, C.POINTER(L.Plane)) for k in keep]), keep

    def run_sequence(name, h, w, nc, prec, frames_per_rank, lossless, **kw):
        """frames sharded over ranks; returns encode / decode MPix/s of the whole job"""
        dt = np.uint8 if prec <= 8 else np.uint16
        mine = shard_frames(frames_per_rank * world, rank, world)
        src = pinned([make_frame(h, w, nc, prec, args.kind, seed=f).astype(dt) for f in mine])
        dst = pinned([np.zeros((h, w, nc), dt) for _ in mine])
        F = len(mine)
        p = L.default_params(width=w, height=h, ncomp=nc, prec=prec, **kw)
        seq = C.c_void_p(lib.j2kcuda_sequence_create(C.byref(p), local, args.workers))
        tin, k1 = table(src); tout, k2 = table(dst)
        outs = (C.POINTER(C.c_uint8) * F)(); lens = (C.c_size_t * F)()
        te = td = 0.0
        for rep in range(args.reps + 1):
            barrier(); t0 = time.perf_counter()
            assert lib.j2kcuda_sequence_encode(seq, tin, F, nc, 0, outs, lens) == 0, L.last_error()
            barrier(); t1 = time.perf_counter()
            assert lib.j2kcuda_sequence_decode(seq, C.cast(outs, C.POINTER(C.POINTER(C.c_uint8))), lens, F, 0, tout, nc, 0) == 0, L.last_error()
            barrier(); t2 = time.perf_counter()
            if rep:                                   # the first pass creates the contexts
                te += maxtime(t1 - t0); td += maxtime(t2 - t1)
            nbytes = sum(lens[i] for i in range(F))
            if rep < args.reps:
                for i in range(F):
                    lib.j2kcuda_free(outs[i])
        ok = True
        if lossless:
            ok = all(bool((a == b).all()) for a, b in zip(src, dst))
        else:
            worst = max(float((a.to(torch.int32) - b.to(torch.int32)).abs().max()) for a, b in zip(src, dst))
            ok = worst < (1 << prec) / 8              # rate-capped: sanity only, parity is tests/test_gpu_configs.py
        for i in range(F):
            lib.j2kcuda_free(outs[i])
        lib.j2kcuda_sequence_destroy(seq)
        mp = args.reps * F * world * w * h / 1e6
        return {"config": name, "n_gpus": world, "frames_per_step": F * world, "encode_mpix_s": mp / te, "decode_mpix_s": mp / td,
                "roundtrip_mpix_s": mp / (te + td), "codestream_bytes_per_frame": nbytes / F, "verified": ok,
                "sharding": "frames over ranks, no collective"}

    def run_tiles(name, h, w, nc, prec, tile, nframes):
        """C4: tiles of one frame sharded over the ranks, fragments gathered and joined on rank 0"""
        frames = [make_frame(h, w, nc, prec, args.kind, seed=100 + f) for f in range(nframes)]
        src = pinned(frames)
        tx, ty = -(-w // tile), -(-h // tile)
        first, last = shard_tiles(tx * ty, rank, world)
        count = last - first
        p = L.default_params(width=w, height=h, ncomp=nc, prec=prec, mct=1, tile_w=tile, tile_h=tile, tile_first=first, tile_count=count if world > 1 else 0)
        whole = L.default_params(width=w, height=h, ncomp=nc, prec=prec, mct=1, tile_w=tile, tile_h=tile)
        lib.j2kcuda_encoder_create.restype = C.c_void_p
        enc = C.c_void_p(lib.j2kcuda_encoder_create(C.byref(p)))
        assert enc.value, L.last_error()
        te = 0.0
        streams = []
        phases = [0.0, 0.0, 0.0, 0.0]
        import queue
        import threading

        def exchange(q, rep, streams):
            """gather + join of finished fragments, on its own thread: the exchange of frame f overlaps the encode of f+1"""
            torch.cuda.set_device(local)
            while True:
                item = q.get()
                if item is None:
                    return
                out, n = item
                tc = time.perf_counter()
                frag = np.ctypeslib.as_array(out, shape=(n,))                          # the library's buffer, not a copy
                parts = gather_bytes(frag, device="cuda", dst=0, as_arrays=True)       # the one exchange step: tile bitstreams to rank 0
                td = time.perf_counter()
                if rank == 0:                                                          # a Python copy of the file only for the final check
                    streams.append(join_tile_streams(whole, parts, length_only=rep < args.reps))
                lib.j2kcuda_free(out)
                phases[2] = td - tc; phases[3] = time.perf_counter() - td

        for rep in range(args.reps + 1):
            streams = []
            q = queue.Queue()
            th = None
            barrier(); t0 = time.perf_counter()
            if world > 1:
                th = threading.Thread(target=exchange, args=(q, rep, streams)); th.start()
            for t in src:
                pl = j2k.planes_of(t.numpy())
                out = C.POINTER(C.c_uint8)(); n = C.c_size_t(0)
                ta = time.perf_counter()
                assert lib.j2kcuda_encoder_upload(enc, pl, nc) == 0, L.last_error()
                assert lib.j2kcuda_encoder_run(enc) == 0, L.last_error()
                tb = time.perf_counter()
                assert lib.j2kcuda_encoder_finish(enc, C.byref(out), C.byref(n)) == 0, L.last_error()
                phases[0] = tb - ta; phases[1] = time.perf_counter() - tb
                if world > 1:
                    q.put((out, n.value))
                else:
                    streams.append(C.string_at(out, n.value) if rep == args.reps else n.value)
                    lib.j2kcuda_free(out)
            if th:
                q.put(None); th.join()
            barrier(); t1 = time.perf_counter()
            if rep:
                te += maxtime(t1 - t0)
        lib.j2kcuda_encoder_destroy(enc)
        ok = True
        if rank == 0:
            ok = bool(np.array_equal(j2k.decode(streams[0]), frames[0])) and bool(np.array_equal(j2k.decode(streams[-1]), frames[-1]))
        mp = args.reps * nframes * w * h / 1e6
        return {"config": name, "n_gpus": world, "frames_per_step": nframes, "encode_mpix_s": mp / te, "verified": ok,
                "codestream_bytes_per_frame": len(streams[0]) if streams else None,
                "rank0_last_frame_ms": {"upload_run": phases[0] * 1e3, "host_tier2": phases[1] * 1e3, "gather": phases[2] * 1e3, "join": phases[3] * 1e3},
                "sharding": "tiles of each frame over ranks (contiguous ranges), fragments gathered to rank 0 and joined while the next frame is encoded"}

    def run_tiles_api(name, h, w, nc, prec, tile, nframes):
        """C4 through j2kcuda_tiles_* (api_multi.cpp): ONE process, tiles of each frame over `--devices` GPUs, encode AND decode"""
        ndev = args.devices or torch.cuda.device_count()
        frames = [make_frame(h, w, nc, prec, args.kind, seed=100 + f) for f in range(nframes)]
        src = pinned(frames)
        dst = pinned([np.zeros_like(f) for f in frames])
        p = L.default_params(width=w, height=h, ncomp=nc, prec=prec, mct=1, tile_w=tile, tile_h=tile)
        devs = (C.c_int * ndev)(*range(ndev))
        lib.j2kcuda_tiles_create.restype = C.c_void_p
        t = C.c_void_p(lib.j2kcuda_tiles_create(C.byref(p), devs, ndev))
        assert t.value, L.last_error()
        te = td = 0.0
        nbytes = 0
        for rep in range(args.reps + 1):
            outs = []
            t0 = time.perf_counter()
            for f in src:
                out = C.POINTER(C.c_uint8)(); n = C.c_size_t(0)
                assert lib.j2kcuda_tiles_encode(t, j2k.planes_of(f.numpy()), nc, C.byref(out), C.byref(n)) == 0, L.last_error()
                outs.append((out, n.value))
            t1 = time.perf_counter()
            for (out, n), g in zip(outs, dst):
                assert lib.j2kcuda_tiles_decode(t, out, C.c_size_t(n), 0, j2k.planes_of(g.numpy()), nc) == 0, L.last_error()
            t2 = time.perf_counter()
            if rep:
                te += t1 - t0; td += t2 - t1
            nbytes = outs[0][1]
            for out, n in outs:
                lib.j2kcuda_free(out)
        lib.j2kcuda_tiles_destroy(t)
        ok = all(bool((a == b).all()) for a, b in zip(src, dst))
        mp = args.reps * nframes * w * h / 1e6
        return {"config": name, "n_gpus": ndev, "frames_per_step": nframes, "encode_mpix_s": mp / te, "decode_mpix_s": mp / td,
                "roundtrip_mpix_s": mp / (te + td), "codestream_bytes_per_frame": nbytes, "verified": ok,
                "sharding": "one process, j2kcuda_tiles_*: contiguous tile runs per device, one host thread per device, joined on the host"}

    def run_sequence_api(name, h, w, nc, prec, nframes, **kw):
        """C5 through j2kcuda_sequence_create_multi: ONE process, frames over `--devices` GPUs"""
        ndev = args.devices or torch.cuda.device_count()
        dt = np.uint8 if prec <= 8 else np.uint16
        F = nframes * ndev
        src = pinned([make_frame(h, w, nc, prec, args.kind, seed=f).astype(dt) for f in range(F)])
        dst = pinned([np.zeros((h, w, nc), dt) for _ in range(F)])
        p = L.default_params(width=w, height=h, ncomp=nc, prec=prec, **kw)
        devs = (C.c_int * ndev)(*range(ndev))
        lib.j2kcuda_sequence_create_multi.restype = C.c_void_p
        cores = os.cpu_count() or 8
        wpd = max(4, min(args.workers, cores // ndev))
        seq = C.c_void_p(lib.j2kcuda_sequence_create_multi(C.byref(p), devs, ndev, wpd))
        assert seq.value, L.last_error()
        tin, k1 = table(src); tout, k2 = table(dst)
        outs = (C.POINTER(C.c_uint8) * F)(); lens = (C.c_size_t * F)()
        te = td = 0.0
        for rep in range(args.reps + 1):
            t0 = time.perf_counter()
            assert lib.j2kcuda_sequence_encode(seq, tin, F, nc, 0, outs, lens) == 0, L.last_error()
            t1 = time.perf_counter()
            assert lib.j2kcuda_sequence_decode(seq, C.cast(outs, C.POINTER(C.POINTER(C.c_uint8))), lens, F, 0, tout, nc, 0) == 0, L.last_error()
            t2 = time.perf_counter()
            if rep:
                te += t1 - t0; td += t2 - t1
            for i in range(F):
                lib.j2kcuda_free(outs[i])
        lib.j2kcuda_sequence_destroy(seq)
        ok = all(bool((a == b).all()) for a, b in zip(src, dst))
        mp = args.reps * F * w * h / 1e6
        return {"config": name, "n_gpus": ndev, "frames_per_step": F, "workers_per_device": wpd, "encode_mpix_s": mp / te, "decode_mpix_s": mp / td,
                "roundtrip_mpix_s": mp / (te + td), "verified": ok,
                "sharding": "one process, j2kcuda_sequence_create_multi: frames to whichever device has a free worker"}

    jobs = [
        ("C4api", lambda: run_tiles_api("C4 8192x4320 16-bit RGB lossless, 1024x1024 tiles, j2kcuda_tiles_* in one process", 4320, 8192, 3, 16, 1024, args.frames8k)),
        ("C5api", lambda: run_sequence_api("C5 4096x2160 16-bit RGB lossless sequence, j2kcuda_sequence_create_multi in one process", 2160, 4096, 3, 16, 8, mct=1)),
        ("C1", lambda: run_sequence("C1 2048x1080 8-bit RGB lossless 5/3", 1080, 2048, 3, 8, 16, True, mct=1)),
        ("C2", lambda: run_sequence("C2 4096x2160 16-bit RGBA lossless 5/3", 2160, 4096, 4, 16, 8, True, mct=1)),
        ("C3", lambda: run_sequence("C3 DCI 4K 12-bit XYZ 9/7 ICT cinema profile (PCRD rate cap)", 2160, 4096, 3, 12, 8, False,
                                    irreversible=1, mct=1, profile=L.PROFILE_CINEMA_4K)),
        ("C4", lambda: run_tiles("C4 8192x4320 16-bit RGB lossless, 1024x1024 tiles", 4320, 8192, 3, 16, 1024, args.frames8k)),
        ("C5", lambda: run_sequence("C5 4096x2160 16-bit RGB lossless sequence", 2160, 4096, 3, 16, 16, True, mct=1)),
    ]
    for key, fn in jobs:
        if (args.only and key not in args.only.split(",")) or (not args.only and key.endswith("api")) or (key.endswith("api") and world > 1):
            continue
        r = fn()
        if rank == 0:
            print(json.dumps(r), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
