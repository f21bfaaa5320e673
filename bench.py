#!/usr/bin/env python
"""bench.py -- 4K lossless JPEG 2000 throughput of libj2kcuda on B200 (BASELINE.json metric).

A step = one pass of the hot path over one 4096x2160 16-bit RGBA frame (BASELINE configs[1]):
lossless 5/3, RCT on components 0-2, 6 resolutions, 64x64 code-blocks, LRCP, one layer, single tile.
  value  MPix/s with the frame already resident in HBM (device stages + result D2H + host Tier-2)
  e2e    MPix/s through the reference-facing C-ABI calls with pinned HOST buffers (H2D and D2H inside)
  roofline  the DWT level-1 kernel: algorithmic bytes (8 B per sample per component) / its CUDA-event time
  cpu_baseline  the reference's own CPU implementation (ext/grok compiled into oracle/_ref) on one frame
`--impl reference` times that CPU implementation instead (rank 0 only).
One process per GPU; frames are sharded across ranks with no collective on the data path (weak scaling).
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

W, H, NC, PREC = 4096, 2160, 4, 16
WORKLOAD = "4096x2160 16-bit RGBA lossless 5/3 (RCT, 6 resolutions, 64x64 code-blocks, 1 tile)"
NROT = 4   # distinct frames rotated through: 4 x 70.8 MB input + 283 MB of planes per frame > 126 MB L2


def peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"], "measured"
    except Exception:
        return 6650.0, "fallback"


class ClockSampler(threading.Thread):
    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop = index, [], False

    def run(self):
        q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
            "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
        while not self.stop:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + q, "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            time.sleep(0.2)

    def summary(self):
        sm = sorted(int(r[0]) for r in self.rows if r[0].isdigit())
        mx = [int(r[1]) for r in self.rows if r[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 2 + i and r[2 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(self.rows)}


def reference_encode_fn():
    """the reference's CPU implementation of the path: oracle/_ref (ext/grok compiled from its own sources)"""
    import refgrok
    if refgrok.available():
        ncores = os.cpu_count() or 1
        def roundtrip(planes):
            cs = refgrok.encode(planes, PREC, mct=1, numresolution=6, threads=ncores)
            refgrok.decode(cs, threads=ncores)
            return cs
        return roundtrip, "reference", ncores
    import hostpipe                       # oracle port: scalar, single thread
    from j2k_b200 import _lib as L

    def enc(planes):
        h, w = planes[0].shape
        p = L.default_params(width=w, height=h, ncomp=len(planes), prec=PREC, mct=1)
        return hostpipe.encode(planes, PREC, p)
    return enc, "port", 1


def run_reference(args, rank):
    if rank != 0:
        return
    from j2k_b200.frames import make_frame
    enc, kind, cores = reference_encode_fn()
    # bounded sample: a port (scalar Python-driven oracle) gets a 1/16 crop, the compiled reference a full frame
    h, w = (H, W) if kind == "reference" else (H // 4, W // 4)
    frames = [make_frame(h, w, NC, PREC, args.kind, seed=s) for s in range(2)]
    planes = [[f[..., c].astype(np.int32) for c in range(NC)] for f in frames]
    for i in range(min(args.warmup, 1)):
        enc(planes[i % 2])
    t0 = time.perf_counter()
    for i in range(args.steps):
        enc(planes[i % 2])
    dt = time.perf_counter() - t0
    v = args.steps * w * h / dt / 1e6
    sample = "%d frame(s) of %dx%d 16-bit RGBA per step, ext/grok CPU encoder + decoder, %d threads" % (1, w, h, cores)
    print(json.dumps({
        "impl": "reference", "metric": "4K lossless encode+decode MPix/s", "value": v, "unit": "MPix/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "int32", "data": "synthetic (%s)" % args.kind,
        "config": {"workload": WORKLOAD, "direction": "encode then decode of the same frame"},
        "cpu_baseline": {"value": v, "unit": "MPix/s", "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": v, "unit": "MPix/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--kind", default="smooth", choices=["smooth", "noise", "ramp"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank)
        return

    import torch
    import torch.distributed as dist
    from j2k_b200 import _lib as L
    from j2k_b200.frames import make_frame
    import j2k
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: libj2kcuda has no CPU path")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    lib = L.lib()
    lib.j2kcuda_encoder_create.restype = C.c_void_p
    p = L.default_params(width=W, height=H, ncomp=NC, prec=PREC, mct=1, numresolution=6, device=local)
    enc = C.c_void_p(lib.j2kcuda_encoder_create(C.byref(p)))
    if not enc.value:
        raise SystemExit("encoder_create failed: " + L.last_error())

    # this rank's frames: pinned host copies (e2e arm) and device copies (kernel arm)
    host, dev, planes = [], [], []
    for i in range(NROT):
        f = make_frame(H, W, NC, PREC, args.kind, seed=rank * NROT + i)
        t = torch.empty((H, W, NC), dtype=torch.uint16).pin_memory()
        t.numpy()[...] = f
        host.append(t)
        dev.append(t.cuda())
        planes.append(j2k.planes_of(t.numpy()))
    torch.cuda.synchronize()
    out = C.POINTER(C.c_uint8)()
    n = C.c_size_t(0)
    ms = (C.c_float * 8)()
    cnt = (C.c_uint64 * 6)()

    lib.j2kcuda_decoder_create.restype = C.c_void_p
    dec = C.c_void_p(lib.j2kcuda_decoder_create(local))
    if not dec.value:
        raise SystemExit("decoder_create failed: " + L.last_error())
    back = torch.empty((H, W, NC), dtype=torch.uint16).pin_memory()       # decoded frame lands here (e2e arm)
    back_planes = j2k.planes_of(back.numpy())
    dms = (C.c_float * 8)()
    tdec = [0.0]

    def step_device(i):
        """encode a frame that is resident in HBM, decode the codestream, leave the pixels in HBM"""
        d = dev[i % NROT]
        rc = lib.j2kcuda_encoder_set_device_input(enc, C.c_void_p(d.data_ptr()), C.c_int64(W * NC * 2), L.USHORT)
        rc = rc or lib.j2kcuda_encoder_run(enc)
        rc = rc or lib.j2kcuda_encoder_finish(enc, C.byref(out), C.byref(n))
        t0 = time.perf_counter()
        rc = rc or lib.j2kcuda_decoder_prepare(dec, out, n, 0)
        rc = rc or lib.j2kcuda_decoder_run(dec)
        tdec[0] += time.perf_counter() - t0
        if rc:
            raise SystemExit("step failed: " + L.last_error())
        lib.j2kcuda_free(out)
        return n.value

    def step_host(i):
        """the same through the reference-facing calls with pinned host buffers on both sides"""
        rc = lib.j2kcuda_encoder_upload(enc, planes[i % NROT], NC)
        rc = rc or lib.j2kcuda_encoder_run(enc)
        rc = rc or lib.j2kcuda_encoder_finish(enc, C.byref(out), C.byref(n))
        t0 = time.perf_counter()
        rc = rc or lib.j2kcuda_decoder_prepare(dec, out, n, 0)
        rc = rc or lib.j2kcuda_decoder_run(dec)
        rc = rc or lib.j2kcuda_decoder_download(dec, back_planes, NC)
        tdec[0] += time.perf_counter() - t0
        if rc:
            raise SystemExit("step failed: " + L.last_error())
        lib.j2kcuda_free(out)
        return n.value

    def timed(step):
        for i in range(args.warmup):
            step(i)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        acc = np.zeros(8)
        dacc = np.zeros(7)
        tdec[0] = 0.0
        launches = nbytes = d2h = 0
        torch.cuda.synchronize()
        e0.record()
        for i in range(args.steps):
            nbytes += step(i)
            lib.j2kcuda_encoder_timings(enc, ms, 8)
            lib.j2kcuda_encoder_counters(enc, cnt, 6)
            acc += np.array(list(ms))
            launches += cnt[0]
            d2h = cnt[5]
            lib.j2kcuda_decoder_timings(dec, dms, 7)
            lib.j2kcuda_decoder_counters(dec, cnt, 6)
            dacc += np.array(list(dms)[:7])
            launches += cnt[0]
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1)], device="cuda")
        if world > 1:
            dist.barrier()
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item()), acc / args.steps, launches, nbytes / args.steps, d2h, dacc / args.steps, tdec[0] / args.steps * 1e3

    sampler = ClockSampler(local)
    sampler.start()
    t_dev, st_dev, launches, cs_bytes, d2h_dev, dst_dev, dec_ms_dev = timed(step_device)
    t_e2e, st_e2e, _, _, d2h_e2e, dst_e2e, dec_ms_e2e = timed(step_host)
    if not np.array_equal(back.numpy(), host[(args.steps - 1) % NROT].numpy()):
        raise SystemExit("round trip is not lossless")
    sampler.stop = True
    sampler.join(timeout=3)

    lib.j2kcuda_encoder_counters(enc, cnt, 6)
    nblocks, nsym = int(cnt[1]), int(cnt[2])
    lib.j2kcuda_encoder_destroy(enc)
    lib.j2kcuda_decoder_destroy(dec)
    if rank != 0:
        return
    mpix = W * H / 1e6
    value = world * args.steps * mpix / (t_dev * 1e-3)
    e2e = world * args.steps * mpix / (t_e2e * 1e-3)
    peak, peak_src = peaks()
    l1_bytes = 8.0 * W * H * NC                      # 4 B read + 4 B written per sample, level 1 of every component
    l1_ms = float(st_dev[7])
    achieved = l1_bytes / (l1_ms * 1e-3) / 1e9
    line = {
        "metric": "4K lossless encode+decode MPix/s", "value": value, "unit": "MPix/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": t_dev / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "int32", "data": "synthetic (%s)" % args.kind,
        "config": {"workload": WORKLOAD, "direction": "encode then decode of the same frame (round trip verified lossless)", "frames_per_step_per_gpu": 1,
                   "l2": "inputs and planes larger than L2: %d rotating frames" % NROT,
                   "codestream_bytes_per_frame": cs_bytes, "code_blocks_per_frame": nblocks, "mq_symbols_per_frame": nsym},
        "e2e": {"value": e2e, "unit": "MPix/s", "h2d_bytes_per_step": int(W * H * NC * 2 + cs_bytes), "d2h_bytes_per_step": int(d2h_e2e + W * H * NC * 2),
                "ms_per_step": t_e2e / args.steps},
        "gpu_launches": int(launches),
        "roofline": {"bound": "hbm", "kernel": "k_dwt_forward<Fwd53> level 1", "achieved": achieved, "peak": peak,
                     "unit": "GB/s", "frac": achieved / peak, "traffic": None, "peak_source": peak_src,
                     "algorithmic_bytes_per_launch": l1_bytes, "avg_launch_ms": l1_ms},
        "stages_ms": {k: float(v) for k, v in zip(["upload", "mct", "dwt", "t1_context", "t1_mq", "compact_d2h", "host_t2", "dwt_l1"], st_dev)},
        "decode_stages_ms": {k: float(v) for k, v in zip(["host_parse", "h2d", "t1", "idwt", "imct_pack", "d2h", "idwt_l1"], dst_dev)},
        "encode_mpix_s": mpix / ((t_dev / args.steps - dec_ms_dev) * 1e-3), "decode_mpix_s": mpix / (dec_ms_dev * 1e-3),
        "t1": {"code_blocks_per_s": nblocks / ((st_dev[3] + st_dev[4]) * 1e-3), "mq_symbols_per_s": nsym / (st_dev[4] * 1e-3)},
        "clocks": sampler.summary(),
    }
    if not args.no_cpu_baseline:
        fn, kind, cores = reference_encode_fn()
        h, w = (H, W) if kind == "reference" else (H // 4, W // 4)
        f = make_frame(h, w, NC, PREC, args.kind, seed=0)
        pl = [f[..., c].astype(np.int32) for c in range(NC)]
        reps = 2 if kind == "reference" else 1
        t0 = time.perf_counter()
        for _ in range(reps):
            fn(pl)
        dt = (time.perf_counter() - t0) / reps
        line["cpu_baseline"] = {"value": w * h / dt / 1e6, "unit": "MPix/s", "cores": cores, "kind": kind,
                                "sample": "%d encode+decode round trip(s) of one %dx%d 16-bit RGBA frame, ext/grok on the host cores" % (reps, w, h)}
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
