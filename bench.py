#!/usr/bin/env python
"""bench.py -- 4K lossless JPEG 2000 throughput of libj2kcuda on B200 (BASELINE.json metric).

A step = one pass of the hot path (encode, then decode what was encoded) over one batch of FRAMES
4096x2160 16-bit RGBA frames (BASELINE configs[1]): lossless 5/3, RCT on components 0-2, 6 resolutions,
64x64 code-blocks, LRCP, one layer, single tile.  Frames go through j2kcuda_sequence_encode / _decode,
which keep several frames in flight (independent frames, no exchange).
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
FRAMES = 32  # frames per step per GPU (one batch of the image sequence): 32 x 70.8 MB of input and 283 MB of planes per
             # frame in flight -- far beyond the 126 MB L2, so no flush is needed between steps.  Round 1 used 16 = one frame per
             # worker and phase; with two, the host Tier-2 of the first wave overlaps the kernels of the second (BASELINE C5 is a
             # 240-frame sequence: the steady state is what the path runs in)
TRAFFIC_L1 = 239819520  # dram__bytes_read.sum 142 357 760 + dram__bytes_write.sum 97 461 760 of one level-1 launch, from the
                        # ncu --set full capture in profiles/r2_frame_full_raw.csv.gz (part of the 141.6 MB written stays in L2)
WORKERS = 16  # contexts (stream + buffers + host thread) the sequence keeps in flight


def peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"], "measured"
    except Exception:
        return 6650.0, "fallback"


class ClockSampler(threading.Thread):
    """SM clock and throttle reasons of this rank's GPU every 200 ms DURING the timed region (B200_PROFILING.md's clocks line).
    Read through NVML inside the process (what nvidia-smi itself calls); if the binding is missing, ONE `nvidia-smi -lms 200`
    loop process.  (Round 1 spawned a fresh nvidia-smi per sample from every rank: at 8 ranks those start-ups, each of which
    opens all 8 GPUs, slowed the CUDA calls of every rank.)"""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index, uuid=None):
        super().__init__(daemon=True)
        self.index, self.uuid, self.rows, self.stop, self.how = index, uuid, [], False, "nvml"

    def run_nvml(self):
        import pynvml as N
        N.nvmlInit()
        h = N.nvmlDeviceGetHandleByUUID(self.uuid) if self.uuid else N.nvmlDeviceGetHandleByIndex(self.index)   # CUDA_VISIBLE_DEVICES-proof
        bits = [N.nvmlClocksThrottleReasonHwSlowdown, N.nvmlClocksThrottleReasonHwThermalSlowdown,
                N.nvmlClocksThrottleReasonSwThermalSlowdown, N.nvmlClocksThrottleReasonSwPowerCap]
        while not self.stop:
            r = N.nvmlDeviceGetCurrentClocksThrottleReasons(h)
            self.rows.append([str(N.nvmlDeviceGetClockInfo(h, N.NVML_CLOCK_SM)), str(N.nvmlDeviceGetMaxClockInfo(h, N.NVML_CLOCK_SM))] +
                             ["Active" if r & b else "Not Active" for b in bits])
            time.sleep(0.2)

    def run_smi(self):
        self.how = "nvidia-smi -lms 200"
        p = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "200"],
                             stdout=subprocess.PIPE, text=True)
        try:
            for line in p.stdout:
                if self.stop:
                    break
                if line.strip():
                    self.rows.append([x.strip() for x in line.split(",")])
        finally:
            p.kill()

    def run(self):
        try:
            self.run_nvml()
        except Exception:
            try:
                self.run_smi()
            except Exception:
                pass

    def summary(self):
        sm = sorted(int(r[0]) for r in self.rows if r[0].isdigit())
        mx = [int(r[1]) for r in self.rows if r[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 2 + i and r[2 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(self.rows), "how": self.how}


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
        "config": {"workload": WORKLOAD, "direction": "encode then decode of the same frame", "frames_per_step_per_gpu": 1},
        "cpu_baseline": {"value": v, "unit": "MPix/s", "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": v, "unit": "MPix/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=6)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--kind", default="smooth", choices=["smooth", "noise", "ramp"])
    ap.add_argument("--frames", type=int, default=FRAMES)
    ap.add_argument("--workers", type=int, default=0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank)
        return
    if args.workers <= 0:
        # frames in flight per GPU: enough to keep the Tier-1 kernels of several frames on the device while other frames are in
        # host Tier-2.  The workers sleep while the device works (blocking event waits), so a rank may have more of them than cores:
        # on 4 cores 16 workers still beat 6 (profiles/r2_cpuexp.log); the cap only bounds the thread count on small hosts
        args.workers = WORKERS if world <= 2 else max(8, min(WORKERS, 3 * (os.cpu_count() or 16) // world))

    import torch
    import torch.distributed as dist
    from j2k_b200 import _lib as L
    from j2k_b200.frames import make_frame, shard_frames
    import j2k
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: libj2kcuda has no CPU path")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    lib = L.lib()
    F = args.frames
    p = L.default_params(width=W, height=H, ncomp=NC, prec=PREC, mct=1, numresolution=6, device=local)

    # this rank's shard of the sequence: frame f of the job goes to rank f % world (no data-path collective)
    mine = shard_frames(F * world, rank, world)
    host, dev = [], []
    for f in mine:
        t = torch.empty((H, W, NC), dtype=torch.uint16).pin_memory()
        t.numpy()[...] = make_frame(H, W, NC, PREC, args.kind, seed=f)
        host.append(t)
        dev.append(t.cuda())
    back = [torch.empty((H, W, NC), dtype=torch.uint16).pin_memory() for _ in range(F)]   # decoded frames, e2e arm
    back_dev = [torch.empty((H, W, NC), dtype=torch.uint16, device="cuda") for _ in range(F)]
    torch.cuda.synchronize()

    def plane_table(arrays):
        keep = [j2k.planes_of(a) for a in arrays]
        tab = (C.POINTER(L.Plane) * len(keep))(*[C.cast(k, C.POINTER(L.Plane)) for k in keep])
        return tab, keep

    def dev_table(tensors):
        keep = []
        for t in tensors:
            pl = (L.Plane * 1)()
            pl[0].buf = t.data_ptr(); pl[0].colbytes = NC * 2; pl[0].rowbytes = W * NC * 2
            pl[0].width = W; pl[0].height = H; pl[0].sample_type = L.USHORT
            keep.append(pl)
        tab = (C.POINTER(L.Plane) * len(keep))(*[C.cast(k, C.POINTER(L.Plane)) for k in keep])
        return tab, keep

    host_in, _k1 = plane_table([t.numpy() for t in host])
    host_out, _k2 = plane_table([t.numpy() for t in back])
    dev_in, _k3 = dev_table(dev)
    dev_out, _k4 = dev_table(back_dev)

    lib.j2kcuda_sequence_create.restype = C.c_void_p
    seq = C.c_void_p(lib.j2kcuda_sequence_create(C.byref(p), local, args.workers))
    if not seq.value:
        raise SystemExit("sequence_create failed: " + L.last_error())
    outs = (C.POINTER(C.c_uint8) * F)()
    lens = (C.c_size_t * F)()

    def step(flags, tin, tout):
        """encode the batch, decode the codestreams it produced"""
        rc = lib.j2kcuda_sequence_encode(seq, tin, F, NC, flags, outs, lens)
        if rc:
            raise SystemExit("sequence_encode failed: " + L.last_error())
        t0 = time.perf_counter()
        rc = lib.j2kcuda_sequence_decode(seq, C.cast(outs, C.POINTER(C.POINTER(C.c_uint8))), lens, F, 0, tout, NC, flags)
        if rc:
            raise SystemExit("sequence_decode failed: " + L.last_error())
        dt = time.perf_counter() - t0
        nbytes = sum(lens[i] for i in range(F))
        for i in range(F):
            lib.j2kcuda_free(outs[i])
        return nbytes, dt

    def stats(reset):
        em, dm = (C.c_double * 8)(), (C.c_double * 8)()
        ec, dc = (C.c_uint64 * 6)(), (C.c_uint64 * 6)()
        fr = (C.c_uint64 * 2)()
        lib.j2kcuda_sequence_stats(seq, em, ec, dm, dc, fr, 1 if reset else 0)
        return np.array(list(em)), list(ec), np.array(list(dm)), list(dc), list(fr)

    per_rank_ms = []          # ms per step of every rank, one list per timed() call (device-resident, then host buffers)

    def timed(flags, tin, tout):
        for _ in range(args.warmup):
            step(flags, tin, tout)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        stats(True)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        nbytes, tdec = 0, 0.0
        torch.cuda.synchronize()
        e0.record()
        for _ in range(args.steps):
            nb, dt = step(flags, tin, tout)
            nbytes += nb
            tdec += dt
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1)], device="cuda")
        if world > 1:
            dist.barrier()
            every = [torch.zeros_like(t) for _ in range(world)]
            dist.all_gather(every, t)
            per_rank_ms.append([round(float(x.item()) / args.steps, 2) for x in every])
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item()), nbytes / (args.steps * F), tdec / args.steps * 1e3, stats(True)

    sampler = ClockSampler(local, "GPU-" + str(torch.cuda.get_device_properties(local).uuid))
    sampler.start()
    t_dev, cs_bytes, dec_ms_dev, st_dev = timed(1, dev_in, dev_out)          # J2KCUDA_SEQ_DEVICE_PIXELS
    for i in range(F):
        if not torch.equal(back_dev[i], dev[i]):
            raise SystemExit("device round trip of frame %d is not lossless" % i)
    t_e2e, _, dec_ms_e2e, st_e2e = timed(0, host_in, host_out)
    for i in range(F):
        if not np.array_equal(back[i].numpy(), host[i].numpy()):
            raise SystemExit("round trip of frame %d is not lossless" % i)
    # the same K steps with the encode of batch i+1 overlapping the decode of batch i (two sequence objects, two host
    # threads): the encoder's Tier-1 is bound by the dependent chain of its longest block and leaves most issue slots
    # free, the decoder's is bound by issue slots, so a transcoding pipeline runs both at once.  Reported beside the
    # sequential figure, not instead of it.
    def pipelined():
        import threading
        seq2 = C.c_void_p(lib.j2kcuda_sequence_create(C.byref(p), local, args.workers))
        outs2 = [(C.POINTER(C.c_uint8) * F)() for _ in range(2)]
        lens2 = [(C.c_size_t * F)() for _ in range(2)]
        err = []

        def enc(slot):
            if lib.j2kcuda_sequence_encode(seq, host_in, F, NC, 0, outs2[slot], lens2[slot]):
                err.append("encode: " + L.last_error())

        def dec(slot):
            if lib.j2kcuda_sequence_decode(seq2, C.cast(outs2[slot], C.POINTER(C.POINTER(C.c_uint8))), lens2[slot], F, 0, host_out, NC, 0):
                err.append("decode: " + L.last_error())
            for i in range(F):
                lib.j2kcuda_free(outs2[slot][i])

        def run(k):
            enc(0)
            for i in range(1, k):
                te = threading.Thread(target=enc, args=(i & 1,)); td = threading.Thread(target=dec, args=((i - 1) & 1,))
                te.start(); td.start(); te.join(); td.join()
            dec((k - 1) & 1)

        run(2)                                         # warm-up: creates the second sequence's contexts
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        run(args.steps)
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1)], device="cuda")
        if world > 1:
            dist.barrier()
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        lib.j2kcuda_sequence_destroy(seq2)
        if err:
            raise SystemExit("pipelined run failed: " + err[0])
        for i in range(F):
            if not np.array_equal(back[i].numpy(), host[i].numpy()):
                raise SystemExit("pipelined round trip of frame %d is not lossless" % i)
        return float(t.item())

    t_pipe = pipelined()

    # the stress content class: uniform noise, every bit-plane of every block coded -- Tier-1 works ~3.4 x longer per
    # frame than on the smooth frames of the headline; two timed steps of the same F-frame batch, device-resident
    def noise_value():
        nf = min(F, 8)
        nd = []
        for f in range(nf):
            t = torch.empty((H, W, NC), dtype=torch.uint16).pin_memory()
            t.numpy()[...] = make_frame(H, W, NC, PREC, "noise", seed=1000 + rank * nf + f)
            nd.append(t.cuda())
        nb = [torch.empty((H, W, NC), dtype=torch.uint16, device="cuda") for _ in range(nf)]
        tin, _ka = dev_table(nd)
        tout, _kb = dev_table(nb)
        o = (C.POINTER(C.c_uint8) * nf)(); l = (C.c_size_t * nf)()

        def one():
            if lib.j2kcuda_sequence_encode(seq, tin, nf, NC, 1, o, l):
                raise SystemExit("noise encode failed: " + L.last_error())
            if lib.j2kcuda_sequence_decode(seq, C.cast(o, C.POINTER(C.POINTER(C.c_uint8))), l, nf, 0, tout, NC, 1):
                raise SystemExit("noise decode failed: " + L.last_error())
            for i in range(nf):
                lib.j2kcuda_free(o[i])
        one()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(2):
            one()
        e1.record()
        torch.cuda.synchronize()
        for i in range(nf):
            if not torch.equal(nb[i], nd[i]):
                raise SystemExit("noise round trip of frame %d is not lossless" % i)
        t = torch.tensor([e0.elapsed_time(e1)], device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return world * 2 * nf * W * H / (float(t.item()) * 1e-3) / 1e6, nf

    v_noise, nf_noise = noise_value()

    # the plug-in's situation: ONE frame at a time through the one-shot calls the j2k::Codec binding makes (j2kcuda_encode /
    # j2kcuda_decode, contexts kept by the library between calls), from and to PAGEABLE memory
    def pageable_single():
        img = np.ascontiguousarray(host[0].numpy().copy())           # pageable copy
        out = np.zeros_like(img)
        pl_in, pl_out = j2k.planes_of(img), j2k.planes_of(out)
        te, td = [], []
        for i in range(4):
            o = C.POINTER(C.c_uint8)(); n = C.c_size_t(0)
            t0 = time.perf_counter()
            if lib.j2kcuda_encode(C.byref(p), pl_in, NC, C.byref(o), C.byref(n)):
                raise SystemExit("one-shot encode failed: " + L.last_error())
            t1 = time.perf_counter()
            if lib.j2kcuda_decode(o, n, 0, pl_out, NC, None, None):
                raise SystemExit("one-shot decode failed: " + L.last_error())
            t2 = time.perf_counter()
            lib.j2kcuda_free(o)
            if i:
                te.append(t1 - t0); td.append(t2 - t1)
        if not np.array_equal(out, img):
            raise SystemExit("one-shot round trip is not lossless")
        return min(te) * 1e3, min(td) * 1e3

    enc_ms_page, dec_ms_page = pageable_single() if rank == 0 else (0.0, 0.0)
    sampler.stop = True
    sampler.join(timeout=3)
    lib.j2kcuda_sequence_destroy(seq)

    # the roofline kernel timed alone: one context, frames one after the other (nothing else on the device),
    # CUDA events recorded by the library on the context's own stream around the level-1 launch
    solo = C.c_void_p(lib.j2kcuda_sequence_create(C.byref(p), local, 1))
    for _ in range(2):
        step_solo = lib.j2kcuda_sequence_encode(solo, dev_in, F, NC, 1, outs, lens)
        if step_solo:
            raise SystemExit("solo pass failed: " + L.last_error())
        lib.j2kcuda_sequence_decode(solo, C.cast(outs, C.POINTER(C.POINTER(C.c_uint8))), lens, F, 0, dev_out, NC, 1)
        for i in range(F):
            lib.j2kcuda_free(outs[i])
        em, ec, dm, dc, fr = (C.c_double * 8)(), (C.c_uint64 * 6)(), (C.c_double * 8)(), (C.c_uint64 * 6)(), (C.c_uint64 * 2)()
        lib.j2kcuda_sequence_stats(solo, em, ec, dm, dc, fr, 1)
    lib.j2kcuda_sequence_destroy(solo)
    solo_enc, solo_dec, solo_n = np.array(list(em)) / fr[0], np.array(list(dm)) / fr[1], int(fr[0])
    if rank != 0:
        return

    em_b, ec_b, dm_b, dc_b, fr_b = st_dev
    em_e, ec_e, dm_e, dc_e, fr_e = st_e2e
    nfr = max(int(fr_b[0]), 1)
    mpix = W * H / 1e6
    value = world * args.steps * F * mpix / (t_dev * 1e-3)
    e2e = world * args.steps * F * mpix / (t_e2e * 1e-3)
    peak, peak_src = peaks()
    l1_bytes = 8.0 * W * H * NC                      # 4 B read + 4 B written per sample, level 1 of every component
    all_bytes = float(sum(8 * NC * (-(-W // (1 << l))) * (-(-H // (1 << l))) for l in range(5)))     # levels 1..5 of a 6-resolution transform
    l1_ms = float(solo_enc[7])
    achieved = l1_bytes / (l1_ms * 1e-3) / 1e9
    l1_ms_batch = float(em_b[7]) / nfr
    il1_ms = float(solo_dec[6])
    nblocks, nsym = int(ec_b[1]) // nfr, int(ec_b[2]) // nfr
    enc_names = ["upload", "mct", "dwt", "t1_context", "t1_mq", "compact_d2h", "host_t2", "dwt_l1"]
    dec_names = ["host_parse", "h2d", "t1", "idwt", "imct_pack", "d2h", "idwt_l1"]
    line = {
        "metric": "4K lossless encode+decode MPix/s", "value": value, "unit": "MPix/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": t_dev / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "int32", "data": "synthetic (%s)" % args.kind,
        "config": {"workload": WORKLOAD, "direction": "encode then decode of the same frames (every round trip verified lossless)",
                   "frames_per_step_per_gpu": F, "frames_in_flight": args.workers,
                   "sharding": "frame f of the job -> rank f %% %d, no collective on the data path" % world,
                   "l2": "inputs larger than L2: %d distinct frames per step (%.0f MB in, %.0f MB of planes each)" % (F, F * W * H * NC * 2 / 1e6, 2 * W * H * NC * 4 / 1e6),
                   "codestream_bytes_per_frame": cs_bytes, "code_blocks_per_frame": nblocks, "mq_symbols_per_frame": nsym},
        "e2e": {"value": e2e, "unit": "MPix/s", "h2d_bytes_per_step": int(F * (W * H * NC * 2 + cs_bytes)),
                "d2h_bytes_per_step": int(ec_e[5] / args.steps + F * W * H * NC * 2), "ms_per_step": t_e2e / args.steps},
        "gpu_launches": int(ec_b[0] + dc_b[0]),
        "e2e_pipelined": {"value": args.steps * F * world * W * H / (t_pipe * 1e-3) / 1e6, "unit": "MPix/s", "ms_per_step": t_pipe / args.steps,
                          "note": "same K steps, host buffers, encode of batch i+1 overlapping decode of batch i"},
        "roofline": {"bound": "hbm", "kernel": "k_dwt_forward<Fwd53> level 1", "achieved": achieved, "peak": peak,
                     "unit": "GB/s", "frac": achieved / peak, "traffic": TRAFFIC_L1, "peak_source": peak_src + " (burst: kernel timed alone)",
                     "algorithmic_bytes_per_launch": l1_bytes, "avg_launch_ms": l1_ms, "launches_timed": solo_n,
                     "how": "library CUDA events on the context's stream around the launch, one frame at a time after the timed region",
                     "in_batch": {"avg_launch_ms": l1_ms_batch, "achieved": l1_bytes / (l1_ms_batch * 1e-3) / 1e9,
                                  "note": "same launch inside the timed region, sharing the device with the Tier-1 kernels of the other frames in flight"},
                     "inverse_level1": {"avg_launch_ms": il1_ms, "achieved": l1_bytes / (il1_ms * 1e-3) / 1e9},
                     "all_levels": {"algorithmic_bytes": all_bytes,
                                    "forward": {"ms": float(solo_enc[2]), "achieved": all_bytes / (float(solo_enc[2]) * 1e-3) / 1e9, "frac": all_bytes / (float(solo_enc[2]) * 1e-3) / 1e9 / peak},
                                    "inverse": {"ms": float(solo_dec[3]), "achieved": all_bytes / (float(solo_dec[3]) * 1e-3) / 1e9, "frac": all_bytes / (float(solo_dec[3]) * 1e-3) / 1e9 / peak},
                                    "note": "BASELINE.md 4: 8 B per sample of every level's input region, all five launches of one frame (CUDA events around them, frame alone)"},
                     "traffic_source": "profiles/r2_frame_full_raw.csv.gz (ncu --set full of tools/stage_times.py roundtrip1, round 2)"},
        "value_noise": {"value": v_noise, "unit": "MPix/s", "frames_per_step_per_gpu": nf_noise,
                        "note": "same round trip on uniform-noise frames (every bit-plane of every code-block coded): the stress figure beside the smooth-frame headline"},
        "e2e_pageable": {"value": W * H / ((enc_ms_page + dec_ms_page) * 1e-3) / 1e6, "unit": "MPix/s", "encode_ms": enc_ms_page, "decode_ms": dec_ms_page,
                         "note": "ONE frame at a time through j2kcuda_encode + j2kcuda_decode (what the j2k::Codec binding calls) from and to pageable host memory; contexts cached by the library"},
        "stages_ms_per_frame_alone": {k: float(v) for k, v in zip(enc_names, solo_enc)},
        "decode_stages_ms_per_frame_alone": {k: float(v) for k, v in zip(dec_names, solo_dec)},
        "stages_ms_per_frame_in_batch": {k: float(v) / nfr for k, v in zip(enc_names, em_b)},
        "decode_stages_ms_per_frame_in_batch": {k: float(v) / nfr for k, v in zip(dec_names, dm_b)},
        "encode_mpix_s": F * mpix / ((t_dev / args.steps - dec_ms_dev) * 1e-3), "decode_mpix_s": F * mpix / (dec_ms_dev * 1e-3),
        "e2e_encode_mpix_s": F * mpix / ((t_e2e / args.steps - dec_ms_e2e) * 1e-3), "e2e_decode_mpix_s": F * mpix / (dec_ms_e2e * 1e-3),
        "t1": {"code_blocks_per_s_alone": nblocks / ((solo_enc[3] + solo_enc[4]) * 1e-3), "mq_symbols_per_s_alone": nsym / (solo_enc[4] * 1e-3),
               "decode_code_blocks_per_s_alone": nblocks / (solo_dec[2] * 1e-3),
               "code_blocks_per_s_job": world * args.steps * F * nblocks / (t_dev * 1e-3)},
        "clocks": sampler.summary(),
        "per_rank_ms_per_step": per_rank_ms,
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
