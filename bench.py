#!/usr/bin/env python
"""bench.py -- BASELINE.json's metric on BASELINE.json's configs.

  python bench.py --gpus N --steps K --warmup W            (N>1: launched by torchrun, one rank per GPU)
  python bench.py --impl reference --gpus N --steps K --warmup W

Headline workload (config.workload) = BASELINE configs[1]: cv::resize 3840x2160 -> 1920x1080 CV_8UC3
INTER_LINEAR (the reference turns this into the 2x2 box filter, resize.cpp:3817).  A step is ONE batched
launch over FRAMES_PER_STEP distinct device-resident frames (1 GB of traffic per step: far larger than the
126 MB L2, so every step streams from HBM).  `value` = output Mpix/s of the whole job, inputs resident in
HBM.  `e2e` = the same metric through the HAL hook b200_hal_resize with pinned HOST buffers (H2D + kernel
+ D2H inside the timed region).  `roofline` = algorithmic bytes (whole source + destination)
per launch / mean launch duration (CUDA events on the launch stream) against MEASURED_PEAKS.json hbm_gbs.
`cpu_baseline` / `--impl reference` = the reference's own CPU code (oracle/_ref/libcvref.so, compiled from
/root/reference by oracle/build_ref.sh; cv::hal::resize, pthreads parallel_for_, all host cores).
`ops` = the other BASELINE configs, device-resident, each with its own roofline figures (informational).
Multi-GPU: frames are independent -> every rank runs the same per-GPU batch on its own GPU, no data-path
collective (weak scaling); torch.distributed (NCCL) is used only for the barrier and the max-over-ranks time.
"""
import argparse
import ctypes as C
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

INTER_NEAREST, INTER_LINEAR, INTER_AREA = 0, 1, 3
B_CONSTANT, B_REPLICATE, B_REFLECT_101 = 0, 1, 4
CV_8UC3, CV_8UC4, CV_32FC1 = 16, 24, 5

SRC_W, SRC_H, DST_W, DST_H, CN = 3840, 2160, 1920, 1080, 3
FRAMES_PER_STEP = 32
WORKLOAD = "cv::resize 3840x2160->1920x1080 CV_8UC3 INTER_LINEAR (BASELINE configs[1])"


# ----------------------------------------------------------------------------------------------
def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """samples nvidia-smi clocks / throttle reasons during the timed region"""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.idx, self.rows, self.proc = gpu_index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.idx), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm = [float(r[1]) for r in self.rows if len(r) >= 9 and r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in self.rows if len(r) >= 9 and r[2].replace(".", "").isdigit()]
        reasons = set()
        for r in self.rows:
            if len(r) >= 9:
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def dist_env():
    return int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))


# ----------------------------------------------------------------------------- reference (CPU) --
def load_cvref():
    p = os.path.join(ROOT, "oracle", "_ref", "libcvref.so")
    if os.path.exists(p):
        lib, kind, fn = C.CDLL(p), "reference", "cvref_hal_resize"
    else:                      # the oracle port always exists (single-threaded restatement)
        so = os.path.join(ROOT, "oracle", "libgeom_oracle.so")
        if not os.path.exists(so):
            subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "-s"])
        lib, kind, fn = C.CDLL(so), "port", "orc_resize"
    f = getattr(lib, fn)
    f.argtypes = [C.c_int, C.c_void_p, C.c_size_t, C.c_int, C.c_int, C.c_void_p, C.c_size_t, C.c_int, C.c_int,
                  C.c_double, C.c_double, C.c_int]
    f.restype = C.c_int
    cores = 1
    if kind == "reference":
        lib.cvref_setNumThreads(-1)
        cores = lib.cvref_getNumThreads()
    return f, kind, cores


def cpu_frames(n_bufs, seed=2026):
    rng = np.random.default_rng(seed)
    srcs = [rng.integers(0, 256, (SRC_H, SRC_W, CN), dtype=np.uint8) for _ in range(n_bufs)]
    dsts = [np.zeros((DST_H, DST_W, CN), np.uint8) for _ in range(n_bufs)]
    return srcs, dsts


def cpu_run(f, srcs, dsts, n_frames):
    t0 = time.perf_counter()
    for i in range(n_frames):
        s, d = srcs[i % len(srcs)], dsts[i % len(dsts)]
        rc = f(CV_8UC3, s.ctypes.data, s.strides[0], SRC_W, SRC_H, d.ctypes.data, d.strides[0], DST_W, DST_H,
               DST_W / SRC_W, DST_H / SRC_H, INTER_LINEAR)
        assert rc == 0
    return time.perf_counter() - t0


def cpu_baseline(budget_s=10.0):
    f, kind, cores = load_cvref()
    srcs, dsts = cpu_frames(8)
    cpu_run(f, srcs, dsts, 8)                                  # warm-up (thread pool, page faults)
    t1 = cpu_run(f, srcs, dsts, 8) / 8
    n = int(max(16, min(4000, budget_s / max(t1, 1e-6))))
    t = cpu_run(f, srcs, dsts, n)
    mpix = n * DST_W * DST_H / t / 1e6
    return {"value": round(mpix, 1), "unit": "Mpix/s", "cores": cores, "kind": kind,
            "sample": f"{n} frames of the headline workload over 8 rotating host buffers, {t:.2f} s, "
                      f"{'cv::hal::resize of oracle/_ref (pthreads parallel_for_)' if kind == 'reference' else 'oracle port'}",
            "ms_per_frame": round(t / n * 1e3, 4)}


def run_reference_arm(args):
    rank, _, world = dist_env()
    if rank != 0:
        return
    f, kind, cores = load_cvref()
    frames = 16                                               # bounded sample per step
    srcs, dsts = cpu_frames(8)
    for _ in range(args.warmup):
        cpu_run(f, srcs, dsts, frames)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        cpu_run(f, srcs, dsts, frames)
    dt = time.perf_counter() - t0
    mpix = args.steps * frames * DST_W * DST_H / dt / 1e6
    sample = f"{frames} frames per step (bounded sample of the {FRAMES_PER_STEP}-frame step) over 8 rotating host buffers"
    print(json.dumps({
        "impl": "reference", "metric": "resize output Mpix/s", "value": round(mpix, 1), "unit": "Mpix/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(dt / args.steps * 1e3, 3),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": WORKLOAD, "frames_per_step": frames, "threads": cores},
        "cpu_baseline": {"value": round(mpix, 1), "unit": "Mpix/s", "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": round(mpix, 1), "unit": "Mpix/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


# ------------------------------------------------------------------------------------- GPU arm --
def affine_inverse(w, h, angle, scale, cx, cy):
    from opencv_b200 import invert_affine
    a = np.deg2rad(angle)
    al, be = scale * np.cos(a), scale * np.sin(a)
    return invert_affine(np.array([[al, be, (1 - al) * cx - be * cy], [-be, al, be * cx + (1 - al) * cy]]))


def affine_footprint_px(sw, sh, dw, dh, M):
    """|F|: distinct in-bounds source pixels touched by any bilinear tap; numpy restatement of
    the fixed-point coordinate scheme, used only to count algorithmic bytes."""
    x = np.arange(dw, dtype=np.float64)
    y = np.arange(dh, dtype=np.float64)[:, None]
    X = (np.rint((M[0, 1] * y + M[0, 2]) * 1024).astype(np.int64) + 16 + np.rint(M[0, 0] * x * 1024).astype(np.int64)) >> 10
    Y = (np.rint((M[1, 1] * y + M[1, 2]) * 1024).astype(np.int64) + 16 + np.rint(M[1, 0] * x * 1024).astype(np.int64)) >> 10
    bm = np.zeros((sh + 2, sw + 2), bool)
    for j in (0, 1):
        for i in (0, 1):
            xx, yy = X + i, Y + j
            ok = (xx >= 0) & (xx < sw) & (yy >= 0) & (yy < sh)
            bm[yy[ok], xx[ok]] = True
    return int(bm.sum())


def run_gpu_arm(args):
    import torch
    import torch.distributed as dist
    from opencv_b200 import capi
    lib = capi.lib
    rank, local_rank, world = dist_env()
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    stream = torch.cuda.current_stream()
    sp = C.c_void_p(stream.cuda_stream)
    peak, peak_src = measured_peak()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---------------- headline: batched device-resident resize 4K -> 1080p 8UC3 INTER_LINEAR ----
    B = FRAMES_PER_STEP
    s_step, d_step = SRC_W * CN, DST_W * CN
    src = torch.empty((B, SRC_H, s_step), dtype=torch.uint8, device="cuda")
    dst = torch.empty((B, DST_H, d_step), dtype=torch.uint8, device="cuda")
    assert lib.b200_synth_frames_dev(src.data_ptr(), SRC_H * s_step, s_step, SRC_W, SRC_H, CN, rank * B, B, 2026, sp) == 0

    def step():
        rc = lib.b200_resize_batch_dev(CV_8UC3, src.data_ptr(), SRC_H * s_step, s_step, SRC_W, SRC_H,
                                       dst.data_ptr(), DST_H * d_step, d_step, DST_W, DST_H, B, 0.0, 0.0, INTER_LINEAR, sp)
        assert rc == 0, capi.last_error()

    for _ in range(max(args.warmup, 3)):
        step()
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
        time.sleep(0.25)
    l0 = lib.b200_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record(stream)
    for _ in range(args.steps):
        step()
    e1.record(stream)
    barrier()
    launches = lib.b200_launch_count() - l0
    ms = e0.elapsed_time(e1)
    if world > 1:
        t = torch.tensor([ms], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    clocks = sampler.stop() if rank == 0 else None
    ms_per_step = ms / args.steps
    mpix = world * B * DST_W * DST_H / (ms_per_step * 1e-3) / 1e6
    alg_bytes = B * (SRC_W * SRC_H * CN + DST_W * DST_H * CN)           # per launch (one launch per step)
    local_ms = e0.elapsed_time(e1) / args.steps
    achieved = alg_bytes / (local_ms * 1e-3) / 1e9

    result = {
        "metric": "resize output Mpix/s", "value": round(mpix, 1), "unit": "Mpix/s", "n_gpus": world,
        "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": round(ms_per_step, 4),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": WORKLOAD, "frames_per_step_per_gpu": B, "launches_per_step": 1,
                   "l2": "inputs larger than L2 (796 MB of distinct source frames per step vs 126 MB L2)",
                   "parallelism": f"{world} GPU(s), independent frames, no collective"},
        "gpu_launches": int(launches),
        "roofline": {"bound": "hbm", "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s",
                     "frac": round(achieved / peak, 4), "traffic": None,
                     "kernel": "resize batch kernel (2x2 box)", "bytes_per_launch": alg_bytes, "peak_source": peak_src},
    }
    if rank == 0:
        result["clocks"] = clocks

    # ---------------- e2e through the HAL hook with pinned host buffers (rank-local, then max) ----
    nb = 4
    e2e_frames = 8
    hs = [torch.empty((SRC_H, SRC_W, CN), dtype=torch.uint8).pin_memory() for _ in range(nb)]
    hd = [torch.empty((DST_H, DST_W, CN), dtype=torch.uint8).pin_memory() for _ in range(nb)]
    for i, h in enumerate(hs):
        h.copy_(src[i].view(SRC_H, SRC_W, CN).cpu())
    lib.b200_hal_set_min_pixels(0)

    def e2e_step():
        for i in range(e2e_frames):
            s_, d_ = hs[i % nb], hd[i % nb]
            rc = lib.b200_hal_resize(CV_8UC3, s_.data_ptr(), SRC_W * CN, SRC_W, SRC_H, d_.data_ptr(), DST_W * CN,
                                     DST_W, DST_H, DST_W / SRC_W, DST_H / SRC_H, INTER_LINEAR)
            assert rc == 0, capi.last_error()
    e2e_step()
    e2e_steps = max(3, min(args.steps, 10))
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_step()
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([dt], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dt = float(t.item())
    checksum = int(hd[0].sum().item())                                  # the result is read on the host
    result["e2e"] = {"value": round(world * e2e_steps * e2e_frames * DST_W * DST_H / dt / 1e6, 1), "unit": "Mpix/s",
                     "h2d_bytes_per_step": e2e_frames * SRC_W * SRC_H * CN, "d2h_bytes_per_step": e2e_frames * DST_W * DST_H * CN,
                     "api": "b200_hal_resize (cv_hal_resize replacement), pinned host buffers", "frames_per_step": e2e_frames,
                     "steps": e2e_steps, "host_checksum": checksum}
    del hs, hd

    # ---------------- other BASELINE configs (rank 0, N=1 only): informational per-op rooflines ----
    if world == 1 and not args.no_ops:
        result["ops"] = run_ops(lib, capi, torch, sp, stream, peak)
        result["cpu_baseline"] = cpu_baseline(args.cpu_budget)
    if rank == 0:
        print(json.dumps(result))
    if world > 1:
        dist.destroy_process_group()


def time_launches(torch, stream, fn, n_sets, iters=48, warm=6):
    for i in range(warm):
        fn(i % n_sets)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for i in range(iters):
        fn(i % n_sets)
    e1.record(stream)
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e-3


def run_ops(lib, capi, torch, sp, stream, peak):
    """device-resident single-image calls over >=16 rotating (src,dst) sets (> 2x L2), SURVEY 8(d)."""
    ops = []
    bv = (C.c_double * 4)(0, 0, 0, 0)

    def synth(n, h, w, cn):
        t = torch.empty((n, h, w * cn), dtype=torch.uint8, device="cuda")
        assert lib.b200_synth_frames_dev(t.data_ptr(), h * w * cn, w * cn, w, h, cn, 0, n, 7, sp) == 0
        return t

    def record(name, secs, out_px, alg_bytes, extra=None):
        r = {"op": name, "us": round(secs * 1e6, 2), "mpix_s": round(out_px / secs / 1e6, 1),
             "gb_s": round(alg_bytes / secs / 1e9, 1), "frac_of_hbm_peak": round(alg_bytes / secs / 1e9 / peak, 4),
             "algorithmic_bytes": alg_bytes}
        if extra:
            r.update(extra)
        ops.append(r)

    # C1 / C1-4K warpAffine 8UC3 LINEAR CONSTANT rot 15 deg, scale 0.9
    for (w, h, n) in ((1920, 1080, 48), (3840, 2160, 16)):
        s, d = synth(n, h, w, 3), torch.empty((n, h, w * 3), dtype=torch.uint8, device="cuda")
        M = affine_inverse(w, h, 15.0, 0.9, w / 2, h / 2)
        Mc = (C.c_double * 6)(*M.ravel())
        fp = affine_footprint_px(w, h, w, h, M)

        def fn(i, s=s, d=d, w=w, h=h, Mc=Mc):
            assert lib.b200_warpAffine_dev(CV_8UC3, s[i].data_ptr(), w * 3, w, h, d[i].data_ptr(), w * 3, w, h, Mc,
                                           INTER_LINEAR, B_CONSTANT, bv, sp) == 0
        record(f"warpAffine {w}x{h} 8UC3 LINEAR CONSTANT", time_launches(torch, stream, fn, n), w * h, fp * 3 + w * h * 3,
               {"footprint_px": fp})
        del s, d
    # C2a INTER_AREA, C2c fractional / 3x3 / generic linear
    s = synth(16, 2160, 3840, 3)
    for (dw, dh, interp, nm) in ((1920, 1080, INTER_AREA, "AREA (2x2 box)"), (1600, 900, INTER_AREA, "AREA fractional"),
                                 (1280, 720, INTER_AREA, "AREA 3x3"), (1600, 900, INTER_LINEAR, "LINEAR generic")):
        d = torch.empty((16, dh, dw * 3), dtype=torch.uint8, device="cuda")

        def fn(i, d=d, dw=dw, dh=dh, interp=interp):
            assert lib.b200_resize_dev(CV_8UC3, s[i].data_ptr(), 3840 * 3, 3840, 2160, d[i].data_ptr(), dw * 3, dw, dh,
                                       0.0, 0.0, interp, sp) == 0
        record(f"resize 3840x2160->{dw}x{dh} 8UC3 {nm}", time_launches(torch, stream, fn, 16), dw * dh,
               3840 * 2160 * 3 + dw * dh * 3)
        del d
    del s
    # C2b resize 1080p -> 4K 32FC1 LINEAR
    s = torch.rand((16, 1080, 1920), dtype=torch.float32, device="cuda")
    d = torch.empty((16, 2160, 3840), dtype=torch.float32, device="cuda")

    def fn(i):
        assert lib.b200_resize_dev(CV_32FC1, s[i].data_ptr(), 1920 * 4, 1920, 1080, d[i].data_ptr(), 3840 * 4, 3840, 2160,
                                   0.0, 0.0, INTER_LINEAR, sp) == 0
    record("resize 1920x1080->3840x2160 32FC1 LINEAR", time_launches(torch, stream, fn, 16), 3840 * 2160,
           1920 * 1080 * 4 + 3840 * 2160 * 4)
    del s, d
    # C3 warpPerspective 4K 8UC4 LINEAR REPLICATE
    s, d = synth(16, 2160, 3840, 4), torch.empty((16, 2160, 3840 * 4), dtype=torch.uint8, device="cuda")
    Hi = np.linalg.inv(np.array([[1.12085816, 0.0203087925, 97], [0.0422569112, 1.05413990, -64],
                                 [4.77793919e-05, -3.76190421e-06, 1]]))
    Hc = (C.c_double * 9)(*Hi.ravel())

    def fn(i):
        assert lib.b200_warpPerspective_dev(CV_8UC4, s[i].data_ptr(), 3840 * 4, 3840, 2160, d[i].data_ptr(), 3840 * 4,
                                            3840, 2160, Hc, INTER_LINEAR, B_REPLICATE, bv, sp) == 0
    record("warpPerspective 3840x2160 8UC4 LINEAR REPLICATE", time_launches(torch, stream, fn, 16), 3840 * 2160,
           32845684 + 3840 * 2160 * 4, {"footprint_px": 8211421, "footprint_source": "SURVEY 8(d) C3"})
    del s, d
    # C4 remap 720p 8UC3 LINEAR REFLECT_101, three map formats (map bytes reported separately)
    w, h = 1280, 720
    s, d = synth(64, h, w, 3), torch.empty((64, h, w * 3), dtype=torch.uint8, device="cuda")
    f = 600.0
    xs, ys = np.meshgrid(np.arange(w, dtype=np.float64), np.arange(h, dtype=np.float64))
    X, Y = (xs - 639.5) / f, (ys - 359.5) / f
    r = np.sqrt(X * X + Y * Y)
    th = np.arctan(r)
    sc = np.where(r < 1e-8, 1.0, th * (1 - 0.05 * th ** 2 + 0.01 * th ** 4) / np.maximum(r, 1e-300))
    mx, my = (f * X * sc + 639.5).astype(np.float32), (f * Y * sc + 359.5).astype(np.float32)
    ix, iy = np.rint(mx * np.float32(32)).astype(np.int64), np.rint(my * np.float32(32)).astype(np.int64)
    m1 = np.stack([ix >> 5, iy >> 5], -1).astype(np.int16)
    m2 = ((iy & 31) * 32 + (ix & 31)).astype(np.uint16)
    t_mx, t_my = torch.from_numpy(mx).cuda(), torch.from_numpy(my).cuda()
    t_mxy = torch.from_numpy(np.ascontiguousarray(np.stack([mx, my], -1))).cuda()
    t_m1, t_m2 = torch.from_numpy(m1).cuda(), torch.from_numpy(m2.view(np.int16)).cuda()
    img = lambda i: (CV_8UC3, s[i].data_ptr(), w * 3, w, h, d[i].data_ptr(), w * 3, w, h)   # noqa: E731
    fns = {
        "2xCV_32FC1": (lambda i: lib.b200_remap32f_dev(*img(i), t_mx.data_ptr(), w * 4, t_my.data_ptr(), w * 4, INTER_LINEAR, B_REFLECT_101, bv, sp), 7372800),
        "CV_32FC2": (lambda i: lib.b200_remap32fc2_dev(*img(i), t_mxy.data_ptr(), w * 8, INTER_LINEAR, B_REFLECT_101, bv, sp), 7372800),
        "CV_16SC2+CV_16UC1": (lambda i: lib.b200_remap16s_dev(*img(i), t_m1.data_ptr(), w * 4, t_m2.data_ptr(), w * 2, INTER_LINEAR, B_REFLECT_101, bv, sp), 5529600),
    }
    for nm, (g, map_bytes) in fns.items():
        def fn(i, g=g):
            assert g(i) == 0
        secs = time_launches(torch, stream, fn, 64, iters=128)
        record(f"remap 1280x720 8UC3 LINEAR REFLECT_101 maps {nm}", secs, w * h, 4428903,
               {"map_bytes_not_in_roofline": map_bytes, "gb_s_incl_maps": round((4428903 + map_bytes) / secs / 1e9, 1)})
    del s, d
    # C5 per-GPU pipeline: 1080p -> 960x540 INTER_LINEAR (2x2 box) -> warpAffine, 64 frames per launch pair
    n, w, h, mw, mh = 64, 1920, 1080, 960, 540
    s = synth(n, h, w, 3)
    mid = torch.empty((n, mh, mw * 3), dtype=torch.uint8, device="cuda")
    d = torch.empty((n, mh, mw * 3), dtype=torch.uint8, device="cuda")
    M = affine_inverse(mw, mh, 15.0, 0.9, 480.0, 270.0)
    Mc = (C.c_double * 6)(*M.ravel())
    fp = affine_footprint_px(mw, mh, mw, mh, M)

    def fn(_):
        assert lib.b200_resize_warpAffine_batch_dev(CV_8UC3, s.data_ptr(), h * w * 3, w * 3, w, h, mid.data_ptr(), mh * mw * 3, mw * 3, mw, mh,
                                                    d.data_ptr(), mh * mw * 3, mw * 3, mw, mh, n, INTER_LINEAR, Mc, INTER_LINEAR,
                                                    B_CONSTANT, bv, sp) == 0
    secs = time_launches(torch, stream, fn, 1, iters=20, warm=3)
    per_frame = (w * h * 3 + mw * mh * 3) + (fp * 3 + mw * mh * 3)
    record("pipeline 1080p resize->960x540 + warpAffine, 64 frames/launch", secs, n * mw * mh, n * per_frame,
           {"frames_per_s": round(n / secs, 1)})
    return ops


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-ops", action="store_true", help="skip the informational per-op table and cpu_baseline")
    ap.add_argument("--cpu-budget", type=float, default=10.0, help="seconds of CPU work for cpu_baseline")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_gpu_arm(args)


if __name__ == "__main__":
    main()
