#!/usr/bin/env python
"""bench.py -- BASELINE.json's metric on BASELINE.json's configs.

  python bench.py --gpus N --steps K --warmup W            (N>1: launched by torchrun, one rank per GPU)
  python bench.py --impl reference --gpus N --steps K --warmup W

Headline workload (config.workload) = BASELINE configs[4], the config the metric's "1/2/4/8 B200" clause is
quoted on: the batched 1080p CV_8UC3 synthetic frame stream through the per-frame pipeline
    cv::resize 1920x1080 -> 960x540 INTER_LINEAR (exact 2x: the reference's 2x2 box, resize.cpp:3817)
    cv::warpAffine 960x540 INTER_LINEAR BORDER_CONSTANT, 15 deg rotation + 0.9 scale about (480,270)
A step is ONE pass over a batch of FRAMES_PER_STEP distinct device-resident frames per GPU (one persistent
launch: box-filter strips and warp tiles from one work queue, half-size frames in an L2-resident ring).  Frames come from a per-GPU ring of RING_CHUNKS batches (1.6 GB, generated on the device before
the timed region), so every step streams from HBM, never from the 126 MB L2.
`value` = final-output Mpix/s of the whole job.  `e2e` = the same pipeline, pinned HOST frames in and pinned
HOST frames out, through the C-ABI frame-stream call b200_hal_resize_warpAffine_batch (H2D + kernels + D2H all
inside the timed region, overlapped chunk by chunk inside the call); `e2e_hooks` = the reference's own call
pattern, one synchronous b200_hal_resize + b200_hal_warpAffine (the cv_hal_* replacements) per frame.  `roofline` = algorithmic bytes per launch (SURVEY
8(d): source footprint + destination) / mean launch duration of the dominant kernel (CUDA events recorded
on the launch stream inside the timed region) against MEASURED_PEAKS.json hbm_gbs.
`cpu_baseline` / `--impl reference` = the reference's own CPU code (oracle/_ref/libcvref.so compiled from
/root/reference by oracle/build_ref.sh: cv::hal::resize + cv::hal::warpAffine, pthreads parallel_for_, all
host cores) on a bounded sample of the same frames.
`ops` = the other BASELINE configs (4K warpAffine / resize / warpPerspective / remap), device-resident over
>=16 rotating buffer sets, each with its own roofline figures (informational; the 70 % targets live here).
Multi-GPU: frames are independent -> every rank runs its own batches on its own GPU, no data-path collective
(weak scaling); torch.distributed (NCCL) is used only for the barrier and the max-over-ranks time.
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
CV_8UC3, CV_8UC4, CV_32FC1, CV_16UC3 = 16, 24, 5, 18

SRC_W, SRC_H, MID_W, MID_H, CN = 1920, 1080, 960, 540, 3
DST_W, DST_H = MID_W, MID_H
FRAMES_PER_STEP = 64
RING_CHUNKS = 4
STREAM_FRAMES = 8192


def ncu_traffic(kernel_prefix):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the headline kernel, read from the NEWEST committed
    `ncu --set full` summary under profiles/ that holds a launch of it (tools/ncu_summary.py output); returns
    (bytes or None, file name or None) -- never a constant, so a kernel change cannot leave a stale figure behind."""
    import csv
    import glob
    best = (None, None)
    for path in sorted(glob.glob(os.path.join(ROOT, "profiles", "r*_ncu_full.csv")), key=lambda q: (int(os.path.basename(q).split("_")[0][1:] or 0), os.path.basename(q))):
        try:
            rows = list(csv.reader(open(path)))
        except OSError:
            continue
        if not rows:
            continue
        hdr = rows[0]
        def col(name):
            for i, h in enumerate(hdr):
                if h.startswith(name):
                    unit = h[h.find("[") + 1:h.find("]")] if "[" in h else "byte"
                    return i, {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1.0)
            return None, 1.0
        (ir, fr), (iw, fw) = col("dram__bytes_read.sum"), col("dram__bytes_write.sum")
        if ir is None or iw is None:
            continue
        vals = [float(r[ir]) * fr + float(r[iw]) * fw for r in rows[1:] if kernel_prefix in r[0]]
        if vals:
            best = (int(statistics.median(vals)), os.path.basename(path))
    return best


WORKLOAD = ("batched 1080p CV_8UC3 synthetic frame stream: cv::resize ->960x540 INTER_LINEAR + cv::warpAffine "
            "INTER_LINEAR BORDER_CONSTANT rot15/scale0.9 (BASELINE configs[4])")


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
_IMG = [C.c_int, C.c_void_p, C.c_size_t, C.c_int, C.c_int, C.c_void_p, C.c_size_t, C.c_int, C.c_int]


def load_cvref():
    """the reference's own CPU code when oracle/_ref is built, else the single-threaded oracle port"""
    p = os.path.join(ROOT, "oracle", "_ref", "libcvref.so")
    if os.path.exists(p):
        lib, kind, names = C.CDLL(p), "reference", ("cvref_hal_resize", "cvref_hal_warpAffine")
    else:
        so = os.path.join(ROOT, "oracle", "libgeom_oracle.so")
        if not os.path.exists(so):
            subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "-s"])
        lib, kind, names = C.CDLL(so), "port", ("orc_resize", "orc_warpAffine")
    f_resize, f_warp = getattr(lib, names[0]), getattr(lib, names[1])
    f_resize.argtypes = _IMG + [C.c_double, C.c_double, C.c_int]
    f_warp.argtypes = _IMG + [C.c_void_p, C.c_int, C.c_int, C.c_void_p]
    cores = 1
    if kind == "reference":
        lib.cvref_setNumThreads(-1)
        cores = lib.cvref_getNumThreads()
    return f_resize, f_warp, kind, cores


class Checker:
    """Parity before timing: the reference's own CPU code (oracle/_ref/libcvref.so) when it was built, else the
    plain-C restatement (oracle/libgeom_oracle.so).  Used ONLY to compare results here; nothing timed on the GPU arm runs it."""
    SIG = {"warpAffine": _IMG + [C.c_void_p, C.c_int, C.c_int, C.c_void_p], "warpPerspective": _IMG + [C.c_void_p, C.c_int, C.c_int, C.c_void_p],
           "resize": _IMG + [C.c_double, C.c_double, C.c_int],
           "remap32f": _IMG + [C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.c_int, C.c_int, C.c_void_p],
           "remap32fc2": _IMG + [C.c_void_p, C.c_size_t, C.c_int, C.c_int, C.c_void_p],
           "remap16s": _IMG + [C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.c_int, C.c_int, C.c_void_p]}

    def __init__(self):
        p = os.path.join(ROOT, "oracle", "_ref", "libcvref.so")
        if os.path.exists(p):
            self.lib, self.kind = C.CDLL(p), "reference (oracle/_ref/libcvref.so)"
            names = {"warpAffine": "cvref_hal_warpAffine", "warpPerspective": "cvref_hal_warpPerspective", "resize": "cvref_hal_resize",
                     "remap32f": "cvref_remap32f", "remap32fc2": "cvref_remap32fc2", "remap16s": "cvref_remap16s"}
        else:
            so = os.path.join(ROOT, "oracle", "libgeom_oracle.so")
            if not os.path.exists(so):
                subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "-s"])
            self.lib, self.kind = C.CDLL(so), "oracle port (oracle/libgeom_oracle.so)"
            names = {k: "orc_" + k for k in self.SIG}
        self.fn = {}
        for k, sym in names.items():
            f = getattr(self.lib, sym)
            f.argtypes, f.restype = self.SIG[k], C.c_int
            self.fn[k] = f
        self.checked = []

    @staticmethod
    def _img(a, typ):
        h, w = a.shape[:2]
        return [typ, a.ctypes.data, a.strides[0], w, h]

    def call(self, op, typ, src, dst, *rest):
        rc = self.fn[op](*(self._img(src, typ) + self._img(dst, typ)[1:]), *rest)
        assert rc == 0, f"checker {op} rc={rc}"

    def time_undistort_map(self, K, d12, w, h, got1, got2, torch):
        """the reference's cv::initUndistortRectifyMap on this host (CV_16SC2): compares the device maps with it, returns seconds per call
        (None when oracle/_ref is absent or this host lacks AVX2, the dispatch variant the device code reproduces)"""
        f = getattr(self.lib, "cvref_initUndistortRectifyMap", None)
        if f is None or not self.lib.cvref_have_avx2():
            return None
        f.argtypes = [C.c_void_p] * 2 + [C.c_int] + [C.c_void_p] * 2 + [C.c_int] * 3 + [C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p]
        f.restype = C.c_int
        Kc, dc = np.ascontiguousarray(K, np.float64), np.ascontiguousarray(d12, np.float64)
        m1, m2 = np.zeros((h, w, 2), np.int16), np.zeros((h, w), np.uint16)
        iR, tilt = np.zeros(9), np.zeros(9)
        args = (Kc.ctypes.data, dc.ctypes.data, 12, None, Kc.ctypes.data, w, h, 11, m1.ctypes.data, w * 4, m2.ctypes.data, w * 2, iR.ctypes.data, tilt.ctypes.data)
        assert f(*args) == 0
        torch.cuda.synchronize()
        self.same(f"initUndistortRectifyMap {w}x{h} map1", got1.cpu().numpy(), m1)
        self.same(f"initUndistortRectifyMap {w}x{h} map2", got2.cpu().numpy().view(np.uint16), m2)
        t0 = time.perf_counter()
        n = 0
        while time.perf_counter() - t0 < 0.5:
            f(*args)
            n += 1
        return (time.perf_counter() - t0) / n

    def time_undistort(self, K, d12, src_t, got_t, torch):
        """the reference's cv::undistort on this host: compares the device result with it, returns seconds per call (None without oracle/_ref / AVX2)"""
        f = getattr(self.lib, "cvref_undistort", None)
        if f is None or not self.lib.cvref_have_avx2():
            return None
        f.argtypes = [C.c_int, C.c_void_p, C.c_size_t, C.c_int, C.c_int, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
        f.restype = C.c_int
        torch.cuda.synchronize()
        src = src_t.cpu().numpy()
        h, wb = src.shape
        w = wb // 3
        want = np.zeros_like(src)
        Kc, dc = np.ascontiguousarray(K, np.float64), np.ascontiguousarray(d12, np.float64)
        args = (16, src.ctypes.data, wb, w, h, want.ctypes.data, wb, Kc.ctypes.data, dc.ctypes.data, 12, None)
        assert f(*args) == 0
        self.same(f"undistort {w}x{h}", got_t.cpu().numpy(), want)
        t0, n = time.perf_counter(), 0
        while time.perf_counter() - t0 < 0.5:
            f(*args)
            n += 1
        return (time.perf_counter() - t0) / n

    def time_pyrdown(self, typ, src_t, got_t, w, h, cn, dt, torch):
        """the reference's cv::pyrDown on this host: compares the device result with it, returns seconds per call (None without oracle/_ref)"""
        f = getattr(self.lib, "cvref_pyrDown", None)
        if f is None:
            return None
        f.argtypes = [C.c_int, C.c_void_p, C.c_size_t, C.c_int, C.c_int, C.c_void_p, C.c_size_t, C.c_int, C.c_int, C.c_int, C.c_void_p]
        f.restype = C.c_int
        torch.cuda.synchronize()
        src = src_t.cpu().numpy().view(dt).reshape(h, w * cn)
        want = np.zeros((h // 2, (w // 2) * cn), dt)
        args = (typ, src.ctypes.data, src.strides[0], w, h, want.ctypes.data, want.strides[0], w // 2, h // 2, 4, None)
        assert f(*args) == 0
        self.same(f"pyrDown {w}x{h} type {typ}", got_t.cpu().numpy().view(dt).reshape(want.shape), want)
        return self._time_host(f, args)

    def time_pyrup(self, typ, src_t, got_t, w, h, cn, dt, torch):
        """the reference's cv::pyrUp on this host, as time_pyrdown"""
        f = getattr(self.lib, "cvref_pyrUp", None)
        if f is None:
            return None
        f.argtypes = [C.c_int, C.c_void_p, C.c_size_t, C.c_int, C.c_int, C.c_void_p, C.c_size_t, C.c_int, C.c_int, C.c_int]
        f.restype = C.c_int
        torch.cuda.synchronize()
        src = src_t.cpu().numpy().view(dt).reshape(h, w * cn)
        want = np.zeros((h * 2, w * 2 * cn), dt)
        args = (typ, src.ctypes.data, src.strides[0], w, h, want.ctypes.data, want.strides[0], w * 2, h * 2, 4)
        assert f(*args) == 0
        self.same(f"pyrUp {w}x{h} type {typ}", got_t.cpu().numpy().view(dt).reshape(want.shape), want)
        return self._time_host(f, args)

    def time_warppolar(self, src_t, got_t, w, h, dw, dh, cx, cy, rad, flags, torch):
        """the reference's cv::warpPolar on this host (maps + remap), as time_pyrdown"""
        f = getattr(self.lib, "cvref_warpPolar", None)
        if f is None:
            return None
        f.argtypes = [C.c_int, C.c_void_p, C.c_size_t, C.c_int, C.c_int, C.c_void_p, C.c_size_t, C.c_int, C.c_int, C.c_float, C.c_float, C.c_double, C.c_int]
        f.restype = C.c_int
        torch.cuda.synchronize()
        src = src_t.cpu().numpy().reshape(h, w * 3)
        want = np.zeros((dh, dw * 3), np.uint8)
        args = (16, src.ctypes.data, src.strides[0], w, h, want.ctypes.data, want.strides[0], dw, dh, cx, cy, rad, flags)
        assert f(*args) == 0
        self.same(f"warpPolar {w}x{h} -> {dw}x{dh}", got_t.cpu().numpy().reshape(want.shape), want)
        return self._time_host(f, args)

    @staticmethod
    def _time_host(f, args):
        t0, n = time.perf_counter(), 0
        while time.perf_counter() - t0 < 0.5:
            f(*args)
            n += 1
        return (time.perf_counter() - t0) / n

    def same(self, what, got, want, rtol=0.0):
        if got.dtype.kind == "f":
            ok = np.allclose(got, want, rtol=rtol, atol=0.0, equal_nan=True) if rtol else bool(((got == want) | (np.isnan(got) & np.isnan(want))).all())
        else:
            ok = bool((got == want).all())
        assert ok, f"parity before timing FAILED: {what} differs from the {self.kind}"
        self.checked.append(what)


def pipeline_matrix():
    return affine_inverse(MID_W, MID_H, 15.0, 0.9, MID_W / 2.0, MID_H / 2.0)


class CpuPipeline:
    def __init__(self, n_bufs=8, seed=2026):
        self.f_resize, self.f_warp, self.kind, self.cores = load_cvref()
        rng = np.random.default_rng(seed)
        self.srcs = [rng.integers(0, 256, (SRC_H, SRC_W, CN), dtype=np.uint8) for _ in range(n_bufs)]
        self.mids = [np.zeros((MID_H, MID_W, CN), np.uint8) for _ in range(n_bufs)]
        self.dsts = [np.zeros((DST_H, DST_W, CN), np.uint8) for _ in range(n_bufs)]
        self.M = (C.c_double * 6)(*pipeline_matrix().ravel())
        self.bv = (C.c_double * 4)(0, 0, 0, 0)

    def run(self, n_frames):
        t0 = time.perf_counter()
        for i in range(n_frames):
            k = i % len(self.srcs)
            s, m, d = self.srcs[k], self.mids[k], self.dsts[k]
            rc = self.f_resize(CV_8UC3, s.ctypes.data, s.strides[0], SRC_W, SRC_H, m.ctypes.data, m.strides[0], MID_W, MID_H,
                               MID_W / SRC_W, MID_H / SRC_H, INTER_LINEAR)
            assert rc == 0
            rc = self.f_warp(CV_8UC3, m.ctypes.data, m.strides[0], MID_W, MID_H, d.ctypes.data, d.strides[0], DST_W, DST_H,
                             self.M, INTER_LINEAR, B_CONSTANT, self.bv)
            assert rc == 0
        return time.perf_counter() - t0

    def describe(self):
        return ("cv::hal::resize + cv::hal::warpAffine of oracle/_ref (reference sources, pthreads parallel_for_)"
                if self.kind == "reference" else "single-threaded oracle port")


def cpu_baseline(budget_s=10.0):
    cpu = CpuPipeline(n_bufs=32)
    cpu.run(32)                                                 # warm-up (thread pool, page faults)
    t1 = cpu.run(32) / 32
    n = int(max(32, min(20000, budget_s / max(t1, 1e-6))))
    t = cpu.run(n)
    out = {"value": round(n * DST_W * DST_H / t / 1e6, 1), "unit": "Mpix/s", "cores": cpu.cores, "kind": cpu.kind,
           "sample": f"{n} frames of the headline pipeline over 32 rotating host buffers (199 MB of source: past the LLC), {t:.2f} s; {cpu.describe()}",
           "frames_per_s": round(n / t, 1)}
    if cpu.kind == "reference":
        # the same code on ONE thread (cv::setNumThreads(1) == OPENCV_FOR_THREADS_NUM=1)
        lib = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libcvref.so"))
        lib.cvref_setNumThreads(1)
        cpu.run(4)
        n1 = int(max(8, min(2000, 0.3 * budget_s / max(cpu.run(4) / 4, 1e-6))))
        t = cpu.run(n1)
        lib.cvref_setNumThreads(-1)
        out["single_thread"] = {"value": round(n1 * DST_W * DST_H / t / 1e6, 1), "unit": "Mpix/s", "cores": 1,
                                "sample": f"{n1} frames, cv::setNumThreads(1), {t:.2f} s"}
    return out


def shared_config():
    """the `config` both arms print: same workload, same frames per step"""
    return {"workload": WORKLOAD, "frames_per_step_per_gpu": FRAMES_PER_STEP,
            "frame": f"{SRC_W}x{SRC_H} CV_8UC3 -> {MID_W}x{MID_H} -> warpAffine {DST_W}x{DST_H}",
            "l2": "inputs larger than the caches: distinct frames per step, >= 32 rotating buffers"}


def run_reference_arm(args):
    rank, _, world = dist_env()
    if rank != 0:
        return
    cpu = CpuPipeline(n_bufs=32)                                # 199 MB of source frames: every step streams from DRAM
    frames = FRAMES_PER_STEP                                    # the GPU arm's step
    for _ in range(max(args.warmup, 1)):
        cpu.run(frames)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        cpu.run(frames)
    dt = time.perf_counter() - t0
    mpix = args.steps * frames * DST_W * DST_H / dt / 1e6
    sample = f"{frames} frames per step (the GPU arm's step) over 32 rotating host buffers; {cpu.describe()}"
    print(json.dumps({
        "impl": "reference", "metric": "pipeline output Mpix/s", "value": round(mpix, 1), "unit": "Mpix/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(dt / args.steps * 1e3, 3),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": shared_config(), "arm": {"threads": cpu.cores, "code": cpu.describe()},
        "frames_per_s": round(args.steps * frames / dt, 1),
        "cpu_baseline": {"value": round(mpix, 1), "unit": "Mpix/s", "cores": cpu.cores, "kind": cpu.kind, "sample": sample},
        "e2e": {"value": round(mpix, 1), "unit": "Mpix/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


# ------------------------------------------------------------------------------------- GPU arm --
def affine_inverse(w, h, angle, scale, cx, cy):
    """getRotationMatrix2D (imgwarp.cpp:2871-2888) inverted the way cv::warpAffine does (:2401-2411)"""
    a = angle * np.pi / 180
    al, be = scale * np.cos(a), scale * np.sin(a)
    m = np.array([al, be, (1 - al) * cx - be * cy, -be, al, be * cx + (1 - al) * cy], np.float64)
    D = m[0] * m[4] - m[1] * m[3]
    D = 1. / D if D != 0 else 0.
    A11, A22 = m[4] * D, m[0] * D
    m[0] = A11; m[1] *= -D
    m[3] *= -D; m[4] = A22
    b1 = -m[0] * m[2] - m[1] * m[5]
    b2 = -m[3] * m[2] - m[4] * m[5]
    m[2] = b1; m[5] = b2
    return m.reshape(2, 3)


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


def bind_to_gpu_cpus(index):
    """Pins this rank to the CPUs NVML reports as local to its GPU, so that the pinned host frames of the e2e arm are
    allocated on that NUMA node and every rank feeds its GPU over its own memory controllers.  Best effort."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(index)
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (os.cpu_count() + 63) // 64)
        cpus = {64 * i + b for i, w in enumerate(words) for b in range(64) if (w >> b) & 1}
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
            return len(cpus)
    except Exception:
        pass
    return 0


def run_gpu_arm(args):
    import torch
    import torch.distributed as dist
    from opencv_b200 import capi
    from opencv_b200 import dist as bdist
    lib = capi.lib
    rank, local_rank, world = dist_env()
    torch.cuda.set_device(local_rank)
    numa = bind_to_gpu_cpus(local_rank)          # before any pinned allocation: host buffers land on the GPU's NUMA node
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    stream = torch.cuda.current_stream()
    sp = C.c_void_p(stream.cuda_stream)
    peak, peak_src = measured_peak()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---------------- headline: batched device-resident pipeline over a ring of distinct frames ----
    B, R = FRAMES_PER_STEP, RING_CHUNKS
    s_step, m_step = SRC_W * CN, MID_W * CN
    s_fs, m_fs = SRC_H * s_step, MID_H * m_step
    src = torch.empty((R * B, SRC_H, s_step), dtype=torch.uint8, device="cuda")
    mid = torch.empty((B, MID_H, m_step), dtype=torch.uint8, device="cuda")
    dst = torch.empty((R * B, DST_H, m_step), dtype=torch.uint8, device="cuda")
    first, _count = bdist.frame_block(STREAM_FRAMES, world, rank)            # this rank's contiguous block of the 8192-frame stream
    assert lib.b200_synth_frames_dev(src.data_ptr(), s_fs, s_step, SRC_W, SRC_H, CN, first, R * B, 2026, sp) == 0
    M = pipeline_matrix()
    Mc = (C.c_double * 6)(*M.ravel())
    bv = (C.c_double * 4)(0, 0, 0, 0)
    steps = args.steps

    def step(i):
        c = i % R
        rc = lib.b200_resize_warpAffine_batch_dev(CV_8UC3, src.data_ptr() + c * B * s_fs, s_fs, s_step, SRC_W, SRC_H,
                                                  mid.data_ptr(), m_fs, m_step, MID_W, MID_H,
                                                  dst.data_ptr() + c * B * m_fs, m_fs, m_step, DST_W, DST_H,
                                                  B, INTER_LINEAR, Mc, INTER_LINEAR, B_CONSTANT, bv, sp)
        assert rc == 0, capi.last_error()

    # parity before timing: frames 0 and 1 of the first batch, through the very call that is timed below,
    # against the reference's CPU code on the same bytes
    chk = Checker() if rank == 0 else None
    if chk:
        step(0)
        torch.cuda.synchronize()
        for f in (0, 1):
            s_np = src[f].cpu().numpy().reshape(SRC_H, SRC_W, CN)
            got = dst[f].cpu().numpy().reshape(DST_H, DST_W, CN)
            m_ref, d_ref = np.zeros((MID_H, MID_W, CN), np.uint8), np.zeros((DST_H, DST_W, CN), np.uint8)
            chk.call("resize", CV_8UC3, s_np, m_ref, MID_W / SRC_W, MID_H / SRC_H, INTER_LINEAR)
            chk.call("warpAffine", CV_8UC3, m_ref, d_ref, Mc, INTER_LINEAR, B_CONSTANT, bv)
            chk.same(f"headline frame {f}", got, d_ref)
    warm = max(args.warmup, 3)
    for i in range(warm):
        step(i)
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
        time.sleep(0.3)
    l0 = lib.b200_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record(stream)
    for i in range(steps):
        step(i)
    e1.record(stream)
    barrier()
    launches = lib.b200_launch_count() - l0
    # Per-kernel durations: the same steps once more with CUDA events recorded on the launch stream directly around
    # every kernel (b200_profile_begin/end).  Kept out of the region that defines `value` because the extra event
    # records cost ~15 % of a 190 us step; the kernels, data and order are identical.
    k_ms, k_n = (C.c_double * 3)(), (C.c_ulonglong * 3)()
    prof_steps = min(steps, 100)
    if not args.no_kernel_events:
        lib.b200_profile_begin()
        for i in range(prof_steps):
            step(i)
        assert lib.b200_profile_end(k_ms, k_n, 3) == 0
        torch.cuda.synchronize()
    ms = bdist.max_over_ranks(e0.elapsed_time(e1))           # device time, max over ranks
    if rank == 0:
        time.sleep(0.2)
    clocks = sampler.stop() if rank == 0 else None
    ms_per_step = ms / steps
    mpix = world * B * DST_W * DST_H / (ms_per_step * 1e-3) / 1e6
    fp = affine_footprint_px(MID_W, MID_H, DST_W, DST_H, M)
    bytes_resize = B * (SRC_W * SRC_H * CN + MID_W * MID_H * CN)         # per step, all launches of the kind
    bytes_warp = B * (fp * CN + DST_W * DST_H * CN)
    n_r, n_w = max(int(k_n[0]), 1), max(int(k_n[1]), 1)
    # mean duration of ONE launch and the algorithmic bytes ONE launch moves (a step issues n/steps launches per kind)
    kernels = {
        "box2x2_u8_kernel<3> (resize 1080p->540p, batched)": {"seconds": k_ms[0] * 1e-3 / n_r, "bytes_per_launch": bytes_resize * prof_steps / n_r,
                                                               "launches_per_step": n_r / prof_steps},
        "warp_affine_tma_u8<3> (960x540 LINEAR CONSTANT, batched)": {"seconds": k_ms[1] * 1e-3 / n_w, "bytes_per_launch": bytes_warp * prof_steps / n_w,
                                                                            "launches_per_step": n_w / prof_steps},
    }
    if int(k_n[0]) == 0 and int(k_n[1]) > 0:
        # the one-launch pipeline (pipe_box_warp_u8): strips and tiles in one persistent kernel, the half-size frames stay
        # in L2, so the algorithmic HBM bytes of a launch are the source frames in and the destination frames out
        kernels = {"pipe_box_warp_u8<3> (resize 1080p->540p + warpAffine 960x540, one persistent launch per batch)": {
            "seconds": k_ms[1] * 1e-3 / n_w, "bytes_per_launch": B * (SRC_W * SRC_H * CN + DST_W * DST_H * CN) * prof_steps / n_w,
            "launches_per_step": n_w / prof_steps}}
    if args.no_kernel_events:
        kernels = {"pipeline (kernel events disabled)": {"seconds": ms_per_step * 1e-3, "bytes_per_launch": bytes_resize + bytes_warp, "launches_per_step": launches / steps}}
    dom = max(kernels, key=lambda k: kernels[k]["seconds"])
    achieved = kernels[dom]["bytes_per_launch"] / kernels[dom]["seconds"] / 1e9
    traffic, traffic_src = ncu_traffic(dom.split("<")[0])

    result = {
        "metric": "pipeline output Mpix/s", "value": round(mpix, 1), "unit": "Mpix/s", "n_gpus": world,
        "steps": steps, "warmup": warm, "ms_per_step": round(ms_per_step, 4),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": shared_config(),
        "arm": {"launches_per_step": launches / steps,
                "l2": f"ring of {R * B} distinct frames per GPU ({R * B * s_fs / 1e6:.0f} MB), {B * s_fs / 1e6:.0f} MB read per step vs 126 MB L2",
                "parallelism": f"{world} GPU(s), independent frames, no collective"},
        "frames_per_s": round(world * B / (ms_per_step * 1e-3), 1),
        "gpu_launches": int(launches),
        "roofline": {"bound": "hbm", "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s",
                     "frac": round(achieved / peak, 4), "traffic": traffic, "traffic_source": traffic_src, "kernel": dom,
                     "bytes_per_launch": kernels[dom]["bytes_per_launch"], "peak_source": peak_src,
                     "per_kernel": {k: {"us": round(v["seconds"] * 1e6, 2), "launches_per_step": v["launches_per_step"],
                                        "gb_s": round(v["bytes_per_launch"] / v["seconds"] / 1e9, 1),
                                        "frac": round(v["bytes_per_launch"] / v["seconds"] / 1e9 / peak, 4),
                                        "bytes_per_launch": v["bytes_per_launch"]} for k, v in kernels.items()},
                     "pipeline_gb_s": round(sum(v["bytes_per_launch"] * v["launches_per_step"] for v in kernels.values()) / (ms_per_step * 1e-3) / 1e9, 1)},
    }
    if rank == 0:
        result["clocks"] = clocks
        result["parity_before_timing"] = {"against": chk.kind, "headline": "frames 0 and 1 of the timed batch call: bit-exact"}

    # ---------------- e2e: host frames in, host frames out ---------------------------------------
    # (1) the frame-stream entry: one C-ABI call per step, B pinned host frames -> B pinned host results; upload,
    #     kernels and download of consecutive chunks overlap inside the call.  This is `e2e`.
    hsrc = torch.empty((B, SRC_H, s_step), dtype=torch.uint8).pin_memory()
    hdst = torch.empty((B, DST_H, m_step), dtype=torch.uint8).pin_memory()
    hsrc.copy_(src[:B].cpu())

    def stream_step():
        rc = lib.b200_hal_resize_warpAffine_batch(CV_8UC3, hsrc.data_ptr(), s_fs, s_step, SRC_W, SRC_H, MID_W, MID_H,
                                                  hdst.data_ptr(), m_fs, m_step, DST_W, DST_H, B, INTER_LINEAR, Mc,
                                                  INTER_LINEAR, B_CONSTANT, bv, 0)
        assert rc == 0, capi.last_error()
    for _ in range(3):
        stream_step()
    e2e_steps = max(5, min(steps, 20))
    barrier()
    l0 = lib.b200_launch_count()
    step_times = []
    for _ in range(e2e_steps):
        t0 = time.perf_counter()
        stream_step()                                                   # returns with every result frame on the host
        step_times.append(time.perf_counter() - t0)
    e2e_launches = lib.b200_launch_count() - l0
    # synchronous host calls on a shared box: report the median step (one descheduled call otherwise moves the total),
    # max over ranks
    dt = bdist.max_over_ranks(statistics.median(step_times))
    checksum = int(hdst.sum(dtype=torch.int64).item())                  # the result is read on the host
    # the device-resident pipeline on the same frames must give the same bytes
    step(0)
    torch.cuda.synchronize()
    same = bool((dst[:B].cpu() == hdst).all().item())
    result["e2e"] = {"value": round(world * B * DST_W * DST_H / dt / 1e6, 1), "unit": "Mpix/s",
                     "timing": "median wall-clock step (the call is synchronous), max over ranks",
                     "step_ms_min_median_max": [round(min(step_times) * 1e3, 3), round(statistics.median(step_times) * 1e3, 3), round(max(step_times) * 1e3, 3)],
                     "h2d_bytes_per_step": B * SRC_W * SRC_H * CN, "d2h_bytes_per_step": B * DST_W * DST_H * CN,
                     "h2d_gb_s": round(B * SRC_W * SRC_H * CN / dt / 1e9, 1),
                     "api": "b200_hal_resize_warpAffine_batch: pinned host frames in, pinned host frames out, 3-stream chunk pipeline",
                     "frames_per_step": B, "steps": e2e_steps, "gpu_launches": int(e2e_launches),
                     "frames_per_s": round(world * B / dt, 1), "host_checksum": checksum, "equals_device_path": same,
                     "cpus_bound_per_rank": numa}
    del hsrc, hdst

    # (2) the reference's own call pattern: one synchronous cv_hal_resize + cv_hal_warpAffine per frame
    nb, hook_frames = 4, 16
    hs = [torch.empty((SRC_H, SRC_W, CN), dtype=torch.uint8).pin_memory() for _ in range(nb)]
    hm = [torch.empty((MID_H, MID_W, CN), dtype=torch.uint8).pin_memory() for _ in range(nb)]
    hd = [torch.empty((DST_H, DST_W, CN), dtype=torch.uint8).pin_memory() for _ in range(nb)]
    for i, h in enumerate(hs):
        h.copy_(src[i].view(SRC_H, SRC_W, CN).cpu())
    lib.b200_hal_set_min_pixels(0)

    def hook_step():
        for i in range(hook_frames):
            s_, m_, d_ = hs[i % nb], hm[i % nb], hd[i % nb]
            rc = lib.b200_hal_resize(CV_8UC3, s_.data_ptr(), s_step, SRC_W, SRC_H, m_.data_ptr(), m_step, MID_W, MID_H,
                                     MID_W / SRC_W, MID_H / SRC_H, INTER_LINEAR)
            assert rc == 0, capi.last_error()
            rc = lib.b200_hal_warpAffine(CV_8UC3, m_.data_ptr(), m_step, MID_W, MID_H, d_.data_ptr(), m_step, DST_W, DST_H,
                                         Mc, INTER_LINEAR, B_CONSTANT, bv)
            assert rc == 0, capi.last_error()
    hook_step()
    hook_step()
    hook_steps = max(5, min(steps, 15))
    barrier()
    step_times = []
    for _ in range(hook_steps):
        t0 = time.perf_counter()
        hook_step()
        step_times.append(time.perf_counter() - t0)
    dth = bdist.max_over_ranks(statistics.median(step_times))
    result["e2e_hooks"] = {"value": round(world * hook_frames * DST_W * DST_H / dth / 1e6, 1), "unit": "Mpix/s",
                           "frames_per_s": round(world * hook_frames / dth, 1),
                           "h2d_bytes_per_step": hook_frames * (SRC_W * SRC_H * CN + MID_W * MID_H * CN),
                           "d2h_bytes_per_step": hook_frames * (MID_W * MID_H * CN + DST_W * DST_H * CN),
                           "api": "b200_hal_resize + b200_hal_warpAffine (cv_hal_* replacements), pinned host buffers, one frame per call",
                           "hook_policy": "all (b200_hal_set_policy(1), set by the Python binding: both calls run on the GPU; under the C default, "
                                          "\"profitable\", the hook declines this 2x2 resize and OpenCV's CPU code does it)",
                           "frames_per_step": hook_frames, "steps": hook_steps}
    # (3) the same calls on PAGEABLE host memory, which is what a cv::Mat holds: the driver stages every copy itself
    ps = [np.ascontiguousarray(h.numpy()) for h in hs]
    pm = [np.empty((MID_H, MID_W, CN), np.uint8) for _ in range(nb)]
    pd = [np.empty((DST_H, DST_W, CN), np.uint8) for _ in range(nb)]

    def hook_step_pageable():
        for i in range(hook_frames):
            s_, m_, d_ = ps[i % nb], pm[i % nb], pd[i % nb]
            rc = lib.b200_hal_resize(CV_8UC3, s_.ctypes.data, s_step, SRC_W, SRC_H, m_.ctypes.data, m_step, MID_W, MID_H,
                                     MID_W / SRC_W, MID_H / SRC_H, INTER_LINEAR)
            assert rc == 0, capi.last_error()
            rc = lib.b200_hal_warpAffine(CV_8UC3, m_.ctypes.data, m_step, MID_W, MID_H, d_.ctypes.data, m_step, DST_W, DST_H,
                                         Mc, INTER_LINEAR, B_CONSTANT, bv)
            assert rc == 0, capi.last_error()
    hook_step_pageable()
    hook_step_pageable()
    barrier()
    step_times = []
    for _ in range(hook_steps):
        t0 = time.perf_counter()
        hook_step_pageable()
        step_times.append(time.perf_counter() - t0)
    dtp = bdist.max_over_ranks(statistics.median(step_times))
    assert all((pd[i] == hd[i].numpy()).all() for i in range(nb)), "pageable and pinned hook results differ"
    result["e2e_hooks_pageable"] = dict(result["e2e_hooks"], value=round(world * hook_frames * DST_W * DST_H / dtp / 1e6, 1),
                                        frames_per_s=round(world * hook_frames / dtp, 1),
                                        api="b200_hal_resize + b200_hal_warpAffine (cv_hal_* replacements), PAGEABLE host buffers (numpy, as a cv::Mat), one frame per call")
    del hs, hm, hd, ps, pm, pd, src, mid, dst

    # ---------------- other BASELINE configs (N=1 only): informational per-op rooflines ----------
    if world == 1 and not args.no_ops:
        result["e2e_streams"] = run_host_streams(lib, capi, torch, chk)
        result["ops"] = run_ops(lib, capi, torch, sp, stream, peak, chk)
        result["parity_before_timing"]["ops"] = chk.checked[2:]
        result["cpu_baseline"] = cpu_baseline(args.cpu_budget)
        cb = result["cpu_baseline"]["value"]
        for k in ("e2e", "e2e_hooks", "e2e_hooks_pageable"):
            result[k]["ratio_vs_cpu_baseline"] = round(result[k]["value"] / cb, 2) if cb else None
    if rank == 0:
        print(json.dumps(result))
    if world > 1:
        dist.destroy_process_group()


def time_launches(torch, stream, fn, n_sets, iters=48, warm=6, lib=None):
    """returns (device seconds per call, launches back to back between two CUDA events; the same with the library's
    per-kernel event brackets, b200_profile_begin/end).  The first is the figure the tables use: measured on B200 the
    bracketed time is ~2 us LONGER per launch (the event records serialise consecutive kernels' ramp-up and tail), so the
    back-to-back loop is GPU-bound, not host-bound, even for 9 us kernels."""
    for i in range(warm):
        fn(i % n_sets)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for i in range(iters):
        fn(i % n_sets)
    e1.record(stream)
    torch.cuda.synchronize()
    wall = e0.elapsed_time(e1) / iters * 1e-3
    if lib is None:
        return wall, wall
    k_ms, k_n = (C.c_double * 3)(), (C.c_ulonglong * 3)()
    lib.b200_profile_begin()
    for i in range(iters):
        fn(i % n_sets)
    assert lib.b200_profile_end(k_ms, k_n, 3) == 0
    torch.cuda.synchronize()
    return wall, sum(k_ms) * 1e-3 / iters


def run_host_streams(lib, capi, torch, chk):
    """BASELINE configs 3 and 4 as frame streams in pinned HOST memory through the chunked host entries (b200_hal_warpPerspective_batch,
    b200_hal_remap_batch), beside one synchronous hook call per frame and beside the reference on the host threads"""
    out = []
    bv = (C.c_double * 4)(0, 0, 0, 0)

    def wall(fn, reps=3):
        fn()
        ts = []
        for _ in range(reps):
            t0 = time.perf_counter(); fn(); ts.append(time.perf_counter() - t0)
        return statistics.median(ts)

    # config 4: remap 720p 8UC3 INTER_LINEAR REFLECT_101 through one CV_16SC2 + CV_16UC1 map, 64 frames
    w, h, n = 1280, 720, 64
    rng = np.random.default_rng(4)
    src = torch.from_numpy(rng.integers(0, 256, (n, h, w * 3), dtype=np.uint8)).pin_memory()
    dst = torch.empty((n, h, w * 3), dtype=torch.uint8).pin_memory()
    dst2 = torch.empty_like(dst).pin_memory()
    xs, ys = np.meshgrid(np.arange(w, dtype=np.float64), np.arange(h, dtype=np.float64))
    X, Y = (xs - 639.5) / 600.0, (ys - 359.5) / 600.0
    r = np.sqrt(X * X + Y * Y); th = np.arctan(r)
    sc = np.where(r < 1e-8, 1.0, th * (1 - 0.05 * th ** 2) / np.maximum(r, 1e-300))
    mx, my = (600.0 * X * sc + 639.5).astype(np.float32), (600.0 * Y * sc + 359.5).astype(np.float32)
    ix, iy = np.rint(mx * np.float32(32)).astype(np.int64), np.rint(my * np.float32(32)).astype(np.int64)
    m1 = np.ascontiguousarray(np.stack([ix >> 5, iy >> 5], -1).astype(np.int16)); m2 = np.ascontiguousarray(((iy & 31) * 32 + (ix & 31)).astype(np.uint16))
    fs = h * w * 3

    def stream():
        assert lib.b200_hal_remap_batch(CV_8UC3, src.data_ptr(), fs, w * 3, w, h, dst.data_ptr(), fs, w * 3, w, h, n, 11, m1.ctypes.data, w * 4,
                                        m2.ctypes.data, w * 2, INTER_LINEAR, B_REFLECT_101, bv, 0) == 0, capi.last_error()

    def per_frame():
        for f in range(n):
            assert lib.b200_hal_remap16s(CV_8UC3, src.data_ptr() + f * fs, w * 3, w, h, dst2.data_ptr() + f * fs, w * 3, w, h, m1.ctypes.data, w * 4,
                                         m2.ctypes.data, w * 2, INTER_LINEAR, B_REFLECT_101, bv) == 0, capi.last_error()
    ts, tp = wall(stream), wall(per_frame)
    assert torch.equal(dst, dst2), "remap host stream differs from the per-frame hooks"
    row = {"op": "remap 1280x720 8UC3 LINEAR REFLECT_101, CV_16SC2 map, 64 host frames", "stream_mpix_s": round(n * w * h / ts / 1e6, 1),
           "hook_per_frame_mpix_s": round(n * w * h / tp / 1e6, 1), "api": "b200_hal_remap_batch vs b200_hal_remap16s per frame, pinned host memory"}
    f_cpu = chk.fn.get("remap16s")
    if f_cpu is not None and "reference" in chk.kind:
        s_np, d_np = src.numpy().reshape(n, h, w, 3), np.zeros((n, h, w, 3), np.uint8)

        def cpu():
            for f in range(n):
                chk.call("remap16s", CV_8UC3, s_np[f], d_np[f], m1.ctypes.data, w * 4, m2.ctypes.data, w * 2, INTER_LINEAR, B_REFLECT_101, bv)
        tc = wall(cpu, 2)
        chk.same("remap host stream (frame 63)", dst.numpy().reshape(n, h, w, 3)[n - 1], d_np[n - 1])
        row.update(cpu_reference_mpix_s=round(n * w * h / tc / 1e6, 1), stream_vs_cpu=round(tc / ts, 2), cpu_threads=os.cpu_count())
    out.append(row)
    del src, dst, dst2
    # config 3: warpPerspective 4K 8UC4 LINEAR REPLICATE, 8 frames
    w, h, n = 3840, 2160, 8
    src = torch.from_numpy(rng.integers(0, 256, (n, h, w * 4), dtype=np.uint8)).pin_memory()
    dst = torch.empty((n, h, w * 4), dtype=torch.uint8).pin_memory()
    dst2 = torch.empty_like(dst).pin_memory()
    Hi = np.linalg.inv(np.array([[1.12085816, 0.0203087925, 97], [0.0422569112, 1.05413990, -64], [4.77793919e-05, -3.76190421e-06, 1]]))
    Hc = (C.c_double * 9)(*Hi.ravel())
    fs = h * w * 4

    def stream_p():
        assert lib.b200_hal_warpPerspective_batch(CV_8UC4, src.data_ptr(), fs, w * 4, w, h, dst.data_ptr(), fs, w * 4, w, h, n, Hc, INTER_LINEAR, B_REPLICATE, bv, 0) == 0, capi.last_error()

    def per_frame_p():
        for f in range(n):
            assert lib.b200_hal_warpPerspective(CV_8UC4, src.data_ptr() + f * fs, w * 4, w, h, dst2.data_ptr() + f * fs, w * 4, w, h, Hc, INTER_LINEAR, B_REPLICATE, bv) == 0
    ts, tp = wall(stream_p), wall(per_frame_p)
    assert torch.equal(dst, dst2), "warpPerspective host stream differs from the per-frame hooks"
    row = {"op": "warpPerspective 3840x2160 8UC4 LINEAR REPLICATE, 8 host frames", "stream_mpix_s": round(n * w * h / ts / 1e6, 1),
           "hook_per_frame_mpix_s": round(n * w * h / tp / 1e6, 1), "api": "b200_hal_warpPerspective_batch vs b200_hal_warpPerspective per frame, pinned host memory"}
    if "reference" in chk.kind:
        s_np, d_np = src.numpy().reshape(n, h, w, 4), np.zeros((n, h, w, 4), np.uint8)

        def cpu_p():
            for f in range(n):
                chk.call("warpPerspective", CV_8UC4, s_np[f], d_np[f], Hc, INTER_LINEAR, B_REPLICATE, bv)
        tc = wall(cpu_p, 2)
        chk.same("warpPerspective host stream (frame 7)", dst.numpy().reshape(n, h, w, 4)[n - 1], d_np[n - 1])
        row.update(cpu_reference_mpix_s=round(n * w * h / tc / 1e6, 1), stream_vs_cpu=round(tc / ts, 2), cpu_threads=os.cpu_count())
    out.append(row)
    return out


def run_ops(lib, capi, torch, sp, stream, peak, chk):
    """device-resident single-image calls over >=16 rotating (src,dst) sets (> 2x L2), SURVEY 8(d)."""
    ops = []
    bv = (C.c_double * 4)(0, 0, 0, 0)

    def synth(n, h, w, cn):
        t = torch.empty((n, h, w * cn), dtype=torch.uint8, device="cuda")
        assert lib.b200_synth_frames_dev(t.data_ptr(), h * w * cn, w * cn, w, h, cn, 0, n, 7, sp) == 0
        return t

    def record(name, times, out_px, alg_bytes, extra=None):
        secs, ev = times       # back-to-back device time per call; same with per-kernel event brackets (adds ~2 us per launch)
        r = {"op": name, "us": round(secs * 1e6, 2), "us_event_bracketed": round(ev * 1e6, 2), "mpix_s": round(out_px / secs / 1e6, 1),
             "gb_s": round(alg_bytes / secs / 1e9, 1), "frac_of_hbm_peak": round(alg_bytes / secs / 1e9 / peak, 4),
             "algorithmic_bytes": alg_bytes}
        if extra:
            r.update(extra)
        ops.append(r)

    def verify(name, op, typ, s_t, d_t, s_shape, d_shape, *rest, dtype=np.uint8):
        """one call of the op that is about to be timed, compared with the reference CPU code on the same bytes"""
        torch.cuda.synchronize()
        s_np = s_t.cpu().numpy().view(dtype).reshape(s_shape)
        got = d_t.cpu().numpy().view(dtype).reshape(d_shape)
        want = np.zeros_like(got)
        chk.call(op, typ, s_np, want, *rest)
        chk.same(name, got, want)

    # C1 / C1-4K warpAffine 8UC3 LINEAR CONSTANT rot 15 deg, scale 0.9
    for (w, h, n) in ((1920, 1080, 48), (3840, 2160, 16)):
        s, d = synth(n, h, w, 3), torch.empty((n, h, w * 3), dtype=torch.uint8, device="cuda")
        M = affine_inverse(w, h, 15.0, 0.9, w / 2, h / 2)
        Mc = (C.c_double * 6)(*M.ravel())
        fp = affine_footprint_px(w, h, w, h, M)

        def fn(i, s=s, d=d, w=w, h=h, Mc=Mc):
            assert lib.b200_warpAffine_dev(CV_8UC3, s[i].data_ptr(), w * 3, w, h, d[i].data_ptr(), w * 3, w, h, Mc,
                                           INTER_LINEAR, B_CONSTANT, bv, sp) == 0
        fn(0)
        verify(f"warpAffine {w}x{h} 8UC3", "warpAffine", CV_8UC3, s[0], d[0], (h, w, 3), (h, w, 3), Mc, INTER_LINEAR, B_CONSTANT, bv)
        record(f"warpAffine {w}x{h} 8UC3 LINEAR CONSTANT", time_launches(torch, stream, fn, n, lib=lib), w * h, fp * 3 + w * h * 3,
               {"footprint_px": fp})
        del s, d
    # same transform at 4K for 8UC4 and 8UC1 (same kernel family; fewer instructions per byte moved than 8UC3)
    for cn, typ in ((4, CV_8UC4), (1, 0)):
        w, h, n = 3840, 2160, 12
        s, d = synth(n, h, w, cn), torch.empty((n, h, w * cn), dtype=torch.uint8, device="cuda")
        M = affine_inverse(w, h, 15.0, 0.9, w / 2, h / 2)
        Mc = (C.c_double * 6)(*M.ravel())
        fp = affine_footprint_px(w, h, w, h, M)

        def fn(i, s=s, d=d, w=w, h=h, Mc=Mc, cn=cn, typ=typ):
            assert lib.b200_warpAffine_dev(typ, s[i].data_ptr(), w * cn, w, h, d[i].data_ptr(), w * cn, w, h, Mc,
                                           INTER_LINEAR, B_CONSTANT, bv, sp) == 0
        fn(0)
        verify(f"warpAffine 4K 8UC{cn}", "warpAffine", typ, s[0], d[0], (h, w, cn), (h, w, cn), Mc, INTER_LINEAR, B_CONSTANT, bv)
        record(f"warpAffine {w}x{h} 8UC{cn} LINEAR CONSTANT", time_launches(torch, stream, fn, n, lib=lib), w * h, fp * cn + w * h * cn,
               {"footprint_px": fp})
        del s, d
    # 16UC3 and 32FC1 at 4K: the generic gather kernel (no staged fast path for these depths)
    for (typ, cn, dt, tdt, nm) in ((CV_16UC3, 3, np.uint16, torch.int16, "16UC3"), (CV_32FC1, 1, np.float32, torch.float32, "32FC1")):
        w, h, n = 3840, 2160, 8
        esz = 2 if dt == np.uint16 else 4
        s = (torch.rand((n, h, w * cn), device="cuda") * 1000).to(tdt)
        d = torch.empty_like(s)
        M = affine_inverse(w, h, 15.0, 0.9, w / 2, h / 2)
        Mc = (C.c_double * 6)(*M.ravel())
        fp = affine_footprint_px(w, h, w, h, M)

        def fn(i, s=s, d=d, w=w, h=h, Mc=Mc, cn=cn, typ=typ, esz=esz):
            assert lib.b200_warpAffine_dev(typ, s[i].data_ptr(), w * cn * esz, w, h, d[i].data_ptr(), w * cn * esz, w, h, Mc,
                                           INTER_LINEAR, B_CONSTANT, bv, sp) == 0
        fn(0)
        verify(f"warpAffine 4K {nm}", "warpAffine", typ, s[0], d[0], (h, w, cn) if cn > 1 else (h, w), (h, w, cn) if cn > 1 else (h, w),
               Mc, INTER_LINEAR, B_CONSTANT, bv, dtype=dt)
        record(f"warpAffine {w}x{h} {nm} LINEAR CONSTANT", time_launches(torch, stream, fn, n, lib=lib), w * h, (fp + w * h) * cn * esz,
               {"footprint_px": fp})
        del s, d
    # C2a INTER_LINEAR / INTER_AREA (both the 2x2 box), C2c fractional / 3x3 / generic linear
    s = synth(16, 2160, 3840, 3)
    for (dw, dh, interp, nm) in ((1920, 1080, INTER_LINEAR, "LINEAR (2x2 box)"), (1920, 1080, INTER_AREA, "AREA (2x2 box)"), (1600, 900, INTER_AREA, "AREA fractional"),
                                 (1280, 720, INTER_AREA, "AREA 3x3"), (1600, 900, INTER_LINEAR, "LINEAR generic")):
        d = torch.empty((16, dh, dw * 3), dtype=torch.uint8, device="cuda")

        def fn(i, d=d, dw=dw, dh=dh, interp=interp):
            assert lib.b200_resize_dev(CV_8UC3, s[i].data_ptr(), 3840 * 3, 3840, 2160, d[i].data_ptr(), dw * 3, dw, dh,
                                       0.0, 0.0, interp, sp) == 0
        fn(0)
        verify(f"resize 4K->{dw}x{dh} {nm}", "resize", CV_8UC3, s[0], d[0], (2160, 3840, 3), (dh, dw, 3), dw / 3840, dh / 2160, interp)
        record(f"resize 3840x2160->{dw}x{dh} 8UC3 {nm}", time_launches(torch, stream, fn, 16, lib=lib), dw * dh,
               3840 * 2160 * 3 + dw * dh * 3)
        del d
    del s
    # C2b resize 1080p -> 4K 32FC1 LINEAR
    s = torch.rand((16, 1080, 1920), dtype=torch.float32, device="cuda")
    d = torch.empty((16, 2160, 3840), dtype=torch.float32, device="cuda")

    def fn(i):
        assert lib.b200_resize_dev(CV_32FC1, s[i].data_ptr(), 1920 * 4, 1920, 1080, d[i].data_ptr(), 3840 * 4, 3840, 2160,
                                   0.0, 0.0, INTER_LINEAR, sp) == 0
    fn(0)
    verify("resize 1080p->4K 32FC1", "resize", CV_32FC1, s[0], d[0], (1080, 1920), (2160, 3840), 2.0, 2.0, INTER_LINEAR, dtype=np.float32)
    record("resize 1920x1080->3840x2160 32FC1 LINEAR", time_launches(torch, stream, fn, 16, lib=lib), 3840 * 2160,
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
    fn(0)
    verify("warpPerspective 4K 8UC4", "warpPerspective", CV_8UC4, s[0], d[0], (2160, 3840, 4), (2160, 3840, 4), Hc, INTER_LINEAR, B_REPLICATE, bv)
    record("warpPerspective 3840x2160 8UC4 LINEAR REPLICATE", time_launches(torch, stream, fn, 16, lib=lib), 3840 * 2160,
           32845684 + 3840 * 2160 * 4, {"footprint_px": 8211421, "footprint_source": "SURVEY 8(d) C3"})
    nbp = 8

    def fn_p(i):
        assert lib.b200_warpPerspective_batch_dev(CV_8UC4, s.data_ptr() + i * nbp * 2160 * 3840 * 4, 2160 * 3840 * 4, 3840 * 4, 3840, 2160,
                                                  d.data_ptr() + i * nbp * 2160 * 3840 * 4, 2160 * 3840 * 4, 3840 * 4, 3840, 2160, nbp, Hc,
                                                  INTER_LINEAR, B_REPLICATE, bv, sp) == 0
    fn_p(0)
    verify("warpPerspective 4K 8UC4 batch (last frame)", "warpPerspective", CV_8UC4, s[nbp - 1], d[nbp - 1], (2160, 3840, 4), (2160, 3840, 4), Hc, INTER_LINEAR, B_REPLICATE, bv)
    record(f"warpPerspective 3840x2160 8UC4 LINEAR REPLICATE, {nbp} frames per launch", time_launches(torch, stream, fn_p, 2, iters=30, lib=lib),
           nbp * 3840 * 2160, nbp * (32845684 + 3840 * 2160 * 4), {"footprint_px": 8211421})
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
    chk_rest = {"2xCV_32FC1": ("remap32f", (mx.ctypes.data, w * 4, my.ctypes.data, w * 4)),
                "CV_32FC2": ("remap32fc2", (np.ascontiguousarray(np.stack([mx, my], -1)), )),
                "CV_16SC2+CV_16UC1": ("remap16s", (m1.ctypes.data, w * 4, m2.ctypes.data, w * 2))}
    mxy_np = chk_rest["CV_32FC2"][1][0]
    chk_rest["CV_32FC2"] = ("remap32fc2", (mxy_np.ctypes.data, w * 8))
    for nm, (g, map_bytes) in fns.items():
        def fn(i, g=g):
            assert g(i) == 0
        fn(0)
        verify(f"remap 720p {nm}", chk_rest[nm][0], CV_8UC3, s[0], d[0], (h, w, 3), (h, w, 3), *chk_rest[nm][1], INTER_LINEAR, B_REFLECT_101, bv)
        times = time_launches(torch, stream, fn, 64, iters=128, lib=lib)
        record(f"remap 1280x720 8UC3 LINEAR REFLECT_101 maps {nm}", times, w * h, 4428903,
               {"map_bytes_not_in_roofline": map_bytes, "gb_s_incl_maps": round((4428903 + map_bytes) / times[0] / 1e9, 1)})
    # a frame stream through one map: 32 frames per launch, the map is read from HBM once and then served by L2
    nbr = 32
    for nm, (mt, p1, st1, p2, st2, map_bytes) in {"CV_32FC2": (13, t_mxy.data_ptr(), w * 8, None, 0, 7372800),
                                                  "CV_16SC2+CV_16UC1": (11, t_m1.data_ptr(), w * 4, t_m2.data_ptr(), w * 2, 5529600)}.items():
        def fn_m(i, mt=mt, p1=p1, st1=st1, p2=p2, st2=st2):
            assert lib.b200_remap_batch_dev(CV_8UC3, s.data_ptr() + i * nbr * h * w * 3, h * w * 3, w * 3, w, h,
                                            d.data_ptr() + i * nbr * h * w * 3, h * w * 3, w * 3, w, h, nbr, mt, p1, st1, p2, st2,
                                            INTER_LINEAR, B_REFLECT_101, bv, sp) == 0
        fn_m(0)
        verify(f"remap 720p {nm} batch (last frame)", chk_rest[nm][0], CV_8UC3, s[nbr - 1], d[nbr - 1], (h, w, 3), (h, w, 3), *chk_rest[nm][1],
               INTER_LINEAR, B_REFLECT_101, bv)
        record(f"remap 1280x720 8UC3 LINEAR REFLECT_101 maps {nm}, {nbr} frames per launch (one map)",
               time_launches(torch, stream, fn_m, 2, iters=40, lib=lib), nbr * w * h, nbr * 4428903 + map_bytes)
    del s, d
    # 8(f) rank 2: the generators of remap's maps, written straight into device memory (the CPU beside it: the reference's own
    # cv::initUndistortRectifyMap / cv::fisheye::initUndistortRectifyMap on the box's host threads, when oracle/_ref was built)
    Kc = np.array([[1000., 0, 640.3], [0, 1010., 359.8], [0, 0, 1]])
    d14 = np.array([-0.28, 0.09, 0.001, -0.0005, -0.012, 0.01, -0.02, 0.003, 0.001, -0.002, 0.0015, 0.0007, 0., 0.])
    for (w, h) in ((1280, 720), (3840, 2160)):
        Kw = Kc * np.array([[w / 1280.], [h / 720.], [1.]])
        Ia = (C.c_double * 9)()
        assert lib.b200_undistort_params((C.c_double * 9)(*Kw.ravel()), None, C.c_double(0.0), C.c_double(0.0), Ia, None) == 0
        gm1 = torch.empty((4, h, w, 2), dtype=torch.int16, device="cuda")
        gm2 = torch.empty((4, h, w), dtype=torch.int16, device="cuda")
        Ka, Da = (C.c_double * 9)(*Kw.ravel()), (C.c_double * 12)(*d14[:12])

        def fn_u(i, gm1=gm1, gm2=gm2, w=w, h=h, Ka=Ka, Da=Da, Ia=Ia):
            assert lib.b200_initUndistortRectifyMap_dev(Ka, Da, Ia, None, w, h, 11, gm1[i].data_ptr(), w * 4, gm2[i].data_ptr(), w * 2, sp) == 0
        fn_u(0)
        extra = {}
        cpu = chk.time_undistort_map(Kw, d14[:12], w, h, gm1[0], gm2[0], torch)
        if cpu:
            extra = {"cpu_reference_us": round(cpu * 1e6, 1), "cpu_threads": os.cpu_count()}
        record(f"initUndistortRectifyMap {w}x{h} CV_16SC2 (12 coefficients)", time_launches(torch, stream, fn_u, 4, iters=40), w * h, w * h * 6, extra)
        Df = (C.c_double * 4)(-0.05, 0.01, 0.002, -0.001)

        def fn_f(i, gm1=gm1, gm2=gm2, w=w, h=h, Ka=Ka, Ia=Ia):
            assert lib.b200_fisheye_initUndistortRectifyMap_dev(Ka, Df, Ia, w, h, 11, gm1[i].data_ptr(), w * 4, gm2[i].data_ptr(), w * 2, sp) == 0
        fn_f(0)
        record(f"fisheye::initUndistortRectifyMap {w}x{h} CV_16SC2", time_launches(torch, stream, fn_f, 4, iters=40), w * h, w * h * 6)
        if (w, h) == (1280, 720):
            # cv::undistort = per-stripe map + remap, per frame (what a caller without cached maps pays), beside the reference on the host
            su, du = synth(8, h, w, 3), torch.empty((8, h, w * 3), dtype=torch.uint8, device="cuda")

            def fn_ud(i, su=su, du=du, gm1=gm1, gm2=gm2, w=w, h=h, Ka=Ka, Da=Da):
                assert lib.b200_undistort_dev(CV_8UC3, su[i].data_ptr(), w * 3, w, h, du[i].data_ptr(), w * 3, Ka, Da, None, None,
                                              gm1[i % 4].data_ptr(), w * 4, gm2[i % 4].data_ptr(), w * 2, sp) == 0
            fn_ud(0)
            extra = {}
            cpu = chk.time_undistort(Kw, d14[:12], su[0], du[0], torch)
            if cpu:
                extra = {"cpu_reference_us": round(cpu * 1e6, 1), "cpu_threads": os.cpu_count()}
            record(f"undistort {w}x{h} 8UC3 (map build + remap, 2 launches)", time_launches(torch, stream, fn_ud, 8, iters=40), w * h,
                   w * h * 3 * 2 + w * h * 6 * 2, extra)
            del su, du
        del gm1, gm2
    # 8(f) rank 4: cv::pyrDown (the cv_hal_pyrdown replacement's kernel), 4K -> 1080p, beside the reference on the host
    for (dt, tdt, cn, nm, typ) in ((np.uint8, torch.uint8, 3, "8UC3", CV_8UC3), (np.float32, torch.float32, 1, "32FC1", CV_32FC1)):
        w, h, n = 3840, 2160, 8
        esz = np.dtype(dt).itemsize * cn
        s = synth(n, h, w, cn) if dt == np.uint8 else torch.rand((n, h, w), dtype=torch.float32, device="cuda") * 255
        d = torch.empty((n, h // 2, (w // 2) * cn), dtype=tdt, device="cuda")

        def fn_pd(i, s=s, d=d, w=w, h=h, esz=esz, typ=typ):
            assert lib.b200_pyrDown_dev(typ, s[i].data_ptr(), w * esz, w, h, d[i].data_ptr(), (w // 2) * esz, w // 2, h // 2, 4, sp) == 0
        fn_pd(0)
        extra = {}
        cpu = chk.time_pyrdown(typ, s[0], d[0], w, h, cn, dt, torch)
        if cpu:
            extra = {"cpu_reference_us": round(cpu * 1e6, 1), "cpu_threads": os.cpu_count()}
        record(f"pyrDown {w}x{h} {nm} -> {w // 2}x{h // 2}", time_launches(torch, stream, fn_pd, n, iters=40), (w // 2) * (h // 2),
               w * h * esz + (w // 2) * (h // 2) * esz, extra)
        if cn == 3:
            def fn_pdb(i, s=s, d=d, w=w, h=h, esz=esz, typ=typ, n=n):
                assert lib.b200_pyrDown_batch_dev(typ, s.data_ptr(), h * w * esz, w * esz, w, h, d.data_ptr(), (h // 2) * (w // 2) * esz, (w // 2) * esz,
                                                  w // 2, h // 2, n, 4, sp) == 0
            record(f"pyrDown {w}x{h} {nm} -> {w // 2}x{h // 2}, {n} frames per launch", time_launches(torch, stream, fn_pdb, 1, iters=20),
                   n * (w // 2) * (h // 2), n * (w * h * esz + (w // 2) * (h // 2) * esz))
        del s, d
    # cv::warpPolar (forward): polar maps on the device + remap, 1080p -> 1024 radii x 2048 angles
    w, h, dw, dh, n = 1920, 1080, 1024, 2048, 8
    s = synth(n, h, w, 3)
    d = torch.zeros((n, dh, dw * 3), dtype=torch.uint8, device="cuda")
    pmx = torch.empty((dh, dw), dtype=torch.float32, device="cuda"); pmy = torch.empty_like(pmx)

    def fn_wp(i, s=s, d=d):
        assert lib.b200_warpPolar_dev(CV_8UC3, s[i].data_ptr(), w * 3, w, h, d[i].data_ptr(), dw * 3, dw, dh, C.c_float(960.0), C.c_float(540.0),
                                      C.c_double(600.0), INTER_LINEAR | 8, pmx.data_ptr(), dw * 4, pmy.data_ptr(), dw * 4, sp) == 0
    fn_wp(0)
    extra = {}
    cpu = chk.time_warppolar(s[0], d[0], w, h, dw, dh, 960.0, 540.0, 600.0, INTER_LINEAR | 8, torch)
    if cpu:
        extra = {"cpu_reference_us": round(cpu * 1e6, 1), "cpu_threads": os.cpu_count()}
    record(f"warpPolar {w}x{h} 8UC3 -> {dw}x{dh} LINEAR (map build + remap, 2 launches)", time_launches(torch, stream, fn_wp, n, iters=40), dw * dh,
           dw * dh * 3 * 2 + dw * dh * 8 * 2, extra)
    del s, d, pmx, pmy
    # cv::pyrUp (the cv_hal_pyrup replacement's kernel), 1080p -> 4K
    for (dt, tdt, cn, nm, typ) in ((np.uint8, torch.uint8, 3, "8UC3", CV_8UC3), (np.float32, torch.float32, 1, "32FC1", CV_32FC1)):
        w, h, n = 1920, 1080, 8
        esz = np.dtype(dt).itemsize * cn
        s = synth(n, h, w, cn) if dt == np.uint8 else torch.rand((n, h, w), dtype=torch.float32, device="cuda") * 255
        d = torch.empty((n, h * 2, w * 2 * cn), dtype=tdt, device="cuda")

        def fn_pu(i, s=s, d=d, w=w, h=h, esz=esz, typ=typ):
            assert lib.b200_pyrUp_dev(typ, s[i].data_ptr(), w * esz, w, h, d[i].data_ptr(), w * 2 * esz, w * 2, h * 2, 4, sp) == 0
        fn_pu(0)
        extra = {}
        cpu = chk.time_pyrup(typ, s[0], d[0], w, h, cn, dt, torch)
        if cpu:
            extra = {"cpu_reference_us": round(cpu * 1e6, 1), "cpu_threads": os.cpu_count()}
        record(f"pyrUp {w}x{h} {nm} -> {w * 2}x{h * 2}", time_launches(torch, stream, fn_pu, n, iters=40), w * 2 * h * 2,
               w * h * esz + w * 2 * h * 2 * esz, extra)
        if cn == 3:
            def fn_pub(i, s=s, d=d, w=w, h=h, esz=esz, typ=typ, n=n):
                assert lib.b200_pyrUp_batch_dev(typ, s.data_ptr(), h * w * esz, w * esz, w, h, d.data_ptr(), 4 * h * w * esz, w * 2 * esz, w * 2, h * 2, n, 4, sp) == 0
            record(f"pyrUp {w}x{h} {nm} -> {w * 2}x{h * 2}, {n} frames per launch", time_launches(torch, stream, fn_pub, 1, iters=20), n * w * 2 * h * 2,
                   n * (w * h * esz + w * 2 * h * 2 * esz))
        del s, d
    # the two 70 %-target ops at 4K, batched (one launch over 16 / 8 distinct frames): a single 4K frame is a 5-10 us kernel,
    # where launch ramp and tail weigh as much as the memory system
    nb = 16
    s = synth(nb * 2, 2160, 3840, 3)
    d = torch.empty((nb * 2, 1080, 1920 * 3), dtype=torch.uint8, device="cuda")

    def fn_b(i):
        assert lib.b200_resize_batch_dev(CV_8UC3, s.data_ptr() + i * nb * 2160 * 3840 * 3, 2160 * 3840 * 3, 3840 * 3, 3840, 2160,
                                         d.data_ptr() + i * nb * 1080 * 1920 * 3, 1080 * 1920 * 3, 1920 * 3, 1920, 1080, nb, 0.0, 0.0, INTER_LINEAR, sp) == 0
    fn_b(0)
    verify("resize 4K->1080p batch (last frame)", "resize", CV_8UC3, s[nb - 1], d[nb - 1], (2160, 3840, 3), (1080, 1920, 3), 0.5, 0.5, INTER_LINEAR)
    record(f"resize 3840x2160->1920x1080 8UC3 LINEAR (2x2 box), {nb} frames per launch", time_launches(torch, stream, fn_b, 2, iters=30, lib=lib),
           nb * 1920 * 1080, nb * (3840 * 2160 * 3 + 1920 * 1080 * 3))
    del d
    nb = 8
    d = torch.empty((nb * 2, 2160, 3840 * 3), dtype=torch.uint8, device="cuda")
    M4 = affine_inverse(3840, 2160, 15.0, 0.9, 1920.0, 1080.0)
    M4c = (C.c_double * 6)(*M4.ravel())
    fp4 = affine_footprint_px(3840, 2160, 3840, 2160, M4)

    def fn_a(i):
        assert lib.b200_warpAffine_batch_dev(CV_8UC3, s.data_ptr() + i * nb * 2160 * 3840 * 3, 2160 * 3840 * 3, 3840 * 3, 3840, 2160,
                                             d.data_ptr() + i * nb * 2160 * 3840 * 3, 2160 * 3840 * 3, 3840 * 3, 3840, 2160, nb, M4c, INTER_LINEAR,
                                             B_CONSTANT, bv, sp) == 0
    fn_a(0)
    verify("warpAffine 4K 8UC3 batch (last frame)", "warpAffine", CV_8UC3, s[nb - 1], d[nb - 1], (2160, 3840, 3), (2160, 3840, 3), M4c, INTER_LINEAR, B_CONSTANT, bv)
    record(f"warpAffine 3840x2160 8UC3 LINEAR CONSTANT, {nb} frames per launch", time_launches(torch, stream, fn_a, 2, iters=30, lib=lib),
           nb * 3840 * 2160, nb * (fp4 * 3 + 3840 * 2160 * 3), {"footprint_px": fp4})
    del s, d
    # the same with 4 channels (one pixel per 32-bit word: fewer instructions per byte moved)
    s = synth(nb * 2, 2160, 3840, 4)
    d = torch.empty((nb * 2, 2160, 3840 * 4), dtype=torch.uint8, device="cuda")

    def fn_a4(i):
        assert lib.b200_warpAffine_batch_dev(CV_8UC4, s.data_ptr() + i * nb * 2160 * 3840 * 4, 2160 * 3840 * 4, 3840 * 4, 3840, 2160,
                                             d.data_ptr() + i * nb * 2160 * 3840 * 4, 2160 * 3840 * 4, 3840 * 4, 3840, 2160, nb, M4c, INTER_LINEAR,
                                             B_CONSTANT, bv, sp) == 0
    fn_a4(0)
    verify("warpAffine 4K 8UC4 batch (last frame)", "warpAffine", CV_8UC4, s[nb - 1], d[nb - 1], (2160, 3840, 4), (2160, 3840, 4), M4c, INTER_LINEAR, B_CONSTANT, bv)
    record(f"warpAffine 3840x2160 8UC4 LINEAR CONSTANT, {nb} frames per launch", time_launches(torch, stream, fn_a4, 2, iters=30, lib=lib),
           nb * 3840 * 2160, nb * (fp4 * 4 + 3840 * 2160 * 4), {"footprint_px": fp4})
    del s, d
    # dnn::blobFromImages fused with the resize (8(f) rank 4): 32 x 1080p 8UC3 -> 640x640 fp32 NCHW, swapRB, mean, scale
    nbb, bw_, bh_ = 32, 640, 640
    sb = synth(nbb * 2, 1080, 1920, 3)
    db = torch.empty((nbb * 2, 3, bh_, bw_), dtype=torch.float32, device="cuda")
    sc_, mn_ = (C.c_double * 4)(1 / 255., 1 / 255., 1 / 255., 1.0), (C.c_double * 4)(104.0, 117.0, 123.0, 0.0)

    def fn_blob(i):
        assert lib.b200_blobFromImages_dev(CV_8UC3, sb.data_ptr() + i * nbb * 1080 * 1920 * 3, 1080 * 1920 * 3, 1920 * 3, 1920, 1080,
                                           db.data_ptr() + i * nbb * 3 * bh_ * bw_ * 4, bw_, bh_, nbb, sc_, mn_, 1, sp) == 0
    record(f"blobFromImages {nbb}x 1920x1080 8UC3 -> 640x640 fp32 NCHW (resize LINEAR + swapRB + mean + scale, one kernel)",
           time_launches(torch, stream, fn_blob, 2, iters=30, lib=lib), nbb * bw_ * bh_, nbb * (1920 * 1080 * 3 + bw_ * bh_ * 3 * 4))
    del sb, db
    # the two batched kernels of the headline pipeline, each alone and back to back (64 frames per launch, ring of 4 chunks)
    n, R, w, h, mw, mh = 64, 4, 1920, 1080, 960, 540
    s = synth(n * R, h, w, 3)
    mid = torch.empty((n * R, mh, mw * 3), dtype=torch.uint8, device="cuda")
    d = torch.empty((n * R, mh, mw * 3), dtype=torch.uint8, device="cuda")
    M = affine_inverse(mw, mh, 15.0, 0.9, 480.0, 270.0)
    Mc = (C.c_double * 6)(*M.ravel())
    fp = affine_footprint_px(mw, mh, mw, mh, M)

    def fn_r(i):
        assert lib.b200_resize_batch_dev(CV_8UC3, s.data_ptr() + i * n * h * w * 3, h * w * 3, w * 3, w, h,
                                         mid.data_ptr() + i * n * mh * mw * 3, mh * mw * 3, mw * 3, mw, mh, n, 0.0, 0.0, INTER_LINEAR, sp) == 0
    record("resize_batch alone: 64x 1080p->960x540 8UC3 (2x2 box)", time_launches(torch, stream, fn_r, R, iters=80, lib=lib), n * mw * mh,
           n * (w * h * 3 + mw * mh * 3))

    def fn_w(i):
        assert lib.b200_warpAffine_batch_dev(CV_8UC3, mid.data_ptr() + i * n * mh * mw * 3, mh * mw * 3, mw * 3, mw, mh,
                                             d.data_ptr() + i * n * mh * mw * 3, mh * mw * 3, mw * 3, mw, mh, n, Mc, INTER_LINEAR,
                                             B_CONSTANT, bv, sp) == 0
    record("warpAffine_batch alone: 64x 960x540 8UC3 LINEAR CONSTANT", time_launches(torch, stream, fn_w, R, iters=80, lib=lib), n * mw * mh,
           n * (fp * 3 + mw * mh * 3))
    return ops


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=400)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-kernel-events", action="store_true", help="do not bracket kernels with CUDA events in the timed region")
    ap.add_argument("--no-ops", action="store_true", help="skip the informational per-op table and cpu_baseline")
    ap.add_argument("--cpu-budget", type=float, default=10.0, help="seconds of CPU work for cpu_baseline")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_gpu_arm(args)


if __name__ == "__main__":
    main()
