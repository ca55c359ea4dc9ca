"""The parameter grids of the reference's own perf tests for this path (modules/imgproc/perf/perf_warp.cpp:22-227, perf_resize.cpp:36-283,
perf_remap.cpp:12-68, perf_pyramids.cpp:8-68), run three ways on the same inputs: the reference's CPU code on the box's host threads (oracle/_ref), the HAL hooks
on pageable host memory (what cv::warpAffine etc. call when the HAL is linked), and the device-resident entries.  Every GPU result is
compared with the reference before it is timed.  Writes a markdown table (profiles/<name>.md).
usage: python tools/perf_view.py profiles/r2_perf_view.md"""
import ctypes as C
import math
import os
import statistics
import sys
import time

sys.path.insert(0, os.path.join(os.path.dirname(__file__), '..'))
import numpy as np
import torch
import bench
from opencv_b200 import capi

lib = capi.lib
chk = bench.Checker()
assert "reference" in chk.kind, "oracle/_ref/libcvref.so is needed for this view"
st = torch.cuda.current_stream(); sp = C.c_void_p(st.cuda_stream)
NP = {0: np.uint8, 2: np.uint16, 3: np.int16, 5: np.float32}
TD = {0: torch.uint8, 2: torch.uint16, 3: torch.int16, 5: torch.float32}
SZ = {"VGA": (640, 480), "qHD": (960, 540), "720p": (1280, 720), "1080p": (1920, 1080), "2160p": (3840, 2160)}
rows = []


def tname(t):
    return {0: "8U", 2: "16U", 3: "16S", 5: "32F"}[t & 7] + f"C{(t >> 3) + 1}"


def gradient(h, w, t):          # cvtest::fillGradient-like content
    cn, dt = (t >> 3) + 1, NP[t & 7]
    x = np.linspace(0, 255, w)[None, :, None]; y = np.linspace(0, 255, h)[:, None, None]
    a = (x * (np.arange(cn)[None, None, :] + 1) / cn + y) / 2
    return np.ascontiguousarray(a.astype(dt)).reshape((h, w, cn) if cn > 1 else (h, w))


def wall(fn, reps):
    fn()
    ts = []
    for _ in range(reps):
        t0 = time.perf_counter(); fn(); ts.append(time.perf_counter() - t0)
    return statistics.median(ts)


def dev_time(fn, iters=30):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(st)
    for _ in range(iters): fn()
    e1.record(st); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e-3


def run(test, label, op, t, src, dshape, rest_cpu, hook, hook_rest, dev, dev_rest):
    """rest_*: the arguments after the destination image for the three entry kinds"""
    dt = NP[t & 7]
    want = np.zeros(dshape, dt); got = np.zeros(dshape, dt)
    h, w = src.shape[:2]; dh, dw = dshape[:2]
    img = lambda a, b: (t, a.ctypes.data, a.strides[0], w, h, b.ctypes.data, b.strides[0], dw, dh)
    cpu = lambda: chk.call(op, t, src, want, *rest_cpu)
    t_cpu = wall(cpu, 5)
    lib.b200_hal_set_policy(0)                       # the C default: would the hook take this cell?
    taken = hook(*img(src, got), *hook_rest) == 0
    lib.b200_hal_set_policy(1)                       # measured anyway
    rc = hook(*img(src, got), *hook_rest)
    if rc != 0:
        rows.append((test, label, t_cpu, None, None, False)); return
    chk.same(f"{test} {label}", got, want)
    t_hook = wall(lambda: hook(*img(src, got), *hook_rest), 7)
    ds = torch.from_numpy(src).cuda(); dd = torch.zeros(dshape, dtype=TD[t & 7], device="cuda")
    dimg = (t, ds.data_ptr(), src.strides[0], w, h, dd.data_ptr(), got.strides[0], dw, dh)
    assert dev(*dimg, *dev_rest, sp) == 0, capi.last_error()
    torch.cuda.synchronize()
    chk.same(f"{test} {label} (device)", dd.cpu().numpy(), want)
    t_dev = dev_time(lambda: dev(*dimg, *dev_rest, sp))
    rows.append((test, label, t_cpu, t_hook, t_dev, taken))


def rotation_inverse(cx, cy, angle, scale):          # cv::getRotationMatrix2D, inverted as cv::warpAffine does
    a = math.radians(angle); al, be = scale * math.cos(a), scale * math.sin(a)
    M = np.array([[al, be, (1 - al) * cx - be * cy], [-be, al, be * cx + (1 - al) * cy]])
    return bench.invert_2x3(M) if hasattr(bench, "invert_2x3") else np.linalg.inv(np.vstack([M, [0, 0, 1]]))[:2]


bv150 = (C.c_double * 4)(150, 150, 150, 150)
# ---- perf_warp.cpp: WarpAffine, WarpPerspective (512x512 source, rotation 30 deg, scale 2.2)
for t in (0, 24):
    for sz in ("VGA", "720p", "1080p"):
        for interp in (0, 1):
            for border in (0, 1):
                w, h = SZ[sz]
                src = gradient(512, 512, t)
                Mi = np.ascontiguousarray(rotation_inverse(256.0, 256.0, 30.0, 2.2)); Mc = (C.c_double * 6)(*Mi.ravel())
                dshape = (h, w) if t == 0 else (h, w, 4)
                run("WarpAffine", f"{tname(t)} {sz} interp{interp} border{border}", "warpAffine", t, src, dshape, (Mc, interp, border, bv150),
                    lib.b200_hal_warpAffine, (Mc, interp, border, bv150), lib.b200_warpAffine_dev, (Mc, interp, border, bv150))
for sz in ("VGA", "720p", "1080p"):
    for interp in (0, 1):
        for border in (0, 1):
            for cn in (1, 3, 4):
                w, h = SZ[sz]; t = (cn - 1) << 3
                src = gradient(512, 512, t)
                a = math.radians(30.0); al, be = 2.2 * math.cos(a), 2.2 * math.sin(a)
                Hm = np.array([[al, be, (1 - al) * 256 - be * 256], [-be, al, be * 256 + (1 - al) * 256], [.3 / w, .3 / h, 1]])
                Hc = (C.c_double * 9)(*np.linalg.inv(Hm).ravel())
                dshape = (h, w) if cn == 1 else (h, w, cn)
                run("WarpPerspective", f"8UC{cn} {sz} interp{interp} border{border}", "warpPerspective", t, src, dshape, (Hc, interp, border, bv150),
                    lib.b200_hal_warpPerspective, (Hc, interp, border, bv150), lib.b200_warpPerspective_dev, (Hc, interp, border, bv150))
# ---- perf_resize.cpp: ResizeAreaFast (scale 2), ResizeArea (0.1 / 0.25 / 0.81), linear up / down
for t in (0, 16, 24, 2, 18, 26):
    for sz in ("VGA", "qHD", "720p", "1080p"):
        w, h = SZ[sz]; dw, dh = w // 2, h // 2; cn = (t >> 3) + 1
        src = gradient(h, w, t); dshape = (dh, dw) if cn == 1 else (dh, dw, cn)
        a3 = (C.c_double(dw / w), C.c_double(dh / h), 3)
        run("ResizeAreaFast", f"{tname(t)} {sz} /2", "resize", t, src, dshape, a3, lib.b200_hal_resize, a3, lib.b200_resize_dev, a3)
for t in (0, 16, 24):
    for sz in ("VGA", "qHD", "720p", "1080p", "2160p"):
        for scale in (0.1, 0.25, 0.81):
            w, h = SZ[sz]; dw, dh = int(round(w * scale)), int(round(h * scale)); cn = (t >> 3) + 1
            src = gradient(h, w, t); dshape = (dh, dw) if cn == 1 else (dh, dw, cn)
            a3 = (C.c_double(dw / w), C.c_double(dh / h), 3)
            run("ResizeArea", f"{tname(t)} {sz} x{scale}", "resize", t, src, dshape, a3, lib.b200_hal_resize, a3, lib.b200_resize_dev, a3)
for t in (0, 16, 2, 5, 21):
    for (a, b) in (("VGA", "720p"), ("VGA", "qHD"), ("1080p", "VGA"), ("720p", "qHD"), ("2160p", "1080p")):
        (w, h), (dw, dh) = SZ[a], SZ[b]; cn = (t >> 3) + 1
        src = gradient(h, w, t); dshape = (dh, dw) if cn == 1 else (dh, dw, cn)
        a3 = (C.c_double(dw / w), C.c_double(dh / h), 1)
        run("resizeLinear", f"{tname(t)} {a}->{b}", "resize", t, src, dshape, a3, lib.b200_hal_resize, a3, lib.b200_resize_dev, a3)
# ---- perf_remap.cpp: Remap (upside-down map with a half-pixel shift so that INTER_LINEAR blends)
for sz in ("VGA", "1080p"):
    for t in (2, 3, 5, 16):
        for interp in (0, 1):
            w, h = SZ[sz]
            src = gradient(h, w, t); cn = (t >> 3) + 1; dshape = (h, w) if cn == 1 else (h, w, cn)
            xs, ys = np.meshgrid(np.arange(w, dtype=np.float32), np.arange(h, dtype=np.float32))
            mx = np.ascontiguousarray(w - 1 - xs + 0.5); my = np.ascontiguousarray(h - 1 - ys + 0.25)
            rest = (mx.ctypes.data, w * 4, my.ctypes.data, w * 4, interp, 1, bv150)
            dmx, dmy = torch.from_numpy(mx).cuda(), torch.from_numpy(my).cuda()
            drest = (dmx.data_ptr(), w * 4, dmy.data_ptr(), w * 4, interp, 1, bv150)
            run("Remap", f"{tname(t)} {sz} 2xCV_32FC1 interp{interp}", "remap32f", t, src, dshape, rest, lib.b200_hal_remap32f, rest, lib.b200_remap32f_dev, drest)

# ---- perf_pyramids.cpp: pyrDown (1080p, 720p, VGA, QVGA, 127x61), pyrUp (720p, VGA, QVGA, 127x61), BORDER_DEFAULT
SZ.update({"QVGA": (320, 240), "ODD": (127, 61)})
PYR_SIG = [C.c_int, C.c_void_p, C.c_size_t, C.c_int, C.c_int, C.c_void_p, C.c_size_t, C.c_int, C.c_int, C.c_int]


def run_pyr(up, sz, t):
    name = "pyrUp" if up else "pyrDown"
    w, h = SZ[sz]; cn = (t >> 3) + 1; dt = NP[t & 7]
    dw, dh = (2 * w, 2 * h) if up else ((w + 1) // 2, (h + 1) // 2)
    src = gradient(h, w, t); dshape = (dh, dw) if cn == 1 else (dh, dw, cn)
    want = np.zeros(dshape, dt); got = np.zeros(dshape, dt)
    ref = getattr(chk.lib, "cvref_" + name); ref.restype = C.c_int
    ref.argtypes = PYR_SIG + ([] if up else [C.c_void_p])
    tail = () if up else (None,)
    cpu = lambda: ref(t, src.ctypes.data, src.strides[0], w, h, want.ctypes.data, want.strides[0], dw, dh, 4, *tail)
    assert cpu() == 0
    t_cpu = wall(cpu, 5)
    hook = lib.b200_hal_pyrup if up else lib.b200_hal_pyrdown
    hk = lambda: hook(src.ctypes.data, src.strides[0], w, h, got.ctypes.data, got.strides[0], dw, dh, t & 7, cn, 4)
    label = f"{tname(t)} {sz}"
    lib.b200_hal_set_policy(0)                       # the C default: would the hook take this cell?
    taken = hk() == 0
    lib.b200_hal_set_policy(1)                       # measured anyway
    if hk() != 0:
        rows.append((name, label, t_cpu, None, None, False)); return
    chk.same(f"{name} {label}", got, want)
    t_hook = wall(hk, 7)
    ds = torch.from_numpy(src).cuda(); dd = torch.zeros(dshape, dtype=TD[t & 7], device="cuda")
    dev = lib.b200_pyrUp_dev if up else lib.b200_pyrDown_dev
    dv = lambda: dev(t, ds.data_ptr(), src.strides[0], w, h, dd.data_ptr(), got.strides[0], dw, dh, 4, sp)
    assert dv() == 0, capi.last_error()
    torch.cuda.synchronize()
    chk.same(f"{name} {label} (device)", dd.cpu().numpy(), want)
    rows.append((name, label, t_cpu, t_hook, dev_time(dv), taken))


for up, sizes in ((False, ("1080p", "720p", "VGA", "QVGA", "ODD")), (True, ("720p", "VGA", "QVGA", "ODD"))):
    for sz in sizes:
        for t in (0, 16, 24, 3, 19, 27, 5, 21, 29):
            run_pyr(up, sz, t)

out = sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/perf_view.md"
with open(out, "w") as f:
    f.write("# The reference's perf-test grids on this box (tools/perf_view.py)\n\n")
    f.write(f"CPU = the reference's own code (oracle/_ref) on {os.cpu_count()} host threads; hooks = `b200_hal_*` on pageable host memory (upload + kernel + download per call,\n"
            "what `cv::warpAffine` etc. run when the HAL is linked); device = `b200_*_dev` on resident images.  Every GPU result was compared with the reference first.\n"
            "`declined` = the hook returned NOT_IMPLEMENTED (below the pixel gate, or a cell left to the CPU): OpenCV would run its CPU path.\n"
            "`default policy`: whether the hook takes the cell under its default policy (`b200_hal_set_policy(0)`); the times are measured with policy `all`.\n\n")
    groups = {}
    for r in rows: groups.setdefault(r[0], []).append(r)
    f.write("| test (perf file) | cases | CPU ms (geo-mean) | hooks ms | device ms | hooks vs CPU | device vs CPU | hook wins / cases | taken by default | of those, wins | losing cells taken by default |\n|---|---|---|---|---|---|---|---|---|---|---|\n")
    gm = lambda v: math.exp(sum(math.log(x) for x in v) / len(v)) if v else float("nan")
    for g, rs in groups.items():
        ok = [r for r in rs if r[3] is not None]
        f.write(f"| {g} | {len(rs)} | {gm([r[2] for r in ok]) * 1e3:.3f} | {gm([r[3] for r in ok]) * 1e3:.3f} | {gm([r[4] for r in ok]) * 1e3:.4f} | "
                f"{gm([r[2] / r[3] for r in ok]):.2f}x | {gm([r[2] / r[4] for r in ok]):.0f}x | {sum(1 for r in ok if r[3] < r[2])} / {len(rs)} | "
                f"{sum(1 for r in ok if r[5])} | {sum(1 for r in ok if r[5] and r[3] < r[2])} | {sum(1 for r in ok if r[5] and r[3] >= r[2] * 1.1)} |\n")
    f.write("\n| test | case | CPU ms | hooks ms | device ms | hooks vs CPU | default policy |\n|---|---|---|---|---|---|---|\n")
    for r in rows:
        if r[3] is None: f.write(f"| {r[0]} | {r[1]} | {r[2] * 1e3:.3f} | declined | | | |\n")
        else: f.write(f"| {r[0]} | {r[1]} | {r[2] * 1e3:.3f} | {r[3] * 1e3:.3f} | {r[4] * 1e3:.4f} | {r[2] / r[3]:.2f}x | {'GPU' if r[5] else 'CPU'} |\n")
print(open(out).read()[:3000])
