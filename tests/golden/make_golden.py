"""Generates tests/golden/hotpath_golden.npz from the REFERENCE ITSELF.

Run in the build container (needs oracle/_ref/libcvref.so, i.e. /root/reference compiled by
oracle/build_ref.sh):   python tests/golden/make_golden.py

Every record stores the exact inputs and the reference's output, so the fixtures are usable on the
GPU box where neither /root/reference nor (necessarily) oracle/_ref exist.
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from common import *  # noqa: E402,F403


def main():
    R = cvref()
    assert R is not None, "build oracle/_ref first (bash oracle/build_ref.sh)"
    rng = np.random.default_rng(20261106)
    out = {}
    n = 0
    dts = {"u8": np.uint8, "u16": np.uint16, "f32": np.float32}
    bv = (17.0, 130.5, 250.0, 3.0)

    shared = {}

    def put(op, params, arrays, dst):
        # inputs are de-duplicated: record stores the name of a shared array
        nonlocal n
        key = f"c{n:03d}"
        out[key + "_op"] = np.array(op)
        out[key + "_params"] = np.array(params, np.float64)
        for k, v in arrays.items():
            sid = shared.setdefault(id(v), f"s{len(shared):03d}")
            out[sid] = v
            out[f"{key}_{k}"] = np.array(sid)
        out[key + "_dst"] = dst
        n += 1

    mx0 = mx = rng.uniform(-8, 37 + 8, (25, 34)).astype(np.float32)
    my0 = my = rng.uniform(-8, 23 + 8, (25, 34)).astype(np.float32)
    m10, m20 = convert_maps_fixed(mx0, my0)

    for dn, dt in dts.items():
        for cn in (1, 3, 4):
            sh, sw, dh, dw = 23, 37, 25, 34
            src = rand_image(rng, sh, sw, cn, dt)
            dshape = (dh, dw) if cn == 1 else (dh, dw, cn)
            for interp in (INTER_NEAREST, INTER_LINEAR):
                border = [BORDER_CONSTANT, BORDER_REPLICATE, BORDER_REFLECT_101, BORDER_WRAP][n % 4]
                M = rotation_inverse(sw, sh, rng.uniform(-180, 180), rng.uniform(0.6, 1.5))
                d = np.zeros(dshape, dt); assert R.warpAffine(src, d, M, interp, border, bv) == 0
                put("warpAffine", [interp, border, *bv, *M.ravel()], {"src": src}, d)
                border = [BORDER_CONSTANT, BORDER_REPLICATE, BORDER_REFLECT_101, BORDER_WRAP][n % 4]
                H = rand_homography_inverse(rng, dw, dh, 0.3)
                d = np.zeros(dshape, dt); assert R.warpPerspective(src, d, H, interp, border, bv) == 0
                put("warpPerspective", [interp, border, *bv, *H.ravel()], {"src": src}, d)
                border = [BORDER_CONSTANT, BORDER_REPLICATE, BORDER_REFLECT_101, BORDER_WRAP][n % 4]
                mx, my = mx0, my0
                d = np.zeros(dshape, dt); assert R.remap32f(src, d, mx, my, interp, border, bv) == 0
                put("remap32f", [interp, border, *bv], {"src": src, "mapx": mx, "mapy": my}, d)
                m1, m2 = m10, m20
                border = [BORDER_CONSTANT, BORDER_REPLICATE, BORDER_REFLECT_101, BORDER_WRAP][n % 4]
                d = np.zeros(dshape, dt); assert R.remap16s(src, d, m1, m2, interp, border, bv) == 0
                put("remap16s", [interp, border, *bv], {"src": src, "map1": m1, "map2": m2}, d)
            for (rh, rw) in ((11, 18), (9, 16), (46, 74), (7, 37), (30, 45)):   # 37x23 -> ...
                for interp in (INTER_NEAREST, INTER_LINEAR, INTER_AREA) + ((5,) if dn in ("u8", "u16") else ()) + (6,):   # 5/6: *_EXACT
                    d = np.zeros((rh, rw) if cn == 1 else (rh, rw, cn), dt)
                    assert R.resize(src, d, interp) == 0
                    put("resize", [interp], {"src": src}, d)
            src2 = rand_image(rng, 24, 36, cn, dt)
            for (rh, rw) in ((12, 18), (8, 12), (12, 12)):                            # exact 2x, 3x, 2x3
                for interp in (INTER_LINEAR, INTER_AREA):
                    d = np.zeros((rh, rw) if cn == 1 else (rh, rw, cn), dt)
                    assert R.resize(src2, d, interp) == 0
                    put("resize", [interp], {"src": src2}, d)
    # INTER_CUBIC (2) / INTER_LANCZOS4 (4) records, appended after the original ones so those keep their numbering
    for dn, dt in dts.items():
        for cn in (1, 3):
            sh, sw, dh, dw = 23, 37, 25, 34
            src = rand_image(rng, sh, sw, cn, dt)
            dshape = (dh, dw) if cn == 1 else (dh, dw, cn)
            for interp in (2, 4):
                border = [BORDER_CONSTANT, BORDER_REPLICATE, BORDER_REFLECT_101, BORDER_WRAP, BORDER_REFLECT][n % 5]
                M = rotation_inverse(sw, sh, rng.uniform(-180, 180), rng.uniform(0.6, 1.5))
                d = np.zeros(dshape, dt); assert R.warpAffine(src, d, M, interp, border, bv) == 0
                put("warpAffine", [interp, border, *bv, *M.ravel()], {"src": src}, d)
                border = [BORDER_CONSTANT, BORDER_REPLICATE, BORDER_REFLECT_101, BORDER_WRAP, BORDER_REFLECT][n % 5]
                H = rand_homography_inverse(rng, dw, dh, 0.3)
                d = np.zeros(dshape, dt); assert R.warpPerspective(src, d, H, interp, border, bv) == 0
                put("warpPerspective", [interp, border, *bv, *H.ravel()], {"src": src}, d)
                border = [BORDER_CONSTANT, BORDER_REPLICATE, BORDER_REFLECT_101, BORDER_WRAP, BORDER_REFLECT][n % 5]
                d = np.zeros(dshape, dt); assert R.remap32f(src, d, mx0, my0, interp, border, bv) == 0
                put("remap32f", [interp, border, *bv], {"src": src, "mapx": mx0, "mapy": my0}, d)
                border = [BORDER_CONSTANT, BORDER_REPLICATE, BORDER_REFLECT_101, BORDER_WRAP, BORDER_REFLECT][n % 5]
                d = np.zeros(dshape, dt); assert R.remap16s(src, d, m10, m20, interp, border, bv) == 0
                put("remap16s", [interp, border, *bv], {"src": src, "map1": m10, "map2": m20}, d)
    # INTER_CUBIC resize records (vector body and scalar tail of the vertical pass both occur: widths 8k+r)
    for dn, dt in dts.items():
        for cn in (1, 3, 4):
            src = rand_image(rng, 23, 37, cn, dt)
            for (rh, rw) in ((11, 18), (46, 74), (30, 45), (9, 7)):
                d = np.zeros((rh, rw) if cn == 1 else (rh, rw, cn), dt)
                assert R.resize(src, d, INTER_CUBIC) == 0
                put("resize", [INTER_CUBIC], {"src": src}, d)
    # INTER_LANCZOS4 resize records
    for dn, dt in dts.items():
        for cn in (1, 3):
            src = rand_image(rng, 23, 37, cn, dt)
            for (rh, rw) in ((11, 18), (46, 74), (9, 7)):
                d = np.zeros((rh, rw) if cn == 1 else (rh, rw, cn), dt)
                assert R.resize(src, d, 4) == 0
                put("resize", [4], {"src": src}, d)
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "hotpath_golden.npz")
    np.savez_compressed(path, n=np.array(n), **out)
    print("wrote", path, n, "records", os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
