"""Reader for tests/golden/hotpath_golden.npz (written by tests/golden/make_golden.py)."""
import os

import numpy as np

_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "hotpath_golden.npz")


def load_golden():
    z = np.load(_PATH)
    recs = []
    for i in range(int(z["n"])):
        key = f"c{i:03d}"
        rec = {"op": str(z[key + "_op"]), "params": z[key + "_params"], "dst": z[key + "_dst"]}
        for k in ("src", "mapx", "mapy", "map1", "map2"):
            if f"{key}_{k}" in z.files:
                rec[k] = np.ascontiguousarray(z[str(z[f"{key}_{k}"])])
        recs.append(rec)
    return recs


def run_record(impl, rec):
    """Runs one golden record through an implementation of the hook ABI (tests.common.HookLib-like)."""
    p = rec["params"]
    dst = np.zeros_like(rec["dst"])
    op = rec["op"]
    if op == "resize":
        rc = impl.resize(rec["src"], dst, int(p[0]))
    else:
        interp, border, bv = int(p[0]), int(p[1]), tuple(p[2:6])
        if op == "warpAffine":
            rc = impl.warpAffine(rec["src"], dst, p[6:12].reshape(2, 3), interp, border, bv)
        elif op == "warpPerspective":
            rc = impl.warpPerspective(rec["src"], dst, p[6:15].reshape(3, 3), interp, border, bv)
        elif op == "remap32f":
            rc = impl.remap32f(rec["src"], dst, rec["mapx"], rec["mapy"], interp, border, bv)
        elif op == "remap16s":
            rc = impl.remap16s(rec["src"], dst, rec["map1"], rec["map2"], interp, border, bv)
        else:
            raise ValueError(op)
    assert rc == 0, (op, rc)
    return dst
