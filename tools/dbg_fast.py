import sys, os
sys.path.insert(0, os.path.join(os.path.dirname(__file__), '..')); sys.path.insert(0, os.path.join(os.path.dirname(__file__), '..', 'tests'))
import numpy as np, torch
from common import *
from opencv_b200 import capi
capi.lib.b200_hal_set_min_pixels(0)
G = HookLib(capi.lib, {k: "b200_hal_" + k for k in SIGS})
O = oracle()
rng = np.random.default_rng(0)
cn = int(sys.argv[1]) if len(sys.argv) > 1 else 3
src = rand_image(rng, 270, 480, cn, np.uint8)
M = rotation_inverse(480, 270, 15, 0.9)
d = np.zeros_like(src); r = np.zeros_like(src)
rc = G.warpAffine(src, d, M, INTER_LINEAR, BORDER_CONSTANT)
print("rc", rc, capi.last_error())
O.warpAffine(src, r, M, INTER_LINEAR, BORDER_CONSTANT)
print("mismatch", int((d != r).sum()))
