import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import jplspatial_b200 as J
from oracle import refpy
from tests import util
from tests.test_gpu_reverb import *
ref = refpy.Ref("sse2"); ctx = J.Context(0)
n, L, frames, blocks = 2, 16, 512, 8
x = np.zeros((blocks, n, frames), np.float32); x[0, :, 0] = 1.0
rts = {}
want = run_reference(ref, x, rts, L, frames)
got, _ = run_gpu(ctx, x, rts, L, frames, J.MODE_PARITY)
d = (got.view(np.uint32) != want.view(np.uint32))
idx = np.argwhere(d)
print("mismatches", len(idx), "first", idx[:5])
if len(idx):
    b, i, f, c = idx[0]
    print("want", want[b, i, f], "\ngot ", got[b, i, f])
    # first nonzero frame
nz = np.argwhere(np.abs(want) > 0)
print("first nonzero", nz[0], want[tuple(nz[0][:3])])
print("got there    ", got[tuple(nz[0][:3])])
