#!/usr/bin/env python
"""tools/batch_ab.py [count] -- A/B of the batch kernel's two factorization paths on the GPU: the
supernodal register path must reproduce the column path bit for bit (same fma sequence per entry)."""
import os
import sys
import time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from qpdo_b200 import problems, batch as qbatch

B = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
b = problems.mpc_batch(B)
res = {}
for mode in ("columns", "supernodes"):
    os.environ["QPDO_BATCH_FACTOR"] = mode
    bt, vals = qbatch.from_mpc(b)
    bt.solve(*vals)
    t0 = time.perf_counter()
    res[mode] = bt.solve(*vals)
    print(mode, bt.info(), "%.1f ms, %.0f QP/s (host buffers)" % ((time.perf_counter() - t0) * 1e3, B / (time.perf_counter() - t0)))
    bt.destroy()
xa, ya, sa, aa = res["columns"]
xb, yb, sb, ab = res["supernodes"]
print("x bitwise equal:", np.array_equal(xa, xb), " y:", np.array_equal(ya, yb), " summaries:", np.array_equal(sa, sb), " active:", np.array_equal(aa, ab))
print("max |dx|", np.abs(xa - xb).max(), "iteration mismatches", int((sa["iterations"] != sb["iterations"]).sum()))
exact = np.array_equal(xa, xb) and np.array_equal(ya, yb) and np.array_equal(sa, sb) and np.array_equal(aa, ab)
if not exact:
    # fast-pivot build (default): the front path rounds differently from the column path; same solver, same answers
    ok = sa["status_val"] == 1
    rel = np.abs(xa[ok] - xb[ok]).max(axis=1) / np.maximum(1.0, np.abs(xa[ok]).max(axis=1))
    print("not bit-identical (fast pivots): max rel |dx| %.3e, status mismatches %d, iteration mismatches %d of %d" %
          (rel.max(), int((sa["status_val"] != sb["status_val"]).sum()), int((sa["iterations"] != sb["iterations"]).sum()), B))
    assert rel.max() <= 1e-8 and (sa["status_val"] != sb["status_val"]).sum() == 0 and (sa["iterations"] != sb["iterations"]).sum() <= max(2, B // 2000)
print("AB_OK", "(bit-identical)" if exact else "(to 1e-8)")

# per-phase cycle accounting (qpdo_batch_timing) of the front paths
import ctypes as C
names = ["setup+scaling", "residuals+outer", "newton prep", "K assembly", "factor+forward", "backward", "direction spmv", "linesearch+update"]
for mode in ("supernodes",):
    os.environ["QPDO_BATCH_FACTOR"] = mode
    bt, vals = qbatch.from_mpc(b)
    bt.lib.qpdo_batch_timing.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_double)]
    bt.lib.qpdo_batch_timing(bt.h, 0, None)
    r = bt.solve(*vals)
    assert np.array_equal(r[0], xb), mode
    cyc = (C.c_double * 8)()
    bt.lib.qpdo_batch_timing(bt.h, B, cyc)
    tot = sum(cyc)
    print(mode, " ".join("%s %.0fk (%.0f%%)" % (nm, c / 1e3, 100 * c / tot) for nm, c in zip(names, cyc)), "| total %.0fk cycles/QP" % (tot / 1e3))
    bt.destroy()
