#!/usr/bin/env python
"""tools/exp_phase.py [count] -- per-phase clock64 accounting of the batch kernel (qpdo_batch_timing) and throughput,
for the current build and environment (QPDO_BATCH_CTAS_PER_SM, ...)."""
import ctypes as C
import os
import sys
import time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from qpdo_b200 import problems, batch as qbatch

B = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
b = problems.mpc_batch(B)
names = ["setup+scaling", "residuals+outer", "newton prep", "K assembly", "factor+forward", "backward", "direction spmv", "linesearch+update"]
bt, vals = qbatch.from_mpc(b)
bt.solve(*vals)
t0 = time.perf_counter()
r = bt.solve(*vals)
dt = time.perf_counter() - t0
print(bt.info(), "%.1f ms, %.0f QP/s (host buffers)" % (dt * 1e3, B / dt), "iters mean %.2f oter mean %.2f" % (r[2]["iterations"].mean(), r[2]["oterations"].mean()),
      "objsum %.10f" % r[2]["objective"][r[2]["status_val"] == 1].sum())
bt.lib.qpdo_batch_timing.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_double)]
bt.lib.qpdo_batch_timing(bt.h, 0, None)
r = bt.solve(*vals)
cyc = (C.c_double * 8)()
bt.lib.qpdo_batch_timing(bt.h, B, cyc)
tot = sum(cyc)
print(" ".join("%s %.0fk (%.0f%%)" % (nm, c / 1e3, 100 * c / tot) for nm, c in zip(names, cyc)), "| total %.0fk cycles/QP" % (tot / 1e3))
bt.destroy()
if hasattr(bt.lib, "qpdo_batch_fine"):
    f8 = (C.c_longlong * 8)()
    bt.lib.qpdo_batch_fine(f8)
    npan = max(1, f8[4])
    print("fine (warp 0 of CTA, per W=4 panel step): phase A %.0f, barrier %.0f, phase B %.0f, barrier %.0f cycles; %d panels" % (f8[0] / npan, f8[1] / npan, f8[2] / npan, f8[3] / npan, npan))
    nf = max(1, f8[7])
    print("fine backward per front: prologue + rows below %.0f, pivot-block steps %.0f cycles; %d fronts" % (f8[5] / nf, f8[6] / nf, nf))
