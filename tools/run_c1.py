#!/usr/bin/env python
"""tools/run_c1.py -- BASELINE configs[0] (random sparse QP n=1000 m=500 1 %) on the GPU and, when oracle/_ref
is present, with the unmodified reference on this box's CPU: solve time, iterations, parity."""
import os
import sys
import time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import qpdo_b200
from qpdo_b200 import problems, capi

p = problems.random_qp(1000, 500, 0.01, seed=1)
s = qpdo_b200.QPDO()
t0 = time.time(); s.setup(p["Q"], p["q"], p["A"], p["l"], p["u"], dict(verbose=0)); ts = time.time() - t0
best = 1e9
for _ in range(5):
    s.warm_start(None, None)
    t0 = time.time(); r = s.solve(); best = min(best, time.time() - t0)
print("gpu: setup %.1f ms, solve %.2f ms (best of 5), status %d iter %d/%d obj %.12g" % (ts * 1e3, best * 1e3, r.status_val, r.iterations, r.oterations, r.objective))
s.delete()
try:
    import oracle
    if oracle.have_ref():
        os.environ.setdefault("OPENBLAS_NUM_THREADS", str(os.cpu_count()))
        q = capi.Solver(oracle.ref_library())
        t0 = time.time(); q.setup(p["Q"], p["q"], p["A"], p["l"], p["u"], dict(verbose=0)); ts = time.time() - t0
        bestr = 1e9
        for _ in range(3):
            q.warm_start(None, None)
            t0 = time.time(); o = q.solve(); bestr = min(bestr, time.time() - t0)
        print("ref: setup %.1f ms, solve %.2f ms (best of 3, %d cores), status %d iter %d/%d obj %.12g" % (ts * 1e3, bestr * 1e3, os.cpu_count(), o.status_val, o.iterations, o.oterations, o.objective))
        print("x rel err %.2e  active set equal %s" % (np.abs(r.x - o.x).max() / max(1, np.abs(o.x).max()), np.array_equal(r.active_set, o.active_set)))
        q.delete()
except ImportError:
    pass
