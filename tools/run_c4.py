#!/usr/bin/env python
"""tools/run_c4.py [npath] -- BASELINE configs[3] at full size on the GPU: Lasso-dual QP n=200k m=100k,
regularisation path with qpdo_update_bounds + qpdo_warm_start + qpdo_solve per step.  Prints one line
per path step and checks the KKT conditions of every solution (size-independent property)."""
import ctypes as C
import os
import sys
import time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import qpdo_b200
from qpdo_b200 import problems, capi

npath = int(sys.argv[1]) if len(sys.argv) > 1 else 4
t0 = time.time()
p = problems.lasso_dual()
print("generated C4 in %.1f s: n=%d m=%d nnzA=%d" % (time.time() - t0, p["q"].size, p["A"].shape[0], p["A"].nnz), flush=True)
s = qpdo_b200.QPDO()
t0 = time.time()
s.setup(p["Q"], p["q"], p["A"], p["l"], p["u"], dict(verbose=0))
print("setup %.2f s" % (time.time() - t0), flush=True)
lib = qpdo_b200.lib()
x = y = None
A = p["A"].tocsr()
for j, lam in enumerate(p["lams"][:npath]):
    if j:
        s.update_bounds(np.full(A.shape[0], -lam), np.full(A.shape[0], lam))
        s.warm_start(x, y)
    t0 = time.time()
    r = s.solve()
    dt = time.time() - t0
    x, y = r.x.copy(), r.y.copy()
    # KKT: Qx + q + A'y = 0, y in normal cone of [l,u] at Ax
    rd = np.abs(x + p["q"] + A.T @ y).max()
    Ax = A @ x
    rp = np.maximum(np.maximum(Ax - lam, -lam - Ax), 0).max()
    comp = np.abs(np.where(y > 0, Ax - lam, np.where(y < 0, Ax + lam, 0.0)) * (np.abs(y) > 1e-9)).max()
    print("path %2d lam %.4g: status %d iter %d/%d obj %.10g  %.2f s  | dual res %.2e prim viol %.2e compl %.2e nnz(y) %d" %
          (j, lam, r.status_val, r.iterations, r.oterations, r.objective, dt, rd, rp, comp, int((np.abs(y) > 1e-9).sum())), flush=True)
    assert r.status_val == 1 and rd < 1e-4 and rp < 1e-4
    info = capi.DevSymbolicInfo()
    lib.qpdo_dev_get_symbolic_info(s.device, C.byref(info))
    print("        symbolic now: nsuper %d nlevel %d max_front %d lnz %.3g flops %.3g, analysis total %.2f s" %
          (info.nsuper, info.nlevel, info.max_front, info.lnz, info.flops, info.analyze_seconds), flush=True)
s.delete()
print("C4_OK")
