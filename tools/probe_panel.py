"""tools/probe_panel.py [c1|contNN] -- needs a library built with EXTRA_NVFLAGS=-DPANEL_PROBE: phase stamps of the first
tile CTA of every k_big_panel launch of one factorization (SM cycles) and the gaps between launches (global timer)."""
import sys, os, ctypes as C
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import qpdo_b200
from qpdo_b200 import problems

name = sys.argv[1] if len(sys.argv) > 1 else "c1"
p = problems.random_qp() if name == "c1" else problems.cont_like(int(name[4:]))
g = qpdo_b200.QPDO()
g.setup(p["Q"], p["q"], p["A"], p["l"], p["u"], dict(verbose=0))
lib, dev = qpdo_b200.lib(), g.device
rng = np.random.default_rng(0)
om = rng.random(g.m) * (rng.random(g.m) < 0.5)
b = rng.standard_normal(g.n); x = np.zeros(g.n)
pd = C.POINTER(C.c_double)
fms, sms = C.c_float(0), C.c_float(0)
buf = np.zeros(16 * 4096, dtype=np.int64)
for rep in range(3):
    lib.qpdo_dev_probe_read(buf.ctypes.data_as(C.c_void_p), 0)       # reset
    lib.qpdo_dev_test_factor_solve(dev, om.ctypes.data_as(pd), C.c_double(1e-3), b.ctypes.data_as(pd), x.ctypes.data_as(pd), C.byref(fms), C.byref(sms))
n = lib.qpdo_dev_probe_read(buf.ctypes.data_as(C.c_void_p), 4096)
P = buf[:16 * n].reshape(n, 16)
P = P[np.argsort(P[:, 14])]
names = ["entry->desc", "desc->prefetch issued", "->sync0", "sync0->LDL' start", "LDL'", "(inv done - LDL' done)", "->sync1 (from LDL' done)", "TRSM", "->sync2", "DMMA update", "stores issued"]
idx = [(0, 1), (1, 2), (2, 3), (3, 4), (4, 5), (5, 6), (5, 8), (8, 9), (9, 10), (10, 11), (11, 12)]
print("%s: factor %.3f ms, %d panel launches probed" % (name, fms.value, n))
for nm, (a, c) in zip(names, idx):
    d = P[:, c] - P[:, a]
    print("  %-26s mean %7.0f cycles  (min %6d max %6d)" % (nm, d.mean(), d.min(), d.max()))
tot = P[:, 12] - P[:, 0]
print("  total entry->stores issued mean %.0f cycles = %.2f us at 1.965 GHz" % (tot.mean(), tot.mean() / 1965.0))
gap = P[1:, 14] - P[:-1, 15]
print("  gap exit->next entry: mean %.2f us median %.2f (min %.2f max %.2f)" % (gap.mean() / 1e3, np.median(gap) / 1e3, gap.min() / 1e3, gap.max() / 1e3))
g.delete()
