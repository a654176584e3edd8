"""tools/probe_front.py [contNN] -- needs a library built with EXTRA_NVFLAGS=-DFRONT_PROBE: mean SM cycles per CTA of
k_front_smem<64|128|256> in its phases (load, extend-add of the children, LDL', write-back) over one factorization."""
import sys, os, ctypes as C
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import qpdo_b200
from qpdo_b200 import problems

name = sys.argv[1] if len(sys.argv) > 1 else "cont300"
p = problems.random_qp() if name == "c1" else problems.cont_like(int(name[4:]))
g = qpdo_b200.QPDO()
g.setup(p["Q"], p["q"], p["A"], p["l"], p["u"], dict(verbose=0))
lib, dev = qpdo_b200.lib(), g.device
rng = np.random.default_rng(0)
om = rng.random(g.m) * (rng.random(g.m) < 0.5)
b = rng.standard_normal(g.n); x = np.zeros(g.n)
pd = C.POINTER(C.c_double)
fms, sms = C.c_float(0), C.c_float(0)
buf = np.zeros(36, dtype=np.uint64)
for rep in range(3):
    lib.qpdo_dev_fprobe_read(buf.ctypes.data_as(C.c_void_p))
    lib.qpdo_dev_test_factor_solve(dev, om.ctypes.data_as(pd), C.c_double(1e-3), b.ctypes.data_as(pd), x.ctypes.data_as(pd), C.byref(fms), C.byref(sms))
lib.qpdo_dev_fprobe_read(buf.ctypes.data_as(C.c_void_p))
P = buf.reshape(3, 12).astype(np.float64)
print("%s: factor %.3f ms" % (name, fms.value))
for w, nt in enumerate((64, 128, 256)):
    c = P[w, 4]
    if c == 0:
        continue
    print("  k_front_smem<%3d>: %5d CTAs, mean N %.0f nc %.0f children %.1f | cycles per CTA: load %.0f, extend-add %.0f, LDL' %.0f, write-back %.0f (total %.1f us)" %
          (nt, c, P[w, 5] / c, P[w, 6] / c, P[w, 7] / c, P[w, 0] / c, P[w, 1] / c, P[w, 2] / c, P[w, 3] / c, P[w, :4].sum() / c / 1965.0))
    print("      inside LDL': panel + barrier %.0f, u registers %.0f, trailing update + barrier %.0f" % (P[w, 8] / c, P[w, 9] / c, P[w, 10] / c))
g.delete()
