"""GPU check of the numeric factorization + solves: K = Q + sigma I + A' diag(omega) A, K x = b."""
import sys, os, ctypes as C
import numpy as np, scipy.sparse as sp, scipy.sparse.linalg as spla
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import qpdo_b200
from qpdo_b200 import capi, problems


def check(name, p, scaling=10, seed=0, omega_mode="random"):
    g = qpdo_b200.QPDO()
    g.setup(p["Q"], p["q"], p["A"], p["l"], p["u"], dict(verbose=0, scaling=scaling))
    lib, dev = qpdo_b200.lib(), g.device
    n, m = g.n, g.m
    Ql = sp.tril(sp.csc_matrix(p["Q"]), format="csc"); Ql.sort_indices()
    A = sp.csc_matrix(p["A"]); A.sort_indices(); A.sum_duplicates()
    Av = capi.dev_vector(lib, dev, "A_values", A.nnz)
    Qv = capi.dev_vector(lib, dev, "Q_values", Ql.nnz)
    As = sp.csc_matrix((Av, A.indices, A.indptr), shape=A.shape)
    Qs = sp.csc_matrix((Qv, Ql.indices, Ql.indptr), shape=Ql.shape)
    Qs = Qs + sp.tril(Qs, -1).T
    rng = np.random.default_rng(seed)
    om = rng.random(m) * (rng.random(m) < 0.5) if omega_mode == "random" else np.zeros(m)
    sigma = 1e-3
    b = rng.standard_normal(n)
    K = (Qs + sigma * sp.identity(n) + As.T @ sp.diags(om) @ As).tocsc()
    x = np.zeros(n)
    fms, sms = C.c_float(0), C.c_float(0)
    pd = C.POINTER(C.c_double)
    for rep in range(3):
        rc = lib.qpdo_dev_test_factor_solve(dev, om.ctypes.data_as(pd), sigma, b.ctypes.data_as(pd), x.ctypes.data_as(pd),
                                            C.byref(fms), C.byref(sms))
        if rc:
            print(name, "FAILED:", lib.qpdo_dev_last_error().decode()); return
    info = capi.DevSymbolicInfo(); lib.qpdo_dev_get_symbolic_info(dev, C.byref(info))
    res = np.abs(K @ x - b).max() / max(1e-300, np.abs(b).max())
    xr = spla.spsolve(K, b)
    err = np.abs(x - xr).max() / np.abs(xr).max()
    print("%-10s n=%d m=%d ns=%d nlev=%d maxN=%d lnz=%d flops=%.3g | resid %.2e err-vs-scipy %.2e nan=%d | factor %.3f ms solve %.3f ms" %
          (name, n, m, info.nsuper, info.nlevel, info.max_front, info.lnz, info.flops, res, err, int(np.isnan(x).sum()), fms.value, sms.value))
    sys.stdout.flush()
    g.delete()


if __name__ == "__main__":
    for w in (sys.argv[1:] or ["trio", "rand200", "cont20", "c1", "cont60"]):
        if w == "trio":
            check("trio", problems.infeasibility_trio()[0][1])
        elif w == "rand200":
            check("rand200", problems.random_qp(200, 100, 0.05, seed=3))
        elif w == "c1":
            check("c1", problems.random_qp())
            check("c1-om0", problems.random_qp(), omega_mode="zero")
        elif w.startswith("cont"):
            check(w, problems.cont_like(int(w[4:])))
        elif w == "lasso":
            check("lasso-small", problems.lasso_dual(20000, 10000))
