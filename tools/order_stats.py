#!/usr/bin/env python
"""tools/order_stats.py [problem] -- host-only: symbolic statistics of the orderings (1 = minimum degree, 2 = nested
dissection) on a problem's K pattern: fill, flops, tree height, largest front and the length of the serial chain of
32-pivot panel steps of the big fronts (what the top of the assembly tree costs on the GPU)."""
import ctypes as C
import os
import sys
import time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import qpdo_b200
from qpdo_b200 import problems

P = C.POINTER


class Sym(C.Structure):
    _fields_ = [("n", C.c_int), ("m", C.c_int), ("nsuper", C.c_int), ("nlevel", C.c_int), ("perm", P(C.c_int)), ("pinv", P(C.c_int)),
                ("sn_start", P(C.c_int)), ("sn_nr", P(C.c_int)), ("sn_rowptr", P(C.c_int64)), ("rows", P(C.c_int)), ("rel", P(C.c_int)),
                ("sn_parent", P(C.c_int)), ("child_ptr", P(C.c_int)), ("child_idx", P(C.c_int)), ("Lp", P(C.c_int64)), ("Up", P(C.c_int64)),
                ("sn_level", P(C.c_int)), ("level_ptr", P(C.c_int)), ("level_sn", P(C.c_int)), ("col2sn", P(C.c_int)), ("nnzK", C.c_int64),
                ("Kp", P(C.c_int)), ("Ki", P(C.c_int)), ("Kdst", P(C.c_int64)), ("Kq", P(C.c_int)), ("Ktp", P(C.c_int64)), ("ntriple", C.c_int64),
                ("Kt_row", P(C.c_int)), ("Kt_a1", P(C.c_int)), ("Kt_a2", P(C.c_int)), ("lnz", C.c_int64), ("lnz_stored", C.c_int64),
                ("usize", C.c_int64), ("flops", C.c_double), ("max_front", C.c_int), ("max_nc", C.c_int)]


def stats(p, ordering, big=160):
    import scipy.sparse as sp
    lib = qpdo_b200.lib()
    lib.qpdo_symbolic_analyze.restype = P(Sym)
    lib.qpdo_symbolic_analyze.argtypes = [C.c_int, C.c_int] + [C.c_void_p] * 4 + [C.c_int, C.c_char_p]
    lib.qpdo_symbolic_free.argtypes = [P(Sym)]
    Q = sp.tril(sp.csc_matrix(p["Q"]), format="csc"); Q.sort_indices()
    A = sp.csc_matrix(p["A"]); A.sort_indices()
    n, m = Q.shape[0], A.shape[0]
    Qp, Qi, Ap, Ai = (np.ascontiguousarray(a, dtype=np.int32) for a in (Q.indptr, Q.indices, A.indptr, A.indices))
    err = C.create_string_buffer(256)
    t0 = time.time()
    Sp = lib.qpdo_symbolic_analyze(n, m, Qp.ctypes.data, Qi.ctypes.data, Ap.ctypes.data, Ai.ctypes.data, ordering, err)
    dt = time.time() - t0
    if not Sp:
        raise RuntimeError(err.value.decode())
    S = Sp.contents
    ns = S.nsuper
    nc = np.array([S.sn_start[s + 1] - S.sn_start[s] for s in range(ns)])
    N = nc + np.array([S.sn_nr[s] for s in range(ns)])
    level = np.array([S.sn_level[s] for s in range(ns)])
    # serial chain of panel steps: per level, the big fronts of a level run side by side -> max over the level
    chain = 0
    per_level = []
    for lv in range(S.nlevel):
        sel = (level == lv) & (N > big)
        steps = int(np.ceil(nc[sel] / 32.0).max()) if sel.any() else 0
        per_level.append((lv, int(sel.sum()), steps))
        chain += steps
    out = dict(ordering=ordering, analyze_s=dt, nsuper=ns, nlevel=S.nlevel, lnz=int(S.lnz), lnz_stored=int(S.lnz_stored), usize=int(S.usize),
               gflop=S.flops / 1e9, max_front=S.max_front, max_nc=S.max_nc, big_fronts=int((N > big).sum()), panel_chain=chain,
               levels_with_big=[x for x in per_level if x[1]])
    lib.qpdo_symbolic_free(Sp)
    return out


if __name__ == "__main__":
    name = sys.argv[1] if len(sys.argv) > 1 else "cont300"
    p = {"cont300": lambda: problems.cont_like(300), "cont100": lambda: problems.cont_like(100), "c1": lambda: problems.random_qp(1000, 500, 0.01, seed=1),
         "lasso_mid": lambda: problems.lasso_dual(n=20000, m=10000)}[name]()
    for o in (1, 2):
        s = stats(p, o)
        print({k: v for k, v in s.items() if k != "levels_with_big"})
        print("   big fronts per level (level, fronts, panel steps):", s["levels_with_big"])
