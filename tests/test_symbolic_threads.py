"""Host symbolic analysis (csrc/ordering.cpp, csrc/symbolic.cpp): the nested dissection forks its two parts onto threads and
the K-product map splits its searches over threads -- the result must not depend on the thread count (the numerics of every
golden depend on the ordering).  CPU only."""
import ctypes as C
import hashlib
import os
import sys
import numpy as np
import pytest
import scipy.sparse as sp

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
from qpdo_b200 import problems


def _digest(lib, p, ordering=2):
    import order_stats as o
    P = C.POINTER
    lib.qpdo_symbolic_analyze.restype = P(o.Sym)
    lib.qpdo_symbolic_analyze.argtypes = [C.c_int, C.c_int] + [C.c_void_p] * 4 + [C.c_int, C.c_char_p]
    lib.qpdo_symbolic_free.argtypes = [P(o.Sym)]
    Q = sp.tril(sp.csc_matrix(p["Q"]), format="csc"); Q.sort_indices()
    A = sp.csc_matrix(p["A"]); A.sort_indices()
    n, m = Q.shape[0], A.shape[0]
    a = [np.ascontiguousarray(v, dtype=np.int32) for v in (Q.indptr, Q.indices, A.indptr, A.indices)]
    err = C.create_string_buffer(256)
    Sp = lib.qpdo_symbolic_analyze(n, m, *[v.ctypes.data for v in a], ordering, err)
    assert Sp, err.value
    S = Sp.contents
    h = hashlib.sha1()
    for name, cnt in (("perm", n), ("Kdst", S.nnzK), ("Ktp", S.nnzK + 1), ("Kt_row", S.ntriple), ("Kt_a1", S.ntriple), ("Kt_a2", S.ntriple),
                      ("rel", S.sn_rowptr[S.nsuper]), ("level_sn", S.nsuper), ("sn_start", S.nsuper + 1)):
        if cnt:
            h.update(np.ctypeslib.as_array(getattr(S, name), shape=(cnt,)).tobytes())
    lib.qpdo_symbolic_free(Sp)
    return h.hexdigest()


@pytest.mark.parametrize("gen", ["cont120", "lasso", "rand"])
def test_symbolic_structure_does_not_depend_on_the_thread_count(cuda_lib, gen, monkeypatch):
    p = {"cont120": lambda: problems.cont_like(120), "lasso": lambda: problems.lasso_dual(n=20000, m=10000),
         "rand": lambda: problems.random_qp(1000, 500, 0.01, seed=1)}[gen]()
    out = []
    for nd, sym in (("1", "1"), ("8", "8"), ("4", "2"), ("8", "8")):
        monkeypatch.setenv("QPDO_ND_THREADS", nd)
        monkeypatch.setenv("QPDO_SYM_THREADS", sym)
        out.append(_digest(cuda_lib, p))
    assert len(set(out)) == 1, out


def test_singleton_columns_analysed_apart_give_the_same_factorization_structure(cuda_lib, monkeypatch):
    """Columns of K without off-diagonal entries are left out of ordering / tree / supernodes and appended as fronts of their
    own (QPDO_SYM_COMPRESS=0: everything through the full analysis).  Same statistics, same relative elimination order of the
    coupled columns, and the K entry maps reproduce P K P' numerically in both modes."""
    import order_stats as o
    P = C.POINTER
    lib = cuda_lib
    lib.qpdo_symbolic_analyze.restype = P(o.Sym)
    lib.qpdo_symbolic_analyze.argtypes = [C.c_int, C.c_int] + [C.c_void_p] * 4 + [C.c_int, C.c_char_p]
    lib.qpdo_symbolic_free.argtypes = [P(o.Sym)]
    p = problems.lasso_dual(n=20000, m=10000)
    rng = np.random.default_rng(0)
    rows = np.sort(rng.choice(10000, 150, replace=False))
    Q = sp.tril(sp.csc_matrix(p["Q"]), format="csc"); Q.sort_indices()
    A = sp.csc_matrix(p["A"])[rows].tocsc(); A.sort_indices()
    n, m = Q.shape[0], A.shape[0]
    a = [np.ascontiguousarray(v, dtype=np.int32) for v in (Q.indptr, Q.indices, A.indptr, A.indices)]
    Ar = A.tocsr()
    coupled = np.zeros(n, bool)
    for r in np.nonzero(np.diff(Ar.indptr) >= 2)[0]:
        coupled[Ar.indices[Ar.indptr[r]:Ar.indptr[r + 1]]] = True
    om = rng.random(m) + 0.1
    K = ((Q + sp.tril(Q, -1).T) + 1e-3 * sp.identity(n) + A.T @ sp.diags(om) @ A).tocsc()
    res = []
    for mode in ("0", "1"):
        monkeypatch.setenv("QPDO_SYM_COMPRESS", mode)
        err = C.create_string_buffer(256)
        Sp = lib.qpdo_symbolic_analyze(n, m, *[v.ctypes.data for v in a], 2, err)
        assert Sp, err.value
        S = Sp.contents
        g = lambda nm, cnt: np.ctypeslib.as_array(getattr(S, nm), shape=(max(int(cnt), 1),)).copy()
        perm, Kp, Ki, Kq, Ktp = g("perm", n), g("Kp", n + 1), g("Ki", S.nnzK), g("Kq", S.nnzK), g("Ktp", S.nnzK + 1)
        Kr, a1, a2 = g("Kt_row", S.ntriple)[:S.ntriple], g("Kt_a1", S.ntriple)[:S.ntriple], g("Kt_a2", S.ntriple)[:S.ntriple]
        cs = np.concatenate([[0.0], np.cumsum(om[Kr] * A.data[a1] * A.data[a2])])
        vals = cs[Ktp[1:]] - cs[Ktp[:-1]] + np.where(Kq >= 0, Q.data[np.maximum(Kq, 0)], 0.0)
        vals += np.where(Ki == np.repeat(np.arange(n), np.diff(Kp)), 1e-3, 0.0)
        mine = sp.csc_matrix((vals, Ki, Kp), shape=(n, n))
        assert abs(mine - sp.tril(K[perm][:, perm], format="csc")).max() < 1e-12
        assert sorted(perm.tolist()) == list(range(n))
        res.append((dict(nsuper=S.nsuper, nlevel=S.nlevel, lnz=S.lnz, stored=S.lnz_stored, usize=S.usize, flops=S.flops, maxf=S.max_front,
                         nnzK=S.nnzK, ntri=S.ntriple), [v for v in perm if coupled[v]]))
        lib.qpdo_symbolic_free(Sp)
    assert res[0][0] == res[1][0], (res[0][0], res[1][0])
    assert res[0][1] == res[1][1]
