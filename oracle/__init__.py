"""oracle -- TEST INFRASTRUCTURE ONLY.

ctypes loader for (1) oracle/libqpdo_oracle.so, our plain-C restatement of QPDO's Newton hot path
(oracle/qpdo_oracle.c) and (2) oracle/_ref/libqpdo_ref*.so, the UNMODIFIED reference compiled from
/root/reference by oracle/Makefile.  Only tests/, __graft_entry__.smoke() and bench.py's CPU
baseline legs may import this package; the product (qpdo_b200) never does.
"""
import ctypes as C
import os
import subprocess
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
RESTATEMENT = os.path.join(HERE, "libqpdo_oracle.so")
REF = os.path.join(HERE, "_ref", "libqpdo_ref.so")
REF_NOSUPER = os.path.join(HERE, "_ref", "libqpdo_ref_nosuper.so")

_pd = C.POINTER(C.c_double)
_pl = C.POINTER(C.c_long)


class OracleSettings(C.Structure):
    _fields_ = [("max_time", C.c_double), ("max_iter", C.c_long), ("inner_max_iter", C.c_long),
                ("eps_abs", C.c_double), ("eps_abs_in", C.c_double), ("eps_prim_inf", C.c_double),
                ("eps_dual_inf", C.c_double), ("rho", C.c_double), ("theta", C.c_double), ("delta", C.c_double),
                ("mu_min", C.c_double), ("proximal", C.c_long), ("sigma_init", C.c_double),
                ("sigma_upd", C.c_double), ("sigma_min", C.c_double), ("scaling", C.c_long), ("verbose", C.c_long),
                ("print_interval", C.c_long), ("reset_newton_iter", C.c_long)]


class OracleTrace(C.Structure):
    _fields_ = [("iter", C.c_long), ("oter", C.c_long), ("kind", C.c_long),
                ("res_prim_norm", C.c_double), ("res_dual_norm", C.c_double), ("res_prim_in_norm", C.c_double),
                ("res_dual_in_norm", C.c_double), ("tau", C.c_double), ("sigma", C.c_double),
                ("n_active", C.c_long), ("n_enter", C.c_long), ("n_leave", C.c_long), ("factor_mode", C.c_long)]


class OracleInfo(C.Structure):
    _fields_ = [("iterations", C.c_long), ("oterations", C.c_long), ("status_val", C.c_long),
                ("res_prim_norm", C.c_double), ("res_dual_norm", C.c_double), ("res_prim_in_norm", C.c_double),
                ("res_dual_in_norm", C.c_double), ("objective", C.c_double), ("solve_seconds", C.c_double),
                ("n_trace", C.c_long)]


def build(ref=True, quiet=True):
    """make the restatement (always) and, when /root/reference is present, oracle/_ref."""
    targets = ["restatement"] + (["ref"] if ref else [])
    subprocess.run(["make", "-C", HERE] + targets, check=True,
                   stdout=subprocess.DEVNULL if quiet else None)


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(RESTATEMENT):
            build(ref=False)
        L = C.CDLL(RESTATEMENT)
        L.qpdo_oracle_solve.argtypes = [C.c_long, C.c_long, _pl, _pl, _pd, _pl, _pl, _pd, _pd, C.c_double, _pd, _pd,
                                        C.POINTER(OracleSettings), _pd, _pd, _pd, _pd, _pd, _pd, _pl,
                                        C.POINTER(OracleInfo), C.POINTER(OracleTrace), C.c_long]
        L.qpdo_oracle_solve.restype = C.c_int
        L.qpdo_oracle_scale_data.argtypes = [C.c_long, C.c_long, _pl, _pl, _pd, _pl, _pl, _pd, _pd, _pd, _pd, C.c_long,
                                             _pd, _pd, _pd]
        L.qpdo_oracle_scale_data.restype = None
        L.qpdo_oracle_spmv.argtypes = [C.c_int, C.c_long, C.c_long, _pl, _pl, _pd, _pd, _pd]
        L.qpdo_oracle_spmv.restype = None
        L.qpdo_oracle_linesearch.argtypes = [C.c_long, C.c_long] + [_pd] * 11
        L.qpdo_oracle_linesearch.restype = C.c_double
        L.oracle_default_settings.argtypes = [C.POINTER(OracleSettings)]
        L.oracle_default_settings.restype = None
        _lib = L
    return _lib


def _csc(M, lower=False):
    import scipy.sparse as sp
    M = sp.csc_matrix(M)
    if lower:
        M = sp.tril(M, format="csc")
    M.sort_indices()
    M.sum_duplicates()
    return (np.ascontiguousarray(M.indptr, dtype=np.int64), np.ascontiguousarray(M.indices, dtype=np.int64),
            np.ascontiguousarray(M.data, dtype=np.float64))


def _p(a, t=_pd):
    return None if a is None else a.ctypes.data_as(t)


class OracleResult(object):
    pass


def settings(**kw):
    s = OracleSettings()
    lib().oracle_default_settings(C.byref(s))
    s.verbose = 0
    for k, v in kw.items():
        if not hasattr(s, k):
            raise KeyError(k)
        setattr(s, k, v)
    return s


def solve(p, settings_kw=None, x0=None, y0=None, trace=True):
    """Solve a problem dict (qpdo_b200.problems) with the C restatement."""
    L = lib()
    q = np.ascontiguousarray(p["q"], dtype=np.float64)
    n = q.size
    m = p["A"].shape[0]
    Qp, Qi, Qx = _csc(p["Q"], lower=True)
    Ap, Ai, Ax = _csc(p["A"])
    l = np.ascontiguousarray(np.maximum(p["l"], -1e20), dtype=np.float64)
    u = np.ascontiguousarray(np.minimum(p["u"], 1e20), dtype=np.float64)
    st = settings(**(settings_kw or {}))
    x = np.zeros(n); y = np.zeros(m); dxc = np.zeros(n); dyc = np.zeros(m); act = np.zeros(max(m, 1), dtype=np.int64)
    info = OracleInfo()
    cap = int(st.max_iter) + 1 if trace else 0
    tr = (OracleTrace * max(cap, 1))()
    x0a = None if x0 is None else np.ascontiguousarray(x0, dtype=np.float64)
    y0a = None if y0 is None else np.ascontiguousarray(y0, dtype=np.float64)
    rc = L.qpdo_oracle_solve(n, m, _p(Qp, _pl), _p(Qi, _pl), _p(Qx), _p(Ap, _pl), _p(Ai, _pl), _p(Ax), _p(q),
                             float(p.get("c", 0.0)), _p(l), _p(u), C.byref(st), _p(x0a), _p(y0a), _p(x), _p(y),
                             _p(dxc), _p(dyc), _p(act, _pl), C.byref(info), tr, cap)
    if rc:
        raise ValueError("oracle: invalid data")
    r = OracleResult()
    r.x, r.y, r.dual_inf_cert, r.prim_inf_cert = x, y, dxc, dyc
    r.active_set = act[:m].astype(np.int8)
    for f, _t in OracleInfo._fields_:
        setattr(r, f, getattr(info, f))
    r.trace = [dict((f, getattr(tr[i], f)) for f, _t in OracleTrace._fields_) for i in range(info.n_trace)]
    return r


def have_ref(nosuper=False):
    return os.path.exists(REF_NOSUPER if nosuper else REF)


def ref_library(nosuper=False):
    """the unmodified reference as a qpdo.h library (driven through qpdo_b200.capi.Solver)."""
    from qpdo_b200 import capi
    return capi.load_library(REF_NOSUPER if nosuper else REF)


REF_BATCH = os.path.join(HERE, "libref_batch.so")
REF_SUMMARY = np.dtype([("status_val", np.int32), ("iterations", np.int32), ("oterations", np.int32),
                        ("n_active", np.int32), ("objective", np.float64), ("res_prim_norm", np.float64),
                        ("res_dual_norm", np.float64)])


def ref_batch_loop(n, m, Qp, Qi, Ap, Ai, Qx, Ax, q, l, u, threads=1, max_iter=0, nosuper=False, want_xy=False):
    """The reference user's loop (qpdo_setup / qpdo_solve / qpdo_cleanup per problem, src/qpdo.c:49,304,591)
    over a packed batch, on `threads` host threads, with the unmodified reference library
    (oracle/ref_batch.c).  Returns (wall seconds, summaries[, x, y])."""
    if not os.path.exists(REF_BATCH):
        build(ref=False)
    os.environ.setdefault("OPENBLAS_NUM_THREADS", "1")       # one workspace per thread; BLAS stays serial
    L = C.CDLL(REF_BATCH)
    L.ref_batch_run.restype = C.c_double
    L.ref_batch_run.argtypes = [C.c_char_p, C.c_long, C.c_long, C.c_long, _pl, _pl, _pl, _pl, _pd, _pd, _pd, _pd, _pd,
                                C.c_long, C.c_int, _pd, _pd, C.c_void_p]
    f = lambda a: np.ascontiguousarray(a, dtype=np.float64)
    g = lambda a: np.ascontiguousarray(a, dtype=np.int64)
    Qp, Qi, Ap, Ai = map(g, (Qp, Qi, Ap, Ai))
    Qx, Ax, q, l, u = map(f, (Qx, Ax, q, l, u))
    count = q.shape[0]
    summ = np.zeros(count, dtype=REF_SUMMARY)
    x = np.zeros((count, n)) if want_xy else None
    y = np.zeros((count, m)) if want_xy else None
    path = REF_NOSUPER if nosuper else REF
    secs = L.ref_batch_run(path.encode(), count, n, m, _p(Qp, _pl), _p(Qi, _pl), _p(Ap, _pl), _p(Ai, _pl), _p(Qx), _p(Ax),
                           _p(q), _p(l), _p(u), max_iter, threads, _p(x), _p(y), summ.ctypes.data)
    if secs < 0:
        raise RuntimeError("ref_batch_run failed (%g)" % secs)
    return (secs, summ, x, y) if want_xy else (secs, summ)
