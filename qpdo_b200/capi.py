"""ctypes mirror of the qpdo.h C ABI (reference: include/qpdo.h:14-56, include/types.h:14-224).

The struct layouts below are the DLONG + PROFILING layout the reference's only build recipe uses
(interfaces/mex/qpdo_make.m:300,331): c_float = double, c_int = SuiteSparse_long = long.

This module is ABI plumbing only: it loads *a* shared library that exports the eight qpdo_*
symbols and drives it.  The product (`qpdo_b200.QPDO`) points it at the CUDA library built from
qpdo_b200/csrc; the tests additionally point it at a build of the unmodified reference so both
are driven through byte-identical inputs.
"""
import ctypes as C
import os
import numpy as np

c_int = C.c_long
c_float = C.c_double
QPDO_INFTY = 1e20

# status codes: reference include/constants.h:17-26
QPDO_SOLVED = 1
QPDO_DUAL_TERMINATED = 2
QPDO_NON_CVX = -1
QPDO_PRIMAL_DUAL_INFEASIBLE = -2
QPDO_PRIMAL_INFEASIBLE = -3
QPDO_DUAL_INFEASIBLE = -4
QPDO_MAX_ITER_REACHED = -5
QPDO_MAX_TIME_REACHED = -6
QPDO_UNSOLVED = -10
QPDO_ERROR = -99

CHOLMOD_REAL = 1
CHOLMOD_DOUBLE = 0
CHOLMOD_LONG = 2

_pd = C.POINTER(c_float)
_pi = C.POINTER(c_int)


class CholmodSparse(C.Structure):
    # SuiteSparse/CHOLMOD/Include/cholmod_core.h:1214-1263
    _fields_ = [("nrow", C.c_size_t), ("ncol", C.c_size_t), ("nzmax", C.c_size_t),
                ("p", C.c_void_p), ("i", C.c_void_p), ("nz", C.c_void_p),
                ("x", C.c_void_p), ("z", C.c_void_p),
                ("stype", C.c_int), ("itype", C.c_int), ("xtype", C.c_int),
                ("dtype", C.c_int), ("sorted", C.c_int), ("packed", C.c_int)]


class QPDOSolution(C.Structure):
    _fields_ = [("x", _pd), ("y", _pd)]


class QPDOScaling(C.Structure):
    _fields_ = [("D", _pd), ("Dinv", _pd), ("E", _pd), ("Einv", _pd),
                ("c", c_float), ("cinv", c_float)]


class QPDOInfo(C.Structure):
    _fields_ = [("iterations", c_int), ("oterations", c_int), ("status", C.c_char * 32),
                ("status_val", c_int),
                ("res_prim_norm", c_float), ("res_dual_norm", c_float),
                ("res_prim_in_norm", c_float), ("res_dual_in_norm", c_float),
                ("objective", c_float),
                ("setup_time", c_float), ("solve_time", c_float), ("run_time", c_float)]


class QPDOData(C.Structure):
    _fields_ = [("n", C.c_size_t), ("m", C.c_size_t),
                ("Q", C.POINTER(CholmodSparse)), ("A", C.POINTER(CholmodSparse)),
                ("q", _pd), ("c", c_float), ("l", _pd), ("u", _pd)]


class QPDOSettings(C.Structure):
    _fields_ = [("max_time", c_float), ("max_iter", c_int), ("inner_max_iter", c_int),
                ("eps_abs", c_float), ("eps_abs_in", c_float),
                ("eps_prim_inf", c_float), ("eps_dual_inf", c_float),
                ("rho", c_float), ("theta", c_float), ("delta", c_float), ("mu_min", c_float),
                ("proximal", c_int), ("sigma_init", c_float), ("sigma_upd", c_float),
                ("sigma_min", c_float), ("scaling", c_int), ("verbose", c_int),
                ("print_interval", c_int), ("reset_newton_iter", c_int)]


class QPDOCholmod(C.Structure):
    _fields_ = [("LD", C.c_void_p), ("LD_Q", C.c_void_p), ("E_temp", C.c_void_p),
                ("D_temp", C.c_void_p), ("linsys_rhs", C.c_void_p), ("dx", C.c_void_p),
                ("dy", C.c_void_p), ("Adx", C.c_void_p), ("Qdx", C.c_void_p),
                ("Atdy", C.c_void_p), ("reset_newton", c_int),
                ("active_set", _pi), ("active_set_old", _pi), ("n_active_set", c_int),
                ("enter", _pi), ("n_enter", c_int), ("leave", _pi), ("n_leave", c_int),
                ("At_scale", C.c_void_p), ("At_sqrt_mu", C.c_void_p)]


class QPDOWorkspace(C.Structure):
    _fields_ = [("data", C.POINTER(QPDOData)),
                ("x", _pd), ("y", _pd), ("Ax", _pd), ("Qx", _pd), ("Aty", _pd),
                ("initialized", c_int),
                ("temp_m", _pd), ("temp_n", _pd), ("temp_2m", _pd),
                ("mu", _pd), ("sqrt_mu", _pd), ("sqrt_mu_min", c_float), ("sqrt_delta", c_float),
                ("n_mu_changed", c_int), ("sigma", c_float), ("sigma_mined", c_int),
                ("norm_q", c_float),
                ("xbar", _pd), ("ybar", _pd), ("dx", _pd), ("dy", _pd), ("tau", c_float),
                ("Qdx", _pd), ("Adx", _pd), ("Atdy", _pd),
                ("w", _pd), ("z", _pd), ("df", _pd), ("res_prim", _pd), ("res_dual", _pd),
                ("res_prim_old", _pd), ("res_prim_in", _pd), ("res_dual_in", _pd),
                ("linsys_rhs", _pd),
                ("res_prim_norm_old", c_float), ("res_dual_norm_old", c_float),
                ("ls_eta", c_float), ("ls_beta", c_float), ("ls_delta", _pd), ("ls_alpha", _pd),
                ("ls_taus", C.c_void_p), ("ls_idx_L", _pi), ("ls_idx_P", _pi), ("ls_idx_J", _pi),
                ("eps_prim", c_float), ("eps_dual", c_float), ("eps_prim_in", c_float),
                ("eps_dual_in", c_float), ("eps_in", c_float),
                ("D_temp", _pd), ("E_temp", _pd),
                ("chol", C.POINTER(QPDOCholmod)), ("settings", C.POINTER(QPDOSettings)),
                ("scaling", C.POINTER(QPDOScaling)), ("solution", C.POINTER(QPDOSolution)),
                ("info", C.POINTER(QPDOInfo)), ("timer", C.c_void_p)]


def load_library(path):
    """dlopen `path` (RTLD_LOCAL) and declare the eight qpdo.h prototypes on it."""
    if not os.path.exists(path):
        raise FileNotFoundError(path)
    lib = C.CDLL(path, mode=os.RTLD_LOCAL | os.RTLD_NOW)
    PW = C.POINTER(QPDOWorkspace)
    lib.qpdo_set_default_settings.argtypes = [C.POINTER(QPDOSettings)]
    lib.qpdo_set_default_settings.restype = None
    lib.qpdo_setup.argtypes = [C.POINTER(QPDOData), C.POINTER(QPDOSettings)]
    lib.qpdo_setup.restype = PW
    lib.qpdo_warm_start.argtypes = [PW, _pd, _pd]
    lib.qpdo_warm_start.restype = None
    lib.qpdo_solve.argtypes = [PW]
    lib.qpdo_solve.restype = None
    lib.qpdo_update_settings.argtypes = [PW, C.POINTER(QPDOSettings)]
    lib.qpdo_update_settings.restype = None
    lib.qpdo_update_bounds.argtypes = [PW, _pd, _pd]
    lib.qpdo_update_bounds.restype = None
    lib.qpdo_update_q.argtypes = [PW, _pd]
    lib.qpdo_update_q.restype = None
    lib.qpdo_cleanup.argtypes = [PW]
    lib.qpdo_cleanup.restype = None
    return lib


def _as_csc_long(M, lower_only=False):
    """scipy sparse -> (indptr, indices, data) int64/float64 CSC with sorted indices."""
    import scipy.sparse as sp
    M = sp.csc_matrix(M)
    if lower_only:
        M = sp.tril(M, format="csc")
    M.sort_indices()
    M.sum_duplicates()
    return (np.ascontiguousarray(M.indptr, dtype=np.int64),
            np.ascontiguousarray(M.indices, dtype=np.int64),
            np.ascontiguousarray(M.data, dtype=np.float64), M.shape)


def _sparse_struct(indptr, indices, data, shape, stype):
    s = CholmodSparse()
    s.nrow, s.ncol = shape
    s.nzmax = max(1, len(data))
    s.p = indptr.ctypes.data
    s.i = indices.ctypes.data
    s.nz = None
    s.x = data.ctypes.data
    s.z = None
    s.stype = stype
    s.itype = CHOLMOD_LONG
    s.xtype = CHOLMOD_REAL
    s.dtype = CHOLMOD_DOUBLE
    s.sorted = 1
    s.packed = 1
    return s


class Result(object):
    __slots__ = ("x", "y", "prim_inf_cert", "dual_inf_cert", "status_val", "status",
                 "iterations", "oterations", "objective", "res_prim_norm", "res_dual_norm",
                 "res_prim_in_norm", "res_dual_in_norm", "setup_time", "solve_time", "run_time",
                 "active_set", "wall_setup", "wall_solve")

    def __repr__(self):
        return ("Result(status=%r(%d), iter=%d, oter=%d, obj=%.15g, rp=%.3e, rd=%.3e)" %
                (self.status, self.status_val, self.iterations, self.oterations, self.objective,
                 self.res_prim_norm, self.res_dual_norm))


class Solver(object):
    """Host-side mirror of the reference's user-facing class (interfaces/mex/qpdo.m:50-233):
    setup / warm_start / update_q / update_bounds / solve / delete over the qpdo.h C ABI."""

    def __init__(self, lib):
        self._lib = lib
        self._work = None
        self._keep = None
        self.n = self.m = 0

    # qpdo.m:36-40 default_settings
    def default_settings(self):
        s = QPDOSettings()
        self._lib.qpdo_set_default_settings(C.byref(s))
        return s

    def _settings(self, overrides):
        if isinstance(overrides, QPDOSettings):
            return overrides
        s = self.default_settings()
        for k, v in (overrides or {}).items():
            if not hasattr(s, k):
                raise KeyError("unknown QPDO setting %r" % k)
            setattr(s, k, v)
        return s

    # qpdo.m:50-160 setup(Q,q,A,bmin,bmax,options)
    def setup(self, Q, q, A, bmin, bmax, settings=None, c=0.0):
        self.delete()
        q = np.ascontiguousarray(q, dtype=np.float64).ravel()
        n = q.size
        if A is None:
            import scipy.sparse as sp
            A = sp.csc_matrix((0, n))
        m = A.shape[0]
        # qpdo.m:138-139: infinities are clipped to +-QPDO_INFTY
        l = np.maximum(np.full(m, -np.inf) if bmin is None else np.asarray(bmin, dtype=np.float64).ravel(), -QPDO_INFTY)
        u = np.minimum(np.full(m, np.inf) if bmax is None else np.asarray(bmax, dtype=np.float64).ravel(), QPDO_INFTY)
        l = np.ascontiguousarray(l)
        u = np.ascontiguousarray(u)
        assert l.size == m and u.size == m, "Incorrect dimension of l/u"
        assert Q.shape == (n, n), "Q must be n x n"
        assert A.shape[1] == n
        Qp, Qi, Qx, Qs = _as_csc_long(Q, lower_only=True)   # mex passes stype=-1 (qpdo_mex.c:150)
        Ap, Ai, Ax, As = _as_csc_long(A)
        Qs_ = _sparse_struct(Qp, Qi, Qx, Qs, -1)
        As_ = _sparse_struct(Ap, Ai, Ax, As, 0)
        d = QPDOData()
        d.n, d.m = n, m
        d.Q, d.A = C.pointer(Qs_), C.pointer(As_)
        d.q = q.ctypes.data_as(_pd)
        d.c = float(c)
        d.l = l.ctypes.data_as(_pd)
        d.u = u.ctypes.data_as(_pd)
        st = self._settings(settings)
        w = self._lib.qpdo_setup(C.byref(d), C.byref(st))
        # setup deep-copies everything (qpdo.c:100-108): inputs may be dropped now
        if not w:
            raise ValueError("qpdo_setup returned NULL (invalid data or settings)")
        self._work = w
        self.n, self.m = n, m
        return self

    def warm_start(self, x=None, y=None):
        xa = None if x is None else np.ascontiguousarray(x, dtype=np.float64).ravel()
        ya = None if y is None else np.ascontiguousarray(y, dtype=np.float64).ravel()
        assert xa is None or xa.size == self.n, "Incorrect dimension of x"
        assert ya is None or ya.size == self.m, "Incorrect dimension of y"
        self._lib.qpdo_warm_start(self._work,
                                  None if xa is None else xa.ctypes.data_as(_pd),
                                  None if ya is None else ya.ctypes.data_as(_pd))

    def update_q(self, q):
        qa = np.ascontiguousarray(q, dtype=np.float64).ravel()
        assert qa.size == self.n
        self._lib.qpdo_update_q(self._work, qa.ctypes.data_as(_pd))

    def update_bounds(self, bmin=None, bmax=None):
        la = None if bmin is None else np.ascontiguousarray(np.maximum(np.asarray(bmin, dtype=np.float64).ravel(), -QPDO_INFTY))
        ua = None if bmax is None else np.ascontiguousarray(np.minimum(np.asarray(bmax, dtype=np.float64).ravel(), QPDO_INFTY))
        self._lib.qpdo_update_bounds(self._work,
                                     None if la is None else la.ctypes.data_as(_pd),
                                     None if ua is None else ua.ctypes.data_as(_pd))

    def update_settings(self, settings):
        st = self._settings(settings)
        self._lib.qpdo_update_settings(self._work, C.byref(st))

    @property
    def info(self):
        return self._work.contents.info.contents

    @property
    def work(self):
        return self._work.contents

    # qpdo.m:220-231 solve -> x, y, prim_inf_cert, dual_inf_cert, info (qpdo_mex.c:237-280)
    def solve(self):
        self._lib.qpdo_solve(self._work)
        w = self._work.contents
        info = w.info.contents
        n, m = self.n, self.m
        r = Result()
        r.status_val = int(info.status_val)
        r.status = info.status.decode()
        r.iterations = int(info.iterations)
        r.oterations = int(info.oterations)
        r.objective = float(info.objective)
        r.res_prim_norm = float(info.res_prim_norm)
        r.res_dual_norm = float(info.res_dual_norm)
        r.res_prim_in_norm = float(info.res_prim_in_norm)
        r.res_dual_in_norm = float(info.res_dual_in_norm)
        r.setup_time = float(info.setup_time)
        r.solve_time = float(info.solve_time)
        r.run_time = float(info.run_time)
        r.x = np.array(np.ctypeslib.as_array(w.solution.contents.x, shape=(n,))) if n else np.zeros(0)
        r.y = np.array(np.ctypeslib.as_array(w.solution.contents.y, shape=(m,))) if m else np.zeros(0)
        r.prim_inf_cert = np.array(np.ctypeslib.as_array(w.dy, shape=(m,))) if m else np.zeros(0)
        r.dual_inf_cert = np.array(np.ctypeslib.as_array(w.dx, shape=(n,))) if n else np.zeros(0)
        aset = w.chol.contents.active_set
        r.active_set = (np.array(np.ctypeslib.as_array(aset, shape=(m,))).astype(np.int8)
                        if (m and aset) else np.zeros(m, dtype=np.int8))
        return r

    def delete(self):
        if self._work is not None:
            self._lib.qpdo_cleanup(self._work)
            self._work = None

    def __del__(self):
        try:
            self.delete()
        except Exception:
            pass


# --------------------------------------------------------------------------- qpdo_cuda.h mirror
class DevScalars(C.Structure):
    _fields_ = [("res_prim_norm", C.c_double), ("res_dual_norm", C.c_double), ("res_prim_in_norm", C.c_double),
                ("res_dual_in_norm", C.c_double), ("tau", C.c_double), ("ls_eta", C.c_double), ("ls_beta", C.c_double),
                ("objective", C.c_double), ("inf_a", C.c_double), ("inf_b", C.c_double), ("inf_c", C.c_double),
                ("sigma", C.c_double), ("n_active", C.c_int), ("n_enter", C.c_int), ("n_leave", C.c_int),
                ("n_mu_changed", C.c_int), ("reset_newton", C.c_int), ("factor_mode", C.c_int), ("ls_nL", C.c_int),
                ("not_posdef", C.c_int), ("flag", C.c_int), ("n_refactor", C.c_int), ("n_update", C.c_int),
                ("n_qonly", C.c_int), ("pad", C.c_int)]


class DevProfile(C.Structure):
    _fields_ = [("assemble_ms", C.c_double), ("factor_ms", C.c_double), ("solve_ms", C.c_double),
                ("spmv_ms", C.c_double), ("linesearch_ms", C.c_double), ("residual_ms", C.c_double),
                ("other_ms", C.c_double), ("n_newton", C.c_long), ("n_residual", C.c_long),
                ("kernel_launches", C.c_long)]


class DevSymbolicInfo(C.Structure):
    _fields_ = [("nsuper", C.c_int), ("nlevel", C.c_int), ("max_front", C.c_int), ("max_nc", C.c_int),
                ("lnz", C.c_int64), ("lnz_stored", C.c_int64), ("usize", C.c_int64), ("nnzK", C.c_int64),
                ("ntriple", C.c_int64), ("flops", C.c_double), ("analyze_seconds", C.c_double)]


def declare_device_layer(lib):
    """prototypes of the diagnostics / test entry points of include/qpdo_cuda.h"""
    vp = C.c_void_p
    lib.qpdo_dev_last_error.restype = C.c_char_p
    lib.qpdo_dev_count.restype = C.c_int
    lib.qpdo_dev_set_profiling.argtypes = [vp, C.c_int]
    lib.qpdo_dev_get_profile.argtypes = [vp, C.POINTER(DevProfile), C.c_int]
    lib.qpdo_dev_get_symbolic_info.argtypes = [vp, C.POINTER(DevSymbolicInfo)]
    lib.qpdo_dev_get_vector.argtypes = [vp, C.c_char_p, _pd, C.c_long]
    lib.qpdo_dev_get_vector.restype = C.c_long
    lib.qpdo_dev_test_factor_solve.argtypes = [vp, _pd, C.c_double, _pd, _pd, C.POINTER(C.c_float), C.POINTER(C.c_float)]
    lib.qpdo_dev_test_spmv.argtypes = [vp, C.c_int, _pd, _pd]
    lib.qpdo_dev_measure_fp64.argtypes = [C.c_int, _pd, _pd]
    lib.qpdo_dev_solve_noalloc.argtypes = [vp, _pd, _pd]
    lib.qpdo_dev_solve_noalloc_max_doubles.restype = C.c_long
    return lib


def dev_vector(lib, dev, name, cap):
    out = np.zeros(max(1, cap))
    n = lib.qpdo_dev_get_vector(dev, name.encode(), out.ctypes.data_as(_pd), cap)
    if n < 0:
        raise RuntimeError(lib.qpdo_dev_last_error().decode())
    return out[:min(n, cap)]
