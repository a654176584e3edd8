"""ctypes mirror of include/qpdo_io.h: read a free-format QPS file into the arrays qpdo_setup takes.

The parsing is host C inside libqpdo_b200.so (csrc/qpdo_io.c); this module only copies the returned
QPDOData (reference include/types.h:81-90) into scipy / numpy objects in the dict layout the rest of
the package uses: {"Q" (symmetric, both triangles), "q", "c", "A", "l", "u", "m_rows"}.
"""
import ctypes as C
import numpy as np
import scipy.sparse as sp
from . import capi


def _csc(S):
    nrow, ncol = int(S.nrow), int(S.ncol)
    p = np.ctypeslib.as_array(C.cast(S.p, C.POINTER(C.c_long)), shape=(ncol + 1,)).copy()
    nnz = int(p[-1])
    if nnz:
        i = np.ctypeslib.as_array(C.cast(S.i, C.POINTER(C.c_long)), shape=(nnz,)).copy()
        x = np.ctypeslib.as_array(C.cast(S.x, C.POINTER(C.c_double)), shape=(nnz,)).copy()
    else:
        i = np.zeros(0, dtype=np.int64); x = np.zeros(0)
    return sp.csc_matrix((x, i, p), shape=(nrow, ncol))


def _raise(what, path, rc, err_line):
    if rc == -1:
        raise OSError("%s: cannot open %s" % (what, path))
    if rc == -2:
        raise ValueError("%s: %s: syntax error or unknown name at line %d" % (what, path, err_line))
    if rc == -4:
        raise ValueError("%s: %s: dimensions of the files do not agree" % (what, path))
    if rc != 0:
        raise MemoryError("%s: allocation failed" % what)


def _data_dict(d):
    n, m = int(d.n), int(d.m)
    Ql = _csc(d.Q.contents)
    Q = (Ql + sp.tril(Ql, -1).T).tocsc()
    A = _csc(d.A.contents)
    q = np.ctypeslib.as_array(d.q, shape=(max(n, 1),))[:n].copy()
    l = np.ctypeslib.as_array(d.l, shape=(max(m, 1),))[:m].copy()
    u = np.ctypeslib.as_array(d.u, shape=(max(m, 1),))[:m].copy()
    return {"Q": Q, "q": q, "c": float(d.c), "A": A, "l": l, "u": u}


def read_mtx(path, lib=None):
    """MatrixMarket file -> (scipy CSC as stored, stype): symmetric files give the lower triangle, stype -1."""
    if lib is None:
        from . import lib as _lib
        lib = _lib()
    lib.qpdo_read_mtx.restype = C.c_int
    lib.qpdo_read_mtx.argtypes = [C.c_char_p, C.POINTER(C.POINTER(capi.CholmodSparse)), C.POINTER(C.c_long)]
    lib.qpdo_free_sparse.restype = None
    lib.qpdo_free_sparse.argtypes = [C.POINTER(capi.CholmodSparse)]
    out = C.POINTER(capi.CholmodSparse)()
    err_line = C.c_long(0)
    rc = lib.qpdo_read_mtx(str(path).encode(), C.byref(out), C.byref(err_line))
    _raise("qpdo_read_mtx", path, rc, err_line.value)
    try:
        return _csc(out.contents), int(out.contents.stype)
    finally:
        lib.qpdo_free_sparse(out)


def read_mtx_problem(A, Q=None, q=None, l=None, u=None, lib=None):
    """Paths of MatrixMarket files -> dict(Q, q, c, A, l, u) (include/qpdo_io.h: qpdo_read_mtx_problem)."""
    if lib is None:
        from . import lib as _lib
        lib = _lib()
    lib.qpdo_read_mtx_problem.restype = C.c_int
    lib.qpdo_read_mtx_problem.argtypes = [C.c_char_p] * 5 + [C.POINTER(C.POINTER(capi.QPDOData)), C.POINTER(C.c_long)]
    lib.qpdo_free_data.restype = None
    lib.qpdo_free_data.argtypes = [C.POINTER(capi.QPDOData)]
    enc = lambda p: None if p is None else str(p).encode()
    out = C.POINTER(capi.QPDOData)()
    err_line = C.c_long(0)
    rc = lib.qpdo_read_mtx_problem(enc(Q), enc(A), enc(q), enc(l), enc(u), C.byref(out), C.byref(err_line))
    _raise("qpdo_read_mtx_problem", A, rc, err_line.value)
    try:
        return _data_dict(out.contents)
    finally:
        lib.qpdo_free_data(out)


def read_qps(path, lib=None):
    """-> dict(Q, q, c, A, l, u, m_rows); raises ValueError with the offending line on a syntax error."""
    if lib is None:
        from . import lib as _lib
        lib = _lib()
    lib.qpdo_read_qps.restype = C.c_int
    lib.qpdo_read_qps.argtypes = [C.c_char_p, C.POINTER(C.POINTER(capi.QPDOData)), C.POINTER(C.c_long), C.POINTER(C.c_long)]
    lib.qpdo_free_data.restype = None
    lib.qpdo_free_data.argtypes = [C.POINTER(capi.QPDOData)]
    out = C.POINTER(capi.QPDOData)()
    m_rows, err_line = C.c_long(0), C.c_long(0)
    rc = lib.qpdo_read_qps(str(path).encode(), C.byref(out), C.byref(m_rows), C.byref(err_line))
    _raise("qpdo_read_qps", path, rc, err_line.value)
    try:
        d = _data_dict(out.contents)
        d["m_rows"] = int(m_rows.value)
        return d
    finally:
        lib.qpdo_free_data(out)
