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
    if rc == -1:
        raise OSError("qpdo_read_qps: cannot open %s" % path)
    if rc == -2:
        raise ValueError("qpdo_read_qps: %s: syntax error or unknown name at line %d" % (path, err_line.value))
    if rc != 0:
        raise MemoryError("qpdo_read_qps: allocation failed")
    try:
        d = out.contents
        n, m = int(d.n), int(d.m)
        Ql = _csc(d.Q.contents)
        Q = (Ql + sp.tril(Ql, -1).T).tocsc()
        A = _csc(d.A.contents)
        q = np.ctypeslib.as_array(d.q, shape=(max(n, 1),))[:n].copy()
        l = np.ctypeslib.as_array(d.l, shape=(max(m, 1),))[:m].copy()
        u = np.ctypeslib.as_array(d.u, shape=(max(m, 1),))[:m].copy()
        return {"Q": Q, "q": q, "c": float(d.c), "A": A, "l": l, "u": u, "m_rows": int(m_rows.value)}
    finally:
        lib.qpdo_free_data(out)
