"""ctypes mirror of include/qpdo_batch.h: many QPs that share one sparsity pattern, one CTA each.

Host-side logic only (packing, partitioning across ranks, the bench legs); all compute is the CUDA
kernel k_batch_solve behind the C ABI.  The reference has no batch entry: its user loops over
qpdo_setup / qpdo_solve / qpdo_cleanup (src/qpdo.c:49,304,591), which is what `reference_loop`
does with a reference build for the CPU baseline.
"""
import ctypes as C
import os
import time

import numpy as np

from . import capi

METRIC = "batched QP solves/s (MPC QPs n=100 m=150, BASELINE configs[4])"
_pd = C.POINTER(C.c_double)
_pi = C.POINTER(C.c_int)


class Summary(C.Structure):           # qpdo_batch_summary, 40 bytes
    _fields_ = [("status_val", C.c_int), ("iterations", C.c_int), ("oterations", C.c_int), ("n_active", C.c_int),
                ("objective", C.c_double), ("res_prim_norm", C.c_double), ("res_dual_norm", C.c_double)]


SUMMARY_DTYPE = np.dtype([("status_val", np.int32), ("iterations", np.int32), ("oterations", np.int32),
                          ("n_active", np.int32), ("objective", np.float64), ("res_prim_norm", np.float64),
                          ("res_dual_norm", np.float64)])
assert SUMMARY_DTYPE.itemsize == C.sizeof(Summary) == 40


def _declare(lib):
    vp = C.c_void_p
    lib.qpdo_batch_create.argtypes = [C.POINTER(vp), C.c_int, C.c_int, _pi, _pi, _pi, _pi, C.c_int,
                                      C.POINTER(capi.QPDOSettings), C.c_int]
    lib.qpdo_batch_destroy.argtypes = [vp]
    lib.qpdo_batch_destroy.restype = None
    lib.qpdo_batch_solve.argtypes = [vp, C.c_int] + [_pd] * 10 + [vp, vp]
    lib.qpdo_batch_solve_device.argtypes = [vp, C.c_int] + [vp] * 12 + [C.c_int, C.POINTER(C.c_float)]
    lib.qpdo_batch_info.argtypes = [vp, _pi, _pi, _pi, _pi]
    lib.qpdo_batch_solve_data.argtypes = [vp, C.c_int, C.POINTER(capi.QPDOSettings), vp, vp, vp, C.c_int]
    lib.qpdo_batch_solve_multi.argtypes = [C.POINTER(vp), C.c_int, C.c_int] + [_pd] * 10 + [vp, vp]
    lib.qpdo_batch_set_adaptive_order.argtypes = [vp, C.c_int]
    lib.qpdo_dev_last_error.restype = C.c_char_p
    return lib


def partition(count, world, rank):
    """contiguous block of instances for `rank`: instance b -> rank floor(b * world / count)"""
    lo = (count * rank + world - 1) // world
    hi = (count * (rank + 1) + world - 1) // world
    return lo, hi


def _dp(a):
    return None if a is None else a.ctypes.data_as(_pd)


class Batch(object):
    """create / solve / destroy over one shared pattern (include/qpdo_batch.h)."""

    def __init__(self, n, m, Qp, Qi, Ap, Ai, capacity, settings=None, device=-1):
        import qpdo_b200
        self.lib = _declare(qpdo_b200.lib())
        self.n, self.m = int(n), int(m)
        self.Qp, self.Qi, self.Ap, self.Ai = (np.ascontiguousarray(a, dtype=np.int32) for a in (Qp, Qi, Ap, Ai))
        self.nnzQ, self.nnzA = int(self.Qp[-1]), int(self.Ap[-1])
        self.capacity = int(capacity)
        st = capi.QPDOSettings()
        self.lib.qpdo_set_default_settings(C.byref(st))
        st.verbose = 0
        for k, v in (settings or {}).items():
            if not hasattr(st, k):
                raise KeyError("unknown QPDO setting %r" % k)
            setattr(st, k, v)
        self.settings = st
        h = C.c_void_p()
        rc = self.lib.qpdo_batch_create(C.byref(h), self.n, self.m, self.Qp.ctypes.data_as(_pi), self.Qi.ctypes.data_as(_pi),
                                        self.Ap.ctypes.data_as(_pi), self.Ai.ctypes.data_as(_pi), self.capacity, C.byref(st), device)
        if rc:
            raise RuntimeError("qpdo_batch_create failed (%d): %s" % (rc, self.lib.qpdo_dev_last_error().decode()))
        self.h = h

    def info(self):
        a, b, c, d = C.c_int(), C.c_int(), C.c_int(), C.c_int()
        self.lib.qpdo_batch_info(self.h, C.byref(a), C.byref(b), C.byref(c), C.byref(d))
        return dict(smem_bytes=a.value, ctas=b.value, lnz=c.value, threads=d.value)

    def solve(self, Qx, Ax, q, l, u, c=None, x0=None, y0=None, want_active=True):
        """HOST buffers in, host results out (copies inside)."""
        f = lambda a: None if a is None else np.ascontiguousarray(a, dtype=np.float64)
        Qx, Ax, q, l, u, c, x0, y0 = map(f, (Qx, Ax, q, l, u, c, x0, y0))
        count = q.shape[0]
        assert Qx.shape == (count, self.nnzQ) and Ax.shape == (count, self.nnzA) and q.shape == (count, self.n)
        assert l.shape == (count, self.m) and u.shape == (count, self.m)
        x = np.empty((count, self.n)); y = np.empty((count, self.m))
        summ = np.zeros(count, dtype=SUMMARY_DTYPE)
        act = np.zeros((count, self.m), dtype=np.uint8) if want_active else None
        rc = self.lib.qpdo_batch_solve(self.h, count, _dp(Qx), _dp(Ax), _dp(q), _dp(l), _dp(u), _dp(c), _dp(x0), _dp(y0),
                                       _dp(x), _dp(y), summ.ctypes.data, None if act is None else act.ctypes.data)
        if rc:
            raise RuntimeError("qpdo_batch_solve failed (%d): %s" % (rc, self.lib.qpdo_dev_last_error().decode()))
        return x, y, summ, act

    def solve_device(self, count, ptrs, out_ptrs, kernel_ms=True):
        """device pointers (ints) in the order of qpdo_batch_solve_device; returns kernel ms"""
        ms = C.c_float(0)
        rc = self.lib.qpdo_batch_solve_device(self.h, count, *[C.c_void_p(p) if p else None for p in ptrs],
                                              *[C.c_void_p(p) if p else None for p in out_ptrs], 1,
                                              C.byref(ms) if kernel_ms else None)
        if rc:
            raise RuntimeError("qpdo_batch_solve_device failed (%d): %s" % (rc, self.lib.qpdo_dev_last_error().decode()))
        return ms.value

    def set_adaptive_order(self, on=True):
        """longest-first queue order from the iteration counts of the previous solve (include/qpdo_batch.h)"""
        self.lib.qpdo_batch_set_adaptive_order(self.h, 1 if on else 0)

    def destroy(self):
        if getattr(self, "h", None):
            self.lib.qpdo_batch_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.destroy()
        except Exception:
            pass


def from_mpc(batch, capacity=None, settings=None, device=-1):
    """Batch object + packed value arrays for a problems.mpc_batch() dict."""
    n, m = batch["n"], batch["m"]
    Qp = np.arange(n + 1, dtype=np.int32); Qi = np.arange(n, dtype=np.int32)
    B = batch["q"].shape[0]
    bt = Batch(n, m, Qp, Qi, batch["Ap"], batch["Ai"], capacity or B, settings, device)
    Qx = np.tile(batch["Qd"], (B, 1))
    return bt, (Qx, batch["Ax"], batch["q"], batch["l"], batch["u"])


def solve_multi(batches, Qx, Ax, q, l, u, c=None, x0=None, y0=None, want_active=True):
    """qpdo_batch_solve_multi: one process, one Batch per GPU, HOST buffers in, host results out."""
    f = lambda a: None if a is None else np.ascontiguousarray(a, dtype=np.float64)
    Qx, Ax, q, l, u, c, x0, y0 = map(f, (Qx, Ax, q, l, u, c, x0, y0))
    b0 = batches[0]
    count = q.shape[0]
    x = np.empty((count, b0.n)); y = np.empty((count, b0.m))
    summ = np.zeros(count, dtype=SUMMARY_DTYPE)
    act = np.zeros((count, b0.m), dtype=np.uint8) if want_active else None
    hs = (C.c_void_p * len(batches))(*[bt.h for bt in batches])
    rc = b0.lib.qpdo_batch_solve_multi(hs, len(batches), count, _dp(Qx), _dp(Ax), _dp(q), _dp(l), _dp(u), _dp(c), _dp(x0), _dp(y0),
                                       _dp(x), _dp(y), summ.ctypes.data, None if act is None else act.ctypes.data)
    if rc:
        raise RuntimeError("qpdo_batch_solve_multi failed (%d): %s" % (rc, b0.lib.qpdo_dev_last_error().decode()))
    return x, y, summ, act
