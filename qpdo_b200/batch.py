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


# ------------------------------------------------------------------------------------------ bench legs
C5_BATCH = 65536


def algorithmic_bytes_per_qp(n, m, nnzQ, nnzA):
    """HBM bytes one instance must move: its values in (Qx, Ax, q, l, u), its results out (x, y, summary,
    active set).  The whole solver loop in between lives in shared memory (batch kernel)."""
    return 8 * (nnzQ + nnzA + n + 2 * m) + 8 * (n + m) + 40 + m


def bench(args, dist, ClockSampler, peaks):
    """C5 arm of bench.py: the whole batch (BASELINE configs[4]) partitioned in contiguous blocks over
    the ranks; a step = every rank solving its block once.  `value`: inputs resident in HBM, device
    pointers through qpdo_batch_solve_device.  `e2e`: qpdo_batch_solve with pinned HOST buffers."""
    import torch
    from . import problems
    B = args.size or C5_BATCH
    lo, hi = partition(B, dist.world, dist.rank)
    cnt = hi - lo
    torch.cuda.set_device(dist.local)
    b = problems.mpc_batch(cnt, first=lo)
    bt, vals = from_mpc(b, capacity=cnt, device=dist.local)
    n, m, nnzQ, nnzA = bt.n, bt.m, bt.nnzQ, bt.nnzA
    info = bt.info()
    # pinned host buffers (the e2e leg) and device-resident copies (the value leg)
    host = [torch.from_numpy(np.ascontiguousarray(v)).pin_memory() for v in vals]
    dev = [h.cuda(non_blocking=False) for h in host]
    dx = torch.empty((cnt, n), dtype=torch.float64, device="cuda")
    dy = torch.empty((cnt, m), dtype=torch.float64, device="cuda")
    dsum = torch.zeros((cnt, 40), dtype=torch.uint8, device="cuda")
    dact = torch.zeros((cnt, m), dtype=torch.uint8, device="cuda")
    hx = torch.empty((cnt, n), dtype=torch.float64).pin_memory()
    hy = torch.empty((cnt, m), dtype=torch.float64).pin_memory()
    hsum = torch.zeros((cnt, 40), dtype=torch.uint8).pin_memory()
    hact = torch.zeros((cnt, m), dtype=torch.uint8).pin_memory()
    torch.cuda.synchronize()
    ptrs = [t.data_ptr() for t in dev] + [0, 0, 0]                       # c, x0, y0 = NULL
    outs = [dx.data_ptr(), dy.data_ptr(), dsum.data_ptr(), dact.data_ptr()]

    def step_device():
        return bt.solve_device(cnt, ptrs, outs, kernel_ms=True)

    def step_host():
        rc = bt.lib.qpdo_batch_solve(bt.h, cnt, *[C.cast(h.data_ptr(), _pd) for h in host], None, None, None,
                                     C.cast(hx.data_ptr(), _pd), C.cast(hy.data_ptr(), _pd), hsum.data_ptr(), hact.data_ptr())
        if rc:
            raise RuntimeError("qpdo_batch_solve failed: %s" % bt.lib.qpdo_dev_last_error().decode())

    for _ in range(max(args.warmup, 3)):
        step_device()
    torch.cuda.synchronize(); dist.barrier()
    kern_ms = []
    with ClockSampler(dist.local) as clk:
        t0 = time.perf_counter()
        for _ in range(args.steps):
            kern_ms.append(step_device())
        torch.cuda.synchronize()
        wall = time.perf_counter() - t0
    dist.barrier()
    dev_ms = float(np.sum(kern_ms))                       # CUDA events on the batch's own stream
    step_host()                                           # e2e warm-up
    torch.cuda.synchronize(); dist.barrier()
    e2e_steps = max(1, min(args.steps, 3))
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        step_host()
    e2e_wall = time.perf_counter() - t0
    dist.barrier()
    # status / objective summaries: the only inter-rank traffic, 40 B per QP
    summ = dsum.cpu().numpy().view(SUMMARY_DTYPE).reshape(-1)
    pad = np.zeros(((B + dist.world - 1) // dist.world) * 40, dtype=np.uint8)
    pad[:cnt * 40] = summ.view(np.uint8)
    allsum = dist.gather_bytes(pad)
    parts = []
    for r, a in enumerate(allsum):
        l2, h2 = partition(B, dist.world, r)
        parts.append(a[:(h2 - l2) * 40].view(SUMMARY_DTYPE))
    allsum = np.concatenate(parts)
    # the host-buffer leg must reproduce the device leg bit for bit
    same = bool(np.array_equal(hsum.numpy().view(SUMMARY_DTYPE).reshape(-1)["iterations"], summ["iterations"]))
    worst = dist.reduce([dev_ms, wall * 1e3, e2e_wall * 1e3], "max")
    total_iter = dist.reduce([float(summ["iterations"].sum())], "sum")[0]
    pk, pk_src = peaks()
    bytes_qp = algorithmic_bytes_per_qp(n, m, nnzQ, nnzA)
    k_ms = worst[0] / args.steps
    rl = {"bound": "hbm", "kernel": "k_batch_solve (one CTA per QP, solver loop in shared memory)",
          "achieved": bytes_qp * cnt / (k_ms * 1e-3) / 1e9, "peak": pk["hbm_gbs"], "peak_source": pk_src, "unit": "GB/s",
          "traffic": None, "bytes_per_qp": bytes_qp,
          "note": "HBM is touched only to load an instance and store its result; the %d loop passes per QP run in shared memory "
                  "(latency/issue-bound, see profiles/)" % round(total_iter / B)}
    rl["frac"] = rl["achieved"] / rl["peak"]
    h2d = 8 * cnt * (nnzQ + nnzA + n + 2 * m)
    d2h = cnt * (8 * (n + m) + 40 + m)
    out = {"metric": METRIC, "value": B * args.steps / (worst[1] * 1e-3), "unit": "QP/s", "n_gpus": dist.world, "steps": args.steps,
           "warmup": max(args.warmup, 3), "ms_per_step": worst[1] / args.steps, "higher_is_better": True, "scaling": "strong",
           "vs_baseline": None, "dtype": "f64", "data": "synthetic",
           "config": {"workload": "c5: %d MPC QPs (n=%d, m=%d, nnzA=%d, common pattern), contiguous blocks over %d rank(s); "
                                  "inputs %.0f MB per rank > 126 MB L2 (no flush needed)" % (B, n, m, nnzA, dist.world, h2d / 1e6),
                      "batch": B, "n": n, "m": m, "nnzA": nnzA, "per_rank": cnt},
           "kernel_ms_per_step": k_ms, "kernel_qp_per_s": B / (k_ms * 1e-3),
           "batch_kernel": dict(info, mean_iterations=total_iter / B),
           "result": {"solved": int((allsum["status_val"] == 1).sum()), "primal_infeasible": int((allsum["status_val"] == -3).sum()),
                      "dual_infeasible": int((allsum["status_val"] == -4).sum()), "max_iter_reached": int((allsum["status_val"] == -2).sum()),
                      "objective_sum": float(allsum["objective"][allsum["status_val"] == 1].sum()), "host_leg_matches_device_leg": same},
           "roofline": rl,
           "e2e": {"value": B * e2e_steps / (worst[2] * 1e-3), "unit": "QP/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                   "call": "qpdo_batch_solve (pinned host buffers in, x / y / summary / active set out)"},
           "gpu_launches": args.steps, "clocks": clk.summary()}
    bt.destroy()
    return out


def bench_reference(args, bounded=True):
    """The reference's own CPU path for C5: the per-problem qpdo_setup / qpdo_solve / qpdo_cleanup loop
    (oracle/ref_batch.c over oracle/_ref/libqpdo_ref.so) on all host cores, over a bounded sample
    of the same instances."""
    import oracle
    from . import problems
    ncores = os.cpu_count() or 1
    sample = min(args.size or C5_BATCH, 2048 * max(1, ncores // 4))
    b = problems.mpc_batch(sample)
    n, m = b["n"], b["m"]
    Qp = np.arange(n + 1); Qi = np.arange(n); Qx = np.tile(b["Qd"], (sample, 1))
    if oracle.have_ref():
        kind = "reference"
        best = None
        for _ in range(max(1, min(getattr(args, "steps", 1) or 1, 3))):
            secs, summ = oracle.ref_batch_loop(n, m, Qp, Qi, b["Ap"], b["Ai"], Qx, b["Ax"], b["q"], b["l"], b["u"], threads=ncores)
            best = secs if best is None else min(best, secs)
        secs = best
    else:
        kind, ncores = "port", 1
        sample = min(sample, 512)
        t0 = time.perf_counter()
        for i in range(sample):
            oracle.solve(problems.mpc_instance(b, i), trace=False)
        secs = time.perf_counter() - t0
    return {"value": sample / secs, "unit": "QP/s", "cores": ncores, "kind": kind,
            "sample": "first %d instances of the C5 batch, one qpdo_setup/qpdo_solve/qpdo_cleanup per QP, %d host thread(s)" % (sample, ncores),
            "wall_s": secs}


def smoke():
    """a few MPC QPs through the batch C ABI on cuda:0, checked against the oracle"""
    import oracle
    from . import problems
    B = 12
    b = problems.mpc_batch(B, first=100)
    bt, vals = from_mpc(b)
    x, y, summ, act = bt.solve(*vals)
    bt.destroy()
    for i in range(B):
        o = oracle.solve(problems.mpc_instance(b, i), trace=False)
        ok = (int(summ[i]["status_val"]), int(summ[i]["iterations"]), int(summ[i]["oterations"])) == (o.status_val, o.iterations, o.oterations)
        ok = ok and np.array_equal(act[i].astype(np.int8), o.active_set)
        if o.status_val == 1:
            ok = ok and np.abs(x[i] - o.x).max() <= 1e-8 * max(1.0, np.abs(o.x).max()) and np.abs(y[i] - o.y).max() <= 1e-8 * max(1.0, np.abs(o.y).max())
        if not ok:
            raise AssertionError("smoke(): batch kernel differs from the oracle on MPC instance %d" % i)
    print("smoke batch: %d MPC QPs, statuses %s, iterations %s -> OK" % (B, list(summ["status_val"]), list(summ["iterations"])))
