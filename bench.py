#!/usr/bin/env python
"""bench.py -- QPDO Newton hot path on B200: KKT factor+solve ms/iter (C3) and batched QP solves/s (C5).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c5|c3|c1|c4] [--impl reference]

Prints ONE JSON line (rank 0).  The product arm calls only the CUDA library through the qpdo.h /
qpdo_batch.h C ABI; oracle/ is executed only by the `cpu_baseline` leg and by `--impl reference`.
A "step" is one pass of the hot path over one batch of synthetic input:
  c5: solving the whole batch of MPC QPs (BASELINE.json configs[4]);  value = QP solves/s, all ranks.
      At N=1 the default line also carries `kkt` (the other half of BASELINE's metric: the c3 record
      with the reference timed on the same box) and `parity` (all 65,536 GPU summaries against the
      reference's, instance by instance).
  c3: one full solve of the CONT-300-shaped QP (configs[2]);          value = KKT factor+solve ms per Newton iteration
  c1: one solve of the random n=1000 m=500 QP (configs[0]);           value = ms per solve (latency path)
  c4: the 20-step Lasso regularisation path at n=200k m=100k (configs[3]); value = KKT ms per Newton iteration
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))), "measured"
    except Exception:
        return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0}, "fallback"


class ClockSampler(object):
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.stop = index, [], False
        self.th = threading.Thread(target=self._run, daemon=True)

    def _run(self):
        while not self.stop:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                self.rows.append([c.strip() for c in out.strip().split(",")])
            except Exception:
                pass
            time.sleep(0.2)

    def __enter__(self):
        self.th.start()
        return self

    def __exit__(self, *a):
        self.stop = True
        self.th.join(timeout=6)

    def summary(self):
        sm = [float(r[0]) for r in self.rows if len(r) >= 6 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 6 and r[1].replace(".", "").isdigit()]
        reasons = set()
        for r in self.rows:
            if len(r) >= 6:
                for nm, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[2:6]):
                    if v.lower().startswith("active"):
                        reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


class Dist(object):
    """one process per GPU; torch.distributed is plumbing (barrier + max/sum over ranks) only"""

    def __init__(self, gpus, backend=None):
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        self.td = None
        if self.world > 1:
            import torch
            import torch.distributed as td
            self.torch, self.td = torch, td
            if backend is None:
                backend = "nccl" if torch.cuda.is_available() else "gloo"
            self.backend = backend
            if backend == "nccl":
                torch.cuda.set_device(self.local)
            td.init_process_group(backend=backend)
            self.dev = torch.device("cuda", self.local) if backend == "nccl" else torch.device("cpu")

    def barrier(self):
        if self.td:
            self.td.barrier()

    def reduce(self, vals, op="max"):
        if not self.td:
            return list(vals)
        t = self.torch.tensor(list(vals), dtype=self.torch.float64, device=self.dev)
        self.td.all_reduce(t, op=self.td.ReduceOp.MAX if op == "max" else self.td.ReduceOp.SUM)
        return t.cpu().tolist()

    def gather_bytes(self, arr):
        """all-gather equal-sized uint8 numpy arrays (status / objective summaries)"""
        if not self.td:
            return [arr]
        t = self.torch.from_numpy(arr.copy()).to(self.dev)
        outs = [self.torch.empty_like(t) for _ in range(self.world)]
        self.td.all_gather(outs, t)
        return [o.cpu().numpy() for o in outs]

    def close(self):
        if self.td:
            self.td.destroy_process_group()


# ------------------------------------------------------------------------------------------ C3
C3_N = 300


def c3_problem(N):
    from qpdo_b200 import problems
    return problems.cont_like(N)


def c3_algorithmic(info, n, m, nnzA, nnzQ):
    """algorithmic bytes / flops per Newton iteration of the KKT phases"""
    lnz = float(info.lnz)
    return {"solve_bytes": 2 * (12 * lnz + 8 * n) + 16 * n,            # triangular solves, both sweeps
            "factor_flops": float(info.flops),                        # sum_j colcount_j^2
            "spmv_bytes": 3 * (12 * nnzA + 4 * (n + 1) + 8 * (n + m)) + (12 * nnzQ + 4 * (n + 1) + 16 * n)}


def run_c3(args, dist):
    import qpdo_b200
    from qpdo_b200 import capi
    lib = qpdo_b200.lib()
    N = args.size or C3_N
    p = c3_problem(N)
    n, m = p["q"].size, p["A"].shape[0]
    st = dict(verbose=0)
    if args.max_iter:
        st["max_iter"] = args.max_iter
    s = qpdo_b200.QPDO()
    t0 = time.time()
    s.setup(p["Q"], p["q"], p["A"], p["l"], p["u"], st)
    setup_s = time.time() - t0
    for _ in range(max(args.warmup, 1)):
        r = s.solve()
    info = capi.DevSymbolicInfo()                      # after a solve: large problems analyse lazily (pattern management)
    lib.qpdo_dev_get_symbolic_info(s.device, C.byref(info))
    prof = capi.DevProfile()
    lib.qpdo_dev_set_profiling(s.device, 1)
    lib.qpdo_dev_get_profile(s.device, C.byref(prof), 1)
    dist.barrier()
    with ClockSampler(dist.local) as clk:
        t0 = time.time()
        for _ in range(args.steps):
            r = s.solve()
        wall = time.time() - t0
    dist.barrier()
    lib.qpdo_dev_get_profile(s.device, C.byref(prof), 1)
    lib.qpdo_dev_set_profiling(s.device, 0)
    s.delete()
    # end to end through the public API with host buffers: setup (H2D of Q, A, q, l, u + analysis) +
    # solve (D2H of x, y, certificates) per step
    e2e_steps = max(1, min(args.steps, 2))
    t0 = time.time()
    for _ in range(e2e_steps):
        s2 = qpdo_b200.QPDO()
        s2.setup(p["Q"], p["q"], p["A"], p["l"], p["u"], st)
        r2 = s2.solve()
        s2.delete()
    e2e_wall = (time.time() - t0) / e2e_steps
    nnzA, nnzQ = p["A"].nnz, int(n)
    newton = max(1, prof.n_newton)
    kkt_ms = (prof.assemble_ms + prof.factor_ms + prof.solve_ms) / newton
    alg = c3_algorithmic(info, n, m, nnzA, nnzQ)
    worst = dist.reduce([kkt_ms, wall / args.steps * 1e3], "max")
    pk, pk_src = peaks()
    solve_ms = prof.solve_ms / newton
    rl = {"bound": "hbm", "kernel": "supernodal triangular solves (k_fwd / k_bwd and the cluster kernels k_fwd_cl / k_bwd_cl, one graph)", "achieved": alg["solve_bytes"] / (solve_ms * 1e-3) / 1e9,
          "peak": pk["hbm_gbs"], "peak_source": pk_src, "unit": "GB/s", "traffic": None}
    rl["frac"] = rl["achieved"] / rl["peak"]
    rec3 = None
    try:
        rec3 = json.load(open(os.path.join(ROOT, "profiles", "c3_ncu.json")))
        rl["traffic"] = rec3.get("solve_dram_bytes_per_newton")
        rl["traffic_source"] = "sum of dram__bytes_read+write over the k_fwd* / k_bwd* launches of one Newton step, ncu: profiles/c3_ncu.json @ %s" % rec3.get("commit")
    except Exception:
        pass
    # numeric factorization against the FP64 peaks measured on this device in this run (register-resident DMMA / DFMA loops)
    dm, df = C.c_double(0), C.c_double(0)
    lib.qpdo_dev_measure_fp64(dist.local, C.byref(dm), C.byref(df))
    fac_tf = alg["factor_flops"] / (prof.factor_ms / newton * 1e-3) / 1e12
    rlf = {"bound": "tensor", "kernel": "multifrontal LDL' (k_front_smem, k_big_panel: row-tile solve and rank-32 update by mma.sync.m8n8k4.f64)", "achieved": fac_tf,
           "peak": dm.value, "peak_source": "qpdo_dev_measure_fp64: mma.sync.m8n8k4.f64 loop over all SMs, this run", "dfma_peak": df.value,
           "unit": "TFLOP/s (f64)", "frac": fac_tf / dm.value if dm.value > 0 else None, "frac_of_dfma_peak": fac_tf / df.value if df.value > 0 else None,
           "flops_per_factorization": alg["factor_flops"]}
    out = {"metric": "KKT factor+solve ms per Newton iteration (CONT-%d-shaped QP, n=%d m=%d)" % (N, n, m),
           "value": worst[0], "unit": "ms/iter", "n_gpus": dist.world, "steps": args.steps, "warmup": args.warmup,
           "ms_per_step": worst[1], "higher_is_better": False, "scaling": "replicas", "vs_baseline": None, "dtype": "f64",
           "data": "synthetic", "config": {"workload": "c3: CONT-%d-like banded PDE-constraint QP, one solve per step; "
                                           "working set (L panels + update matrices) %.0f MB vs 126 MB L2" %
                                           (N, 8e-6 * (info.lnz_stored + info.usize)),
                                           "N": N, "n": n, "m": m, "nnzA": int(nnzA)},
           "phases_ms_per_newton": {"assemble": prof.assemble_ms / newton, "factor": prof.factor_ms / newton, "solve": solve_ms,
                                    "spmv": prof.spmv_ms / newton, "linesearch": prof.linesearch_ms / newton,
                                    "residual": prof.residual_ms / max(1, prof.n_residual), "other": prof.other_ms / newton},
           "symbolic": {"nsuper": info.nsuper, "nlevel": info.nlevel, "max_front": info.max_front, "lnz": int(info.lnz),
                        "lnz_stored": int(info.lnz_stored), "flops": info.flops, "analyze_s": info.analyze_seconds, "setup_s": setup_s},
           "result": {"status": r.status_val, "iterations": r.iterations, "oterations": r.oterations, "objective": r.objective},
           "factor_gflops": alg["factor_flops"] / (prof.factor_ms / newton * 1e-3) / 1e9,
           "roofline": rl, "roofline_factor": rlf,
           "e2e": {"value": e2e_wall * 1e3 / max(1, r2.iterations), "unit": "ms/iter (setup + solve wall clock / loop passes)",
                   "h2d_bytes_per_step": int(12 * (nnzA + nnzQ) + 8 * (n + 2 * m) + 8 * (n + 1)),
                   "d2h_bytes_per_step": int(8 * 3 * (n + m) + 8 * m), "setup_plus_solve_s": e2e_wall},
           "gpu_launches": int(prof.kernel_launches), "clocks": clk.summary()}
    return out


def ref_c3(args, sample_iters=12):
    """the unmodified reference (oracle/_ref, CHOLMOD + OpenBLAS on all host cores) on a bounded
    sample of C3: the first `sample_iters` loop passes of the same solve"""
    import oracle
    from qpdo_b200 import capi
    ncores = os.cpu_count() or 1
    os.environ.setdefault("OPENBLAS_NUM_THREADS", str(ncores))
    N = args.size or C3_N
    p = c3_problem(N)
    kind = "reference" if oracle.have_ref() else "port"
    t0 = time.time()
    if kind == "reference":
        s = capi.Solver(oracle.ref_library())
        s.setup(p["Q"], p["q"], p["A"], p["l"], p["u"], dict(verbose=0, max_iter=sample_iters))
        r = s.solve(); s.delete()
        secs, iters = r.solve_time, r.iterations
    else:
        r = oracle.solve(p, dict(max_iter=sample_iters), trace=False)
        secs, iters = r.solve_seconds, r.iterations
    return {"value": secs * 1e3 / max(1, iters), "unit": "ms/iter (whole loop pass incl. factor+solve)", "cores": ncores if kind == "reference" else 1,
            "kind": kind, "sample": "first %d loop passes of the C3 (N=%d) solve, %s" %
            (iters, N, "CHOLMOD supernodal + OpenBLAS" if kind == "reference" else "C restatement"), "wall_s": time.time() - t0}


# ------------------------------------------------------------------------------------------ C5
from qpdo_b200 import batch as qbatch      # ctypes mirror of include/qpdo_batch.h (no compute of its own)
import ctypes as C

C5_BATCH = 65536


def ncu_record():
    """figures of the last `ncu --set full` capture of k_batch_solve (DRAM bytes and executed warp instructions per
    QP), written by tools/ncu_batch_record.py with the commit they were taken at: profiles/c5_ncu.json"""
    try:
        return json.load(open(os.path.join(ROOT, "profiles", "c5_ncu.json")))
    except Exception:
        return None


def knife_edge_ids():
    """C5 instances on which the UNMODIFIED reference changes its own iteration total under one-ulp changes of q
    (profiles/r2_knife_edge.json, tools/knife_edge.py)"""
    try:
        cases = json.load(open(os.path.join(ROOT, "profiles", "r2_knife_edge.json")))["cases"]
        return sorted(int(c["case"].split()[-1]) for c in cases if c["case"].startswith("c5 instance"))
    except Exception:
        return []


def parity_record(gpu, ref):
    """all GPU summaries against the reference's, instance by instance (north_star: status and iteration counts exact,
    objective to 1e-8 relative)"""
    n = min(gpu.size, ref.size)
    g, r = gpu[:n], ref[:n]
    st = np.nonzero(g["status_val"] != r["status_val"])[0]
    it = np.nonzero((g["iterations"] != r["iterations"]) | (g["oterations"] != r["oterations"]))[0]
    ok = (g["status_val"] == 1) & (r["status_val"] == 1)
    rel = np.abs(g["objective"][ok] - r["objective"][ok]) / np.maximum(1.0, np.abs(r["objective"][ok]))
    knife = set(knife_edge_ids())
    return {"instances": int(n), "status_mismatch": int(st.size), "iteration_mismatch": int(it.size),
            "first_ids": [int(i) for i in it[:16]],
            "iteration_deltas": [int(g["iterations"][i]) - int(r["iterations"][i]) for i in it[:16]],
            "unexplained": [int(i) for i in it if int(i) not in knife][:16],
            "documented_knife_edge_ids": sorted(knife),
            "max_rel_objective": float(rel.max()) if rel.size else 0.0,
            "reference_iterations_total": int(r["iterations"].sum()), "gpu_iterations_total": int(g["iterations"].sum()),
            "against": "oracle/_ref (unmodified reference) on the whole batch, same run",
            "note": "a mismatch listed under documented_knife_edge_ids is a decision the reference itself flips under a "
                    "one-ulp change of q (profiles/r2_knife_edge.json)"}


def algorithmic_bytes_per_qp(n, m, nnzQ, nnzA):
    """HBM bytes one instance must move: its values in (Qx, Ax, q, l, u), its results out (x, y, summary,
    active set).  The whole solver loop in between lives in shared memory (batch kernel)."""
    return 8 * (nnzQ + nnzA + n + 2 * m) + 8 * (n + m) + 40 + m


def c5_config(B, n, m, nnzA, world):
    per_rank = (B + world - 1) // world
    return {"workload": "c5: %d MPC QPs (n=%d, m=%d, nnzA=%d, common pattern), contiguous blocks over %d rank(s); "
                        "inputs %.0f MB per rank vs 126 MB L2" % (B, n, m, nnzA, world, 8e-6 * per_rank * (n + nnzA + n + 2 * m)),
            "batch": B, "n": n, "m": m, "nnzA": nnzA, "per_rank": per_rank}


def run_c5(args, dist):
    """C5 arm of bench.py: the whole batch (BASELINE configs[4]) partitioned in contiguous blocks over
    the ranks; a step = every rank solving its block once.  `value`: inputs resident in HBM, device
    pointers through qpdo_batch_solve_device.  `e2e`: qpdo_batch_solve with pinned HOST buffers."""
    import torch
    from qpdo_b200 import problems
    B = args.size or C5_BATCH
    lo, hi = qbatch.partition(B, dist.world, dist.rank)
    cnt = hi - lo
    torch.cuda.set_device(dist.local)
    b = problems.mpc_batch(cnt, first=lo)
    bt, vals = qbatch.from_mpc(b, capacity=cnt, device=dist.local)
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
        rc = bt.lib.qpdo_batch_solve(bt.h, cnt, *[C.cast(h.data_ptr(), qbatch._pd) for h in host], None, None, None,
                                     C.cast(hx.data_ptr(), qbatch._pd), C.cast(hy.data_ptr(), qbatch._pd), hsum.data_ptr(), hact.data_ptr())
        if rc:
            raise RuntimeError("qpdo_batch_solve failed: %s" % bt.lib.qpdo_dev_last_error().decode())

    # work-queue order: longest instance first by the iteration counts of the previous solve of the same batch (the
    # receding-horizon pattern the batch entry exists for); the warm-up steps leave that order on the device.  The
    # natural-order figure of the same launch is reported beside it (`kernel_ms_natural_order`).
    nat_ms = [step_device() for _ in range(2)][-1]
    bt.set_adaptive_order(not args.natural_order)
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
    summ = dsum.cpu().numpy().view(qbatch.SUMMARY_DTYPE).reshape(-1).copy()

    def gather_summaries(local):
        """status / objective summaries: the only inter-rank traffic, 40 B per QP, one all-gather"""
        pad = np.zeros(((B + dist.world - 1) // dist.world) * 40, dtype=np.uint8)
        pad[:cnt * 40] = local.view(np.uint8).reshape(-1)
        parts = []
        for r, a in enumerate(dist.gather_bytes(pad)):
            l2, h2 = qbatch.partition(B, dist.world, r)
            parts.append(a[:(h2 - l2) * 40].view(qbatch.SUMMARY_DTYPE))
        return np.concatenate(parts)

    def step_e2e():
        step_host()
        return gather_summaries(hsum.numpy())             # the NCCL all-gather is part of the end-to-end step

    step_e2e()                                            # e2e warm-up (also NCCL's first-call setup)
    torch.cuda.synchronize(); dist.barrier()
    e2e_steps = max(1, min(args.steps, 3))
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        allsum = step_e2e()
    e2e_wall = time.perf_counter() - t0
    dist.barrier()
    # the host-buffer leg must reproduce the device leg bit for bit
    same = bool(np.array_equal(hsum.numpy().view(qbatch.SUMMARY_DTYPE).reshape(-1)["iterations"], summ["iterations"]))
    worst = dist.reduce([dev_ms, wall * 1e3, e2e_wall * 1e3], "max")
    total_iter = dist.reduce([float(summ["iterations"].sum())], "sum")[0]
    pk, pk_src = peaks()
    rec = ncu_record()
    bytes_qp = algorithmic_bytes_per_qp(n, m, nnzQ, nnzA)
    k_ms = worst[0] / args.steps
    rl = {"bound": "hbm", "kernel": "k_batch_solve (one CTA per QP, solver loop in shared memory)",
          "achieved": bytes_qp * cnt / (k_ms * 1e-3) / 1e9, "peak": pk["hbm_gbs"], "peak_source": pk_src, "unit": "GB/s",
          "traffic": (rec["dram_bytes_per_qp"] * cnt) if rec else None,
          "traffic_source": ("dram__bytes_read+write of one ncu --set full launch, per QP x QPs per launch: profiles/c5_ncu.json, captured at commit %s" % rec["commit"]) if rec else "no ncu record",
          "bytes_per_qp": bytes_qp,
          "note": "HBM is touched only to load an instance and store its result; the %d loop passes per QP run in shared memory "
                  "(latency/issue-bound, see profiles/)" % round(total_iter / B)}
    rl["frac"] = rl["achieved"] / rl["peak"]
    # the resource that binds: warp-instruction issue (4 schedulers x SMs x SM clock); executed warp instructions per
    # QP from the ncu record
    issue = None
    if rec and rec.get("warp_inst_per_qp"):
        sms = 148
        f_sm = float(pk.get("sm_max_mhz", 1965.0)) * 1e6
        issue = {"bound": "issue slots", "achieved": rec["warp_inst_per_qp"] * cnt / (k_ms * 1e-3) / 1e12, "peak": 4 * sms * f_sm / 1e12,
                 "unit": "T warp-inst/s", "warp_inst_per_qp": rec["warp_inst_per_qp"], "source": "profiles/c5_ncu.json @ %s" % rec["commit"]}
        issue["frac"] = issue["achieved"] / issue["peak"]
    h2d = 8 * cnt * (nnzQ + nnzA + n + 2 * m)
    d2h = cnt * (8 * (n + m) + 40 + m)
    out = {"metric": qbatch.METRIC, "value": B * args.steps / (worst[1] * 1e-3), "unit": "QP/s", "n_gpus": dist.world, "steps": args.steps,
           "warmup": max(args.warmup, 3), "ms_per_step": worst[1] / args.steps, "higher_is_better": True, "scaling": "strong",
           "vs_baseline": None, "dtype": "f64", "data": "synthetic",
           "config": c5_config(B, n, m, nnzA, dist.world),
           "kernel_ms_per_step": k_ms, "kernel_qp_per_s": B / (k_ms * 1e-3),
           "kernel_ms_natural_order": dist.reduce([nat_ms], "max")[0],
           "queue_order": "natural" if args.natural_order else "longest first, from the iteration counts of the previous step (qpdo_batch_set_adaptive_order); results do not depend on it",
           "batch_kernel": dict(info, mean_iterations=total_iter / B),
           "result": {"solved": int((allsum["status_val"] == 1).sum()), "primal_infeasible": int((allsum["status_val"] == -3).sum()),
                      "dual_infeasible": int((allsum["status_val"] == -4).sum()), "max_iter_reached": int((allsum["status_val"] == -2).sum()),
                      "objective_sum": float(allsum["objective"][allsum["status_val"] == 1].sum()), "host_leg_matches_device_leg": same},
           "roofline": rl, "roofline_issue": issue,
           "e2e": {"value": B * e2e_steps / (worst[2] * 1e-3), "unit": "QP/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                   "call": "qpdo_batch_solve (pinned host buffers in, x / y / summary / active set out) + the all-gather of the 40-byte summaries over %s, all inside the timed region" % (dist.backend if dist.td else "no collective (one rank)")},
           "gpu_launches": args.steps, "clocks": clk.summary()}
    out["result"]["share_primal_infeasible"] = out["result"]["primal_infeasible"] / float(B)
    bt.destroy()
    return out, allsum


def ref_c5(args, bounded=True, want_summaries=False):
    """The reference's own CPU path for C5: the per-problem qpdo_setup / qpdo_solve / qpdo_cleanup loop
    (oracle/ref_batch.c over oracle/_ref/libqpdo_ref.so) on all host cores, over a bounded sample
    of the same instances."""
    import oracle
    from qpdo_b200 import problems
    ncores = os.cpu_count() or 1
    sample = min(args.size or C5_BATCH, C5_BATCH)          # the whole batch: ~3 s on 16 cores per pass
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
    out = {"value": sample / secs, "unit": "QP/s", "cores": ncores, "kind": kind,
           "sample": "first %d instances of the C5 batch, one qpdo_setup/qpdo_solve/qpdo_cleanup per QP, %d host thread(s)" % (sample, ncores),
           "wall_s": secs}
    return (out, summ) if (want_summaries and kind == "reference") else (out, None) if want_summaries else out



# ------------------------------------------------------------------------------------------ C1 (latency path)
def c1_problem():
    from qpdo_b200 import problems
    return problems.random_qp(1000, 500, 0.01, seed=1)


def run_c1(args, dist):
    """BASELINE configs[0]: one cold solve of the random n=1000 m=500 QP per step (qpdo_warm_start(NULL, NULL) +
    qpdo_solve on a set-up workspace); e2e = qpdo_setup + qpdo_solve + qpdo_cleanup with host buffers."""
    import qpdo_b200
    from qpdo_b200 import capi
    lib = qpdo_b200.lib()
    p = c1_problem()
    n, m = p["q"].size, p["A"].shape[0]
    s = qpdo_b200.QPDO()
    s.setup(p["Q"], p["q"], p["A"], p["l"], p["u"], dict(verbose=0))
    for _ in range(max(args.warmup, 3)):
        s.warm_start(None, None); r = s.solve()
    prof = capi.DevProfile()
    lib.qpdo_dev_set_profiling(s.device, 1)
    lib.qpdo_dev_get_profile(s.device, C.byref(prof), 1)
    s.warm_start(None, None); s.solve()
    lib.qpdo_dev_get_profile(s.device, C.byref(prof), 1)
    lib.qpdo_dev_set_profiling(s.device, 0)           # the phase timers synchronise: the timed steps run without them
    sinfo = capi.DevSymbolicInfo()
    lib.qpdo_dev_get_symbolic_info(s.device, C.byref(sinfo))
    dist.barrier()
    ts = []
    with ClockSampler(dist.local) as clk:
        for _ in range(args.steps):
            s.warm_start(None, None)
            t0 = time.perf_counter(); r = s.solve(); ts.append(time.perf_counter() - t0)
    dist.barrier()
    s.delete()
    te = []
    for _ in range(max(1, min(args.steps, 5))):
        t0 = time.perf_counter()
        s2 = qpdo_b200.QPDO(); s2.setup(p["Q"], p["q"], p["A"], p["l"], p["u"], dict(verbose=0)); r2 = s2.solve(); s2.delete()
        te.append(time.perf_counter() - t0)
    newton = max(1, prof.n_newton)
    worst = dist.reduce([float(np.mean(ts)) * 1e3], "max")
    return {"metric": "ms per solve (random sparse QP n=%d m=%d density 1 %%, BASELINE configs[0])" % (n, m), "value": worst[0], "unit": "ms/solve",
            "n_gpus": dist.world, "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": worst[0], "higher_is_better": False,
            "scaling": "replicas", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": "c1: random sparse convex QP n=1000 m=500 density 1 % through qpdo_setup/qpdo_solve; latency-bound "
                                   "(working set in L2; no L2 flush: every pass re-reads what the previous one wrote)", "n": n, "m": m, "nnzA": int(p["A"].nnz)},
            "best_ms": float(np.min(ts)) * 1e3,
            "kkt_ms_per_newton": (prof.assemble_ms + prof.factor_ms + prof.solve_ms) / newton,
            "launches_per_newton_step": prof.kernel_launches / float(newton), "newton_steps": int(prof.n_newton),
            "phases_ms_per_newton": {"assemble": prof.assemble_ms / newton, "factor": prof.factor_ms / newton, "solve": prof.solve_ms / newton,
                                     "spmv": prof.spmv_ms / newton, "linesearch": prof.linesearch_ms / newton,
                                     "residual": prof.residual_ms / max(1, prof.n_residual), "other": prof.other_ms / newton},
            "symbolic": {"nsuper": sinfo.nsuper, "nlevel": sinfo.nlevel, "max_front": sinfo.max_front, "lnz": int(sinfo.lnz), "flops": sinfo.flops},
            "result": {"status": r.status_val, "iterations": r.iterations, "oterations": r.oterations, "objective": r.objective},
            "roofline": {"bound": "latency", "kernel": "whole loop pass (launch / dependency latency; HBM and FP64 pipes idle at n=1000)",
                         "achieved": None, "peak": None, "unit": "GB/s", "frac": None, "traffic": None},
            "e2e": {"value": float(np.mean(te)) * 1e3, "unit": "ms/solve (qpdo_setup + qpdo_solve + qpdo_cleanup, host buffers)",
                    "h2d_bytes_per_step": int(12 * (p["A"].nnz + p["Q"].nnz) + 8 * (n + 2 * m)), "d2h_bytes_per_step": int(8 * 3 * (n + m))},
            "gpu_launches": int(prof.kernel_launches), "clocks": clk.summary()}


def ref_c1(args):
    import oracle
    from qpdo_b200 import capi
    ncores = os.cpu_count() or 1
    os.environ.setdefault("OPENBLAS_NUM_THREADS", str(ncores))
    p = c1_problem()
    if oracle.have_ref():
        q = capi.Solver(oracle.ref_library())
        q.setup(p["Q"], p["q"], p["A"], p["l"], p["u"], dict(verbose=0))
        ts = []
        for _ in range(max(3, min(args.steps or 3, 10))):
            q.warm_start(None, None)
            t0 = time.perf_counter(); o = q.solve(); ts.append(time.perf_counter() - t0)
        q.delete()
        kind, it = "reference", o.iterations
    else:
        t0 = time.perf_counter(); o = oracle.solve(p, trace=False); ts = [time.perf_counter() - t0]
        kind, ncores, it = "port", 1, o.iterations
    return {"value": float(np.mean(ts)) * 1e3, "unit": "ms/solve", "cores": ncores, "kind": kind, "iterations": int(it),
            "sample": "%d cold solves of the C1 QP on a set-up workspace (qpdo_warm_start(NULL, NULL) + qpdo_solve), %s" %
            (len(ts), "CHOLMOD + OpenBLAS" if kind == "reference" else "C restatement")}


# ------------------------------------------------------------------------------------------ C4 (regularisation path)
def c4_problem(args):
    from qpdo_b200 import problems
    if args.size:                                         # reduced sizes for quick checks: n = size, m = size / 2
        return problems.lasso_dual(n=args.size, m=args.size // 2)
    return problems.lasso_dual()


def c4_path(solver, p, npath):
    """qpdo_update_bounds + qpdo_warm_start + qpdo_solve per regularisation step; returns per-step
    (seconds, iterations, status)"""
    m = p["A"].shape[0]
    out, x, y = [], None, None
    for j, lam in enumerate(p["lams"][:npath]):
        if j:
            solver.update_bounds(np.full(m, -lam), np.full(m, lam))
            solver.warm_start(x, y)
        t0 = time.perf_counter(); r = solver.solve(); dt = time.perf_counter() - t0
        x, y = r.x.copy(), r.y.copy()
        out.append((dt, r.iterations, r.status_val, r.objective))
    return out


def run_c4(args, dist):
    import qpdo_b200
    from qpdo_b200 import capi
    lib = qpdo_b200.lib()
    p = c4_problem(args)
    n, m = p["q"].size, p["A"].shape[0]
    npath = len(p["lams"])
    # the path is one sequential run without warm-up steps: create the CUDA context with a throwaway tiny QP first, so that the
    # timed qpdo_setup is the setup of this problem and not the first CUDA call of the process (c3 / c1 time theirs warm too)
    from qpdo_b200 import problems as _pr
    _w = _pr.random_qp(40, 20, 0.2, seed=1); _s = qpdo_b200.QPDO(); _s.setup(_w["Q"], _w["q"], _w["A"], _w["l"], _w["u"], dict(verbose=0)); _s.solve(); _s.delete()
    s = qpdo_b200.QPDO()
    t0 = time.perf_counter(); s.setup(p["Q"], p["q"], p["A"], p["l"], p["u"], dict(verbose=0)); setup_s = time.perf_counter() - t0
    prof = capi.DevProfile()
    lib.qpdo_dev_set_profiling(s.device, 1)
    lib.qpdo_dev_get_profile(s.device, C.byref(prof), 1)
    dist.barrier()
    with ClockSampler(dist.local) as clk:
        steps = c4_path(s, p, npath)
    dist.barrier()
    lib.qpdo_dev_get_profile(s.device, C.byref(prof), 1)
    sinfo = capi.DevSymbolicInfo(); lib.qpdo_dev_get_symbolic_info(s.device, C.byref(sinfo))
    s.delete()
    newton = max(1, prof.n_newton)
    path_s = sum(t for t, _, _, _ in steps)
    worst = dist.reduce([(prof.assemble_ms + prof.factor_ms + prof.solve_ms) / newton, path_s], "max")
    return {"metric": "KKT factor+solve ms per Newton iteration (Lasso-as-QP n=%d m=%d, %d-step regularisation path with warm starts)" % (n, m, npath),
            "value": worst[0], "unit": "ms/iter", "n_gpus": dist.world, "steps": 1, "warmup": 0, "ms_per_step": worst[1] * 1e3,
            "higher_is_better": False, "scaling": "replicas", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": "c4: Lasso dual QP, qpdo_update_bounds + qpdo_warm_start + qpdo_solve per path step; a step = the whole "
                                   "%d-step path (the path is sequential: one run, no warm-up; A and L exceed L2)" % npath, "n": n, "m": m, "nnzA": int(p["A"].nnz)},
            "path_seconds": worst[1], "setup_s": setup_s, "path_iterations": [it for _, it, _, _ in steps], "path_status": [st for _, _, st, _ in steps],
            "newton_steps": int(prof.n_newton),
            "host_analysis_seconds": float(sinfo.analyze_seconds),   # symbolic (re-)analyses on the support-following pattern, all of setup + path
            "phases_ms_per_newton": {"assemble": prof.assemble_ms / newton, "factor": prof.factor_ms / newton, "solve": prof.solve_ms / newton,
                                     "spmv": prof.spmv_ms / newton, "linesearch": prof.linesearch_ms / newton},
            "roofline": {"bound": "hbm", "kernel": "k_spmv_rows (A x, A' y; the active-set factor is tiny on this path)",
                         "achieved": (3 * (12.0 * p["A"].nnz + 4 * (n + 1) + 8 * (n + m)) + 28.0 * n) / max(1e-9, prof.spmv_ms / newton * 1e-3) / 1e9,
                         "peak": peaks()[0]["hbm_gbs"], "unit": "GB/s", "traffic": None},
            "e2e": {"value": (setup_s + worst[1]) * 1e3 / max(1, sum(it for _, it, _, _ in steps)), "unit": "ms per loop pass (qpdo_setup + the whole path, wall clock)",
                    "h2d_bytes_per_step": int(12 * p["A"].nnz + 20 * n + 16 * m * npath + 8 * (n + m) * (npath - 1)),
                    "d2h_bytes_per_step": int(8 * 3 * (n + m) * npath), "setup_plus_path_s": setup_s + worst[1]},
            "gpu_launches": int(prof.kernel_launches), "clocks": clk.summary()}


def ref_c4(args):
    import oracle
    from qpdo_b200 import capi
    ncores = os.cpu_count() or 1
    os.environ.setdefault("OPENBLAS_NUM_THREADS", str(ncores))
    if not oracle.have_ref():
        return {"value": None, "error": "oracle/_ref not built: the C restatement has no sparse path for n = 200k"}
    p = c4_problem(args)
    npath = min(len(p["lams"]), args.ref_path_steps or len(p["lams"]))
    q = capi.Solver(oracle.ref_library())
    t0 = time.perf_counter(); q.setup(p["Q"], p["q"], p["A"], p["l"], p["u"], dict(verbose=0)); setup_s = time.perf_counter() - t0
    steps = c4_path(q, p, npath)
    q.delete()
    path_s = sum(t for t, _, _, _ in steps)
    iters = sum(it for _, it, _, _ in steps)
    return {"value": path_s * 1e3 / max(1, iters), "unit": "ms per loop pass (whole pass incl. factor/update + solve)", "cores": ncores, "kind": "reference",
            "path_seconds": path_s, "setup_s": setup_s, "path_iterations": [it for _, it, _, _ in steps],
            "sample": "the first %d of %d path steps of C4, CHOLMOD supernodal + OpenBLAS" % (npath, len(p["lams"]))}


# ------------------------------------------------------------------------------------------ main
def kkt_record(args, dist):
    """the other half of BASELINE's metric in the default line: the c3 record (KKT assemble+factor+solve ms per Newton
    step on the CONT-300-shaped QP) with the unmodified reference timed on the same box"""
    sub = argparse.Namespace(**vars(args))
    sub.steps, sub.warmup, sub.size, sub.max_iter = 2, 1, 0, 0
    rec = run_c3(sub, dist)
    keep = ("metric", "value", "unit", "ms_per_step", "phases_ms_per_newton", "symbolic", "result", "factor_gflops", "roofline",
            "roofline_factor", "e2e", "gpu_launches", "config")
    out = dict((k, rec[k]) for k in keep if k in rec)
    if not args.no_cpu_baseline:
        try:
            out["cpu_baseline"] = ref_c3(sub)
        except Exception as e:
            out["cpu_baseline"] = {"value": None, "error": repr(e)}
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=None)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default=None, choices=["c1", "c3", "c4", "c5"])
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--size", type=int, default=0, help="c3: grid N (default 300); c5: batch size (default 65536); c4: n (default 200000)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--natural-order", action="store_true", help="c5: pull the work queue in instance order (no adaptive order)")
    ap.add_argument("--no-kkt", action="store_true", help="c5 at one GPU: leave the c3 record (`kkt`) out of the line")
    ap.add_argument("--ref-path-steps", type=int, default=0, help="c4 reference leg: only the first k path steps")
    ap.add_argument("--max-iter", type=int, default=0, help="c3: cap the loop passes per solve (profiling runs)")
    args = ap.parse_args()
    if args.workload is None:
        args.workload = "c5"
    if args.steps is None:
        args.steps = {"c5": 5, "c3": 3, "c1": 10, "c4": 1}[args.workload]
    args.warmup = max(args.warmup, 0)
    refs = {"c1": ref_c1, "c3": ref_c3, "c4": ref_c4}

    if args.impl == "reference":
        # rank 0 alone runs the CPU reference; the other ranks exit 0 without work
        if int(os.environ.get("RANK", "0")) != 0:
            return 0
        if args.workload == "c5":
            b = ref_c5(args)
            line = {"impl": "reference", "metric": qbatch.METRIC, "value": b["value"], "unit": "QP/s", "higher_is_better": True}
            from qpdo_b200 import problems
            pn, pm, prow = problems.mpc_pattern()[:3]
            cfg = c5_config(args.size or C5_BATCH, pn, pm, int(prow.size), max(1, args.gpus))
        else:
            if args.workload == "c3":
                vals = [ref_c3(args) for _ in range(max(1, min(args.steps, 2)))]
                b = min(vals, key=lambda v: v["value"])
                metric = "KKT factor+solve ms per Newton iteration (CONT-%d-shaped QP)" % (args.size or C3_N)
            else:
                b = refs[args.workload](args)
                metric = {"c1": "ms per solve (random sparse QP n=1000 m=500 density 1 %, BASELINE configs[0])",
                          "c4": "KKT factor+solve ms per Newton iteration (Lasso-as-QP regularisation path)"}[args.workload]
            line = {"impl": "reference", "metric": metric, "value": b["value"], "unit": b["unit"].split(" ")[0], "higher_is_better": False}
            cfg = {"workload": "%s: bounded sample on the host cores (see cpu_baseline.sample)" % args.workload}
        line.update({"n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "scaling": "strong" if args.workload == "c5" else "replicas",
                     "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": cfg,
                     "cpu_baseline": b, "e2e": {"value": b["value"], "unit": line["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}})
        print(json.dumps(line))
        return 0

    dist = Dist(args.gpus)
    try:
        gpu_summaries = None
        if args.workload == "c5":
            out, gpu_summaries = run_c5(args, dist)
        else:
            out = {"c1": run_c1, "c3": run_c3, "c4": run_c4}[args.workload](args, dist)
        if dist.rank == 0:
            if dist.world == 1 and not args.no_cpu_baseline:
                try:
                    if args.workload == "c5":
                        out["cpu_baseline"], ref_summaries = ref_c5(args, bounded=True, want_summaries=True)
                        if ref_summaries is not None:
                            out["parity"] = parity_record(gpu_summaries, ref_summaries)
                    else:
                        out["cpu_baseline"] = refs[args.workload](args)
                except Exception as e:          # the baseline is a reported number, never a reason to fail the bench
                    out["cpu_baseline"] = {"value": None, "error": repr(e)}
            if dist.world == 1 and args.workload == "c5" and not args.no_kkt and not args.size:
                try:
                    out["kkt"] = kkt_record(args, dist)
                except Exception as e:
                    out["kkt"] = {"value": None, "error": repr(e)}
            print(json.dumps(out))
    finally:
        dist.close()
    return 0


if __name__ == "__main__":
    sys.exit(main())
