#!/usr/bin/env python
"""tools/knife_edge.py -- where and why the unmodified reference and a refactoring solver part ways.

For every case the parity tests treat as a knife edge this script
  1. steps the UNMODIFIED reference (oracle/_ref) and the restatement pass by pass (the loop is
     deterministic, so `max_iter = k` exposes the state after pass k-1: src/qpdo.c:343-449),
  2. finds the first pass whose control decision differs, the comparison that decided it
     (check_inner_optimality src/termination.c:22-30, is_dual_infeasible :197-209), both operands
     and the margin, in units of the rounding noise that produced the operand,
  3. perturbs ONE entry of q by ONE ulp, 100 times, and counts the reference's own iteration totals:
     a decision that flips under a 1-ulp change of the data is not pinned by the reference itself.

Writes profiles/r2_knife_edge.json (read by tests/test_knife_edge.py).  Runs in the build container
only (needs oracle/_ref); nothing on the GPU box imports it.
"""
import collections
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import oracle                                   # noqa: E402
from qpdo_b200 import capi, problems            # noqa: E402

C5_KNIFE = (8364, 11054, 64095)                 # the three C5 instances (all status -3)


def ref_pass(lib, p, st, k):
    """state of the reference after k loop passes"""
    s = capi.Solver(lib)
    s.setup(p["Q"], p["q"], p["A"], p["l"], p["u"], dict(dict(verbose=0), **dict(st, max_iter=k)))
    r = s.solve()
    w = s.work
    g = lambda ptr, sz: np.array(np.ctypeslib.as_array(ptr, shape=(sz,)))
    out = dict(status=r.status_val, iterations=r.iterations, oterations=r.oterations, tau=float(w.tau), sigma=float(w.sigma),
               eps_in=float(w.eps_in), rp=r.res_prim_norm, rd=r.res_dual_norm, rpi=r.res_prim_in_norm, rdi=r.res_dual_in_norm,
               ymax=float(np.abs(g(w.y, s.m)).max()), dxmax=float(np.abs(g(w.dx, s.n)).max()))
    s.delete()
    return out


def ref_counts(lib, p, st):
    s = capi.Solver(lib)
    s.setup(p["Q"], p["q"], p["A"], p["l"], p["u"], dict(dict(verbose=0), **st))
    r = s.solve()
    s.delete()
    return (r.status_val, r.iterations, r.oterations)


def perturbation_study(lib, p, st, trials=100):
    """reference and restatement totals under 1-ulp changes of single q entries"""
    n = p["q"].size
    ref, ours = collections.Counter(), collections.Counter()
    for t in range(trials):
        q = p["q"].copy()
        i = t % n
        q[i] = np.nextafter(q[i], np.inf)
        ref["%d/%d/%d" % ref_counts(lib, dict(p, q=q), st)] += 1
        o = oracle.solve(dict(p, q=q), st, trace=False)
        ours["%d/%d/%d" % (o.status_val, o.iterations, o.oterations)] += 1
    return dict(ref), dict(ours)


def first_divergence(lib, p, st):
    r = ref_counts(lib, p, st)
    o = oracle.solve(p, st, trace=False)
    K = max(r[1], o.iterations) + 1
    rows = []
    for k in range(1, K + 1):
        a = ref_pass(lib, p, st, k)
        b = oracle.solve(p, dict(st, max_iter=k))
        rows.append(dict(k=k, ref=a, restatement=dict(status=b.status_val, iterations=b.iterations, oterations=b.oterations,
                                                      rp=b.res_prim_norm, rd=b.res_dual_norm, rpi=b.res_prim_in_norm,
                                                      rdi=b.res_dual_in_norm, tau=b.trace[-1]["tau"] if b.trace else None)))
        if (a["status"], a["iterations"], a["oterations"]) != (b.status_val, b.iterations, b.oterations):
            break
    return r, (o.status_val, o.iterations, o.oterations), rows


def ulp(x):
    return float(np.spacing(abs(x)))


def c5_case(lib, inst):
    p = problems.mpc_instance(problems.mpc_batch(1, first=inst), 0)
    r, o, rows = first_divergence(lib, p, {})
    last = rows[-1]
    a, b = last["ref"], last["restatement"]
    # the pass before the split: both have just taken a Newton step; the reference's inner dual residual is
    # above eps_in (another Newton step), ours is below (subproblem solved -> infeasibility certificate)
    dec = dict(comparison="res_dual_in_norm <= eps_in  (check_inner_optimality, src/termination.c:22-30)",
               eps_in=a["eps_in"], ref_operand=a["rdi"], restatement_operand=b["rdi"],
               y_inf_norm=a["ymax"], ulp_of_y=ulp(a["ymax"]),
               ref_margin=a["rdi"] - a["eps_in"], ref_margin_in_ulp_of_y=(a["rdi"] - a["eps_in"]) / ulp(a["ymax"]),
               operand_in_ulp_of_y=a["rdi"] / ulp(a["ymax"]),
               why="res_dual_in = Qx + q + A'y is a sum of terms of size |y| ~ %.3g that cancels to ~1e-7: the operand is "
                   "%.0f ulp(|y|) of rounding residue of the Newton solve, compared with a threshold %.1f ulp(|y|) away" %
                   (a["ymax"], a["rdi"] / ulp(a["ymax"]), abs(a["rdi"] - a["eps_in"]) / ulp(a["ymax"])))
    pr, po = perturbation_study(lib, p, {})
    return dict(case="c5 instance %d" % inst, reference=r, restatement=o, first_differing_pass=last["k"] - 1, decision=dec,
                one_ulp_perturbations_of_q=dict(trials=100, reference_totals=pr, restatement_totals=po), trace_tail=rows[-4:])


def dual_infeasible_case(lib):
    p = dict((n, q) for n, q, _ in problems.infeasibility_trio())["dual_infeasible"]
    st = dict(max_iter=100)
    r, o, rows = first_divergence(lib, p, st)
    # the deciding test: ||Q dx - sigma tau dx|| <= eps_dual_inf ||dx||  with Q dx = 0 in the unbounded direction,
    # i.e. sigma * tau <= eps_dual_inf with sigma == eps_dual_inf == 1e-6  <=>  tau <= 1
    taus = [(row["k"], row["ref"]["tau"], row["restatement"]["tau"], row["ref"]["sigma"]) for row in rows]
    mu, sig0 = 0.1, 1e-3
    d_ref = (sig0 + 1.0 / mu) - 1.0 / mu             # K22 after "row 3 enters, row 3 leaves" by rank-1 update + downdate
    dec = dict(comparison="||Qdx - sigma tau dx||_inf <= eps_dual_inf ||dx||_inf (is_dual_infeasible, src/termination.c:197-209): "
                          "with Q dx = 0 along x2 this reads sigma*tau <= eps_dual_inf, and sigma == eps_dual_inf == 1e-6 "
                          "at outer iteration 4, i.e. tau <= 1",
               ref_tau=[t[1] for t in taus if t[3] <= 1.0000000000000002e-06 and t[3] > 5e-7][:1],
               restatement_tau=[t[2] for t in taus if t[3] <= 1.0000000000000002e-06 and t[3] > 5e-7][:1],
               K22_by_downdate="fl(fl(sigma + 1/mu) - 1/mu) = %.17g" % d_ref, K22_by_refactor="sigma = %.17g" % sig0,
               relative_difference=(d_ref - sig0) / sig0, ulp_of_pre_downdate_pivot=ulp(sig0 + 1 / mu),
               why="K = Q + sigma I + A' diag(omega) A is diagonal here. Row 3 (x2 >= 1) is active at pass 0 (K22 = sigma + 1/mu = 10.001) "
                   "and leaves at pass 1: the reference DOWNDATES the pivot 10.001 by 10 (ldldowndate_leave_constraints, "
                   "src/cholmod_interface.c:66-72) and is left with sigma +- one ulp of 10.001 (%.3g, a relative ~1e-12 of sigma); a "
                   "refactor has K22 = sigma exactly. dx2 = rhs2/K22 is off by that 1e-12, the exact linesearch answers with "
                   "tau = 1 - 9.4e-13 instead of 1 (pass 1 in `taus`), which leaves y3 = 1.0e-12 instead of ~0. Row 3 is never active "
                   "again, so that y3 stays in res_dual_in for the rest of the solve and every later Newton step has tau = 1 + 1.0e-12 "
                   "(later refactors, triggered by update_sigma, do not remove it: the residue is in the iterate, not in the factor). At "
                   "outer iteration 4 sigma == eps_dual_inf, the test reads tau <= 1, the reference fails it by 4.5e3 ulp of tau = 1e-12 "
                   "and needs one more outer iteration (sigma = 1e-7); a refactoring solver has tau = 1 - 1 ulp and passes." %
                   ulp(sig0 + 1 / mu))
    return dict(case="dual_infeasible (examples/infeasibility_tests.m:60-75)", reference=r, restatement=o,
                first_differing_pass=rows[-1]["k"] - 1, decision=dec, taus=taus, trace_tail=rows[-3:])


def cont40_bounds_case(lib, trials=10):
    """CONT-40 with its equality rows relaxed into ranges, warm-started from the cold solution: a degenerate
    ~440-pass solve.  The reference's own totals under 1-ulp changes of q."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from conftest import golden_cases, load_golden
    gen, st = golden_cases()[0]["cont40"]
    p = gen()
    g = load_golden("cont40")
    pb = dict(p, l=p["l"] - 0.1, u=p["u"] + 0.05)
    tot = []
    for i in range(trials):
        q = pb["q"].copy()
        q[i] = np.nextafter(q[i], np.inf)
        s = capi.Solver(lib)
        s.setup(pb["Q"], q, pb["A"], pb["l"], pb["u"], dict(dict(verbose=0), **st))
        s.warm_start(g["cold_x"], g["cold_y"])
        r = s.solve()
        s.delete()
        tot.append("%d/%d/%d" % (r.status_val, r.iterations, r.oterations))
    o = oracle.solve(pb, st, x0=g["cold_x"], y0=g["cold_y"], trace=False)
    return dict(case="cont40 bounds sequence (tests/golden/cont40.npz bounds_*)",
                reference=(int(g["bounds_status_val"]), int(g["bounds_iterations"]), int(g["bounds_oterations"])),
                restatement=(o.status_val, o.iterations, o.oterations),
                one_ulp_perturbations_of_q=dict(trials=trials, reference_totals=tot),
                decision=dict(why="a degenerate several-hundred-pass solve: the reference gives %d different iteration totals in %d "
                                  "runs that differ by one ulp in one entry of q" % (len(set(tot)), trials)))


def main():
    if not oracle.have_ref():
        raise SystemExit("needs oracle/_ref (build container only)")
    lib = oracle.ref_library()
    out = dict(generated_by="tools/knife_edge.py", cases=[dual_infeasible_case(lib)] + [c5_case(lib, i) for i in C5_KNIFE] + [cont40_bounds_case(lib)])
    # a control: an ordinary instance is stable under the same perturbations
    p = problems.mpc_instance(problems.mpc_batch(1, first=100), 0)
    pr, po = perturbation_study(lib, p, {})
    out["control"] = dict(case="c5 instance 100", one_ulp_perturbations_of_q=dict(trials=100, reference_totals=pr, restatement_totals=po))
    path = os.path.join(ROOT, "profiles", "r2_knife_edge.json")
    json.dump(out, open(path, "w"), indent=1)
    for c in out["cases"]:
        print(c["case"], "ref", c["reference"], "restatement", c["restatement"], "first differing pass", c.get("first_differing_pass"))
        print("   ", c["decision"]["why"])
        if "one_ulp_perturbations_of_q" in c:
            print("    reference under 1-ulp perturbations:", c["one_ulp_perturbations_of_q"]["reference_totals"])
    print("control:", out["control"])


if __name__ == "__main__":
    main()
