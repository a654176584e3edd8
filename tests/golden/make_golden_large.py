"""Full-size golden fixtures from the UNMODIFIED reference (minutes of CPU each; build container only).
   python tests/golden/make_golden_large.py cont300 | lasso
Stores status / iteration counts / objective / active-set digest, and x, y rounded-trip exact (float64)."""
import os, sys, time, hashlib
import numpy as np
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import oracle
from qpdo_b200 import capi, problems


def run(name, p, st, path=False, stride=1):
    ref = oracle.ref_library()
    s = capi.Solver(ref)
    t0 = time.time()
    s.setup(p["Q"], p["q"], p["A"], p["l"], p["u"], dict(dict(verbose=0), **st))
    r = s.solve()
    out = dict(status_val=r.status_val, iterations=r.iterations, oterations=r.oterations, objective=r.objective,
               res_prim_norm=r.res_prim_norm, res_dual_norm=r.res_dual_norm, x=r.x, y=r.y,
               active_set=np.packbits(r.active_set.astype(np.uint8)), n_active=int(r.active_set.sum()),
               ref_solve_seconds=r.solve_time, ref_setup_seconds=r.setup_time)
    print(name, r, "n_active", out["n_active"], "solve %.2fs" % r.solve_time, flush=True)
    if path:
        m = p["A"].shape[0]
        xs, ys = r.x, r.y
        for j, lam in enumerate(p["lams"][1:], 1):
            s.update_bounds(np.full(m, -lam), np.full(m, lam))
            s.warm_start(xs, ys)
            rj = s.solve()
            out.update({"path%d_status_val" % j: rj.status_val, "path%d_iterations" % j: rj.iterations,
                        "path%d_oterations" % j: rj.oterations, "path%d_objective" % j: rj.objective,
                        "path%d_n_active" % j: int(rj.active_set.sum()), "path%d_solve_seconds" % j: rj.solve_time,
                        "path%d_active_set" % j: np.packbits(rj.active_set.astype(np.uint8))})
            if j == len(p["lams"]) - 1:
                out["pathlast_x"] = rj.x; out["pathlast_y"] = rj.y
            print(name, "path", j, rj, "n_active", int(rj.active_set.sum()), "solve %.2fs" % rj.solve_time, flush=True)
            xs, ys = rj.x, rj.y
    s.delete()
    out["total_seconds"] = time.time() - t0
    if stride > 1:                      # keep the fixture small: every stride-th entry of the solution vectors
        for k in ("x", "y", "pathlast_x", "pathlast_y"):
            if k in out:
                out[k] = out[k][::stride]
        out["stride"] = stride
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)


if __name__ == "__main__":
    which = sys.argv[1]
    if which == "cont300":
        run("cont300", problems.cont_like(300), {})
    elif which == "cont120":
        run("cont120", problems.cont_like(120), {})
    elif which == "lasso":
        run("lasso_c4", problems.lasso_dual(), {}, path=True, stride=16)
    elif which == "lasso_mid":
        run("lasso_mid", problems.lasso_dual(20000, 10000, window=200, npath=6), {}, path=True)
