"""Generate tests/golden/*.npz from the UNMODIFIED reference (oracle/_ref/libqpdo_ref.so, built from
/root/reference by oracle/Makefile).  Run in the build container only (the GPU box has no
/root/reference):   python tests/golden/make_golden.py

Each fixture stores the reference's outputs for a seeded problem of qpdo_b200.problems (inputs are
regenerated at test time; a checksum of the inputs is stored to catch generator drift).
"""
import os, sys, hashlib
import numpy as np
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import oracle
from qpdo_b200 import capi, problems


def checksum(p):
    h = hashlib.sha256()
    import scipy.sparse as sp
    for M in (sp.tril(sp.csc_matrix(p["Q"]), format="csc"), sp.csc_matrix(p["A"])):
        M.sort_indices()
        h.update(M.indptr.astype(np.int64).tobytes()); h.update(M.indices.astype(np.int64).tobytes())
        h.update(np.ascontiguousarray(M.data, dtype=np.float64).tobytes())
    for k in ("q", "l", "u"):
        h.update(np.ascontiguousarray(p[k], dtype=np.float64).tobytes())
    return h.hexdigest()


CASES = {
    # name: (generator, settings)
    "degenerate": (lambda: problems.infeasibility_trio()[0][1], dict(max_iter=100)),
    "primal_infeasible": (lambda: problems.infeasibility_trio()[1][1], dict(max_iter=100)),
    "dual_infeasible": (lambda: problems.infeasibility_trio()[2][1], dict(max_iter=100)),
    "degenerate_noscale": (lambda: problems.infeasibility_trio()[0][1], dict(max_iter=100, scaling=0)),
    "primal_infeasible_x10": (lambda: problems.replicate_block_diag(problems.infeasibility_trio()[1][1], 10), dict(max_iter=100)),
    "degenerate_x10": (lambda: problems.replicate_block_diag(problems.infeasibility_trio()[0][1], 10), dict(max_iter=100)),
    "rand200": (lambda: problems.random_qp(200, 100, 0.05, seed=3), {}),
    "rand300": (lambda: problems.random_qp(300, 150, 0.03, seed=2), {}),
    "rand300_noprox": (lambda: problems.random_qp(300, 150, 0.03, seed=2), dict(proximal=0)),
    "rand300_noscale": (lambda: problems.random_qp(300, 150, 0.03, seed=2), dict(scaling=0)),
    "c1": (lambda: problems.random_qp(), {}),
    "cont20": (lambda: problems.cont_like(20), {}),
    "cont40": (lambda: problems.cont_like(40), {}),
    "mpc0": (lambda: problems.mpc_instance(problems.mpc_batch(4), 0), {}),
    "mpc3": (lambda: problems.mpc_instance(problems.mpc_batch(4), 3), {}),
    "lasso_small": (lambda: problems.lasso_dual(2000, 1000, window=100, npath=4), {}),
}


def record(r):
    return dict(status_val=r.status_val, iterations=r.iterations, oterations=r.oterations, objective=r.objective,
                res_prim_norm=r.res_prim_norm, res_dual_norm=r.res_dual_norm, x=r.x, y=r.y, active_set=r.active_set,
                prim_inf_cert=r.prim_inf_cert, dual_inf_cert=r.dual_inf_cert)


def main():
    ref = oracle.ref_library()
    for name, (gen, st) in CASES.items():
        p = gen()
        s = capi.Solver(ref)
        s.setup(p["Q"], p["q"], p["A"], p["l"], p["u"], dict(dict(verbose=0), **st))
        out = dict(("cold_" + k, v) for k, v in record(s.solve()).items())
        out["checksum"] = checksum(p)
        # parametric re-solves through the same workspace: warm start, update_bounds, update_q
        r0x, r0y = out["cold_x"], out["cold_y"]
        if out["cold_status_val"] == 1:
            s.warm_start(r0x * (1 + 1e-3), r0y * (1 - 1e-3))          # demo_mex.m:45-53 style perturbation
            out.update(("warm_" + k, v) for k, v in record(s.solve()).items())
            if "lams" in p:
                xs, ys = r0x, r0y
                for j, lam in enumerate(p["lams"][1:], 1):
                    m = p["A"].shape[0]
                    s.update_bounds(np.full(m, -lam), np.full(m, lam))
                    s.warm_start(xs, ys)
                    rj = s.solve()
                    out.update(("path%d_" % j + k, v) for k, v in record(rj).items())
                    xs, ys = rj.x, rj.y
            else:
                s.update_bounds(p["l"] - 0.1, p["u"] + 0.05)
                s.warm_start(r0x, r0y)
                out.update(("bounds_" + k, v) for k, v in record(s.solve()).items())
                s.update_q(p["q"] * 1.5 + 0.01)
                out.update(("newq_" + k, v) for k, v in record(s.solve()).items())
        s.delete()
        np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
        print(name, out["cold_status_val"], out["cold_iterations"], out["cold_oterations"],
              [(k, int(out[k])) for k in out if k.endswith("_iterations") and not k.startswith("cold")])


if __name__ == "__main__":
    main()
