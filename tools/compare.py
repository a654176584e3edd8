"""Run the same QP through the unmodified reference (oracle/_ref) and the CUDA library; print both."""
import sys, os, time, json
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import qpdo_b200
from qpdo_b200 import capi, problems

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def ref_lib(nosuper=False):
    return capi.load_library(os.path.join(ROOT, "oracle", "_ref", "libqpdo_ref_nosuper.so" if nosuper else "libqpdo_ref.so"))


def run(solver, p, settings=None, warm=None):
    s = dict(verbose=0); s.update(settings or {})
    t0 = time.time()
    solver.setup(p["Q"], p["q"], p["A"], p["l"], p["u"], s)
    t1 = time.time()
    if warm is not None:
        solver.warm_start(*warm)
    r = solver.solve()
    t2 = time.time()
    r.wall_setup, r.wall_solve = t1 - t0, t2 - t1
    return r


def compare(name, p, settings=None, ref=None):
    ref = ref or ref_lib()
    a = run(capi.Solver(ref), p, settings)
    g = qpdo_b200.QPDO()
    b = run(g, p, settings)
    rel = lambda u, v: float(np.abs(u - v).max() / max(1e-300, np.abs(u).max())) if u.size else 0.0
    out = dict(name=name, ref=repr(a), gpu=repr(b), same_status=a.status_val == b.status_val,
               same_iter=(a.iterations, a.oterations) == (b.iterations, b.oterations),
               same_active=bool(np.array_equal(a.active_set, b.active_set)),
               rel_x=rel(a.x, b.x), rel_y=rel(a.y, b.y), rel_obj=abs(a.objective - b.objective) / max(1.0, abs(a.objective)),
               ref_solve_s=a.solve_time, gpu_solve_s=b.solve_time, ref_setup_s=a.wall_setup, gpu_setup_s=b.wall_setup)
    print(json.dumps(out))
    sys.stdout.flush()
    return out, g


if __name__ == "__main__":
    which = sys.argv[1:] or ["c2", "small", "c1", "cont20", "cont60"]
    for w in which:
        if w == "c2":
            for nm, p, st in problems.infeasibility_trio():
                compare(nm, p, dict(max_iter=100))
        elif w == "small":
            compare("rand200", problems.random_qp(200, 100, 0.05, seed=3))
        elif w == "c1":
            compare("c1", problems.random_qp())
        elif w.startswith("cont"):
            compare(w, problems.cont_like(int(w[4:])))
        elif w == "mpc":
            b = problems.mpc_batch(4)
            for i in range(4):
                compare("mpc%d" % i, problems.mpc_instance(b, i))
