#!/usr/bin/env python
"""tests/golden/make_golden_rescale.py -- golden vectors of qpdo_update_settings with MORE scaling sweeps
(reference src/qpdo.c:496-512), produced by the UNMODIFIED reference (oracle/_ref/libqpdo_ref.so) in the build
container: setup with scaling = 2, solve, update_settings(scaling = 10), warm start from the first solution,
solve again.  Writes tests/golden/rescale_seq.npz."""
import os
import sys
import numpy as np
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import oracle
from qpdo_b200 import capi, problems


def sequence(lib, p):
    s = capi.Solver(lib)
    s.setup(p["Q"], p["q"], p["A"], p["l"], p["u"], dict(verbose=0, scaling=2))
    a = s.solve()
    s.update_settings(dict(verbose=0, scaling=10))
    w = s.work
    D, E = (np.ctypeslib.as_array(w.scaling.contents.D, shape=(s.n,)).copy(), np.ctypeslib.as_array(w.scaling.contents.E, shape=(s.m,)).copy())
    c = float(w.scaling.contents.c)
    s.warm_start(a.x, a.y)
    b = s.solve()
    s.delete()
    return a, b, D, E, c


def problem():
    return problems.random_qp(120, 80, 0.05, seed=11)


if __name__ == "__main__":
    a, b, D, E, c = sequence(oracle.ref_library(), problem())
    np.savez_compressed(os.path.join(HERE, "rescale_seq.npz"),
                        x1=a.x, y1=a.y, info1=np.array([a.status_val, a.iterations, a.oterations]), obj1=a.objective,
                        x2=b.x, y2=b.y, info2=np.array([b.status_val, b.iterations, b.oterations]), obj2=b.objective,
                        active2=np.asarray(b.active_set), D=D, E=E, c=c)
    print("first solve", a.status_val, a.iterations, a.oterations, a.objective)
    print("after rescale", b.status_val, b.iterations, b.oterations, b.objective, "c", c)
