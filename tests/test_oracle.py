"""CPU tests: the C restatement (oracle/qpdo_oracle.c) is pinned against the golden vectors made by
the unmodified reference (tests/golden/make_golden.py) and, when present, against oracle/_ref."""
import numpy as np
import pytest
import scipy.sparse as sp
from conftest import rel_err, load_golden, golden_cases
from qpdo_b200 import problems

CASES, checksum = golden_cases()
TOL = 1e-8          # north_star: x, y, objective within 1e-8 relative

# The reference's dual-infeasibility test sits on a knife edge for this fixture: at outer iteration 4 it
# reads sigma*tau <= eps_dual_inf with sigma == eps_dual_inf, i.e. tau <= 1, and the reference's tau is
# 1 + 1.0e-12 -- the residue of ONE rank-1 downdate (10.001 - 10) at pass 1 (trace and arithmetic:
# profiles/r2_knife_edge.json, made by tools/knife_edge.py).  A refactoring solver has tau = 1 - 1 ulp,
# passes that test and stops exactly one outer iteration (two loop passes) earlier: asserted exactly.
KNIFE_EDGE = {"dual_infeasible": (-2, -1)}      # (iterations, oterations) relative to the reference
SLOW = {"c1"}
# CONT-40 with its equality rows relaxed into 0.15-wide ranges is a degenerate ~440-pass solve on which the
# UNMODIFIED reference gives 10 different iteration totals (408..445) in 10 runs that differ by one ulp in
# one entry of q (profiles/r2_knife_edge.json): status + objective only.  CONT-20 is stable and exact.
CHAOTIC_SEQ = {"cont40"}


TOL_SEQ = 1e-6      # long parametric re-solves of (near-)degenerate problems: x is only determined to eps_abs


def check(tag, g, r, prefix, knife=None, TOL=TOL):
    assert r.status_val == int(g[prefix + "status_val"]), tag
    if knife is None:
        assert r.iterations == int(g[prefix + "iterations"]), tag
        assert r.oterations == int(g[prefix + "oterations"]), tag
        assert np.array_equal(r.active_set, g[prefix + "active_set"]), tag
        if r.status_val == 1:
            assert rel_err(g[prefix + "x"], r.x) <= TOL, tag
            assert rel_err(g[prefix + "y"], r.y) <= TOL, tag
            assert abs(r.objective - float(g[prefix + "objective"])) <= TOL * max(1.0, abs(float(g[prefix + "objective"]))), tag
        if r.status_val == -3:
            assert rel_err(g[prefix + "prim_inf_cert"], r.prim_inf_cert) <= 1e-6, tag
    else:
        assert r.iterations == int(g[prefix + "iterations"]) + knife[0], tag
        assert r.oterations == int(g[prefix + "oterations"]) + knife[1], tag
        assert np.array_equal(r.active_set, g[prefix + "active_set"]), tag
        if r.status_val == -4:
            a = g[prefix + "dual_inf_cert"]; b = r.dual_inf_cert
            assert rel_err(a / np.abs(a).max(), b / np.abs(b).max()) <= 1e-6, tag


@pytest.mark.parametrize("name", [n for n in CASES if n not in SLOW])
def test_restatement_matches_reference_golden(oracle_mod, name):
    gen, st = CASES[name]
    p = gen()
    g = load_golden(name)
    assert checksum(p) == str(g["checksum"]), "problem generator drifted from the golden fixture"
    r = oracle_mod.solve(p, st)
    check(name + "/cold", g, r, "cold_", knife=KNIFE_EDGE.get(name))
    if "warm_status_val" in g.files:
        rw = oracle_mod.solve(p, st, x0=g["cold_x"] * (1 + 1e-3), y0=g["cold_y"] * (1 - 1e-3))
        check(name + "/warm", g, rw, "warm_")
    if "lams" in p:                       # C4-style regularisation path: new bounds + warm start
        xs, ys = g["cold_x"], g["cold_y"]
        m = p["A"].shape[0]
        for j, lam in enumerate(p["lams"][1:], 1):
            pj = dict(p, l=np.full(m, -lam), u=np.full(m, lam))
            rj = oracle_mod.solve(pj, st, x0=xs, y0=ys)
            check("%s/path%d" % (name, j), g, rj, "path%d_" % j)
            xs, ys = g["path%d_x" % j], g["path%d_y" % j]
    elif "bounds_status_val" in g.files:
        pb = dict(p, l=p["l"] - 0.1, u=p["u"] + 0.05)
        rb = oracle_mod.solve(pb, st, x0=g["cold_x"], y0=g["cold_y"])
        if name in CHAOTIC_SEQ:
            assert rb.status_val == int(g["bounds_status_val"])
            assert abs(rb.objective - float(g["bounds_objective"])) <= 1e-6 * max(1.0, abs(float(g["bounds_objective"])))
        else:
            check(name + "/bounds", g, rb, "bounds_", TOL=TOL_SEQ)


def test_known_answer_statuses(oracle_mod):
    """examples/infeasibility_tests.m:30,48,75: status 1 / -3 / -4"""
    for name, p, expected in problems.infeasibility_trio():
        assert oracle_mod.solve(p, dict(max_iter=100)).status_val == expected, name


def test_restatement_against_live_reference(oracle_mod):
    if not oracle_mod.have_ref():
        pytest.skip("oracle/_ref not built (no /root/reference on this machine)")
    from qpdo_b200 import capi
    ref = oracle_mod.ref_library()
    for seed in (11, 12, 13):
        p = problems.random_qp(120, 80, 0.06, seed=seed)
        s = capi.Solver(ref)
        s.setup(p["Q"], p["q"], p["A"], p["l"], p["u"], dict(verbose=0))
        a = s.solve(); s.delete()
        b = oracle_mod.solve(p)
        assert (a.status_val, a.iterations, a.oterations) == (b.status_val, b.iterations, b.oterations)
        assert np.array_equal(a.active_set, b.active_set)
        assert rel_err(a.x, b.x) <= TOL and rel_err(a.y, b.y) <= TOL


def test_spmv_order_and_values(oracle_mod):
    import ctypes as C
    L = oracle_mod.lib()
    rng = np.random.default_rng(0)
    A = sp.random(40, 30, 0.2, random_state=1, format="csc"); A.sort_indices()
    Ap, Ai, Ax = A.indptr.astype(np.int64), A.indices.astype(np.int64), A.data.copy()
    pl, pd = C.POINTER(C.c_long), C.POINTER(C.c_double)
    x = rng.standard_normal(30); y = np.zeros(40)
    L.qpdo_oracle_spmv(0, 40, 30, Ap.ctypes.data_as(pl), Ai.ctypes.data_as(pl), Ax.ctypes.data_as(pd), x.ctypes.data_as(pd), y.ctypes.data_as(pd))
    assert np.allclose(y, A @ x, rtol=1e-14, atol=1e-14)
    z = rng.standard_normal(40); yt = np.zeros(30)
    L.qpdo_oracle_spmv(1, 40, 30, Ap.ctypes.data_as(pl), Ai.ctypes.data_as(pl), Ax.ctypes.data_as(pd), z.ctypes.data_as(pd), yt.ctypes.data_as(pd))
    assert np.allclose(yt, A.T @ z, rtol=1e-14, atol=1e-14)
    S = sp.random(30, 30, 0.2, random_state=2); S = (S + S.T).tocsc()
    Lo = sp.tril(S, format="csc"); Lo.sort_indices()
    yq = np.zeros(30)
    L.qpdo_oracle_spmv(2, 30, 30, Lo.indptr.astype(np.int64).ctypes.data_as(pl), Lo.indices.astype(np.int64).ctypes.data_as(pl),
                       Lo.data.ctypes.data_as(pd), x.ctypes.data_as(pd), yq.ctypes.data_as(pd))
    assert np.allclose(yq, S @ x, rtol=1e-13, atol=1e-13)


def test_ruiz_scaling_equilibrates(oracle_mod):
    import ctypes as C
    L = oracle_mod.lib()
    p = problems.random_qp(60, 40, 0.1, seed=5)
    Q = sp.tril(p["Q"], format="csc"); Q.sort_indices(); A = sp.csc_matrix(p["A"]); A.sort_indices()
    pl, pd = C.POINTER(C.c_long), C.POINTER(C.c_double)
    Qp, Qi, Qx = Q.indptr.astype(np.int64), Q.indices.astype(np.int64), Q.data.copy()
    Ap, Ai, Ax = A.indptr.astype(np.int64), A.indices.astype(np.int64), A.data.copy()
    q, l, u = p["q"].copy(), p["l"].copy(), p["u"].copy()
    D = np.zeros(60); E = np.zeros(40); c = C.c_double(0)
    L.qpdo_oracle_scale_data(60, 40, Qp.ctypes.data_as(pl), Qi.ctypes.data_as(pl), Qx.ctypes.data_as(pd),
                             Ap.ctypes.data_as(pl), Ai.ctypes.data_as(pl), Ax.ctypes.data_as(pd), q.ctypes.data_as(pd),
                             l.ctypes.data_as(pd), u.ctypes.data_as(pd), 10, D.ctypes.data_as(pd), E.ctypes.data_as(pd), C.byref(c))
    As = sp.csc_matrix((Ax, Ai, Ap), shape=A.shape)
    assert np.allclose(As.toarray(), np.diag(E) @ A.toarray() @ np.diag(D), rtol=1e-12)
    rn = np.abs(As).max(axis=1).toarray().ravel(); cn = np.abs(As).max(axis=0).toarray().ravel()
    assert 0.5 < rn[rn > 0].min() and rn.max() < 2.0 and 0.5 < cn[cn > 0].min() and cn.max() < 2.0
    assert np.allclose(l, E * p["l"]) and np.allclose(q, c.value * D * p["q"])


def test_linesearch_is_a_root_of_the_merit_derivative(oracle_mod):
    import ctypes as C
    L = oracle_mod.lib()
    rng = np.random.default_rng(3)
    n, m = 20, 30
    pd = C.POINTER(C.c_double)
    dx, Qdx_raw, df, dy, y, Adx, w = (rng.standard_normal(k) for k in (n, n, n, m, m, m, m))
    Qdx = dx * rng.uniform(0.5, 2.0, n)              # dx'Qdx > 0
    mu = rng.uniform(0.01, 10, m); sqrt_mu = 1 / np.sqrt(mu)
    l = -rng.random(m); u = rng.random(m)
    w = l + (u - l) * rng.random(m)                  # alpha >= 0
    df = -10.0 * dx; y = -dy                         # psi'(0) = beta < 0: a root tau > 0 exists
    args = [np.ascontiguousarray(a) for a in (dx, Qdx, df, dy, y, mu, sqrt_mu, Adx, w, l, u)]
    tau = L.qpdo_oracle_linesearch(n, m, *[a.ctypes.data_as(pd) for a in args])
    tm = 0.5 * dy * mu
    eta = 0.5 * (dy @ tm + dx @ Qdx); beta = 0.5 * (y @ tm + dx @ df)
    c0 = (Adx - tm) * sqrt_mu
    delta = np.concatenate([-c0, c0]); alpha = np.concatenate([(w - l) * sqrt_mu, (u - w) * sqrt_mu])
    psi = lambda t: eta * t + beta + delta @ np.maximum(delta * t - alpha, 0)
    assert abs(psi(tau)) <= 1e-9 * (abs(eta) + abs(beta) + np.abs(delta).sum() ** 2)
