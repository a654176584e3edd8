"""GPU parity tests proper (-m gpu): the CUDA library, driven through the qpdo.h C ABI, against
(1) the golden vectors of the unmodified reference, (2) the C restatement on fresh seeded inputs,
(3) oracle/_ref itself when its .so travelled to the box, and (4) size-independent properties at
BASELINE.json's full sizes."""
import ctypes as C
import numpy as np
import pytest
import scipy.sparse as sp
from conftest import rel_err, load_golden, golden_cases
from test_oracle import check, KNIFE_EDGE, CHAOTIC_SEQ, TOL, TOL_SEQ
from qpdo_b200 import capi, problems

pytestmark = pytest.mark.gpu
CASES, checksum = golden_cases()
_pd = C.POINTER(C.c_double)


def gpu_solver():
    import qpdo_b200
    return qpdo_b200.QPDO()


def gpu_solve(p, st=None, warm=None):
    s = gpu_solver()
    s.setup(p["Q"], p["q"], p["A"], p["l"], p["u"], dict(dict(verbose=0), **(st or {})))
    if warm is not None:
        s.warm_start(*warm)
    r = s.solve()
    s.delete()
    return r


@pytest.mark.parametrize("name", list(CASES))
def test_cuda_matches_reference_golden(name):
    gen, st = CASES[name]
    p = gen()
    g = load_golden(name)
    assert checksum(p) == str(g["checksum"])
    s = gpu_solver()
    s.setup(p["Q"], p["q"], p["A"], p["l"], p["u"], dict(dict(verbose=0), **st))
    r = s.solve()
    check(name + "/cold", g, r, "cold_", knife=KNIFE_EDGE.get(name))
    if "warm_status_val" in g.files:
        # same workspace, like the reference's demo (examples/demo_mex.m:45-53)
        s.warm_start(g["cold_x"] * (1 + 1e-3), g["cold_y"] * (1 - 1e-3))
        check(name + "/warm", g, s.solve(), "warm_")
        m = p["A"].shape[0]
        if "lams" in p:
            xs, ys = g["cold_x"], g["cold_y"]
            for j, lam in enumerate(p["lams"][1:], 1):
                s.update_bounds(np.full(m, -lam), np.full(m, lam))
                s.warm_start(xs, ys)
                check("%s/path%d" % (name, j), g, s.solve(), "path%d_" % j)
                xs, ys = g["path%d_x" % j], g["path%d_y" % j]
        else:
            s.update_bounds(p["l"] - 0.1, p["u"] + 0.05)
            s.warm_start(g["cold_x"], g["cold_y"])
            rb = s.solve()
            if name in CHAOTIC_SEQ:
                assert rb.status_val == int(g["bounds_status_val"])
                assert abs(rb.objective - float(g["bounds_objective"])) <= 1e-6 * max(1.0, abs(float(g["bounds_objective"])))
            else:
                check(name + "/bounds", g, rb, "bounds_", TOL=TOL_SEQ)
                s.update_q(p["q"] * 1.5 + 0.01)
                check(name + "/newq", g, s.solve(), "newq_", TOL=TOL_SEQ)
    s.delete()


@pytest.mark.parametrize("seed", [21, 22, 23, 24])
def test_cuda_matches_restatement_on_fresh_inputs(oracle_mod, seed):
    p = problems.random_qp(150 + 10 * seed, 90 + seed, 0.05, seed=seed)
    a = oracle_mod.solve(p)
    b = gpu_solve(p)
    assert (a.status_val, a.iterations, a.oterations) == (b.status_val, b.iterations, b.oterations)
    assert np.array_equal(a.active_set, b.active_set)
    assert rel_err(a.x, b.x) <= TOL and rel_err(a.y, b.y) <= TOL
    assert abs(a.objective - b.objective) <= TOL * max(1.0, abs(a.objective))


def test_cuda_matches_restatement_mpc_family(oracle_mod):
    batch = problems.mpc_batch(12, first=100)
    for i in range(12):
        p = problems.mpc_instance(batch, i)
        a = oracle_mod.solve(p)
        b = gpu_solve(p)
        assert (a.status_val, a.iterations, a.oterations) == (b.status_val, b.iterations, b.oterations), i
        assert np.array_equal(a.active_set, b.active_set), i
        assert rel_err(a.x, b.x) <= TOL and rel_err(a.y, b.y) <= TOL, i


def test_cuda_matches_live_reference_when_present(oracle_mod):
    if not oracle_mod.have_ref():
        pytest.skip("oracle/_ref/libqpdo_ref.so did not travel to this machine")
    try:
        ref = oracle_mod.ref_library()
    except OSError as e:
        pytest.skip("reference build not loadable here: %s" % e)
    for nm, p in (("cont30", problems.cont_like(30)), ("rand400", problems.random_qp(400, 200, 0.02, seed=9))):
        s = capi.Solver(ref)
        s.setup(p["Q"], p["q"], p["A"], p["l"], p["u"], dict(verbose=0))
        a = s.solve(); s.delete()
        b = gpu_solve(p)
        assert (a.status_val, a.iterations, a.oterations) == (b.status_val, b.iterations, b.oterations), nm
        assert np.array_equal(a.active_set, b.active_set), nm
        assert rel_err(a.x, b.x) <= TOL and rel_err(a.y, b.y) <= TOL, nm


def test_spmv_bit_exact_against_restatement(cuda_lib, oracle_mod):
    """Ax, A'y, Qx follow cholmod_sdmult's accumulation order: bit-identical to the restatement."""
    import qpdo_b200
    L = oracle_mod.lib()
    p = problems.random_qp(300, 200, 0.05, seed=7)
    s = qpdo_b200.QPDO()
    s.setup(p["Q"], p["q"], p["A"], p["l"], p["u"], dict(verbose=0, scaling=0))
    n, m = s.n, s.m
    Q = sp.tril(sp.csc_matrix(p["Q"]), format="csc"); Q.sort_indices()
    A = sp.csc_matrix(p["A"]); A.sort_indices()
    pl = C.POINTER(C.c_long)
    rng = np.random.default_rng(0)
    for which, M, nin, nout in ((0, A, n, m), (1, A, m, n), (2, Q, n, n)):
        x = rng.standard_normal(nin); y = np.zeros(nout); yr = np.zeros(nout)
        assert cuda_lib.qpdo_dev_test_spmv(s.device, which, x.ctypes.data_as(_pd), y.ctypes.data_as(_pd)) == 0
        ip, ii = M.indptr.astype(np.int64), M.indices.astype(np.int64)
        L.qpdo_oracle_spmv(which, M.shape[0], M.shape[1], ip.ctypes.data_as(pl), ii.ctypes.data_as(pl),
                           M.data.ctypes.data_as(_pd), x.ctypes.data_as(_pd), yr.ctypes.data_as(_pd))
        assert np.array_equal(y, yr), "spmv %d not bit-exact (max diff %.3e)" % (which, np.abs(y - yr).max())
    s.delete()


def test_ruiz_scaling_bit_exact_against_restatement(cuda_lib, oracle_mod):
    import qpdo_b200
    L = oracle_mod.lib()
    p = problems.random_qp(200, 150, 0.05, seed=8)
    s = qpdo_b200.QPDO()
    s.setup(p["Q"], p["q"], p["A"], p["l"], p["u"], dict(verbose=0))
    n, m = s.n, s.m
    Q = sp.tril(sp.csc_matrix(p["Q"]), format="csc"); Q.sort_indices()
    A = sp.csc_matrix(p["A"]); A.sort_indices()
    pl = C.POINTER(C.c_long)
    Qx, Ax, q, l, u = Q.data.copy(), A.data.copy(), p["q"].copy(), p["l"].copy(), p["u"].copy()
    D = np.zeros(n); E = np.zeros(m); c = C.c_double()
    Qp, Qi, Ap, Ai = (a.astype(np.int64) for a in (Q.indptr, Q.indices, A.indptr, A.indices))
    L.qpdo_oracle_scale_data(n, m, Qp.ctypes.data_as(pl), Qi.ctypes.data_as(pl), Qx.ctypes.data_as(_pd), Ap.ctypes.data_as(pl),
                             Ai.ctypes.data_as(pl), Ax.ctypes.data_as(_pd), q.ctypes.data_as(_pd), l.ctypes.data_as(_pd),
                             u.ctypes.data_as(_pd), 10, D.ctypes.data_as(_pd), E.ctypes.data_as(_pd), C.byref(c))
    for nm, want in (("A_values", Ax), ("Q_values", Qx), ("q", q), ("l", l), ("u", u), ("D", D), ("E", E)):
        got = capi.dev_vector(cuda_lib, s.device, nm, want.size)
        assert np.array_equal(got, want), nm
    assert s.work.scaling.contents.c == c.value
    s.delete()


@pytest.mark.parametrize("gen", ["rand", "cont", "lasso"])
def test_factor_solve_residual(cuda_lib, gen):
    """K x = b with K = Q + sigma I + A' diag(omega) A: backward error of the multifrontal LDL' + solves."""
    import qpdo_b200
    p = {"rand": lambda: problems.random_qp(600, 300, 0.02, seed=3), "cont": lambda: problems.cont_like(50),
         "lasso": lambda: problems.lasso_dual(6000, 3000, window=200)}[gen]()
    s = qpdo_b200.QPDO()
    s.setup(p["Q"], p["q"], p["A"], p["l"], p["u"], dict(verbose=0))
    n, m = s.n, s.m
    A = sp.csc_matrix(p["A"]); A.sort_indices(); A.sum_duplicates()
    Ql = sp.tril(sp.csc_matrix(p["Q"]), format="csc"); Ql.sort_indices()
    As = sp.csc_matrix((capi.dev_vector(cuda_lib, s.device, "A_values", A.nnz), A.indices, A.indptr), shape=A.shape)
    Qs = sp.csc_matrix((capi.dev_vector(cuda_lib, s.device, "Q_values", Ql.nnz), Ql.indices, Ql.indptr), shape=Ql.shape)
    Qs = Qs + sp.tril(Qs, -1).T
    rng = np.random.default_rng(1)
    om = rng.random(m) * (rng.random(m) < 0.5) * 10
    sigma = 1e-3
    K = (Qs + sigma * sp.identity(n) + As.T @ sp.diags(om) @ As).tocsc()
    b = rng.standard_normal(n); x = np.zeros(n)
    f, t = C.c_float(), C.c_float()
    assert cuda_lib.qpdo_dev_test_factor_solve(s.device, om.ctypes.data_as(_pd), sigma, b.ctypes.data_as(_pd), x.ctypes.data_as(_pd),
                                               C.byref(f), C.byref(t)) == 0, cuda_lib.qpdo_dev_last_error()
    resid = np.abs(K @ x - b).max() / (np.abs(K).sum(axis=1).max() * np.abs(x).max() + np.abs(b).max())
    assert resid <= 1e-12, resid
    s.delete()


@pytest.mark.parametrize("gen", ["cont100", "c1"])
def test_factor_solve_is_deterministic_across_repeats_and_stream_modes(cuda_lib, monkeypatch, gen):
    """The same factorization + solve, 12 times over, and once more with the independent launches of a tree level on ONE
    stream (QPDO_B200_CONCURRENT=0), and with the one-CTA solves of the big fronts (QPDO_B200_SOLVE=cta): every x bit-identical.  A race between warps of a front kernel or between the
    concurrent branches of the captured graphs shows up here as a differing bit."""
    import qpdo_b200
    # cont100: 1,013 fronts, every size bucket and the blocked path; c1: a dense 815-row root front (26 panel steps chained by
    # programmatic dependent launch, 26-step cluster solves)
    p = problems.cont_like(100) if gen == "cont100" else problems.random_qp(1000, 500, 0.01, seed=1)
    rng = np.random.default_rng(7)
    xs = []
    for mode in ("1", "0", "cta"):                # "cta": big fronts solved by one CTA each instead of a cluster (same bits by design)
        monkeypatch.setenv("QPDO_B200_CONCURRENT", "0" if mode == "0" else "1")
        monkeypatch.setenv("QPDO_B200_SOLVE", "cta" if mode == "cta" else "cluster")
        s = qpdo_b200.QPDO()
        s.setup(p["Q"], p["q"], p["A"], p["l"], p["u"], dict(verbose=0))
        n, m = s.n, s.m
        if not xs:
            om = rng.random(m) * (rng.random(m) < 0.5) * 10
            b = rng.standard_normal(n)
        f, t = C.c_float(), C.c_float()
        for rep in range(12 if mode == "1" else 2):
            x = np.zeros(n)
            assert cuda_lib.qpdo_dev_test_factor_solve(s.device, om.ctypes.data_as(_pd), 1e-3, b.ctypes.data_as(_pd), x.ctypes.data_as(_pd),
                                                       C.byref(f), C.byref(t)) == 0, cuda_lib.qpdo_dev_last_error()
            xs.append(x)
        s.delete()
    for x in xs[1:]:
        assert np.array_equal(x, xs[0])


def test_singleton_leaf_fronts_elementwise_path_is_bit_identical(cuda_lib, monkeypatch):
    """Columns of K without off-diagonal entries (nearly all of a Lasso-type problem with few weighted rows) are handled by
    element-wise kernels instead of one CTA per front (QPDO_B200_SINGLETONS=0 switches that off): same x, bit for bit."""
    import qpdo_b200
    p = problems.lasso_dual(6000, 3000, window=200)
    rng = np.random.default_rng(11)
    xs = []
    for mode in ("1", "0"):
        monkeypatch.setenv("QPDO_B200_SINGLETONS", mode)
        monkeypatch.setenv("QPDO_PATTERN", "dynamic")       # pattern = the weighted rows only: most columns are isolated
        s = qpdo_b200.QPDO()
        s.setup(p["Q"], p["q"], p["A"], p["l"], p["u"], dict(verbose=0))
        n, m = s.n, s.m
        if not xs:
            om = np.zeros(m); om[rng.choice(m, 40, replace=False)] = rng.random(40) * 5
            b = rng.standard_normal(n)
        f, t = C.c_float(), C.c_float()
        x = np.zeros(n)
        assert cuda_lib.qpdo_dev_test_factor_solve(s.device, om.ctypes.data_as(_pd), 1e-3, b.ctypes.data_as(_pd), x.ctypes.data_as(_pd),
                                                   C.byref(f), C.byref(t)) == 0, cuda_lib.qpdo_dev_last_error()
        info = capi.DevSymbolicInfo(); cuda_lib.qpdo_dev_get_symbolic_info(s.device, C.byref(info))
        assert info.nsuper > 4000                            # thousands of one-pivot fronts
        xs.append(x)
        s.delete()
    assert np.isfinite(xs[0]).all() and np.array_equal(xs[0], xs[1])


def test_full_size_cont300_against_reference_golden(cuda_lib):
    """BASELINE config 3 at full size (n = 90,300, m = 180,300) against the unmodified reference's
    result (tests/golden/cont300.npz, make_golden_large.py) + KKT conditions recomputed on the host."""
    p = problems.cont_like(300)
    g = load_golden("cont300")
    r = gpu_solve(p)
    assert r.status_val == int(g["status_val"]) == 1
    assert (r.iterations, r.oterations) == (int(g["iterations"]), int(g["oterations"]))
    assert np.array_equal(np.packbits(r.active_set.astype(np.uint8)), g["active_set"])
    assert rel_err(g["x"], r.x) <= TOL and rel_err(g["y"], r.y) <= TOL
    assert abs(r.objective - float(g["objective"])) <= TOL * max(1.0, abs(float(g["objective"])))
    A, Q = p["A"], p["Q"]
    Ax = A @ r.x
    assert np.abs(Ax - np.clip(Ax + r.y, p["l"], p["u"])).max() <= 1e-5
    assert np.abs(Q @ r.x + p["q"] + A.T @ r.y).max() <= 1e-5


@pytest.mark.parametrize("name", ["lasso_small", "rand300", "cont40", "mpc0", "c1"])
def test_dynamic_pattern_matches_golden_and_superset(name, monkeypatch):
    """pattern management (factor.cu): the row set of A behind the symbolic structure follows the support of
    omega (re-analysis on demand, what the reference does at every refactor, cholmod_interface.c:10) or is
    the fixed superset; both must reproduce the reference's cold solve"""
    gen, st = CASES[name]
    p = gen()
    g = load_golden(name)
    res = {}
    for mode in ("superset", "dynamic"):
        monkeypatch.setenv("QPDO_PATTERN", mode)
        res[mode] = gpu_solve(p, st)
        check("%s[%s]" % (name, mode), g, res[mode], "cold_", knife=KNIFE_EDGE.get(name))
    a, b = res["superset"], res["dynamic"]
    assert (a.status_val, a.iterations, a.oterations) == (b.status_val, b.iterations, b.oterations)
    assert np.array_equal(a.active_set, b.active_set)
    assert rel_err(a.x, b.x) <= 1e-9 and rel_err(a.y, b.y) <= 1e-9


def test_full_size_lasso_path_against_reference_golden():
    """BASELINE config 4 at full size (n = 200,000, m = 100,000): the whole 20-step regularisation path
    (qpdo_update_bounds + qpdo_warm_start + qpdo_solve per step) against the unmodified reference
    (tests/golden/lasso_c4.npz, make_golden_large.py lasso; x, y stored with a stride)."""
    p = problems.lasso_dual()
    g = load_golden("lasso_c4")
    stride = int(g["stride"])
    m = p["A"].shape[0]
    s = gpu_solver()
    s.setup(p["Q"], p["q"], p["A"], p["l"], p["u"], dict(verbose=0))
    r = s.solve()
    assert (r.status_val, r.iterations, r.oterations) == (int(g["status_val"]), int(g["iterations"]), int(g["oterations"]))
    assert np.array_equal(np.packbits(r.active_set.astype(np.uint8)), g["active_set"])
    assert rel_err(g["x"], r.x[::stride]) <= TOL and rel_err(g["y"], r.y[::stride]) <= TOL
    assert abs(r.objective - float(g["objective"])) <= TOL * max(1.0, abs(float(g["objective"])))
    x, y = r.x.copy(), r.y.copy()
    last = len(p["lams"]) - 1
    for j, lam in enumerate(p["lams"][1:], 1):
        s.update_bounds(np.full(m, -lam), np.full(m, lam))
        s.warm_start(x, y)
        r = s.solve()
        tag = "path %d" % j
        assert (r.status_val, r.iterations, r.oterations) == (int(g["path%d_status_val" % j]), int(g["path%d_iterations" % j]),
                                                              int(g["path%d_oterations" % j])), tag
        assert np.array_equal(np.packbits(r.active_set.astype(np.uint8)), g["path%d_active_set" % j]), tag
        assert abs(r.objective - float(g["path%d_objective" % j])) <= TOL * max(1.0, abs(float(g["path%d_objective" % j]))), tag
        x, y = r.x.copy(), r.y.copy()
        if j == last:
            assert rel_err(g["pathlast_x"], x[::stride]) <= TOL and rel_err(g["pathlast_y"], y[::stride]) <= TOL
    s.delete()


def test_edge_cases_no_constraints_and_scalar_problem(oracle_mod):
    """m = 0 (the Newton matrix is Q + sigma I only, newton.c:21,31-33) and n = m = 1, against the restatement
    (which the live reference agrees with: ref m=0 -> 6 loop passes / 2 outer, n=1 -> 10 / 4)"""
    cases = [dict(Q=sp.csc_matrix(np.array([[2.0, 0.5], [0.5, 1.0]])), q=np.array([1.0, -1.0]), A=sp.csc_matrix((0, 2)),
                  l=np.zeros(0), u=np.zeros(0)),
             dict(Q=sp.csc_matrix(np.array([[1.0]])), q=np.array([-3.0]), A=sp.csc_matrix(np.array([[1.0]])),
                  l=np.array([-1.0]), u=np.array([1.0]))]
    expect = [(1, 6, 2), (1, 10, 4)]
    for p, e in zip(cases, expect):
        o = oracle_mod.solve(p)
        r = gpu_solve(p)
        assert (o.status_val, o.iterations, o.oterations) == e
        assert (r.status_val, r.iterations, r.oterations) == e
        assert rel_err(o.x, r.x) <= TOL and (o.y.size == 0 or rel_err(o.y, r.y) <= TOL)
        assert abs(r.objective - o.objective) <= TOL * max(1.0, abs(o.objective))


@pytest.mark.parametrize("n,m,case", [(20, 30, "root"), (40, 1000, "root"), (40, 1500, "root"), (30, 40000, "root"),
                                      (30, 40000, "ties"), (30, 6000, "no_breakpoint"), (30, 6000, "inf_bounds")])
def test_linesearch_kernels_against_restatement(cuda_lib, oracle_mod, n, m, case):
    """exact_linesearch alone (linesearch.c:8-158): the one-CTA bitonic path (2m <= 2048) and the breakpoint build ->
    8-pass radix sort -> block sweep path against the restatement's qsort + sequential sweep on the same vectors.
    Tolerance 1e-10 relative: the sweep's prefix sums are blocked on the GPU, sequential in the reference."""
    import qpdo_b200
    L = oracle_mod.lib()
    p = problems.random_qp(n, m, min(1.0, 4.0 / n), seed=5)
    s = qpdo_b200.QPDO()
    s.setup(p["Q"], p["q"], p["A"], p["l"], p["u"], dict(verbose=0, scaling=0))
    rng = np.random.default_rng(n + m)
    dx, dy, Adx = rng.standard_normal(n), rng.standard_normal(m), rng.standard_normal(m)
    Qdx = dx * rng.uniform(0.5, 2.0, n)
    mu = rng.uniform(0.01, 10, m); sqrt_mu = 1 / np.sqrt(mu)
    l = -rng.random(m); u = rng.random(m)
    w = l + (u - l) * rng.random(m)
    df = -10.0 * dx; y = -dy
    if case == "ties":                               # many equal breakpoints: the sort order of equal keys must not matter
        mu[:] = 1.0; sqrt_mu[:] = 1.0
        dy[:] = 0.0; y[:] = 0.0
        Adx = rng.choice([-1.0, 1.0, 2.0], m); l[:] = -1.0; u[:] = 1.0; w = rng.choice([-0.5, 0.0, 0.5], m)
    elif case == "no_breakpoint":                    # the root lies before the first breakpoint: tau = -beta / eta
        df = -1e-6 * dx
        y = -1e-6 * dy
    elif case == "inf_bounds":
        l[::2] = -1e20; u[1::3] = 1e20
    args = [np.ascontiguousarray(a, dtype=np.float64) for a in (dx, Qdx, df, dy, y, mu, sqrt_mu, Adx, w, l, u)]
    ptrs = [a.ctypes.data_as(_pd) for a in args]
    tau_ref = L.qpdo_oracle_linesearch(n, m, *ptrs)
    tau = C.c_double(-1.0)
    cuda_lib.qpdo_dev_test_linesearch.argtypes = [C.c_void_p] + [_pd] * 11 + [C.POINTER(C.c_double)]
    cuda_lib.qpdo_dev_test_linesearch.restype = C.c_int
    assert cuda_lib.qpdo_dev_test_linesearch(s.device, *ptrs, C.byref(tau)) == 0
    s.delete()
    assert tau_ref > 0
    assert abs(tau.value - tau_ref) <= 1e-10 * abs(tau_ref), (tau.value, tau_ref)


def _rescale_sequence(compat, monkeypatch):
    import os
    import qpdo_b200
    monkeypatch.setenv("QPDO_B200_COMPAT_RESCALE", "1" if compat else "0")
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "rescale_seq.npz"))
    p = problems.random_qp(120, 80, 0.05, seed=11)
    s = qpdo_b200.QPDO()
    s.setup(p["Q"], p["q"], p["A"], p["l"], p["u"], dict(verbose=0, scaling=2))
    a = s.solve()
    assert [a.status_val, a.iterations, a.oterations] == list(g["info1"])
    assert rel_err(a.x, g["x1"]) <= TOL and rel_err(a.y, g["y1"]) <= TOL
    D_old = np.ctypeslib.as_array(s.work.scaling.contents.D, shape=(s.n,)).copy()
    s.update_settings(dict(verbose=0, scaling=10))
    w = s.work
    assert w.info.contents.status_val != -99          # QPDO_ERROR
    D = np.ctypeslib.as_array(w.scaling.contents.D, shape=(s.n,)).copy()
    E = np.ctypeslib.as_array(w.scaling.contents.E, shape=(s.m,)).copy()
    assert rel_err(E, g["E"]) <= 1e-12 and abs(w.scaling.contents.c - float(g["c"])) <= TOL * float(g["c"])
    return s, p, g, a, D_old, D, E


def test_update_settings_more_scaling_sweeps_reference_compatible(monkeypatch):
    """qpdo_update_settings with more Ruiz sweeps (qpdo.c:496-512) exactly as the reference behaves, including its
    clobbered temp_n (total D = D_increment o (Qx + q)): golden sequence from the unmodified reference
    (tests/golden/make_golden_rescale.py): scaling 2, solve, scaling 10, warm start, solve."""
    s, p, g, a, D_old, D, E = _rescale_sequence(True, monkeypatch)
    assert rel_err(D, g["D"]) <= TOL
    s.warm_start(a.x, a.y)
    b = s.solve()
    assert [b.status_val, b.iterations, b.oterations] == list(g["info2"])
    assert np.array_equal(np.asarray(b.active_set), g["active2"])
    assert rel_err(b.x, g["x2"]) <= TOL and rel_err(b.y, g["y2"]) <= TOL
    assert abs(b.objective - float(g["obj2"])) <= TOL * abs(float(g["obj2"]))
    # decreasing is refused with QPDO_ERROR and leaves the settings alone (qpdo.c:490-495)
    s.update_settings(dict(verbose=0, scaling=3))
    assert s.work.info.contents.status_val == -99 and s.work.settings.contents.scaling == 10
    s.delete()


def test_update_settings_more_scaling_sweeps_intended_total_scaling(cuda_lib, monkeypatch):
    """Default behaviour: E and c as the reference computes them, D = D_increment o D_old (the product the reference's
    comment announces): the scaled matrix on the device is E A D, and the warm-started solve returns the same solution."""
    s, p, g, a, D_old, D, E = _rescale_sequence(False, monkeypatch)
    assert np.all(D > 0) and not np.allclose(D, D_old)
    A = sp.csc_matrix(p["A"])
    v = np.random.default_rng(1).standard_normal(s.n); out = np.zeros(s.m)
    assert cuda_lib.qpdo_dev_test_spmv(s.device, 0, v.ctypes.data_as(_pd), out.ctypes.data_as(_pd)) == 0
    assert rel_err(out, E * (A @ (D * v))) <= 1e-12
    s.warm_start(a.x, a.y)
    b = s.solve()
    assert b.status_val == 1
    assert rel_err(b.x, a.x) <= 1e-4 and abs(b.objective - a.objective) <= 1e-6 * abs(a.objective)
    s.delete()


def test_host_mirror_of_the_scaled_data_matches_the_reference(oracle_mod):
    """work->data->{q, l, u} hold the SCALED vectors after qpdo_setup, qpdo_update_bounds and qpdo_update_q, as in the
    reference (src/scaling.c:40-91, src/qpdo.c:538-543, 560-571): same values as the live reference's workspace"""
    if not oracle_mod.have_ref():
        pytest.skip("oracle/_ref not built")
    p = problems.random_qp(120, 80, 0.08, seed=11)
    g = gpu_solver()
    r = capi.Solver(oracle_mod.ref_library())
    n, m = 120, 80

    def vecs(s):
        d = s.work.data.contents
        f = lambda ptr, k: np.array(np.ctypeslib.as_array(ptr, shape=(k,)))
        return f(d.q, n), f(d.l, m), f(d.u, m)
    for s in (g, r):
        s.setup(p["Q"], p["q"], p["A"], p["l"], p["u"], dict(verbose=0))
    for a, b in zip(vecs(g), vecs(r)):
        assert rel_err(b, a) <= 1e-13
    for s in (g, r):
        s.update_bounds(p["l"] - 0.25, p["u"] + 0.5)
        s.update_q(p["q"] * 1.25 + 0.125)
    for a, b in zip(vecs(g), vecs(r)):
        assert rel_err(b, a) <= 1e-13
    g.delete(); r.delete()
