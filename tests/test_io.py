"""QPS reader (include/qpdo_io.h): host-side parsing checks on the CPU, and on the GPU the
read problem solved through qpdo_setup / qpdo_solve against the C restatement of the reference."""
import numpy as np
import pytest
from conftest import rel_err
from qpdo_b200 import io as qio

# the example of the QPS format description (Maros & Meszaros 1999), plus a range, a free and a fixed variable
QPS = """NAME          QPEX
ROWS
 N  obj
 G  r1
 L  r2
 E  r3
COLUMNS
    c1        r1        2.0        r2        -1.0
    c1        obj       1.5
    c2        r1        1.0        r2        2.0
    c2        obj       -2.0       r3        1.0
    c3        r3        1.0        obj       0.5
    c4        r3        -1.0
RHS
    rhs1      obj       -4.0
    rhs1      r1        2.0        r2        6.0
    rhs1      r3        1.0
RANGES
    rng1      r1        8.0
BOUNDS
 UP bnd1      c1        20.0
 FR bnd1      c3
 FX bnd1      c4        0.25
QUADOBJ
    c1        c1        8.0
    c1        c2        2.0
    c2        c2        10.0
    c3        c3        1.0
    c4        c4        1.0
ENDATA
"""
INF = 1e20


@pytest.fixture()
def qps_file(tmp_path):
    p = tmp_path / "qpex.qps"
    p.write_text(QPS)
    return p


def test_read_qps_arrays(cuda_lib, qps_file):
    d = qio.read_qps(qps_file, cuda_lib)
    assert d["m_rows"] == 3 and d["A"].shape == (6, 4) and d["Q"].shape == (4, 4)      # c1, c2 [0, ..), c4 fixed: 3 bound rows; c3 free: none
    np.testing.assert_array_equal(d["q"], [1.5, -2.0, 0.5, 0.0])
    assert d["c"] == 4.0
    np.testing.assert_array_equal(d["Q"].toarray(), [[8, 2, 0, 0], [2, 10, 0, 0], [0, 0, 1, 0], [0, 0, 0, 1]])
    np.testing.assert_array_equal(d["A"].toarray(), [[2, 1, 0, 0], [-1, 2, 0, 0], [0, 1, 1, -1],
                                                      [1, 0, 0, 0], [0, 1, 0, 0], [0, 0, 0, 1]])
    np.testing.assert_array_equal(d["l"], [2.0, -INF, 1.0, 0.0, 0.0, 0.25])
    np.testing.assert_array_equal(d["u"], [10.0, 6.0, 1.0, 20.0, INF, 0.25])


def test_read_qps_errors(cuda_lib, tmp_path):
    with pytest.raises(OSError):
        qio.read_qps(tmp_path / "missing.qps", cuda_lib)
    bad = tmp_path / "bad.qps"
    bad.write_text(QPS.replace("    c2        obj       -2.0       r3        1.0", "    c2        obj       -2.0       r9        1.0"))
    with pytest.raises(ValueError, match="line 11"):
        qio.read_qps(bad, cuda_lib)


def test_read_qps_objsense(cuda_lib, tmp_path):
    """OBJSENSE MIN (either spelling) is accepted; a maximisation file is rejected with its line number instead of
    being read -- and solved -- as a minimisation"""
    ok1 = tmp_path / "min1.qps"
    ok1.write_text(QPS.replace("ROWS\n", "OBJSENSE\n    MIN\nROWS\n", 1))
    ok2 = tmp_path / "min2.qps"
    ok2.write_text(QPS.replace("ROWS\n", "OBJSENSE MIN\nROWS\n", 1))
    for f in (ok1, ok2):
        np.testing.assert_array_equal(qio.read_qps(f, cuda_lib)["q"], [1.5, -2.0, 0.5, 0.0])
    bad1 = tmp_path / "max1.qps"
    bad1.write_text(QPS.replace("ROWS\n", "OBJSENSE\n    MAX\nROWS\n", 1))
    with pytest.raises(ValueError, match="line 3"):
        qio.read_qps(bad1, cuda_lib)
    bad2 = tmp_path / "max2.qps"
    bad2.write_text(QPS.replace("ROWS\n", "OBJSENSE MAX\nROWS\n", 1))
    with pytest.raises(ValueError, match="line 2"):
        qio.read_qps(bad2, cuda_lib)


def test_read_qps_solves_on_the_oracle(cuda_lib, oracle_mod, qps_file):
    d = qio.read_qps(qps_file, cuda_lib)
    o = oracle_mod.solve(d, trace=False)
    assert o.status_val == 1
    x = np.asarray(o.x)
    assert np.all(d["A"] @ x >= d["l"] - 1e-5) and np.all(d["A"] @ x <= d["u"] + 1e-5)
    # KKT stationarity: Qx + q + A'y = 0
    assert np.abs(d["Q"] @ x + d["q"] + d["A"].T @ np.asarray(o.y)).max() < 1e-5


@pytest.mark.gpu
def test_read_qps_cuda_matches_restatement(cuda_lib, oracle_mod, qps_file):
    from qpdo_b200 import QPDO
    d = qio.read_qps(qps_file, cuda_lib)
    o = oracle_mod.solve(d, trace=False)
    s = QPDO()
    s.setup(d["Q"], d["q"], d["A"], d["l"], d["u"], dict(verbose=0))
    r = s.solve()
    s.delete()
    assert (r.status_val, r.iterations, r.oterations) == (o.status_val, o.iterations, o.oterations)
    assert rel_err(r.x, o.x) < 1e-8 and rel_err(r.y, o.y) < 1e-8


# ---------------------------------------------------------------- MatrixMarket (qpdo_read_mtx*)
def _write_mtx_problem(tmp_path, seed=3, n=12, m=9):
    import scipy.io as sio
    import scipy.sparse as sp
    rng = np.random.default_rng(seed)
    B = sp.random(n, n, 0.2, random_state=seed, format="csc")
    Q = (B @ B.T + sp.identity(n)).tocsc()
    A = sp.random(m, n, 0.4, random_state=seed + 1, format="csc")
    q = rng.standard_normal(n)
    l = -rng.random(m); u = rng.random(m)
    l[0] = -np.inf; u[1] = 1e30                      # infinite bounds spelled two ways
    sio.mmwrite(str(tmp_path / "Q.mtx"), Q, symmetry="symmetric")
    sio.mmwrite(str(tmp_path / "Qg.mtx"), Q, symmetry="general")
    sio.mmwrite(str(tmp_path / "A.mtx"), A)
    sio.mmwrite(str(tmp_path / "q.mtx"), q.reshape(-1, 1))                    # array format
    sio.mmwrite(str(tmp_path / "l.mtx"), sp.csc_matrix(np.where(np.isinf(l), -1e21, l).reshape(-1, 1)))   # coordinate n x 1
    (tmp_path / "u.mtx").write_text("%%MatrixMarket matrix array real general\n% upper bounds\n1 {}\n".format(m)
                                    + "\n".join("inf" if v > 1e20 else repr(float(v)) for v in u) + "\n")
    return Q, A, q, l, u


def test_read_mtx_problem_arrays(cuda_lib, tmp_path):
    import scipy.sparse as sp
    Q, A, q, l, u = _write_mtx_problem(tmp_path)
    S, stype = qio.read_mtx(tmp_path / "Q.mtx", cuda_lib)
    assert stype == -1 and abs(S - sp.tril(Q)).max() < 1e-15 and sp.triu(S, 1).nnz == 0
    S, stype = qio.read_mtx(tmp_path / "A.mtx", cuda_lib)
    assert stype == 0 and S.has_sorted_indices and abs(S - A).max() < 1e-15
    for qfile in ("Q.mtx", "Qg.mtx"):
        d = qio.read_mtx_problem(tmp_path / "A.mtx", tmp_path / qfile, tmp_path / "q.mtx", tmp_path / "l.mtx", tmp_path / "u.mtx", cuda_lib)
        assert abs(d["Q"] - Q).max() < 1e-15 and abs(d["A"] - A).max() < 1e-15
        assert np.allclose(d["q"], q, rtol=0, atol=1e-15)
        assert d["l"][0] == -INF and d["u"][1] == INF
        assert np.allclose(d["l"][1:], l[1:], rtol=0, atol=1e-15) and np.allclose(np.delete(d["u"], 1), np.delete(u, 1), rtol=0, atol=1e-15)
    d = qio.read_mtx_problem(tmp_path / "A.mtx", lib=cuda_lib)          # defaults: LP with free rows
    assert d["Q"].nnz == 0 and not d["q"].any() and np.all(d["l"] == -INF) and np.all(d["u"] == INF)


def test_read_mtx_pattern_skew_and_errors(cuda_lib, tmp_path):
    (tmp_path / "p.mtx").write_text("%%MatrixMarket matrix coordinate pattern symmetric\n3 3 3\n1 1\n3 1\n2 3\n")
    S, stype = qio.read_mtx(tmp_path / "p.mtx", cuda_lib)
    assert stype == -1 and np.array_equal(S.toarray(), [[1, 0, 0], [0, 0, 0], [1, 1, 0]])
    (tmp_path / "k.mtx").write_text("%%MatrixMarket matrix coordinate real skew-symmetric\n2 2 1\n2 1 3.5\n")
    S, stype = qio.read_mtx(tmp_path / "k.mtx", cuda_lib)
    assert stype == 0 and np.array_equal(S.toarray(), [[0, -3.5], [3.5, 0]])
    (tmp_path / "d.mtx").write_text("%%MatrixMarket matrix coordinate real general\n2 2 3\n1 1 1.0\n1 1 2.0\n2 1 -1\n")
    S, _ = qio.read_mtx(tmp_path / "d.mtx", cuda_lib)                   # duplicates are summed
    assert np.array_equal(S.toarray(), [[3, 0], [-1, 0]])
    with pytest.raises(OSError):
        qio.read_mtx(tmp_path / "missing.mtx", cuda_lib)
    (tmp_path / "bad.mtx").write_text("%%MatrixMarket matrix coordinate real general\n2 2 2\n1 1 1.0\n3 1 2.0\n")
    with pytest.raises(ValueError, match="line 4"):
        qio.read_mtx(tmp_path / "bad.mtx", cuda_lib)
    (tmp_path / "cplx.mtx").write_text("%%MatrixMarket matrix coordinate complex general\n1 1 1\n1 1 1.0 0.0\n")
    with pytest.raises(ValueError, match="line 1"):
        qio.read_mtx(tmp_path / "cplx.mtx", cuda_lib)
    Q, A, q, l, u = _write_mtx_problem(tmp_path)
    with pytest.raises(ValueError, match="dimensions"):
        qio.read_mtx_problem(tmp_path / "A.mtx", tmp_path / "Q.mtx", tmp_path / "l.mtx", lib=cuda_lib)   # q of length m


def test_read_mtx_problem_solves_on_the_oracle(cuda_lib, oracle_mod, tmp_path):
    _write_mtx_problem(tmp_path)
    d = qio.read_mtx_problem(tmp_path / "A.mtx", tmp_path / "Q.mtx", tmp_path / "q.mtx", tmp_path / "l.mtx", tmp_path / "u.mtx", cuda_lib)
    o = oracle_mod.solve(d, trace=False)
    assert o.status_val == 1
    x = np.asarray(o.x)
    assert np.abs(d["Q"] @ x + d["q"] + d["A"].T @ np.asarray(o.y)).max() < 1e-5


@pytest.mark.gpu
def test_read_mtx_problem_cuda_matches_restatement(cuda_lib, oracle_mod, tmp_path):
    from qpdo_b200 import QPDO
    _write_mtx_problem(tmp_path)
    d = qio.read_mtx_problem(tmp_path / "A.mtx", tmp_path / "Q.mtx", tmp_path / "q.mtx", tmp_path / "l.mtx", tmp_path / "u.mtx", cuda_lib)
    o = oracle_mod.solve(d, trace=False)
    s = QPDO()
    s.setup(d["Q"], d["q"], d["A"], d["l"], d["u"], dict(verbose=0))
    r = s.solve()
    s.delete()
    assert (r.status_val, r.iterations, r.oterations) == (o.status_val, o.iterations, o.oterations)
    assert rel_err(r.x, o.x) < 1e-8 and rel_err(r.y, o.y) < 1e-8
