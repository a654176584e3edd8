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
