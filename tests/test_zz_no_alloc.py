"""qpdo_dev_solve_noalloc (include/qpdo_cuda.h): the triangular solves with their scratch in a static device array,
no allocation.  Checked through the C ABI against the existing entry point (qpdo_dev_test_factor_solve) on two
calls in a row, and for the rejection of a problem beyond the array's stated size.  Collected last on
purpose."""
import ctypes as C
import numpy as np
import pytest
from qpdo_b200 import problems

pytestmark = pytest.mark.gpu
_pd = C.POINTER(C.c_double)


def _setup(p):
    import qpdo_b200
    s = qpdo_b200.QPDO()
    s.setup(p["Q"], p["q"], p["A"], p["l"], p["u"], dict(verbose=0))
    return s


@pytest.mark.parametrize("gen", ["rand", "cont"])
def test_noalloc_solve_matches_the_allocating_entry_twice_in_a_row(cuda_lib, gen):
    p = {"rand": lambda: problems.random_qp(600, 300, 0.02, seed=3), "cont": lambda: problems.cont_like(60)}[gen]()
    s = _setup(p)
    n, m = s.n, s.m
    rng = np.random.default_rng(5)
    om = rng.random(m) * (rng.random(m) < 0.5) * 10
    f, t = C.c_float(), C.c_float()
    for call in range(2):                                     # second call finds the first call's data in the scratch
        b = rng.standard_normal(n) * (10.0 ** call)
        x_ref, x_na = np.zeros(n), np.full(n, np.nan)
        assert cuda_lib.qpdo_dev_test_factor_solve(s.device, om.ctypes.data_as(_pd), 1e-3, b.ctypes.data_as(_pd), x_ref.ctypes.data_as(_pd),
                                                   C.byref(f), C.byref(t)) == 0, cuda_lib.qpdo_dev_last_error()
        assert cuda_lib.qpdo_dev_solve_noalloc(s.device, b.ctypes.data_as(_pd), x_na.ctypes.data_as(_pd)) == 0, cuda_lib.qpdo_dev_last_error()
        assert np.isfinite(x_na).all()
        # same kernels, same factor, same order of operations: within the 1e-8 of the other parity tests (in fact equal)
        assert np.abs(x_na - x_ref).max() <= 1e-8 * max(1.0, np.abs(x_ref).max())
        assert np.array_equal(x_na, x_ref)
    s.delete()


def test_noalloc_solve_rejects_a_problem_beyond_the_static_scratch(cuda_lib):
    cap = cuda_lib.qpdo_dev_solve_noalloc_max_doubles()
    assert cap == 1 << 19
    p = problems.cont_like(300)                               # needs 574,388 doubles > 524,288
    s = _setup(p)
    n, m = s.n, s.m
    b, x = np.ones(n), np.zeros(n)
    f, t = C.c_float(), C.c_float()
    om = np.ones(m)                                           # full support: the superset pattern is installed (pattern management)
    assert cuda_lib.qpdo_dev_test_factor_solve(s.device, om.ctypes.data_as(_pd), 1e-3, b.ctypes.data_as(_pd), x.ctypes.data_as(_pd),
                                               C.byref(f), C.byref(t)) == 0, cuda_lib.qpdo_dev_last_error()
    y = np.zeros(n)
    assert cuda_lib.qpdo_dev_solve_noalloc(s.device, b.ctypes.data_as(_pd), y.ctypes.data_as(_pd)) == -2
    msg = cuda_lib.qpdo_dev_last_error().decode()
    assert "524288" in msg and "scratch" in msg, msg
    assert not y.any()                                        # nothing was written
    s.delete()
