"""CPU tests of the drop-in boundary: the C-ABI library loads, exports every symbol include/*.h
declares, keeps the reference's defaults / validation / error behaviour, and fails loudly (no CPU
fallback) when there is no CUDA device.  No compute call is made here."""
import ctypes as C
import os
import re
import numpy as np
import pytest
import scipy.sparse as sp
from conftest import ROOT
from qpdo_b200 import capi, problems


def declared_functions(header):
    src = open(os.path.join(ROOT, "include", header)).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    names = re.findall(r"^\s*(?:const\s+)?[A-Za-z_][\w\s\*]*?[\s\*]([a-z_][a-z0-9_]*)\s*\(", src, flags=re.M)
    return sorted(set(n for n in names if n.startswith(("qpdo_", "copy_settings", "update_status"))))


@pytest.mark.parametrize("header", sorted(h for h in os.listdir(os.path.join(ROOT, "include")) if h.endswith(".h")))
def test_library_exports_every_declared_symbol(cuda_lib, header):
    names = declared_functions(header)
    assert len(names) >= 2, names
    missing = [n for n in names if not hasattr(cuda_lib, n)]
    assert not missing, "declared in include/%s but not exported: %s" % (header, missing)


def test_qpdo_h_has_the_eight_reference_entry_points():
    names = declared_functions("qpdo.h")
    for n in ("qpdo_set_default_settings", "qpdo_setup", "qpdo_warm_start", "qpdo_solve", "qpdo_update_settings",
              "qpdo_update_bounds", "qpdo_update_q", "qpdo_cleanup"):       # reference include/qpdo.h:14-56
        assert n in names


def test_default_settings_match_reference_constants(cuda_lib):
    s = capi.QPDOSettings()
    cuda_lib.qpdo_set_default_settings(C.byref(s))
    # reference include/constants.h:44-67
    assert (s.max_iter, s.inner_max_iter, s.proximal, s.scaling, s.verbose, s.print_interval, s.reset_newton_iter) == \
           (10000, 1000, 1, 10, 1, 1, 1000)
    assert (s.eps_abs, s.eps_abs_in, s.eps_prim_inf, s.eps_dual_inf) == (1e-6, 1.0, 1e-6, 1e-6)
    assert (s.rho, s.theta, s.delta, s.mu_min) == (0.1, 0.25, 1e-2, 1e-9)
    assert (s.sigma_init, s.sigma_upd, s.sigma_min, s.max_time) == (1e-3, 1e-1, 1e-7, 1e20)


def test_struct_sizes_match_the_dlong_layout():
    # reference include/types.h with -DDLONG -DPROFILING on LP64
    assert C.sizeof(capi.QPDOSettings) == 19 * 8
    assert C.sizeof(capi.QPDOInfo) == 8 + 8 + 32 + 8 + 8 * 8
    assert C.sizeof(capi.QPDOData) == 8 * 8
    assert C.sizeof(capi.CholmodSparse) == 8 * 8 + 6 * 4 + 0 or C.sizeof(capi.CholmodSparse) == 88


def test_update_status_strings(cuda_lib):
    cuda_lib.update_status.argtypes = [C.POINTER(capi.QPDOInfo), C.c_long]
    info = capi.QPDOInfo()
    for val, text in ((1, "solved"), (-3, "primal infeasible"), (-4, "dual infeasible"),
                      (-5, "maximum iterations reached"), (-6, "max time exceeded"), (-10, "unsolved"), (-99, "error")):
        cuda_lib.update_status(C.byref(info), val)           # reference src/util.c:50-91
        assert info.status.decode() == text and info.status_val == val


@pytest.mark.parametrize("bad", [dict(max_iter=0), dict(rho=1.0), dict(delta=0.0), dict(proximal=2), dict(sigma_min=1.0),
                                 dict(scaling=-1), dict(theta=1.5), dict(mu_min=0.0)])
def test_setup_rejects_invalid_settings_before_touching_the_gpu(bad, capfd):
    import qpdo_b200
    p = problems.infeasibility_trio()[0][1]
    with pytest.raises(ValueError):       # qpdo_setup returns QPDO_NULL (reference src/qpdo.c:54-67)
        qpdo_b200.QPDO().setup(p["Q"], p["q"], p["A"], p["l"], p["u"], dict(dict(verbose=0), **bad))


def test_setup_rejects_crossed_bounds(capfd):
    import qpdo_b200
    p = problems.infeasibility_trio()[0][1]
    with pytest.raises(ValueError):       # validate_data, reference src/validate.c:9-31
        qpdo_b200.QPDO().setup(p["Q"], p["q"], p["A"], np.array([1.0, 1.0, 1.0]), np.array([0.0, 3.0, 3.0]), dict(verbose=0))


def test_no_cpu_fallback_without_a_device(cuda_lib, capfd):
    if cuda_lib.qpdo_dev_count() > 0:
        pytest.skip("a CUDA device is present")
    import qpdo_b200
    p = problems.infeasibility_trio()[0][1]
    with pytest.raises(ValueError):
        qpdo_b200.QPDO().setup(p["Q"], p["q"], p["A"], p["l"], p["u"], dict(verbose=0))
    assert b"no CUDA device" in cuda_lib.qpdo_dev_last_error() or "CUDA" in capfd.readouterr().err


def test_product_does_not_import_the_oracle():
    for dirpath, _dirs, files in os.walk(os.path.join(ROOT, "qpdo_b200")):
        for f in files:
            if f.endswith((".py", ".c", ".cu", ".cpp", ".h")):
                txt = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in txt and "qpdo_oracle" not in txt and "libqpdo_ref" not in txt, f
