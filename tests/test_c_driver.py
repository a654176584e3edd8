"""BASELINE.json configs[0] verbatim: a C driver calling qpdo_setup / qpdo_solve.  tests/c_driver/driver.c is compiled
once against the REFERENCE's own headers (include/qpdo.h, include/types.h + SuiteSparse, -DDLONG -DPROFILING) and the
one object is linked against oracle/_ref/libqpdo_ref.so and against qpdo_b200/lib/libqpdo_b200.so (oracle/Makefile
`c_driver`, built in the build container; the binaries travel to the GPU box).  Same report from both binaries =
compile-time and run-time proof of the drop-in boundary."""
import os
import re
import subprocess
import pytest
from conftest import ROOT

REF_BIN = os.path.join(ROOT, "oracle", "_ref", "c_driver_ref")
B200_BIN = os.path.join(ROOT, "oracle", "_ref", "c_driver_b200")
NUM = re.compile(r"[-+]?\d+\.\d+e[-+]\d+")


def run(binary, *args):
    return subprocess.run([binary] + [str(a) for a in args], capture_output=True, text=True, timeout=300)


def have_binaries():
    return os.path.exists(REF_BIN) and os.path.exists(B200_BIN)


def test_driver_is_built_against_the_reference_headers_only():
    src = open(os.path.join(ROOT, "tests", "c_driver", "driver.c")).read()
    includes = re.findall(r'#include\s+"([^"]+)"', src)
    assert includes == ["qpdo.h"]                       # resolved by -I/root/reference/include in oracle/Makefile
    mk = open(os.path.join(ROOT, "oracle", "Makefile")).read()
    rule = mk[mk.index("$(OUT)/c_driver.o:"):]
    rule = rule[:rule.index("\n$(OUT)/c_driver_ref")]
    assert "$(INCS)" in rule and "../include" not in rule and "-DDLONG" in rule and "-DPROFILING" in rule


def test_reference_binary_solves_c1():
    if not have_binaries():
        pytest.skip("oracle/_ref/c_driver_* not built (needs /root/reference)")
    out = run(REF_BIN).stdout.splitlines()
    assert out[0].startswith("problem n 1000 m 500") and out[1].startswith("status 1 (solved)")


def test_b200_binary_fails_loudly_without_a_gpu(cuda_lib):
    if not have_binaries():
        pytest.skip("oracle/_ref/c_driver_* not built")
    if cuda_lib.qpdo_dev_count() > 0:
        pytest.skip("a CUDA device is present")
    r = run(B200_BIN)
    assert r.returncode == 2 and "qpdo_setup returned NULL" in r.stdout and "no CUDA device" in (r.stdout + r.stderr)


@pytest.mark.gpu
@pytest.mark.parametrize("args", [(), (400, 300, 0.02, 7), (60, 90, 0.2, 3)])
def test_same_object_linked_twice_gives_the_same_report(args):
    if not have_binaries():
        pytest.skip("oracle/_ref/c_driver_* not built")
    a, b = run(REF_BIN, *args), run(B200_BIN, *args)
    assert a.returncode == 0 and b.returncode == 0, (a.stderr, b.stderr, b.stdout)
    la, lb = a.stdout.splitlines(), b.stdout.splitlines()
    assert len(la) == len(lb) and len(la) > 10
    # problem line, status / iteration counts, active-set count and hash: exact
    assert la[:3] == lb[:3], (la[:3], lb[:3])
    for x, y in zip(la[3:], lb[3:]):
        assert NUM.sub("#", x) == NUM.sub("#", y), (x, y)                # same text apart from the floating-point fields
        if x.startswith("resolve"):
            assert x.split("objective")[0] == y.split("objective")[0]     # status / iterations of the warm-started re-solve: exact
        for u, v in zip(NUM.findall(x), NUM.findall(y)):
            u, v = float(u), float(v)
            if x.startswith("objective") and "res_prim" in x and abs(u) < 1e-5:
                continue                                                  # residual norms at convergence: both below eps_abs
            assert abs(u - v) <= 1e-8 * max(1.0, abs(u)), (x, y)
