"""-m gpu: a host C++ program drives the boundary (north_star: "reached from host C++ through a thin C-ABI shim").

tests/cxx/test_bridge.cpp is shaped like the reference's own driver program (Code/C/test_bridge.cpp:76-93): it includes
include/bayesbridge_b200.hpp (class BridgeRegression + the four driver functions over the C ABI), links
libbayesbridge_b200.so, runs all four drivers, the conditionals one by one, the in-library multi-GPU fan-out and the
interrupt hook, and exits 0 only if its own checks hold.  Built by __graft_entry__.build() (g++; no nvcc needed)."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "tests", "cxx", "test_bridge")


def _build():
    src = os.path.join(ROOT, "tests", "cxx", "test_bridge.cpp")
    hdrs = [os.path.join(ROOT, "include", h) for h in ("bayesbridge_b200.h", "bayesbridge_b200.hpp")]
    lib = os.path.join(ROOT, "bayesbridge_b200", "libbayesbridge_b200.so")
    if os.path.exists(EXE) and all(os.path.getmtime(EXE) >= os.path.getmtime(f) for f in [src, lib] + hdrs):
        return
    subprocess.run(["g++", "-O2", "-std=c++17", "-I", os.path.join(ROOT, "include"), "-o", EXE, src, "-L", os.path.dirname(lib),
                    "-lbayesbridge_b200", "-Wl,-rpath,$ORIGIN/../../bayesbridge_b200"], check=True)


def test_cxx_caller_compiles_against_the_header():
    """CPU: the C++ face of the boundary compiles and links; without a GPU the program says so and exits 2"""
    _build()
    import bayesbridge_b200 as bb
    if bb._lib.load().bb_device_count() > 0:
        pytest.skip("a GPU is present: the gpu test runs the program")
    r = subprocess.run([EXE], capture_output=True, text=True, timeout=120)
    assert r.returncode == 2 and "no CPU fallback" in r.stdout


@pytest.mark.gpu
def test_cxx_test_bridge_program():
    _build()
    import bayesbridge_b200 as bb
    ndev = bb._lib.load().bb_device_count()
    assert ndev >= 1
    r = subprocess.run([EXE, "--devices", str(min(ndev, 8))], capture_output=True, text=True, timeout=600)
    print(r.stdout[-3000:]); print(r.stderr[-2000:])
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "all checks passed" in r.stdout
