"""The C++ host mirror (include/ppx_planner.hpp): compiled and linked against libppx.so on CPU, run
on the GPU (tests/cpp/test_planner.cpp follows the reference's test_hybrid_a_star.cpp and
test_reeds_shepp.cpp)."""
import os
import subprocess

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
CPP = os.path.join(HERE, "cpp")
BIN = os.path.join(CPP, "test_planner")


def test_cpp_host_builds_and_links():
    r = subprocess.run(["make", "-C", CPP], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    assert os.path.exists(BIN)
    # the host layer binds the C ABI only: every undefined ppx_ symbol of the binary is exported by libppx.so
    need = {l.split()[-1] for l in subprocess.run(["nm", "-u", BIN], capture_output=True, text=True).stdout.splitlines()
            if " ppx_" in l or l.strip().startswith("U ppx_")}
    lib = os.path.join(os.path.dirname(HERE), "pathplanning_b200", "libppx.so")
    have = {l.split()[-1] for l in subprocess.run(["nm", "-D", "--defined-only", lib], capture_output=True, text=True).stdout.splitlines()}
    assert need and need <= have, need - have


@pytest.mark.gpu
def test_cpp_host_runs():
    if not os.path.exists(BIN):
        subprocess.run(["make", "-C", CPP], check=True)
    r = subprocess.run([BIN], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "all checks passed" in r.stdout
