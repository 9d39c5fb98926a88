"""Builds tests/cpp/test_garb200.cpp (the reference's gar tests on the C++ mirror
include/garb200.hpp) with g++ and runs it on the GPU."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIBDIR = os.path.join(ROOT, "aligator-swk_b200")
SRC = os.path.join(ROOT, "tests", "cpp", "test_garb200.cpp")


def _build(tmp):
    exe = os.path.join(tmp, "test_garb200")
    subprocess.run(["g++", "-std=c++17", "-O2", "-I", os.path.join(ROOT, "include"), SRC, "-o", exe,
                    "-L", LIBDIR, "-lgarb200", f"-Wl,-rpath,{LIBDIR}"], check=True, capture_output=True, text=True)
    return exe


def test_cpp_mirror_compiles(tmp_path):
    assert os.path.exists(_build(str(tmp_path)))


@pytest.mark.gpu
def test_cpp_mirror_reference_cases(tmp_path):
    r = subprocess.run([_build(str(tmp_path))], capture_output=True, text=True, timeout=600)
    print(r.stdout, r.stderr)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "all C++ mirror tests passed" in r.stdout


ADAPTER_SRC = os.path.join(ROOT, "tests", "cpp", "test_adapter_stub.cpp")
STUB = os.path.join(ROOT, "tests", "cpp", "stub")


def _build_adapter(tmp):
    """include/garb200_aligator.hpp (the aligator-side adapter) against the stub of
    <aligator/gar/riccati-base.hpp>: syntax and override signatures (riccati-base.hpp:18-31) are checked
    in a container without Eigen."""
    exe = os.path.join(tmp, "test_adapter_stub")
    subprocess.run(["g++", "-std=c++17", "-O2", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), "-I", STUB,
                    ADAPTER_SRC, "-o", exe, "-L", LIBDIR, "-lgarb200", f"-Wl,-rpath,{LIBDIR}"],
                   check=True, capture_output=True, text=True)
    return exe


def test_adapter_header_compiles_against_stub(tmp_path):
    assert os.path.exists(_build_adapter(str(tmp_path)))


def test_adapter_header_refuses_to_compile_without_aligator(tmp_path):
    r = subprocess.run(["g++", "-std=c++17", "-fsyntax-only", "-I", os.path.join(ROOT, "include"), ADAPTER_SRC],
                       capture_output=True, text=True)
    assert r.returncode != 0 and "riccati-base.hpp> not found" in r.stderr


@pytest.mark.gpu
def test_adapter_through_the_reference_interface(tmp_path):
    r = subprocess.run([_build_adapter(str(tmp_path))], capture_output=True, text=True, timeout=600)
    print(r.stdout, r.stderr)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "all adapter (stub) tests passed" in r.stdout
