"""CPU checks of the drop-in boundary: libgarb200.so loads without a GPU and exports every
function include/garb200.h declares; argument validation that needs no device."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "garb200.h")


def declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    src = re.sub(r"//[^\n]*", "", src)
    return sorted(set(re.findall(r"\b(garb200_[a-z0-9_]+)\s*\(", src)))


def test_header_declares_the_abi():
    syms = declared_symbols()
    for must in ("garb200_create", "garb200_backward", "garb200_forward", "garb200_cycle_append",
                 "garb200_collapse_feedback", "garb200_get_gains", "garb200_kkt_error"):
        assert must in syms


def test_library_exports_every_declared_symbol():
    from aligator_swk_b200 import capi
    L = capi.lib()
    missing = [s for s in declared_symbols() if not hasattr(L, s)]
    assert not missing, missing
    # and the ctypes binding lists exactly the header's functions
    assert sorted(capi.SYMBOLS) == declared_symbols()


def test_version_and_error_strings_need_no_gpu():
    from aligator_swk_b200 import capi
    L = capi.lib()
    assert b"sm_100a" in L.garb200_version() or L.garb200_version()
    assert isinstance(L.garb200_last_error(), bytes)


def test_create_rejects_bad_config_before_touching_cuda():
    from aligator_swk_b200 import capi
    L = capi.lib()
    h = C.c_void_p()
    cfg = capi.Config(nx=0, nu=1, nc=0, nth=0, horizon=3, nc0=0, batch=1, num_legs=0, device=0, flags=0)
    assert L.garb200_create(C.byref(cfg), C.byref(h)) == capi.ERR_INVALID
    assert not h.value
    # parallel-in-time mode overwrites the parameterization (parallel-solver.hxx:51-61): nth must be 0
    cfg = capi.Config(nx=4, nu=1, nc=0, nth=2, horizon=8, nc0=4, batch=1, num_legs=2, device=0, flags=0)
    assert L.garb200_create(C.byref(cfg), C.byref(h)) == capi.ERR_INVALID
    cfg = capi.Config(nx=4, nu=1, nc=0, nth=0, horizon=2, nc0=4, batch=1, num_legs=8, device=0, flags=0)
    assert L.garb200_create(C.byref(cfg), C.byref(h)) == capi.ERR_INVALID     # more legs than knots
    assert L.garb200_create(None, C.byref(h)) == capi.ERR_INVALID


def test_no_cpu_fallback_without_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("only meaningful on a box without a GPU")
    from aligator_swk_b200 import capi
    with pytest.raises(capi.GarB200Error):
        capi.Solver(4, 2, 0, 0, horizon=3, batch=1)


def test_product_sources_never_touch_the_oracle():
    pkg = os.path.join(ROOT, "aligator-swk_b200")
    bad = []
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".hpp", ".cpp")):
                txt = open(os.path.join(dp, f), errors="ignore").read()
                if re.search(r"pyoracle|libgar_oracle|oracle/", txt):
                    bad.append(os.path.join(dp, f))
    assert not bad, bad


def test_headers_are_plain_c(tmp_path):
    """The drop-in boundary is a C ABI: include/garb200.h and garb200_layout.h must compile as C99 (no C++ types,
    no default arguments, no references) with every warning on."""
    import shutil
    import subprocess
    gcc = shutil.which("gcc")
    if gcc is None:
        pytest.skip("no gcc")
    src = tmp_path / "hdr.c"
    src.write_text('#include "garb200.h"\n#include "garb200_layout.h"\n'
                   "int use(void) {\n  garb200_config cfg; garb200_solver *s = 0; garb200_multi *m = 0;\n"
                   "  cfg.nx = 2; (void)m;\n"
                   "  return garb200_create(&cfg, &s) + (int)sizeof(garb200_knot_view) + (int)sizeof(garb200_device_problem_view);\n}\n")
    inc = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "include")
    r = subprocess.run([gcc, "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", "-I", inc, "-c", str(src), "-o", str(tmp_path / "hdr.o")],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
