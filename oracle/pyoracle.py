"""ctypes binding of the CPU oracle (oracle/libgar_oracle.so) — TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this module.  The product package never does.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int)


def _cpu_stamp() -> str:
    """The .so is built -march=native and travels with the repo snapshot to other hosts:
    rebuild when the host CPU (model + ISA flags) differs from the one it was built on."""
    import hashlib
    model = flags = ""
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("model name") and not model:
                model = line
            elif line.startswith("flags") and not flags:
                flags = line
            if model and flags:
                break
    except OSError:
        pass
    return hashlib.sha1((model + flags).encode()).hexdigest()


def build(force: bool = False) -> str:
    so = os.path.join(_HERE, "libgar_oracle.so")
    stamp_file = os.path.join(_HERE, "libgar_oracle.so.cpu")
    srcs = [os.path.join(_HERE, f) for f in os.listdir(_HERE) if f.endswith((".cpp", ".hpp"))]
    srcs.append(os.path.join(_HERE, "..", "include", "garb200_layout.h"))
    stale = (not os.path.exists(so)) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs)
    stamp = _cpu_stamp()
    try:
        stale = stale or open(stamp_file).read().strip() != stamp
    except OSError:
        stale = True
    if force or stale:
        subprocess.check_call(["make", "-C", _HERE, "-B", "libgar_oracle.so"],
                              stdout=subprocess.DEVNULL)
        with open(stamp_file, "w") as f:
            f.write(stamp)
    return so


def lib():
    global _LIB
    if _LIB is None:
        _LIB = C.CDLL(build())
        _LIB.gar_oracle_bench.restype = C.c_double
    return _LIB


def _d(a):
    if a is None:
        return None
    assert a.dtype == np.float64 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(_dp)


def _i(a):
    assert a.dtype == np.int32 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(_ip)


def layouts(nx, nu, nc, nth):
    k = np.zeros(17, np.int32); g = np.zeros(5, np.int32); v = np.zeros(8, np.int32)
    lib().gar_oracle_layouts(nx, nu, nc, nth, _i(k), _i(g), _i(v))
    return k, g, v


def _dims(pp):
    L = pp.layout
    return L.nx, L.nu, L.nc, L.nth, pp.horizon, pp.nc0


def _strides(nx, nu, nc, nth):
    k, g, v = layouts(nx, nu, nc, nth)
    return int(k[16]), int(g[4]), int(v[7])


def prox_solve(pp, mudyn, mueq, theta=None, b=0):
    """ProximalRiccatiSolver backward+forward on problem ``b`` of a PackedProblems batch.
    Returns a dict of packed outputs (same layouts as the C-ABI)."""
    nx, nu, nc, nth, T, nc0 = _dims(pp)
    _, gs, vs_ = _strides(nx, nu, nc, nth)
    out = {
        "gains": np.zeros((T + 1, gs)), "values": np.zeros((T + 1, vs_)),
        "kkt0_ff": np.zeros(nx + nc0), "kkt0_fth": np.zeros((nx + nc0, nth)),
        "thGrad": np.zeros(nth), "thHess": np.zeros((nth, nth)),
        "xs": np.zeros((T + 1, nx)), "us": np.zeros((T + 1, nu)), "vs": np.zeros((T + 1, nc)),
        "lbdas": np.zeros(nc0 + T * nx), "kkt": np.zeros(3),
    }
    th = None if theta is None else np.ascontiguousarray(theta, dtype=np.float64)
    knots = np.ascontiguousarray(pp.knots[b]); G0 = np.ascontiguousarray(pp.G0[b])
    g0 = np.ascontiguousarray(pp.g0[b])
    rc = lib().gar_oracle_prox_solve(
        nx, nu, nc, nth, T, nc0, _i(pp.nu_t), _i(pp.nc_t), _d(knots), _d(G0), _d(g0),
        C.c_double(mudyn), C.c_double(mueq), _d(th), _d(out["gains"]), _d(out["values"]),
        _d(out["kkt0_ff"]), _d(out["kkt0_fth"]), _d(out["thGrad"]), _d(out["thHess"]),
        _d(out["xs"]), _d(out["us"]), _d(out["vs"]), _d(out["lbdas"]), _d(out["kkt"]))
    assert rc == 0
    out["thHess"] = out["thHess"].T.copy()  # stored column-major by the oracle
    return out


def dense_solve(pp, mudyn, mueq, theta=None, gv=None, b=0):
    """RiccatiSolverDense (dense-kernel.hpp:56-212, dense-riccati.hxx:46-118) on problem ``b``.
    gv: [T+1][nth][nc] array (column-major nc x nth blocks) or None; same outputs as prox_solve."""
    nx, nu, nc, nth, T, nc0 = _dims(pp)
    _, gs, vs_ = _strides(nx, nu, nc, nth)
    out = {
        "gains": np.zeros((T + 1, gs)), "values": np.zeros((T + 1, vs_)),
        "kkt0_ff": np.zeros(nx + nc0), "kkt0_fth": np.zeros((nx + nc0, nth)),
        "thGrad": np.zeros(nth), "thHess": np.zeros((nth, nth)),
        "xs": np.zeros((T + 1, nx)), "us": np.zeros((T + 1, nu)), "vs": np.zeros((T + 1, nc)),
        "lbdas": np.zeros(nc0 + T * nx), "kkt": np.zeros(3),
    }
    th = None if theta is None else np.ascontiguousarray(theta, dtype=np.float64)
    g = None if gv is None else np.ascontiguousarray(gv, dtype=np.float64)
    knots = np.ascontiguousarray(pp.knots[b]); G0 = np.ascontiguousarray(pp.G0[b])
    g0 = np.ascontiguousarray(pp.g0[b])
    rc = lib().gar_oracle_dense_solve(
        nx, nu, nc, nth, T, nc0, _i(pp.nu_t), _i(pp.nc_t), _d(knots), _d(G0), _d(g0), _d(g),
        C.c_double(mudyn), C.c_double(mueq), _d(th), _d(out["gains"]), _d(out["values"]),
        _d(out["kkt0_ff"]), _d(out["kkt0_fth"]), _d(out["thGrad"]), _d(out["thHess"]),
        _d(out["xs"]), _d(out["us"]), _d(out["vs"]), _d(out["lbdas"]), _d(out["kkt"]))
    assert rc == 0
    out["thHess"] = out["thHess"].T.copy()
    return out


def parallel_solve(pp, mudyn, mueq, num_legs, collapse=False, b=0):
    nx, nu, nc, nth, T, nc0 = _dims(pp)
    assert nth == 0
    _, gs, vs_ = _strides(nx, nu, nc, nx)
    out = {
        "gains": np.zeros((T + 1, gs)), "values": np.zeros((T + 1, vs_)),
        "xs": np.zeros((T + 1, nx)), "us": np.zeros((T + 1, nu)), "vs": np.zeros((T + 1, nc)),
        "lbdas": np.zeros(nc0 + T * nx), "kkt": np.zeros(3), "info": np.zeros(2, np.int32),
    }
    knots = np.ascontiguousarray(pp.knots[b]); G0 = np.ascontiguousarray(pp.G0[b])
    g0 = np.ascontiguousarray(pp.g0[b])
    rc = lib().gar_oracle_parallel_solve(
        nx, nu, nc, T, nc0, _i(pp.nu_t), _i(pp.nc_t), _d(knots), _d(G0), _d(g0),
        C.c_double(mudyn), C.c_double(mueq), int(num_legs), int(bool(collapse)),
        _d(out["gains"]), _d(out["values"]), _d(out["xs"]), _d(out["us"]), _d(out["vs"]),
        _d(out["lbdas"]), _d(out["kkt"]), _i(out["info"]))
    assert rc == 0
    return out


def kkt_error(pp, xs, us, vs, lbdas, mudyn, mueq, theta=None, b=0):
    nx, nu, nc, nth, T, nc0 = _dims(pp)
    out = np.zeros(3)
    th = None if theta is None else np.ascontiguousarray(theta, dtype=np.float64)
    knots = np.ascontiguousarray(pp.knots[b]); G0 = np.ascontiguousarray(pp.G0[b])
    g0 = np.ascontiguousarray(pp.g0[b])
    f = lambda a: np.ascontiguousarray(a, dtype=np.float64).reshape(-1)
    rc = lib().gar_oracle_kkt_error(
        nx, nu, nc, nth, T, nc0, _i(pp.nu_t), _i(pp.nc_t), _d(knots), _d(G0), _d(g0),
        C.c_double(mudyn), C.c_double(mueq), _d(th), _d(f(xs)), _d(f(us)), _d(f(vs)),
        _d(f(lbdas)), _d(out))
    assert rc == 0
    return out


def dense_kkt(pp, mudyn, mueq, b=0):
    """lqrDenseMatrix (utils.hxx:194-267): returns (mat, rhs); the KKT system is mat z = -rhs."""
    nx, nu, nc, nth, T, nc0 = _dims(pp)
    assert nth == 0
    n = lib().gar_oracle_num_rows(nx, T, nc0, _i(pp.nu_t), _i(pp.nc_t))
    mat = np.zeros((n, n)); rhs = np.zeros(n)
    knots = np.ascontiguousarray(pp.knots[b]); G0 = np.ascontiguousarray(pp.G0[b])
    g0 = np.ascontiguousarray(pp.g0[b])
    rc = lib().gar_oracle_dense_kkt(nx, nu, nc, T, nc0, _i(pp.nu_t), _i(pp.nc_t), _d(knots),
                                    _d(G0), _d(g0), C.c_double(mudyn), C.c_double(mueq),
                                    _d(mat), _d(rhs))
    assert rc == 0
    return mat, rhs


def dense_solution_to_traj(pp, z):
    """utils.hpp:106-137 ordering [lbd0 | x0 u0 v0 lbd1 | x1 ...] -> packed trajectory."""
    nx, nu, nc, nth, T, nc0 = _dims(pp)
    xs = np.zeros((T + 1, nx)); us = np.zeros((T + 1, nu)); vs = np.zeros((T + 1, nc))
    lb = np.zeros(nc0 + T * nx)
    lb[:nc0] = z[:nc0]
    idx = nc0
    for t in range(T + 1):
        m, c = int(pp.nu_t[t]), int(pp.nc_t[t])
        xs[t] = z[idx: idx + nx]; idx += nx
        us[t, :m] = z[idx: idx + m]; idx += m
        vs[t, :c] = z[idx: idx + c]; idx += c
        if t < T:
            lb[nc0 + t * nx: nc0 + (t + 1) * nx] = z[idx: idx + nx]; idx += nx
    return xs, us, vs, lb


def bench(pp, mudyn, mueq, theta=None, num_legs=0, reps=1, nthreads=0):
    """Timed CPU baseline over the whole PackedProblems batch. -> (seconds, threads)."""
    nx, nu, nc, nth, T, nc0 = _dims(pp)
    th = None if theta is None else np.ascontiguousarray(theta, dtype=np.float64)
    used = C.c_int(0)
    knots = np.ascontiguousarray(pp.knots); G0 = np.ascontiguousarray(pp.G0)
    g0 = np.ascontiguousarray(pp.g0)
    secs = lib().gar_oracle_bench(
        nx, nu, nc, nth, T, nc0, _i(pp.nu_t), _i(pp.nc_t), pp.batch, _d(knots), _d(G0),
        _d(g0), C.c_double(mudyn), C.c_double(mueq), _d(th), int(num_legs), int(reps),
        int(nthreads), C.byref(used))
    return float(secs), int(used.value)
