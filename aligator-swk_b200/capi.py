"""ctypes binding of the C ABI (include/garb200.h -> libgarb200.so).

There is no CPU fallback: if the shared library is missing, or no CUDA device is present,
the calls raise.  ``lib()`` itself only needs the .so (so symbol checks run on CPU boxes).
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libgarb200.so")
_LIB = None

OK, ERR_INVALID, ERR_CUDA, ERR_NUMERIC, ERR_STATE = 0, 1, 2, 3, 4
STORE_ALL_VALUES = 1
FORCE_GENERIC = 2
PHASE_TIMERS = 4
EXPLICIT_DYNAMICS = 8
DENSE_KERNEL = 16
NO_PADDING = 32

_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int)


class GarB200Error(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"garb200 error {code}: {msg}")
        self.code = code


class Config(C.Structure):
    _fields_ = [("nx", C.c_int), ("nu", C.c_int), ("nc", C.c_int), ("nth", C.c_int),
                ("horizon", C.c_int), ("nc0", C.c_int), ("batch", C.c_int),
                ("num_legs", C.c_int), ("device", C.c_int), ("flags", C.c_int),
                ("nu_t", _ip), ("nc_t", _ip), ("nu_term", C.c_int), ("nc_term", C.c_int)]


class DeviceField(C.Structure):
    _fields_ = [("ptr", C.c_void_p), ("stride_b", C.c_longlong), ("stride_t", C.c_longlong), ("ld", C.c_int)]


class DeviceProblemView(C.Structure):
    _fields_ = [(n, DeviceField) for n in ("Q", "S", "R", "q", "r", "A", "B", "E", "f", "C", "D", "d",
                                           "Gth", "Gx", "Gu", "gamma", "G0", "g0")]


class KnotView(C.Structure):
    _fields_ = [("nu", C.c_int), ("nc", C.c_int)] + [
        (n, _dp) for n in ("Q", "S", "R", "q", "r", "A", "B", "E", "f", "C", "D", "d",
                           "Gth", "Gx", "Gu", "gamma")]


# every symbol include/garb200.h declares (tests/test_capi_symbols.py checks the export list)
SYMBOLS = [
    "garb200_create", "garb200_destroy", "garb200_last_error", "garb200_version",
    "garb200_set_stream", "garb200_synchronize", "garb200_get_layouts", "garb200_layouts",
    "garb200_upload_packed", "garb200_stage_knot", "garb200_stage_initial", "garb200_commit",
    "garb200_device_problem", "garb200_generate", "garb200_download_packed",
    "garb200_backward", "garb200_forward", "garb200_forward_device",
    "garb200_device_trajectory", "garb200_get_gains", "garb200_download_gains",
    "garb200_device_gains", "garb200_get_value", "garb200_get_initial", "garb200_get_flags",
    "garb200_get_refinement_steps", "garb200_collapse_feedback", "garb200_cycle_append",
    "garb200_kkt_error", "garb200_upload_trajectory", "garb200_kernel_launches", "garb200_last_timings",
    "garb200_fp64_peak", "garb200_timing_log_start", "garb200_timing_log_read",
    "garb200_used_fast_path", "garb200_phase_cycles", "garb200_set_tuning",
    "garb200_kkt_error_noalloc", "garb200_used_explicit_path", "garb200_padded_shape",
    "garb200_pack_from_device", "garb200_linear_step", "garb200_rollout_device", "garb200_device_theta",
    "garb200_upload_gv",
    "garb200_multi_create", "garb200_multi_destroy", "garb200_multi_num_shards", "garb200_multi_shard",
    "garb200_multi_upload_packed", "garb200_multi_generate", "garb200_multi_backward", "garb200_multi_forward",
    "garb200_multi_synchronize", "garb200_multi_kkt_error",
]


def lib():
    """Load libgarb200.so (built by ``__graft_entry__.build()``); raise loudly if absent."""
    global _LIB
    if _LIB is None:
        if not os.path.exists(LIB_PATH):
            raise GarB200Error(ERR_CUDA, f"{LIB_PATH} not built: run __graft_entry__.build(); "
                                         "there is no CPU fallback")
        L = C.CDLL(LIB_PATH)
        L.garb200_last_error.restype = C.c_char_p
        L.garb200_version.restype = C.c_char_p
        L.garb200_kernel_launches.restype = C.c_int64
        L.garb200_destroy.restype = None
        L.garb200_layouts.restype = None
        L.garb200_backward.argtypes = [C.c_void_p, C.c_double, C.c_double]
        L.garb200_kkt_error.argtypes = [C.c_void_p, C.c_double, C.c_double, C.c_void_p, C.c_void_p]
        L.garb200_generate.argtypes = [C.c_void_p, C.c_uint64, C.c_int]
        for name in ("garb200_forward", "garb200_forward_device"):
            getattr(L, name).argtypes = [C.c_void_p] * 6 if name == "garb200_forward_device" else [C.c_void_p] * 6
        L.garb200_upload_packed.argtypes = [C.c_void_p] * 4
        L.garb200_set_stream.argtypes = [C.c_void_p, C.c_void_p]
        L.garb200_kkt_error_noalloc.argtypes = [C.c_void_p] * 9 + [C.c_double, C.c_double, C.c_void_p, C.c_void_p]
        L.garb200_pack_from_device.argtypes = [C.c_void_p, C.c_void_p, C.c_double]
        L.garb200_linear_step.argtypes = [C.c_void_p, C.c_double] + [C.c_void_p] * 8
        L.garb200_rollout_device.argtypes = [C.c_void_p, C.c_double] + [C.c_void_p] * 4
        L.garb200_device_theta.argtypes = [C.c_void_p, C.c_void_p]
        L.garb200_upload_gv.argtypes = [C.c_void_p, C.c_void_p]
        L.garb200_multi_create.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
        L.garb200_multi_destroy.argtypes = [C.c_void_p]
        L.garb200_multi_destroy.restype = None
        L.garb200_multi_num_shards.argtypes = [C.c_void_p]
        L.garb200_multi_shard.argtypes = [C.c_void_p, C.c_int] + [C.c_void_p] * 4
        L.garb200_multi_upload_packed.argtypes = [C.c_void_p] * 4
        L.garb200_multi_generate.argtypes = [C.c_void_p, C.c_uint64, C.c_int]
        L.garb200_multi_backward.argtypes = [C.c_void_p, C.c_double, C.c_double]
        L.garb200_multi_forward.argtypes = [C.c_void_p, C.c_void_p, C.c_int] + [C.c_void_p] * 4
        L.garb200_multi_synchronize.argtypes = [C.c_void_p]
        L.garb200_multi_kkt_error.argtypes = [C.c_void_p, C.c_double, C.c_double, C.c_void_p, C.c_void_p]
        L.garb200_set_tuning.argtypes = [C.c_void_p, C.c_char_p, C.c_int]
        L.garb200_synchronize.argtypes = [C.c_void_p]
        L.garb200_destroy.argtypes = [C.c_void_p]
        L.garb200_get_gains.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        L.garb200_download_gains.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        L.garb200_get_value.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        L.garb200_get_initial.argtypes = [C.c_void_p, C.c_int] + [C.c_void_p] * 4
        L.garb200_get_flags.argtypes = [C.c_void_p, C.c_void_p]
        L.garb200_get_refinement_steps.argtypes = [C.c_void_p, C.c_void_p]
        L.garb200_collapse_feedback.argtypes = [C.c_void_p]
        L.garb200_cycle_append.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
        L.garb200_download_packed.argtypes = [C.c_void_p, C.c_int, C.c_int] + [C.c_void_p] * 3
        L.garb200_device_problem.argtypes = [C.c_void_p] * 4
        L.garb200_device_trajectory.argtypes = [C.c_void_p] * 5
        L.garb200_device_gains.argtypes = [C.c_void_p, C.c_void_p]
        L.garb200_stage_knot.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        L.garb200_stage_initial.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        L.garb200_commit.argtypes = [C.c_void_p]
        L.garb200_kernel_launches.argtypes = [C.c_void_p]
        L.garb200_last_timings.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.garb200_fp64_peak.argtypes = [C.c_int, C.c_int, C.c_void_p]
        L.garb200_used_fast_path.argtypes = [C.c_void_p]
        L.garb200_used_explicit_path.argtypes = [C.c_void_p]
        L.garb200_padded_shape.argtypes = [C.c_void_p, C.c_void_p]
        L.garb200_phase_cycles.argtypes = [C.c_void_p, C.c_void_p]
        L.garb200_timing_log_start.argtypes = [C.c_void_p, C.c_int]
        L.garb200_timing_log_read.argtypes = [C.c_void_p] * 4
        L.garb200_create.argtypes = [C.c_void_p, C.c_void_p]
        L.garb200_upload_trajectory.argtypes = [C.c_void_p] * 5
        L.garb200_get_layouts.argtypes = [C.c_void_p] * 4
        _LIB = L
    return _LIB


def check(rc):
    if rc != OK:
        raise GarB200Error(rc, lib().garb200_last_error().decode())


def _ptr(a):
    """numpy array / int device address / None -> c_void_p"""
    if a is None:
        return None
    if isinstance(a, np.ndarray):
        assert a.flags["C_CONTIGUOUS"], "array must be C-contiguous"
        return a.ctypes.data_as(C.c_void_p)
    return C.c_void_p(int(a))


def c_layouts(nx, nu, nc, nth):
    k = np.zeros(17, np.int32); g = np.zeros(5, np.int32); v = np.zeros(8, np.int32)
    lib().garb200_layouts(nx, nu, nc, nth, k.ctypes.data_as(_ip), g.ctypes.data_as(_ip),
                          v.ctypes.data_as(_ip))
    return k, g, v


def fp64_peak(device=0, kind=0) -> float:
    out = C.c_double(0.0)
    check(lib().garb200_fp64_peak(device, kind, C.byref(out)))
    return out.value


NOALLOC_SCRATCH_BYTES = 1 << 20


def kkt_error_noalloc(nx, nu, nc, nth, horizon, nc0, batch, d_knots, d_G0, d_g0, d_xs, d_us, d_vs, d_lbdas,
                      mudyn, mueq, d_theta=None, nu_t=None, nc_t=None, nu_term=0, nc_term=0, device=0,
                      stream=None):
    """garb200_kkt_error_noalloc: lqrComputeKktError on DEVICE pointers, no handle, no allocation
    (static device scratch of NOALLOC_SCRATCH_BYTES).  -> host array [batch][3]."""
    cfg = Config()
    cfg.nx, cfg.nu, cfg.nc, cfg.nth = nx, nu, nc, nth
    cfg.horizon, cfg.nc0, cfg.batch, cfg.device = horizon, nc0, batch, device
    a = None if nu_t is None else np.ascontiguousarray(nu_t, dtype=np.int32)
    b = None if nc_t is None else np.ascontiguousarray(nc_t, dtype=np.int32)
    cfg.nu_t = None if a is None else a.ctypes.data_as(_ip)
    cfg.nc_t = None if b is None else b.ctypes.data_as(_ip)
    cfg.nu_term, cfg.nc_term = nu_term, nc_term
    out = np.zeros((batch, 3))
    check(lib().garb200_kkt_error_noalloc(C.byref(cfg), _ptr(d_knots), _ptr(d_G0), _ptr(d_g0), _ptr(d_xs),
                                          _ptr(d_us), _ptr(d_vs), _ptr(d_lbdas), _ptr(d_theta),
                                          float(mudyn), float(mueq), _ptr(stream), _ptr(out)))
    return out


class Solver:
    """Thin RAII wrapper over a ``garb200_solver*`` handle."""

    def __init__(self, nx, nu, nc=0, nth=0, horizon=1, nc0=None, batch=1, num_legs=0, device=0,
                 flags=0, nu_t=None, nc_t=None, nu_term=0, nc_term=0):
        self._h = None
        cfg = Config()
        cfg.nx, cfg.nu, cfg.nc, cfg.nth = nx, nu, nc, nth
        cfg.horizon, cfg.nc0, cfg.batch = horizon, (nx if nc0 is None else nc0), batch
        cfg.num_legs, cfg.device, cfg.flags = num_legs, device, flags
        self._nu_t = None if nu_t is None else np.ascontiguousarray(nu_t, dtype=np.int32)
        self._nc_t = None if nc_t is None else np.ascontiguousarray(nc_t, dtype=np.int32)
        cfg.nu_t = None if self._nu_t is None else self._nu_t.ctypes.data_as(_ip)
        cfg.nc_t = None if self._nc_t is None else self._nc_t.ctypes.data_as(_ip)
        cfg.nu_term, cfg.nc_term = nu_term, nc_term
        h = C.c_void_p()
        check(lib().garb200_create(C.byref(cfg), C.byref(h)))
        self._h = h
        self.cfg = cfg
        T = horizon
        self.nu_t = (np.array([nu] * T + [nu_term], np.int32) if nu_t is None else self._nu_t)
        self.nc_t = (np.array([nc] * T + [nc_term], np.int32) if nc_t is None else self._nc_t)
        self.nx, self.nu, self.nc, self.nth = nx, nu, nc, nth
        self.horizon, self.nc0, self.batch, self.num_legs = T, cfg.nc0, batch, num_legs
        self.nth_eff = nx if num_legs >= 2 else nth

    @classmethod
    def for_problems(cls, pp, **kw):
        L = pp.layout
        return cls(L.nx, L.nu, L.nc, L.nth, pp.horizon, pp.nc0, pp.batch,
                   nu_t=pp.nu_t, nc_t=pp.nc_t, **kw)

    def close(self):
        if self._h is not None:
            lib().garb200_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- plumbing
    def set_tuning(self, **knobs):
        """diagnostic launcher knobs (garb200_set_tuning), e.g. ``set_tuning(group=1)``"""
        for k, v in knobs.items():
            check(lib().garb200_set_tuning(self._h, k.encode(), int(v)))

    def set_stream(self, stream_ptr):
        check(lib().garb200_set_stream(self._h, C.c_void_p(int(stream_ptr))))

    def synchronize(self):
        check(lib().garb200_synchronize(self._h))

    # ---- problem data
    def upload_packed(self, knots, G0, g0):
        check(lib().garb200_upload_packed(self._h, _ptr(knots), _ptr(G0), _ptr(g0)))

    def upload(self, pp):
        self._keep = (np.ascontiguousarray(pp.knots), np.ascontiguousarray(pp.G0),
                      np.ascontiguousarray(pp.g0))
        self.upload_packed(*self._keep)
        self.synchronize_stream_only()

    def synchronize_stream_only(self):
        rc = lib().garb200_synchronize(self._h)
        if rc not in (OK, ERR_NUMERIC):
            check(rc)

    def generate(self, seed=42, dist=0):
        check(lib().garb200_generate(self._h, C.c_uint64(seed), int(dist)))

    def upload_gv(self, gv):
        """Gv for the dense solver: [batch][T+1][nth][nc] (column-major nc x nth blocks) or None"""
        g = None if gv is None else np.ascontiguousarray(gv, dtype=np.float64)
        check(lib().garb200_upload_gv(self._h, _ptr(g)))

    def download_packed(self, first, count):
        from .layout import KnotLayout
        L = KnotLayout(self.nx, self.nu, self.nc, self.nth)
        knots = np.zeros((count, self.horizon + 1, L.stride))
        G0 = np.zeros((count, self.nc0 * self.nx)); g0 = np.zeros((count, self.nc0))
        check(lib().garb200_download_packed(self._h, first, count, _ptr(knots), _ptr(G0), _ptr(g0)))
        return knots, G0, g0

    def device_problem(self):
        a, b, c = C.c_void_p(), C.c_void_p(), C.c_void_p()
        check(lib().garb200_device_problem(self._h, C.byref(a), C.byref(b), C.byref(c)))
        return a.value, b.value, c.value

    def pack_from_device(self, view: DeviceProblemView, preg=0.0):
        """device-resident producer (updateLQSubproblem, solver-proxddp.hxx:802-873)"""
        check(lib().garb200_pack_from_device(self._h, C.byref(view), float(preg)))

    def pack_from_tensors(self, fields, preg=0.0):
        """Zero-copy producer input from device tensors (torch CUDA tensors, or anything with
        ``__cuda_array_interface__``): ``fields`` maps the knot field names (Q S R q r A B E f C D d Gth Gx Gu gamma,
        plus G0 g0) to float64 arrays of shape [B, T+1, rows, cols] (matrices), [B, T+1, rows] (vectors),
        [B, c0, nx] / [B, c0] (G0 / g0).  Any batch / knot strides are taken as they are; inside a block the
        ROW stride must be 1 (the solver's blocks are column-major: element (i, j) at i + j * ld), i.e. a tensor
        made as ``x.transpose(-1, -2).contiguous().transpose(-1, -2)`` or a [B, T+1, cols, rows] buffer viewed
        transposed; anything else is refused rather than copied.  A missing ``E`` means explicit dynamics (E = -I).
        Returns the view (keep the tensors alive until the stream has run the packing kernel)."""
        view = DeviceProblemView()
        for name, a in fields.items():
            if a is None:
                continue
            if hasattr(a, "data_ptr"):                       # torch
                if str(a.dtype) not in ("torch.float64",) or not a.is_cuda:
                    raise TypeError(f"{name}: float64 CUDA tensor expected")
                ptr, shape, strides = a.data_ptr(), tuple(a.shape), tuple(a.stride())
            else:
                cai = a.__cuda_array_interface__
                if cai["typestr"] not in ("<f8", "=f8"):
                    raise TypeError(f"{name}: float64 expected")
                ptr, shape = cai["data"][0], tuple(cai["shape"])
                st = cai.get("strides")
                if st is None:
                    st, acc = [], 1
                    for d in reversed(shape):
                        st.insert(0, acc); acc *= d
                    strides = tuple(st)
                else:
                    strides = tuple(x // 8 for x in st)
            f = getattr(view, name)
            if name in ("G0", "g0"):
                if name == "G0":
                    if len(shape) != 3 or (shape[1] > 1 and strides[1] != 1):
                        raise ValueError("G0: [B, c0, nx] with unit stride along c0 (column-major block)")
                    f.ptr, f.stride_b, f.stride_t, f.ld = ptr, strides[0], 0, strides[2]
                else:
                    if len(shape) != 2 or (shape[1] > 1 and strides[1] != 1):
                        raise ValueError("g0: [B, c0] contiguous along c0")
                    f.ptr, f.stride_b, f.stride_t, f.ld = ptr, strides[0], 0, max(shape[1], 1)
                continue
            if len(shape) == 4:
                if shape[2] > 1 and strides[2] != 1:
                    raise ValueError(f"{name}: the row stride inside a block must be 1 (column-major blocks); got strides {strides}")
                f.ptr, f.stride_b, f.stride_t, f.ld = ptr, strides[0], strides[1], strides[3] if shape[3] > 1 else max(shape[2], 1)
            elif len(shape) == 3:
                if shape[2] > 1 and strides[2] != 1:
                    raise ValueError(f"{name}: vectors must be contiguous along their last axis")
                f.ptr, f.stride_b, f.stride_t, f.ld = ptr, strides[0], strides[1], max(shape[2], 1)
            else:
                raise ValueError(f"{name}: [B, T+1, rows, cols] or [B, T+1, rows] expected, got {shape}")
        self.pack_from_device(view, preg)
        return view

    def linear_step(self, alpha, cur, trial):
        """trial = cur + alpha * (solver-owned direction); cur / trial: 4-tuples of device pointers"""
        check(lib().garb200_linear_step(self._h, float(alpha), *[_ptr(a) for a in cur], *[_ptr(a) for a in trial]))

    def rollout_device(self, alpha, d_dxs, d_dus, d_dvs, d_dlbdas):
        check(lib().garb200_rollout_device(self._h, float(alpha), _ptr(d_dxs), _ptr(d_dus), _ptr(d_dvs), _ptr(d_dlbdas)))

    def stage_knot(self, problem, t, view: KnotView):
        check(lib().garb200_stage_knot(self._h, problem, t, C.byref(view)))

    def stage_initial(self, problem, G0, g0):
        check(lib().garb200_stage_initial(self._h, problem, _ptr(G0), _ptr(g0)))

    def commit(self):
        check(lib().garb200_commit(self._h))

    # ---- hot path
    def backward(self, mudyn, mueq):
        check(lib().garb200_backward(self._h, float(mudyn), float(mueq)))

    def traj_arrays(self):
        B, T = self.batch, self.horizon
        return (np.zeros((B, T + 1, self.nx)), np.zeros((B, T + 1, self.nu)),
                np.zeros((B, T + 1, self.nc)), np.zeros((B, self.nc0 + T * self.nx)))

    def forward(self, theta=None, out=None):
        xs, us, vs, lb = out if out is not None else self.traj_arrays()
        th = None if theta is None else np.ascontiguousarray(
            np.broadcast_to(np.asarray(theta, dtype=np.float64), (self.batch, self.nth)))
        check(lib().garb200_forward(self._h, _ptr(th), _ptr(xs), _ptr(us), _ptr(vs), _ptr(lb)))
        return xs, us, vs, lb

    def forward_device(self, d_theta=None, d_xs=None, d_us=None, d_vs=None, d_lbdas=None):
        check(lib().garb200_forward_device(self._h, _ptr(d_theta), _ptr(d_xs), _ptr(d_us),
                                           _ptr(d_vs), _ptr(d_lbdas)))

    def device_trajectory(self):
        p = [C.c_void_p() for _ in range(4)]
        check(lib().garb200_device_trajectory(self._h, *[C.byref(q) for q in p]))
        return tuple(q.value for q in p)

    # ---- readers
    def gain_stride(self):
        from .layout import GainLayout
        return GainLayout(self.nx, self.nu, self.nc, self.nth_eff).stride

    def download_gains(self, first=0, count=None):
        count = self.batch - first if count is None else count
        g = np.zeros((count, self.horizon + 1, self.gain_stride()))
        check(lib().garb200_download_gains(self._h, first, count, _ptr(g)))
        return g

    def get_gains(self, problem, t):
        nr = int(self.nu_t[t] + self.nc_t[t] + 2 * self.nx)
        ff = np.zeros(nr); fb = np.zeros((nr, self.nx)); fth = np.zeros((nr, self.nth_eff))
        check(lib().garb200_get_gains(self._h, problem, t, _ptr(ff), _ptr(fb), _ptr(fth)))
        return ff, fb, fth

    def get_value(self, problem, t):
        from .layout import ValueLayout
        V = ValueLayout(self.nx, self.nth_eff)
        rec = np.zeros(V.stride)
        check(lib().garb200_get_value(self._h, problem, t, _ptr(rec)))
        return V.unpack(rec)

    def get_initial(self, problem):
        s0, h = self.nx + self.nc0, self.nth
        ff = np.zeros(s0); fth = np.zeros((s0, h)); g = np.zeros(h); H = np.zeros((h, h))
        check(lib().garb200_get_initial(self._h, problem, _ptr(ff), _ptr(fth), _ptr(g), _ptr(H)))
        return ff, fth, g, H.T.copy()

    def get_flags(self):
        f = np.zeros(self.batch, np.int32)
        check(lib().garb200_get_flags(self._h, _ptr(f)))
        return f

    def get_refinement_steps(self):
        f = np.zeros(self.batch, np.int32)
        check(lib().garb200_get_refinement_steps(self._h, _ptr(f)))
        return f

    def collapse_feedback(self):
        check(lib().garb200_collapse_feedback(self._h))

    def cycle_append(self, views, shared=False):
        arr = (KnotView * len(views))(*views)
        check(lib().garb200_cycle_append(self._h, arr, int(shared)))

    def kkt_error(self, mudyn, mueq, theta=None):
        out = np.zeros((self.batch, 3))
        th = None if theta is None else np.ascontiguousarray(
            np.broadcast_to(np.asarray(theta, dtype=np.float64), (self.batch, self.nth)))
        check(lib().garb200_kkt_error(self._h, float(mudyn), float(mueq), _ptr(th), _ptr(out)))
        return out

    def upload_trajectory(self, xs, us, vs, lbdas):
        check(lib().garb200_upload_trajectory(self._h, _ptr(xs), _ptr(us), _ptr(vs), _ptr(lbdas)))

    def kernel_launches(self) -> int:
        return int(lib().garb200_kernel_launches(self._h))

    def used_fast_path(self) -> bool:
        return bool(lib().garb200_used_fast_path(self._h))

    def used_explicit_path(self) -> bool:
        return bool(lib().garb200_used_explicit_path(self._h))

    def padded_shape(self):
        """(nx, nu, nc, nth) of the instantiated sweep this problem shape is embedded in (exact zero /
        identity padding, csrc/pad.cu), or None if the shape runs natively / on the generic kernel."""
        out = (C.c_int * 4)()
        return tuple(out) if lib().garb200_padded_shape(self._h, out) else None

    def phase_cycles(self):
        out = np.zeros(16, np.int64)
        check(lib().garb200_phase_cycles(self._h, _ptr(out)))
        return out

    def timing_log_start(self, capacity):
        check(lib().garb200_timing_log_start(self._h, int(capacity)))
        self._tcap = int(capacity)

    def timing_log_read(self):
        """-> (backward_ms[count], forward_ms[count]) device-event durations of the logged calls."""
        b = np.zeros(self._tcap, np.float32); f = np.zeros(self._tcap, np.float32)
        n = C.c_int(0)
        check(lib().garb200_timing_log_read(self._h, _ptr(b), _ptr(f), C.byref(n)))
        return b[: n.value].astype(np.float64), f[: n.value].astype(np.float64)

    def last_timings(self):
        a, b = C.c_float(0), C.c_float(0)
        check(lib().garb200_last_timings(self._h, C.byref(a), C.byref(b)))
        return a.value, b.value


class MultiSolver:
    """garb200_multi_*: one logical batched solver over several GPUs of one box (contiguous batch shards,
    no communication inside the solve, gather by device->host or peer copies)."""

    def __init__(self, nx, nu, nc=0, nth=0, horizon=1, nc0=None, batch=1, devices=(0,), flags=0, nu_term=0, nc_term=0):
        self._h = None
        cfg = Config()
        cfg.nx, cfg.nu, cfg.nc, cfg.nth = nx, nu, nc, nth
        cfg.horizon, cfg.nc0, cfg.batch, cfg.flags = horizon, (nx if nc0 is None else nc0), batch, flags
        cfg.nu_term, cfg.nc_term = nu_term, nc_term
        dev = np.ascontiguousarray(devices, dtype=np.int32)
        h = C.c_void_p()
        check(lib().garb200_multi_create(C.byref(cfg), dev.ctypes.data_as(_ip), len(dev), C.byref(h)))
        self._h = h
        self.nx, self.nu, self.nc, self.nth, self.horizon, self.nc0, self.batch = nx, nu, nc, nth, horizon, cfg.nc0, batch

    def close(self):
        if self._h is not None:
            lib().garb200_multi_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def shards(self):
        out = []
        for i in range(lib().garb200_multi_num_shards(self._h)):
            s, d, f, c = C.c_void_p(), C.c_int(), C.c_int(), C.c_int()
            check(lib().garb200_multi_shard(self._h, i, C.byref(s), C.byref(d), C.byref(f), C.byref(c)))
            out.append((d.value, f.value, c.value))
        return out

    def upload(self, pp):
        self._keep = (np.ascontiguousarray(pp.knots), np.ascontiguousarray(pp.G0), np.ascontiguousarray(pp.g0))
        check(lib().garb200_multi_upload_packed(self._h, *[_ptr(a) for a in self._keep]))

    def generate(self, seed=42, dist=0):
        check(lib().garb200_multi_generate(self._h, C.c_uint64(seed), int(dist)))

    def backward(self, mudyn, mueq):
        check(lib().garb200_multi_backward(self._h, float(mudyn), float(mueq)))

    def forward(self, theta=None):
        B, T = self.batch, self.horizon
        xs, us = np.zeros((B, T + 1, self.nx)), np.zeros((B, T + 1, self.nu))
        vs, lb = np.zeros((B, T + 1, self.nc)), np.zeros((B, self.nc0 + T * self.nx))
        th = None if theta is None else np.ascontiguousarray(np.broadcast_to(np.asarray(theta, dtype=np.float64), (B, self.nth)))
        check(lib().garb200_multi_forward(self._h, _ptr(th), -1, _ptr(xs), _ptr(us), _ptr(vs), _ptr(lb)))
        self.synchronize()
        return xs, us, vs, lb

    def forward_to_device(self, root_device, d_xs, d_us, d_vs, d_lbdas, theta=None):
        th = None if theta is None else np.ascontiguousarray(np.broadcast_to(np.asarray(theta, dtype=np.float64), (self.batch, self.nth)))
        self._th_keep = th
        check(lib().garb200_multi_forward(self._h, _ptr(th), int(root_device), _ptr(d_xs), _ptr(d_us), _ptr(d_vs), _ptr(d_lbdas)))

    def synchronize(self):
        check(lib().garb200_multi_synchronize(self._h))

    def kkt_error(self, mudyn, mueq, theta=None):
        out = np.zeros((self.batch, 3))
        th = None if theta is None else np.ascontiguousarray(np.broadcast_to(np.asarray(theta, dtype=np.float64), (self.batch, self.nth)))
        check(lib().garb200_multi_kkt_error(self._h, float(mudyn), float(mueq), _ptr(th), _ptr(out)))
        return out
