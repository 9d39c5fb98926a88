"""Host-side mirror of the reference's ``aligator.gar`` surface, backed by the CUDA C-ABI.

Same names, argument meaning and error behaviour as the reference bindings
(/root/reference/bindings/python/src/gar/expose-gar.cpp:55-130, expose-prox-riccati.cpp:14-56,
expose-parallel.cpp:16-24, expose-utils.cpp:33-53) so the parity tests read like
/root/reference/tests/gar/*.cpp.  One solver object drives ONE problem here (batch = 1
through the same kernels); batched use goes through ``capi.Solver`` with packed arrays.

Nothing in this module computes on the CPU: backward / forward / KKT residuals all run
the sm_100a kernels; numpy is only used to hold the caller's matrices.
"""
from __future__ import annotations

import numpy as np

from . import capi
from .layout import GainLayout, KnotLayout
from .synth import PackedProblems


class LqrKnot:
    """lqr-problem.hpp:32-196.  Cost 1/2 [x;u]^T [Q S; S^T R] [x;u] + q^T x + r^T u,
    dynamics E x' + A x + B u + f = 0, constraint C x + D u + d = 0."""

    _MATS = ("Q", "S", "R", "q", "r", "A", "B", "E", "f", "C", "D", "d", "Gth", "Gx", "Gu", "Gv", "gamma")

    def __init__(self, nx, nu, nc, nx2=None, nth=0):
        nx2 = nx if nx2 is None else nx2
        if nx2 != nx:
            raise ValueError("garb200 supports nx2 == nx only")
        self.nx, self.nu, self.nc, self.nx2, self.nth = nx, nu, nc, nx2, nth
        z = np.zeros
        self.Q, self.S, self.R = z((nx, nx)), z((nx, nu)), z((nu, nu))
        self.q, self.r = z(nx), z(nu)
        self.A, self.B, self.E, self.f = z((nx, nx)), z((nx, nu)), z((nx, nx)), z(nx)
        self.C, self.D, self.d = z((nc, nx)), z((nc, nu)), z(nc)
        self._alloc_param(nth)

    def _alloc_param(self, nth):
        z = np.zeros
        self.nth = nth
        self.Gth, self.Gx, self.Gu = z((nth, nth)), z((self.nx, nth)), z((self.nu, nth))
        self.Gv, self.gamma = z((self.nc, nth)), z(nth)

    def addParameterization(self, nth):        # lqr-problem.hxx:215-223
        self._alloc_param(nth)
        return self

    def copy(self):
        k = LqrKnot(self.nx, self.nu, self.nc, self.nx2, self.nth)
        for n in self._MATS:
            setattr(k, n, np.array(getattr(self, n), dtype=np.float64, copy=True))
        return k

    def isApprox(self, other, prec=np.finfo(float).eps):
        return all(np.allclose(getattr(self, n), getattr(other, n), rtol=prec, atol=0) for n in self._MATS)

    def fields(self):
        return {n: np.asarray(getattr(self, n), dtype=np.float64) for n in self._MATS if n != "Gv"}


class LqrProblem:
    """lqr-problem.hpp:198-289."""

    def __init__(self, stages, nc0):
        self.stages = [k.copy() for k in stages]   # "constructor creates a copy"
        nx0 = self.stages[0].nx if self.stages else 0
        self.G0 = np.zeros((nc0, nx0))
        self.g0 = np.zeros(nc0)

    @property
    def horizon(self):
        return len(self.stages) - 1

    @property
    def nc0(self):
        return int(np.asarray(self.g0).shape[0])

    @property
    def ntheta(self):
        return self.stages[0].nth

    @property
    def isInitialized(self):
        return len(self.stages) > 0

    @property
    def isParameterized(self):
        return self.isInitialized and self.stages[0].nth > 0

    def addParameterization(self, nth):
        for k in self.stages:
            k.addParameterization(nth)

    def pack(self) -> PackedProblems:
        st = self.stages
        nx = st[0].nx
        if any(k.nx != nx for k in st):
            raise ValueError("garb200 needs the same nx at every knot")
        nth = st[0].nth
        if any(k.nth != nth for k in st):
            raise ValueError("garb200 needs the same nth at every knot")
        nu, nc = max(k.nu for k in st), max(k.nc for k in st)
        L = KnotLayout(nx, nu, nc, nth)
        T = self.horizon
        knots = np.zeros((1, T + 1, L.stride))
        for t, k in enumerate(st):
            L.pack(knots[0, t], k.fields(), nu=k.nu, nc=k.nc)
        G0 = np.asarray(self.G0, dtype=np.float64).reshape(self.nc0, nx)
        return PackedProblems(L, T, self.nc0,
                              np.array([k.nu for k in st], np.int32),
                              np.array([k.nc for k in st], np.int32),
                              knots, G0.reshape(1, -1, order="F").copy(),
                              np.asarray(self.g0, dtype=np.float64).reshape(1, -1).copy())


def lqrInitializeSolution(problem: LqrProblem):
    """utils.hpp:139-168."""
    st = problem.stages
    N = problem.horizon
    xs = [np.zeros(k.nx) for k in st]
    us = [np.zeros(k.nu) for k in st]
    vs = [np.zeros(k.nc) for k in st]
    lbdas = [np.zeros(problem.nc0)] + [np.zeros(st[i].nx2) for i in range(N)]
    if st[-1].nu == 0:
        us.pop()
    return xs, us, vs, lbdas


def lqrNumRows(problem: LqrProblem) -> int:
    """utils.hpp (lqrNumRows): nc0 + sum_t (nx + nu + nc) + sum_{t<N} nx2."""
    st = problem.stages
    return problem.nc0 + sum(k.nx + k.nu + k.nc for k in st) + sum(k.nx2 for k in st[:-1])


def lqrDenseMatrix(problem: LqrProblem, mudyn: float, mueq: float):
    """utils.hxx:194-267 / expose-utils.cpp:33-53: the dense KKT system ``mat z = -rhs`` of the whole problem,
    unknowns ordered [lbd0 | x0 u0 v0 lbd1 | x1 u1 v1 lbd2 | ...].  Pure assembly of the caller's matrices
    (host utility, no arithmetic beyond copies): what ``examples/gar-elqr.py`` feeds a dense LDLT."""
    st, N = problem.stages, problem.horizon
    n = lqrNumRows(problem)
    mat, rhs = np.zeros((n, n)), np.zeros(n)
    nc0, nx0 = problem.nc0, st[0].nx
    G0 = np.asarray(problem.G0, dtype=np.float64).reshape(nc0, nx0)
    mat[nc0:nc0 + nx0, :nc0] = G0.T
    mat[:nc0, nc0:nc0 + nx0] = G0
    mat[np.arange(nc0), np.arange(nc0)] = -mudyn
    rhs[:nc0] = problem.g0
    idx = nc0
    for t, k in enumerate(st):
        nx, nu, nc = k.nx, k.nu, k.nc
        nb = nx + nu + nc
        blk = mat[idx:idx + nb, idx:idx + nb]
        blk[:nx, :nx] = k.Q
        blk[nx:nx + nu, :nx] = np.asarray(k.S).T
        blk[nx:nx + nu, nx:nx + nu] = k.R
        blk[nx + nu:, :nx] = k.C
        blk[nx + nu:, nx:nx + nu] = k.D
        blk[nx + nu + np.arange(nc), nx + nu + np.arange(nc)] = -mueq
        low = np.tril(blk)                                # selfadjointView<Lower>
        blk[:] = low + np.tril(blk, -1).T
        rhs[idx:idx + nx] = k.q
        rhs[idx + nx:idx + nx + nu] = k.r
        rhs[idx + nx + nu:idx + nb] = k.d
        if t != N:
            r0 = idx + nb
            row = np.zeros((nx, nb + 2 * nx))             # [A B 0 -mu I E]
            row[:, :nx] = k.A
            row[:, nx:nx + nu] = k.B
            row[np.arange(nx), nb + np.arange(nx)] = -mudyn
            row[:, nb + nx:] = k.E
            mat[r0:r0 + nx, idx:idx + nb + 2 * nx] = row
            mat[idx:idx + nb + 2 * nx, r0:r0 + nx] = row.T
            rhs[r0:r0 + k.nx2] = k.f
            idx += nb + k.nx2
    return mat, rhs


def lqrDenseSolutionToTraj(problem: LqrProblem, solution):
    """utils.hpp:106-137."""
    st, N = problem.stages, problem.horizon
    nc0 = problem.nc0
    z = np.asarray(solution, dtype=np.float64)
    xs, us, vs, lbdas = [], [], [], [z[:nc0].copy()]
    idx = nc0
    for t, k in enumerate(st):
        xs.append(z[idx:idx + k.nx].copy()); idx += k.nx
        us.append(z[idx:idx + k.nu].copy()); idx += k.nu
        vs.append(z[idx:idx + k.nc].copy()); idx += k.nc
        if t < N:
            lbdas.append(z[idx:idx + k.nx2].copy()); idx += k.nx2
    return xs, us, vs, lbdas


class RiccatiSolverBase:
    """riccati-base.hpp:12-34 — the drop-in boundary."""

    def backward(self, mudyn, mueq):  # pragma: no cover - interface
        raise NotImplementedError

    def forward(self, xs, us, vs, lbdas, theta=None):  # pragma: no cover - interface
        raise NotImplementedError

    def cycleAppend(self, knot):  # pragma: no cover - interface
        raise NotImplementedError

    def collapseFeedback(self):
        pass

    def getFeedforward(self, i):  # pragma: no cover - interface
        raise NotImplementedError

    def getFeedback(self, i):  # pragma: no cover - interface
        raise NotImplementedError


class _GpuSolver(RiccatiSolverBase):
    def __init__(self, problem: LqrProblem, num_legs=0, device=0, store_all_values=True, extra_flags=0):
        self.problem_ = problem
        pp = problem.pack()
        self._pp = pp
        # datas[t].vm of every knot is part of the serial solver's public state (proximal-riccati.hpp:18);
        # the parallel-in-time solver only needs the leg heads, and keeping every value function would
        # take it off the specialised leg sweep
        self._solver = capi.Solver.for_problems(pp, num_legs=num_legs, device=device,
                                                flags=(capi.STORE_ALL_VALUES if store_all_values else 0) | extra_flags)

    def _upload(self):
        # the reference re-reads problem_->stages on every backward() (proximal-riccati.hxx:37)
        self._pp = self.problem_.pack()
        self._solver.upload(self._pp)

    def backward(self, mudyn, mueq):
        self._upload()
        self._solver.backward(mudyn, mueq)
        try:
            self._solver.synchronize()
        except capi.GarB200Error as e:
            if e.code == capi.ERR_NUMERIC:
                return False
            raise
        return True

    def _scatter(self, xs, us, vs, lbdas, X, U, V, Lb):
        st = self.problem_.stages
        N, nx, nc0 = self.problem_.horizon, st[0].nx, self.problem_.nc0
        for t, k in enumerate(st):
            xs[t][:] = X[0, t]
            if t < len(us):
                us[t][:] = U[0, t, :k.nu]
            vs[t][:] = V[0, t, :k.nc]
        lbdas[0][:] = Lb[0, :nc0]
        for t in range(1, N + 1):
            lbdas[t][:] = Lb[0, nc0 + (t - 1) * nx: nc0 + t * nx]

    def getFeedforward(self, i):
        return self._solver.get_gains(0, i)[0]

    def getFeedback(self, i):
        return self._solver.get_gains(0, i)[1]

    def getParamFeedback(self, i):
        return self._solver.get_gains(0, i)[2]

    def value(self, i):
        """datas[i].vm (riccati-kernel.hpp:39-60)."""
        return self._solver.get_value(0, i)

    def kktError(self, mudyn, mueq, theta=None):
        return self._solver.kkt_error(mudyn, mueq, theta)[0]


class ProximalRiccatiSolver(_GpuSolver):
    """proximal-riccati.hpp:12-45, GPU-backed."""

    def __init__(self, problem: LqrProblem, device=0):
        super().__init__(problem, 0, device)

    def forward(self, xs, us, vs, lbdas, theta=None):
        X, U, V, Lb = self._solver.forward(theta)
        self._scatter(xs, us, vs, lbdas, X, U, V, Lb)
        return True

    def cycleAppend(self, knot: LqrKnot):
        pk = KnotLayout(self._pp.layout.nx, knot.nu, knot.nc, knot.nth)
        f = {n: np.ascontiguousarray(np.asarray(v, dtype=np.float64).reshape(-1, order="F"))
             for n, v in knot.fields().items()}
        view = capi.KnotView()
        view.nu, view.nc = knot.nu, knot.nc
        self._keep = f
        for n, a in f.items():
            setattr(view, n, a.ctypes.data_as(capi._dp) if a.size else None)
        self._solver.cycle_append([view], shared=True)

    @property
    def kkt0(self):
        ff, fth, _, _ = self._solver.get_initial(0)
        return {"ff": ff, "fth": fth}

    @property
    def thGrad(self):
        return self._solver.get_initial(0)[2]

    @property
    def thHess(self):
        return self._solver.get_initial(0)[3]


class RiccatiSolverDense(ProximalRiccatiSolver):
    """dense-riccati.hpp:28-58, GPU-backed (csrc/dense.cu): one Bunch-Kaufman factorization of the
    (nu + nc + 2 nx)^2 stage KKT matrix per knot (dense-kernel.hpp:56-212) instead of the proximal sweep.
    Same constructor, virtuals and gains layout; unlike the proximal kernel it reads Gv
    (dense-kernel.hpp:76,134 vs riccati-kernel.hxx:128,272)."""

    def __init__(self, problem: LqrProblem, device=0):
        _GpuSolver.__init__(self, problem, 0, device, store_all_values=True, extra_flags=capi.DENSE_KERNEL)

    def _upload(self):
        super()._upload()
        st = self.problem_.stages
        nc, nth = max(k.nc for k in st), st[0].nth
        if nc * nth:
            gv = np.zeros((1, len(st), nth, nc))
            for t, k in enumerate(st):
                gv[0, t, :, :k.nc] = np.asarray(k.Gv, dtype=np.float64).reshape(k.nc, nth).T
            self._solver.upload_gv(gv)


class ParallelRiccatiSolver(_GpuSolver):
    """parallel-solver.hpp:24-108, GPU-backed: legs are CTAs instead of OpenMP threads."""

    def __init__(self, problem: LqrProblem, num_threads: int, device=0):
        if num_threads < 2:   # ALIGATOR_RUNTIME_ERROR, parallel-solver.hxx:25-29
            raise RuntimeError(f"(ParallelRiccatiSolver) num_threads (={num_threads}) should be "
                               "greater than or equal to 2.")
        self.numThreads = num_threads
        super().__init__(problem, num_threads, device, store_all_values=False)

    def forward(self, xs, us, vs, lbdas, theta=None):   # theta ignored (parallel-solver.hxx:184)
        X, U, V, Lb = self._solver.forward(None)
        self._scatter(xs, us, vs, lbdas, X, U, V, Lb)
        return True

    def collapseFeedback(self):
        self._solver.collapse_feedback()

    def refinementSteps(self):
        return int(self._solver.get_refinement_steps()[0])


def lqrComputeKktError(problem: LqrProblem, xs, us, vs, lbdas, mudyn, mueq, theta=None,
                       verbose=False, device=0):
    """utils.hxx:94-192 on the GPU: uploads the trajectory next to the problem and runs the
    residual kernel.  Returns [dyn, cstr, dual]."""
    pp = problem.pack()
    s = capi.Solver.for_problems(pp, device=device)
    s.upload(pp)
    L = pp.layout
    T, nx, nc0 = pp.horizon, L.nx, pp.nc0
    X, U, V, Lb = s.traj_arrays()
    for t, k in enumerate(problem.stages):
        X[0, t] = xs[t]
        if t < len(us):
            U[0, t, :k.nu] = us[t]
        V[0, t, :k.nc] = vs[t]
    Lb[0, :nc0] = lbdas[0]
    for t in range(1, T + 1):
        Lb[0, nc0 + (t - 1) * nx: nc0 + t * nx] = lbdas[t]
    s.upload_trajectory(X, U, V, Lb)
    return s.kkt_error(mudyn, mueq, theta)[0]
