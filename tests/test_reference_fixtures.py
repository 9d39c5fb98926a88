"""The reference's own deterministic inputs for this path, reproduced (VERDICT r1 "missing" #7, "next" 2(c,d)):

* ``short_problem`` (/root/reference/tests/gar/cholmod.cpp:45-72): the ONLY fully RNG-free fixture the
  reference holds — every entry a constant, E = -I, G0 = -3.14, g0 = 42.  The reference uses it to pin the
  KKT-matrix assembly (cholmod.cpp:76-107, mudyn = 1e-8, mueq = 1e-6); its stage KKT is singular (R and B
  are constant-filled, rank 1), so it is a known-answer test of ``lqrDenseMatrix`` and not of a solve.
  The expected 18 x 18 matrix below is written out by hand from utils.hxx:194-267.
* ``randomlyModifyProblem`` x 10 then re-solve (tests/gar/parallel.cpp:186-197, 262-270).
* an independent-precision yardstick (oracle/extended_precision.py) for oracle and GPU.
"""
import numpy as np
import pytest

from aligator_swk_b200.synth import generate_problems
from conftest import relerr
from extended_precision import refined_solve


def short_problem(horz=1, nx=3, nu=2, nc=1, nc0=3):
    """tests/gar/cholmod.cpp:45-72 (x0 only fixes nc0 = x0.size(): its values are never read)."""
    from aligator_swk_b200 import gar
    knots = []
    for _ in range(horz + 1):
        k = gar.LqrKnot(nx, nu, nc)
        k.Q[:] = +0.11; k.R[:] = +0.12; k.S[:] = -0.2; k.q[:] = 13.0; k.r[:] = -13.0
        k.C[:] = -2.0; k.D[:] = -3.0; k.d[:] = -4.0
        k.f[:] = -0.3; k.A[:] = 0.2; k.B[:] = 0.3; k.E[:] = -np.eye(nx)
        knots.append(k)
    prob = gar.LqrProblem(knots, nc0)
    prob.G0[:] = -3.14
    prob.g0[:] = 42.0
    return prob


def short_problem_expected(mudyn, mueq):
    """Hand-written KKT matrix / rhs of short_problem(horz = 1, nx = 3, nu = 2, nc = 1, nc0 = 3),
    unknowns [lbd0(3) | x0(3) u0(2) v0(1) lbd1(3) | x1(3) u1(2) v1(1)]  (utils.hxx:194-267)."""
    l0, x0, u0, v0, l1, x1, u1, v1 = (slice(0, 3), slice(3, 6), slice(6, 8), slice(8, 9), slice(9, 12),
                                      slice(12, 15), slice(15, 17), slice(17, 18))
    K, r = np.zeros((18, 18)), np.zeros(18)          # entries are written on / below the diagonal only
    K[l0, l0] = -mudyn * np.eye(3); K[x0, l0] = -3.14                      # G0^T  (G0 x0 + g0 - mudyn lbd0)
    for x, u, v in ((x0, u0, v0), (x1, u1, v1)):
        K[x, x] = 0.11; K[u, x] = -0.2; K[u, u] = 0.12                     # [Q S; S^T R]
        K[v, x] = -2.0; K[v, u] = -3.0; K[v, v] = -mueq                    # C, D, -mueq
        r[x] = 13.0; r[u] = -13.0; r[v] = -4.0
    K[l1, x0] = 0.2; K[l1, u0] = 0.3; K[l1, l1] = -mudyn * np.eye(3)       # A, B, -mudyn I
    K[x1, l1] = -np.eye(3)                                                 # E^T = -I
    r[l0] = 42.0; r[l1] = -0.3
    K = np.tril(K)
    K = K + np.tril(K, -1).T
    return K, r


def test_short_problem_dense_kkt_known_answer(oracle):
    from aligator_swk_b200 import gar
    prob = short_problem()
    Kexp, rexp = short_problem_expected(1e-8, 1e-6)
    assert np.array_equal(Kexp, Kexp.T) and gar.lqrNumRows(prob) == 18
    K, r = gar.lqrDenseMatrix(prob, 1e-8, 1e-6)                 # host mirror (expose-utils.cpp:33-53)
    assert np.array_equal(K, Kexp) and np.array_equal(r, rexp)
    Ko, ro = oracle.dense_kkt(prob.pack(), 1e-8, 1e-6)          # the oracle's restatement
    assert np.array_equal(Ko, Kexp) and np.array_equal(ro, rexp)
    # the reference then modifies g0, Q0, R0 and rebuilds (cholmod.cpp:98-105): same three-way agreement
    rng = np.random.default_rng(0)
    prob.g0[:] = rng.uniform(-1, 1, 3)
    W = rng.standard_normal((3, 4)); prob.stages[0].Q[:] = W @ W.T
    W = rng.standard_normal((2, 3)); prob.stages[0].R[:] = W @ W.T
    K2, r2 = gar.lqrDenseMatrix(prob, 1e-8, 1e-6)
    Ko2, ro2 = oracle.dense_kkt(prob.pack(), 1e-8, 1e-6)
    assert np.array_equal(K2, Ko2) and np.array_equal(r2, ro2) and not np.array_equal(K2, Kexp)
    # the stage KKT of this fixture is singular (rank-1 R and B): it is an assembly fixture only
    assert np.linalg.matrix_rank(Kexp) < 18


def test_dense_matrix_mirror_matches_oracle_on_ragged_parametric_free_problem(oracle):
    from test_oracle import short_horz_pb
    from aligator_swk_b200 import gar
    prob = short_horz_pb()
    K, r = gar.lqrDenseMatrix(prob, 1e-9, 1e-7)
    Ko, ro = oracle.dense_kkt(prob.pack(), 1e-9, 1e-7)
    assert np.array_equal(K, Ko) and np.array_equal(r, ro)
    z = np.linalg.solve(K, -r)
    xs, us, vs, lb = gar.lqrDenseSolutionToTraj(prob, z)
    X, U, V, Lb = oracle.dense_solution_to_traj(prob.pack(), z)
    assert np.array_equal(np.array(xs), X) and np.array_equal(np.concatenate(lb), Lb)


def _refined_traj(oracle, pp, mu, b=0):
    mat, rhs = oracle.dense_kkt(pp, mu, mu, b)
    z, hist = refined_solve(mat, rhs)
    return oracle.dense_solution_to_traj(pp, z), hist


@pytest.mark.parametrize("dims", [(12, 4, 0, 50, 1e-11), (36, 12, 0, 100, 1e-11), (36, 12, 6, 40, 1e-9)],
                         ids=lambda d: "nx%d_nu%d_nc%d_T%d_mu%g" % d)
def test_oracle_against_extended_precision_reference(oracle, dims):
    """The CPU port vs the refined dense solve on the bench distribution (E ~ 1000 N(0,1), singular Q
    blocks) at the BASELINE shapes and the bench mu."""
    nx, nu, nc, T, mu = dims
    pp = generate_problems(2, T, nx, nu, nc, seed=nx + T)
    for b in range(2):
        (xs, us, vs, lb), hist = _refined_traj(oracle, pp, mu, b)
        assert min(hist) < 1e-17, hist           # backward error far below double precision
        out = oracle.prox_solve(pp, mu, mu, b=b)
        errs = {k: relerr(out[k], ref) for k, ref in (("xs", xs), ("us", us), ("vs", vs), ("lbdas", lb))}
        print("oracle vs extended-precision reference", dims, b, errs)
        assert max(errs.values()) < 1e-11, errs


# ---- GPU ------------------------------------------------------------------------------------------
@pytest.fixture(scope="module")
def capi():
    from aligator_swk_b200 import capi as c
    c.lib()
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    return c


@pytest.mark.gpu
@pytest.mark.parametrize("dims", [(12, 4, 0, 50, 1e-11), (36, 12, 0, 100, 1e-11), (36, 12, 6, 40, 1e-9)],
                         ids=lambda d: "nx%d_nu%d_nc%d_T%d_mu%g" % d)
def test_gpu_against_extended_precision_reference(capi, oracle, dims):
    """Which one moved?  GPU and CPU port both measured against the refined reference: the GPU must be as
    close to the true KKT solution as the port is (within a factor 10, floor 1e-10)."""
    nx, nu, nc, T, mu = dims
    pp = generate_problems(3, T, nx, nu, nc, seed=nx + T + 1)
    s = capi.Solver.for_problems(pp)
    s.upload(pp)
    s.backward(mu, mu)
    X, U, V, Lb = s.forward()
    assert (s.get_flags() == 0).all()
    for b in range(3):
        (xs, us, vs, lb), hist = _refined_traj(oracle, pp, mu, b)
        assert min(hist) < 1e-17, hist
        out = oracle.prox_solve(pp, mu, mu, b=b)
        e_gpu = max(relerr(X[b], xs), relerr(U[b], us), relerr(V[b], vs), relerr(Lb[b], lb))
        e_cpu = max(relerr(out["xs"], xs), relerr(out["us"], us), relerr(out["vs"], vs), relerr(out["lbdas"], lb))
        print(f"nx={nx} b={b}: GPU vs refined {e_gpu:.2e}, CPU port vs refined {e_cpu:.2e}")
        assert e_gpu <= max(10 * e_cpu, 1e-10), (e_gpu, e_cpu)
    s.close()


def _problem_from_packed(gar, pp, b=0):
    L = pp.layout
    knots = []
    for t in range(pp.horizon + 1):
        m, c = int(pp.nu_t[t]), int(pp.nc_t[t])
        k = gar.LqrKnot(L.nx, m, c, nth=L.nth)
        for name, a in L.unpack(pp.knots[b, t], m, c).items():
            dst = getattr(k, name)
            dst[...] = np.asarray(a).reshape(dst.shape)
        knots.append(k)
    prob = gar.LqrProblem(knots, pp.nc0)
    prob.G0[:] = pp.G0[b].reshape(pp.nc0, L.nx, order="F")
    prob.g0[:] = pp.g0[b]
    return prob


def randomly_modify_problem(prob, rng):
    """tests/gar/parallel.cpp:186-197: simulates updating the LQ problem in SQP."""
    N = prob.horizon
    for i in (0, N // 3, N // 2, N // 2 + 1, N // 2 + 2, N):
        kn = prob.stages[i]
        kn.A[:] = rng.standard_normal((kn.nx, kn.nx))
        kn.B[:] = rng.uniform(-1, 1, (kn.nx, kn.nu))
        kn.q[:] = rng.standard_normal(kn.nx)
        kn.R[:] = np.eye(kn.nu)
        kn.S[:] = 0.0


@pytest.mark.gpu
def test_parallel_solver_class_modify_and_resolve(capi, oracle):
    """tests/gar/parallel.cpp:199-271: nx = 32, nu = 12, N = 50, mu = 1e-9; serial == parallel to 1e-10, then
    ten rounds of randomlyModifyProblem + re-solve through the reference-shaped host API (the solver
    re-reads the problem on every backward)."""
    from aligator_swk_b200 import gar
    nx, nu, N, J, mu, tol = 32, 12, 50, 6, 1e-9, 1e-10
    pp = generate_problems(1, N, nx, nu, seed=42, x0=np.zeros(nx))
    prob = _problem_from_packed(gar, pp)
    par = gar.ParallelRiccatiSolver(prob, J)
    rng = np.random.default_rng(7)
    for it in range(11):
        if it:
            randomly_modify_problem(prob, rng)
        assert par.backward(mu, mu)
        xs, us, vs, lbdas = gar.lqrInitializeSolution(prob)
        par.forward(xs, us, vs, lbdas)
        ref = oracle.prox_solve(prob.pack(), mu, mu)                  # the SERIAL solver is the yardstick
        sx, sl = max(np.abs(ref["xs"]).max(), 1.0), max(np.abs(ref["lbdas"]).max(), 1.0)
        assert np.abs(np.array(xs) - ref["xs"]).max() <= tol * sx, it
        assert np.abs(np.concatenate(lbdas) - ref["lbdas"]).max() <= 10 * tol * sl, it
        e = par.kktError(mu, mu)
        assert e.max() <= max(tol, 100 * ref["kkt"].max()), (it, e, ref["kkt"])
