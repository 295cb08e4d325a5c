"""oracle/nlopt_kraft -- the INDEPENDENT restatement of Kraft's SLSQP behind the oracle's NLopt entry points -- against
scipy (scipy.optimize.minimize(method="SLSQP") is Kraft's algorithm as well, scipy.optimize.nnls Lawson & Hanson's).

What is pinned here: the building blocks (NNLS, LDP, the LDL^T rank-one update) and the whole SQP TRAJECTORY: on smooth
problems of the purchasing managers' shape every iterate of nlopt_kraft equals scipy's to ~1e-12, i.e. both are the same
algorithm.  NLopt's driver logic around it (xtol_abs stopping rule, evaluation counting, best-point rule) follows NLopt's
documentation; no NLopt binary is available offline to compare with (DESIGN.md "Parity")."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest
from scipy import optimize as so

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
FUNC = C.CFUNCTYPE(C.c_double, C.c_uint, C.POINTER(C.c_double), C.POINTER(C.c_double), C.c_void_p)
P = C.POINTER(C.c_double)
NLOPT_LD_SLSQP = 40  # position in the enum of oracle/include/nlopt.h (NLopt 2.7's order)


def _arr(a):
    return a.ctypes.data_as(P)


@pytest.fixture(scope="module")
def kraft():
    lib_path = os.path.join(ROOT, "oracle", "_ref", "libnlopt_kraft.so")
    src = [os.path.join(ROOT, "oracle", "nlopt_kraft", f) for f in ("nlopt_kraft.cpp", "slsqp_kraft.cpp", "slsqp_kraft.h")]
    if not os.path.exists(lib_path) or any(os.path.getmtime(s) > os.path.getmtime(lib_path) for s in src):
        subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "oracle"), "_ref/libnlopt_kraft.so"], check=True)
    lib = C.CDLL(lib_path)
    lib.nlopt_create.restype = C.c_void_p
    lib.nlopt_create.argtypes = [C.c_int, C.c_uint]
    lib.nlopt_destroy.argtypes = [C.c_void_p]
    for name in ("nlopt_set_min_objective", "nlopt_set_max_objective"):
        getattr(lib, name).argtypes = [C.c_void_p, FUNC, C.c_void_p]
    for name in ("nlopt_add_equality_constraint", "nlopt_add_inequality_constraint"):
        getattr(lib, name).argtypes = [C.c_void_p, FUNC, C.c_void_p, C.c_double]
    for name in ("nlopt_set_lower_bounds", "nlopt_set_upper_bounds", "nlopt_set_xtol_abs"):
        getattr(lib, name).argtypes = [C.c_void_p, P]
    lib.nlopt_set_maxeval.argtypes = [C.c_void_p, C.c_int]
    lib.nlopt_optimize.argtypes = [C.c_void_p, P, P]
    lib.nlopt_shim_last_evals.argtypes = [C.c_void_p]
    lib.nlopt_shim_last_iterations.argtypes = [C.c_void_p]
    lib.kraft_nnls.argtypes = [P, C.c_int, C.c_int, C.c_int, P, P, P, C.POINTER(C.c_int)]
    lib.kraft_ldp.argtypes = [P, C.c_int, C.c_int, C.c_int, P, P, P, P, C.POINTER(C.c_int)]
    lib.kraft_ldl.argtypes = [C.c_int, P, P, C.c_double]
    return lib


def solve(lib, fun, grad, cons, lb, ub, x0, xtol, maximize=False, maxeval=100000, trace=None):
    """Through the NLopt entry points; `trace` collects the points at which gradients were asked for (start + iterates)."""
    n = len(x0)
    arr = lambda v: (C.c_double * n)(*v)
    keep = []

    def wrap(f, g, tr=None):
        def cb(_n, x, gr, _d):
            xv = np.array([x[i] for i in range(n)])
            if gr:
                gv = g(xv)
                for i in range(n):
                    gr[i] = gv[i]
                if tr is not None:
                    tr.append(xv.copy())
            return float(f(xv))
        c = FUNC(cb)
        keep.append(c)
        return c

    opt = lib.nlopt_create(NLOPT_LD_SLSQP, n)
    (lib.nlopt_set_max_objective if maximize else lib.nlopt_set_min_objective)(opt, wrap(fun, grad, trace), None)
    for kind, cf, cg in cons:
        (lib.nlopt_add_equality_constraint if kind == "eq" else lib.nlopt_add_inequality_constraint)(opt, wrap(cf, cg), None, 1e-9)
    lib.nlopt_set_lower_bounds(opt, arr(lb))
    lib.nlopt_set_upper_bounds(opt, arr(ub))
    lib.nlopt_set_xtol_abs(opt, arr([xtol] * n))
    lib.nlopt_set_maxeval(opt, maxeval)
    x = arr(x0)
    val = C.c_double()
    rc = lib.nlopt_optimize(opt, x, C.byref(val))
    evals, iters = lib.nlopt_shim_last_evals(opt), lib.nlopt_shim_last_iterations(opt)
    lib.nlopt_destroy(opt)
    return np.array(list(x)), val.value, rc, evals, iters


def test_nnls_matches_lawson_hanson(kraft):
    rng = np.random.default_rng(1)
    for _ in range(40):
        m, n = int(rng.integers(2, 40)), int(rng.integers(2, 40))
        A, b = rng.normal(size=(m, n)), rng.normal(size=m)
        Af, bb, x = np.asfortranarray(A.copy()), b.copy(), np.zeros(n)
        rn, mode = C.c_double(), C.c_int()
        kraft.kraft_nnls(_arr(Af), m, m, n, _arr(bb), _arr(x), C.byref(rn), C.byref(mode))
        xs, rs = so.nnls(A, b)
        assert mode.value == 1
        np.testing.assert_allclose(x, xs, atol=1e-10)
        assert abs(rn.value - rs) < 1e-10


def test_ldp_least_distance(kraft):
    rng = np.random.default_rng(2)
    for _ in range(30):
        m, n = int(rng.integers(1, 14)), int(rng.integers(2, 9))
        G, x0 = rng.normal(size=(m, n)), rng.normal(size=n)
        h = G @ x0 - rng.uniform(0, 1, size=m)  # x0 is feasible
        Gf, x, mult = np.asfortranarray(G.copy()), np.zeros(n), np.zeros(m)
        xn, mode = C.c_double(), C.c_int()
        kraft.kraft_ldp(_arr(Gf), m, m, n, _arr(h.copy()), _arr(x), C.byref(xn), _arr(mult), C.byref(mode))
        assert mode.value == 1
        assert np.all(G @ x - h > -1e-9)
        # KKT of min 1/2 |x|^2 s.t. G x >= h:  x = G^T mult, mult >= 0, complementary slackness
        np.testing.assert_allclose(x, G.T @ mult, atol=1e-9)
        assert np.all(mult >= -1e-12)
        assert np.abs(mult * (G @ x - h)).max() < 1e-8
        assert abs(xn.value - np.linalg.norm(x)) < 1e-12
    # an infeasible system is reported as such (mode 4)
    G = np.asfortranarray(np.array([[1.0, 0.0], [-1.0, 0.0]]))
    h = np.array([1.0, 1.0])  # x >= 1 and -x >= 1
    x, mult, xn, mode = np.zeros(2), np.zeros(2), C.c_double(), C.c_int()
    kraft.kraft_ldp(_arr(G), 2, 2, 2, _arr(h), _arr(x), C.byref(xn), _arr(mult), C.byref(mode))
    assert mode.value == 4


def test_ldl_rank_one_update(kraft):
    rng = np.random.default_rng(3)
    for _ in range(30):
        n = int(rng.integers(1, 12))
        M = rng.normal(size=(n, n))
        B = M @ M.T + n * np.eye(n)
        L, D, Bc = np.eye(n), np.zeros(n), B.copy()
        for j in range(n):
            D[j] = Bc[j, j]
            L[j + 1:, j] = Bc[j + 1:, j] / D[j]
            Bc[j + 1:, j + 1:] -= np.outer(L[j + 1:, j], L[j + 1:, j]) * D[j]
        packed = []
        for j in range(n):
            packed.append(D[j])
            packed += list(L[j + 1:, j])
        z = rng.normal(size=n)
        for sigma in (0.7, -0.05):
            a, zz = np.array(packed), z.copy()
            kraft.kraft_ldl(n, _arr(a), _arr(zz), sigma)
            L2, D2, k = np.eye(n), np.zeros(n), 0
            for j in range(n):
                D2[j] = a[k]
                k += 1
                for i in range(j + 1, n):
                    L2[i, j] = a[k]
                    k += 1
            np.testing.assert_allclose(L2 @ np.diag(D2) @ L2.T, B + sigma * np.outer(z, z), rtol=1e-9, atol=1e-9)


def _smooth_purchasing(seed):
    """Separable convex costs, one linear equality, box bounds, variables scaled by the baseline flows: the purchasing
    managers' problem (reference src/model/PurchasingManager.cpp:202-241) without the kink of the transport penalty."""
    rng = np.random.default_rng(seed)
    n = int(rng.integers(3, 40))
    z = rng.lognormal(0, 1, n)
    price, curv, pen = 1 + 0.2 * rng.random(n), 0.05 + rng.random(n), 0.5
    need, U = z.sum() * (0.8 + 0.5 * rng.random()), z.sum()
    cost = lambda x: float(np.sum(x * z * (price + curv * (x * z - z) / z) + pen * (x * z - z) ** 2 / z) / U)
    cgrad = lambda x: (price + curv * (2 * x * z - z) / z + 2 * pen * (x * z - z) / z) * z / U
    con = lambda x: float((need - np.sum(x * z)) / U)
    cong = lambda x: -z / U
    return n, cost, cgrad, con, cong, np.zeros(n), np.full(n, 3.0), np.ones(n)


@pytest.mark.parametrize("seed", range(6))
def test_trajectory_equals_scipy_slsqp(kraft, seed):
    n, cost, cgrad, con, cong, lb, ub, x0 = _smooth_purchasing(seed)
    tr, ref_tr = [], []
    xs, fs, rc, evals, iters = solve(kraft, cost, cgrad, [("eq", con, cong)], lb, ub, x0, 1e-10, trace=tr)
    ref = so.minimize(cost, x0, jac=cgrad, bounds=list(zip(lb, ub)), method="SLSQP", constraints=[{"type": "eq", "fun": con, "jac": cong}],
                      options={"ftol": 1e-16, "maxiter": 1000}, callback=lambda xk: ref_tr.append(xk.copy()))
    assert rc == 4  # NLOPT_XTOL_REACHED
    # iterate k of nlopt_kraft (the k-th point after the start at which gradients are evaluated) == scipy's iterate k
    m = min(len(tr) - 1, len(ref_tr), 40)
    assert m >= 10
    diffs = [float(np.abs(tr[i + 1] - ref_tr[i]).max()) for i in range(m)]
    assert max(diffs) < 1e-10, diffs
    np.testing.assert_allclose(xs, ref.x, atol=1e-6)
    assert abs(fs - ref.fun) < 1e-12
    assert abs(con(xs)) < 1e-9


@pytest.mark.parametrize("seed", range(4))
def test_inequality_constrained(kraft, seed):
    rng = np.random.default_rng(50 + seed)
    n = int(rng.integers(3, 12))
    Q = rng.normal(size=(n, n))
    Q = Q @ Q.T + np.eye(n)
    c = rng.normal(size=n)
    a1, a2 = rng.normal(size=n), rng.normal(size=n)
    f = lambda x: float(0.5 * x @ Q @ x + c @ x)
    g = lambda x: Q @ x + c
    # NLopt convention fc(x) <= 0
    cons = [("ineq", lambda x: float(a1 @ x - 0.3), lambda x: a1), ("ineq", lambda x: float(np.sum(x ** 2) - 2.0), lambda x: 2 * x),
            ("eq", lambda x: float(a2 @ x - 0.1), lambda x: a2)]
    lb, ub, x0 = np.full(n, -1.0), np.full(n, 1.0), np.zeros(n)
    xs, fs, rc, evals, iters = solve(kraft, f, g, cons, lb, ub, x0, 1e-10)
    ref = so.minimize(f, x0, jac=g, bounds=list(zip(lb, ub)), method="SLSQP", options={"ftol": 1e-15, "maxiter": 500},
                      constraints=[{"type": "ineq", "fun": lambda x: -(a1 @ x - 0.3), "jac": lambda x: -a1},
                                   {"type": "ineq", "fun": lambda x: -(np.sum(x ** 2) - 2.0), "jac": lambda x: -2 * x},
                                   {"type": "eq", "fun": lambda x: a2 @ x - 0.1, "jac": lambda x: a2}])
    assert rc in (1, 4)
    np.testing.assert_allclose(xs, ref.x, atol=1e-6)
    assert abs(fs - ref.fun) < 1e-9


def test_nlopt_driver_rules(kraft):
    n, cost, cgrad, con, cong, lb, ub, x0 = _smooth_purchasing(3)
    # maxeval: stops with NLOPT_MAXEVAL_REACHED (5) after exactly that many objective evaluations, returns the best point seen
    xs, fs, rc, evals, iters = solve(kraft, cost, cgrad, [("eq", con, cong)], lb, ub, x0, 1e-10, maxeval=7)
    assert rc == 5 and evals == 7
    # a start point outside the box is rejected like NLopt does (NLOPT_INVALID_ARGS)
    bad = x0.copy()
    bad[0] = 4.0
    assert solve(kraft, cost, cgrad, [("eq", con, cong)], lb, ub, bad, 1e-10)[2] == -2
    # a loose x tolerance stops earlier than a tight one, never before the second completed iteration
    loose = solve(kraft, cost, cgrad, [("eq", con, cong)], lb, ub, x0, 1e-2)
    tight = solve(kraft, cost, cgrad, [("eq", con, cong)], lb, ub, x0, 1e-12)
    assert loose[2] == 4 and tight[2] == 4 and 2 <= loose[4] < tight[4]
