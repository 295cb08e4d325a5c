"""The optimiser (csrc/sqp_core.h, through oracle/nlopt_shim's NLopt entry points) against an INDEPENDENT SLSQP:
scipy.optimize.minimize(method="SLSQP") wraps Kraft's original Fortran, the same algorithm NLopt's NLOPT_LD_SLSQP is
a C translation of.  NLopt itself is not available offline, so this is the closest pin of the optimiser boundary:
on problems shaped like the reference's two call sites the optimum found by sqp_core must be as good as Kraft's.

  * purchasing-like (reference src/model/PurchasingManager.cpp:149-241): separable convex costs with a kink, one
    linear equality constraint, box bounds;
  * consumer-like (reference src/model/Consumer.cpp:170-253): nested-CES utility (NOT separable), budget constraint.
"""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

scipy_optimize = pytest.importorskip("scipy.optimize")

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
NLOPT_LD_SLSQP = None  # read from the header below
FUNC = C.CFUNCTYPE(C.c_double, C.c_uint, C.POINTER(C.c_double), C.POINTER(C.c_double), C.c_void_p)


@pytest.fixture(scope="module")
def shim():
    global NLOPT_LD_SLSQP
    subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "oracle"), "shim"], check=True)
    lib = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libnlopt_shim.so"))
    lib.nlopt_create.restype = C.c_void_p
    lib.nlopt_create.argtypes = [C.c_int, C.c_uint]
    lib.nlopt_destroy.argtypes = [C.c_void_p]
    for name in ("nlopt_set_min_objective", "nlopt_set_max_objective"):
        getattr(lib, name).argtypes = [C.c_void_p, FUNC, C.c_void_p]
    for name in ("nlopt_add_equality_constraint", "nlopt_add_inequality_constraint"):
        getattr(lib, name).argtypes = [C.c_void_p, FUNC, C.c_void_p, C.c_double]
    for name in ("nlopt_set_lower_bounds", "nlopt_set_upper_bounds", "nlopt_set_xtol_abs"):
        getattr(lib, name).argtypes = [C.c_void_p, C.POINTER(C.c_double)]
    lib.nlopt_set_maxeval.argtypes = [C.c_void_p, C.c_int]
    lib.nlopt_optimize.argtypes = [C.c_void_p, C.POINTER(C.c_double), C.POINTER(C.c_double)]
    lib.nlopt_shim_set_kinks.argtypes = [C.c_void_p] + [C.POINTER(C.c_double)] * 5
    lib.nlopt_shim_last_evals.argtypes = [C.c_void_p]
    # position of NLOPT_LD_SLSQP in the enum of oracle/include/nlopt.h
    names = []
    inside = False
    for line in open(os.path.join(ROOT, "oracle", "include", "nlopt.h")):
        line = line.split("/*")[0].split("//")[0].strip()
        if line.startswith("typedef enum") and not names:
            inside = True
            continue
        if inside:
            if line.startswith("}"):
                break
            names += [t.split("=")[0].strip() for t in line.split(",") if t.strip()]
    NLOPT_LD_SLSQP = names.index("NLOPT_LD_SLSQP")
    return lib


def solve_shim(lib, fun, grad, con, con_grad, lb, ub, x0, xtol, maximize=False, inequality=False, kinks=None):
    n = len(x0)
    arr = lambda v: (C.c_double * n)(*v)

    def wrap(f, g):
        def cb(_n, x, gr, _data):
            xv = np.array([x[i] for i in range(n)])
            if gr:
                gv = g(xv)
                for i in range(n):
                    gr[i] = gv[i]
            return float(f(xv))
        return FUNC(cb)

    fcb, ccb = wrap(fun, grad), wrap(con, con_grad)
    opt = lib.nlopt_create(NLOPT_LD_SLSQP, n)
    (lib.nlopt_set_max_objective if maximize else lib.nlopt_set_min_objective)(opt, fcb, None)
    (lib.nlopt_add_inequality_constraint if inequality else lib.nlopt_add_equality_constraint)(opt, ccb, None, 1e-9)
    lib.nlopt_set_lower_bounds(opt, arr(lb))
    lib.nlopt_set_upper_bounds(opt, arr(ub))
    lib.nlopt_set_xtol_abs(opt, arr([xtol] * n))
    lib.nlopt_set_maxeval(opt, 100000)
    if kinks is not None:  # (scale, target, left slope, right slope, slope at the kink), what the harness tells the shim about the purchasing objective
        lib.nlopt_shim_set_kinks(opt, *[arr(k) for k in kinks])
    x = arr(x0)
    val = C.c_double(0.0)
    rc = lib.nlopt_optimize(opt, x, C.byref(val))
    evals = lib.nlopt_shim_last_evals(opt)
    lib.nlopt_destroy(opt)
    return np.array(list(x)), val.value, rc, evals


@pytest.mark.parametrize("seed", range(8))
def test_purchasing_like_problem(shim, seed):
    rng = np.random.default_rng(seed)
    n = int(rng.integers(3, 40))
    z = rng.lognormal(0.0, 1.0, n)              # baseline flows; variables are scaled by them (x = D / Z*)
    price = 1.0 + 0.2 * rng.random(n)           # price level per supplier
    curv = 0.05 + rng.random(n)                 # price increase with demand
    kink = 0.02                                 # markup * |D - Z*| term
    pen = 0.5
    need = z.sum() * (0.8 + 0.5 * rng.random())
    U = z.sum()

    def cost(x):
        D = x * z
        return float(np.sum(D * (price + curv * np.maximum(D - z, 0.0) / z) + kink * np.abs(D - z) + pen * (D - z) ** 2 / z) / U)

    def cost_grad(x):
        D = x * z
        over = D > z
        g = price + np.where(over, curv * (2 * D - z) / z, 0.0) + kink * np.sign(D - z) + 2 * pen * (D - z) / z
        return g * z / U

    con = lambda x: float((need - np.sum(x * z)) / U)
    con_grad = lambda x: -z / U
    lb, ub = np.zeros(n), np.full(n, 3.0)
    x0 = np.ones(n)
    # kink at D = z (x = 1): slope jumps from -kink to kink + curv (the price increase sets in there as well)
    kinks = (z, z, -kink * z / U, (kink + curv) * z / U, np.zeros(n))
    xs, fs, rc, evals = solve_shim(shim, cost, cost_grad, con, con_grad, lb, ub, x0, 1e-7, kinks=kinks)
    assert rc > 0 or rc == -4
    ref = scipy_optimize.minimize(cost, x0, jac=cost_grad, bounds=list(zip(lb, ub)), method="SLSQP",
                                  constraints=[{"type": "eq", "fun": con, "jac": con_grad}], options={"ftol": 1e-14, "maxiter": 2000})
    assert abs(con(xs)) < 1e-8
    # as good as Kraft's optimum (the kink makes both stop within the same basin; allow 1e-6 relative)
    assert fs <= ref.fun * (1 + 1e-6) + 1e-12, (fs, ref.fun)
    assert evals <= ref.nfev, "the kink-aware solver should not need more evaluations than Kraft's smooth-model SLSQP"


def _ces(shares_goods, shares_baskets, baskets, sub_intra, sub_inter):
    e_in = [(s - 1) / s for s in sub_intra]
    e_out = (sub_inter - 1) / sub_inter

    def util(x):
        total = 0.0
        for b, idx in enumerate(baskets):
            inner = np.sum((x[idx] * shares_goods[idx]) ** e_in[b] * shares_goods[idx] ** (1 / sub_intra[b]))
            total += (inner ** (1 / e_in[b]) * shares_baskets[b]) ** e_out * shares_baskets[b] ** (1 / sub_inter)
        return total ** (1 / e_out)

    def grad(x, h=1e-7):
        g = np.zeros_like(x)
        for i in range(x.size):
            xp, xm = x.copy(), x.copy()
            xp[i] += h
            xm[i] -= h
            g[i] = (util(xp) - util(xm)) / (2 * h)
        return g
    return util, grad


@pytest.mark.parametrize("seed", range(6))
def test_consumer_like_problem(shim, seed):
    rng = np.random.default_rng(100 + seed)
    n = int(rng.integers(4, 27))
    nb = int(rng.integers(1, 4))
    member = rng.integers(0, nb, n)
    member[:nb] = np.arange(nb)
    baskets = [np.nonzero(member == b)[0] for b in range(nb)]
    w = rng.dirichlet(np.ones(n))
    shares_baskets = np.array([w[idx].sum() for idx in baskets])
    shares_goods = np.array([w[i] / shares_baskets[member[i]] for i in range(n)])
    util, ugrad = _ces(shares_goods, shares_baskets, baskets, [2.0] * nb, 1.5)
    prices = 1.0 + 0.3 * (rng.random(n) - 0.3)  # prices moved away from the baseline of 1
    budget = lambda x: float(np.sum(w * prices * x) - 1.0)      # <= 0 / == 0
    budget_grad = lambda x: w * prices
    lb, ub = np.full(n, 0.5), np.full(n, 1.5)
    x0 = np.clip(np.ones(n) / prices.mean(), lb, ub)
    xs, fs, rc, _ = solve_shim(shim, util, ugrad, budget, budget_grad, lb, ub, x0, 1e-8, maximize=True)
    assert rc > 0 or rc == -4
    ref = scipy_optimize.minimize(lambda x: -util(x), x0, jac=lambda x: -ugrad(x), bounds=list(zip(lb, ub)), method="SLSQP",
                                  constraints=[{"type": "eq", "fun": budget, "jac": budget_grad}], options={"ftol": 1e-14, "maxiter": 2000})
    assert abs(budget(xs)) < 1e-7
    assert fs >= -ref.fun * (1 - 2e-6), (fs, -ref.fun)
