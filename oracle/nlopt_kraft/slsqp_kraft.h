/* TEST INFRASTRUCTURE (oracle). Not part of the product path.
 *
 * From-scratch restatement of Dieter Kraft's SLSQP (D. Kraft, "A software package for sequential quadratic
 * programming", DFVLR-FB 88-28, 1988; "Algorithm 733: TOMP", ACM TOMS 20, 1994) as NLopt's NLOPT_LD_SLSQP drives it.
 * NLopt (github.com/stevengj/nlopt, nominal pin v2.7.0 at reference CMakeLists.txt:163) is a build-time download of
 * the reference and is NOT available offline; this file follows the PUBLISHED algorithm, not NLopt's source:
 *
 *   slsqp_minimize   main SQP loop: dense BFGS approximation kept as its LDL^T factors (packed, column-wise), Powell's
 *                    damped update applied as two rank-one modifications (ldl_update), L1 merit function with Kraft's
 *                    inexact Armijo-type line search (factor 0.1, quadratic interpolation floored at 0.1, <= 10 trials),
 *                    at most 5 resets of the BFGS matrix on a non-descent direction, augmented (n+1)-variable
 *                    sub-problem when the linearised constraints are inconsistent;
 *   lsq -> lsei -> lsi -> ldp -> nnls   the QP sub-problem as a least-squares problem with equality and inequality
 *                    constraints (bounds enter as 2n inequality rows), reduced by Householder transformations (h12)
 *                    to a least-distance problem, solved through its dual by Lawson & Hanson's NNLS.
 *
 * NLopt's documented behaviour around it (nlopt.readthedocs.io, "SLSQP"; NLopt's stopping-criteria section):
 * Kraft's own convergence test is switched off (acc = 0) and NLopt's tests are applied after every completed line
 * search: xtol_abs per coordinate (and the exact-repeat test of xtol_rel = 0), maxeval on function evaluations;
 * the first trial point of a line search is evaluated with gradients, later ones without, the accepted one again with
 * gradients; trial points are clipped into the box; the best point seen (lowest objective among feasible points,
 * else least infeasible) is returned.
 *
 * Nothing here includes, links or calls anything under acclimate_b200/ -- this is the INDEPENDENT optimiser the CUDA
 * path is compared against (tests/test_kraft_*.py, DESIGN.md "Parity").
 */
#ifndef ORACLE_SLSQP_KRAFT_H
#define ORACLE_SLSQP_KRAFT_H

#include <functional>
#include <vector>

namespace kraft {

enum Result {
    R_FAILURE = -1,            /* LSQ sub-problem failed: NNLS iteration limit (mode 3) or incompatible constraints (mode 4) */
    R_INVALID_ARGS = -2,
    R_ROUNDOFF_LIMITED = -4,   /* positive directional derivative after 5 resets (mode 8), singular sub-problem (5, 6, 7) */
    R_SUCCESS = 1,
    R_XTOL_REACHED = 4,
    R_MAXEVAL_REACHED = 5,
};

struct Problem {
    int n = 0;
    int meq = 0;   /* equality constraints h_j(x) = 0, j < meq */
    int mineq = 0; /* inequality constraints in Kraft's convention g_j(x) >= 0 */
    /* objective to MINIMISE; grad may be null */
    std::function<double(const double* x, double* grad)> f;
    /* constraint j (equalities first); value returned, gradient written when grad != null */
    std::function<double(int j, const double* x, double* grad)> c;
    std::vector<double> lb, ub, xtol_abs, ctol; /* ctol[j]: feasibility tolerance of constraint j (best-point rule) */
    int maxeval = 0;                             /* <= 0: unlimited */
};

struct Stats {
    int iters = 0;     /* completed SQP iterations (line searches) */
    int evals = 0;     /* objective evaluations */
    int resets = 0;    /* BFGS resets */
    int lsq_mode = 1;  /* mode of the last LSQ call */
    int augmented = 0; /* sub-problems solved in the augmented (inconsistent linearisation) form */
};

/* x: start point in, best point out.  *minf: objective at the returned point. */
Result slsqp_minimize(const Problem& p, double* x, double* minf, Stats* st);

/* building blocks, exposed for the unit tests (tests/test_kraft_slsqp.py through the ctypes entry points of
 * nlopt_kraft.cpp) */
void h12(int mode, int lpivot, int l1, int m, double* u, int iue, double* up, double* c, int ice, int icv, int ncv);
void nnls(double* a, int mda, int m, int n, double* b, double* x, double* rnorm, double* w, double* z, int* index, int* mode);
void ldp(const double* g, int mg, int m, int n, const double* h, double* x, double* xnorm, double* w, int* index, int* mode);
void ldl_update(int n, double* a, double* z, double sigma, double* w);

} /* namespace kraft */

#endif
