/* TEST INFRASTRUCTURE (oracle). Not part of the product path.
 *
 * NLopt provider "kraft" for the oracle build of the reference: the 14 NLopt entry points the reference binds
 * (reference include/optimization.h:137-186, src/model/Consumer.cpp:422) in front of slsqp_kraft.cpp, an independent
 * restatement of Kraft's SLSQP.  Includes NOTHING of the product (acclimate_b200/): this is the optimiser the CUDA
 * path's results are measured against (`make -C oracle kraft`, tests/golden/*.kraft.acc.gz).
 * Only NLOPT_LD_SLSQP is provided (the reference's default for purchasing and utility optimisation,
 * src/input/ModelInitializer.cpp:1034,1048).
 */
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "nlopt.h"
#include "slsqp_kraft.h"

struct nlopt_opt_s {
    nlopt_algorithm algorithm;
    unsigned n;
    nlopt_func f = nullptr;
    void* f_data = nullptr;
    bool maximize = false;
    struct Con { nlopt_func fn; void* data; double tol; };
    std::vector<Con> eq, ineq;
    std::vector<double> lb, ub, xtol;
    bool have_xtol = false;
    int maxeval = 0;
    double maxtime = 0.0;
    const char* errmsg = nullptr;
    int last_iters = 0, last_evals = 0;
};

static long long g_calls = 0, g_iters = 0, g_evals = 0, g_resets = 0, g_max_iters = 0;

extern "C" {

nlopt_opt nlopt_create(nlopt_algorithm algorithm, unsigned n) {
    auto* o = new nlopt_opt_s;
    o->algorithm = algorithm;
    o->n = n;
    o->lb.assign(n, -HUGE_VAL);
    o->ub.assign(n, HUGE_VAL);
    o->xtol.assign(n, 0.0);
    return o;
}
void nlopt_destroy(nlopt_opt opt) { delete opt; }
nlopt_result nlopt_set_max_objective(nlopt_opt opt, nlopt_func f, void* d) { opt->f = f; opt->f_data = d; opt->maximize = true; return NLOPT_SUCCESS; }
nlopt_result nlopt_set_min_objective(nlopt_opt opt, nlopt_func f, void* d) { opt->f = f; opt->f_data = d; opt->maximize = false; return NLOPT_SUCCESS; }
nlopt_result nlopt_set_lower_bounds(nlopt_opt opt, const double* lb) { opt->lb.assign(lb, lb + opt->n); return NLOPT_SUCCESS; }
nlopt_result nlopt_set_upper_bounds(nlopt_opt opt, const double* ub) { opt->ub.assign(ub, ub + opt->n); return NLOPT_SUCCESS; }
nlopt_result nlopt_add_inequality_constraint(nlopt_opt opt, nlopt_func fc, void* d, double tol) { opt->ineq.push_back({fc, d, tol}); return NLOPT_SUCCESS; }
nlopt_result nlopt_add_equality_constraint(nlopt_opt opt, nlopt_func h, void* d, double tol) { opt->eq.push_back({h, d, tol}); return NLOPT_SUCCESS; }
nlopt_result nlopt_set_xtol_abs(nlopt_opt opt, const double* tol) { opt->xtol.assign(tol, tol + opt->n); opt->have_xtol = true; return NLOPT_SUCCESS; }
nlopt_result nlopt_set_maxeval(nlopt_opt opt, int maxeval) { opt->maxeval = maxeval; return NLOPT_SUCCESS; }
nlopt_result nlopt_set_maxtime(nlopt_opt opt, double maxtime) { opt->maxtime = maxtime; return NLOPT_SUCCESS; }
nlopt_result nlopt_set_local_optimizer(nlopt_opt opt, const nlopt_opt) {
    opt->errmsg = "nlopt_kraft: subsidiary optimizers are not provided";
    return NLOPT_INVALID_ARGS;
}
nlopt_result nlopt_set_population(nlopt_opt, unsigned) { return NLOPT_SUCCESS; }
const char* nlopt_get_errmsg(nlopt_opt opt) { return opt->errmsg; }
int nlopt_shim_last_iterations(nlopt_opt opt) { return opt->last_iters; }
int nlopt_shim_last_evals(nlopt_opt opt) { return opt->last_evals; }
void nlopt_provider_stats(long long* out) { out[0] = g_calls; out[1] = g_iters; out[2] = g_evals; out[3] = g_resets; out[4] = g_max_iters; }
void nlopt_provider_set_group(int) {}
const char* nlopt_provider_name(void) { return "kraft"; }

nlopt_result nlopt_optimize(nlopt_opt opt, double* x, double* opt_f) {
    if (opt->algorithm != NLOPT_LD_SLSQP) {
        opt->errmsg = "nlopt_kraft: only NLOPT_LD_SLSQP is provided";
        return NLOPT_INVALID_ARGS;
    }
    if (opt->f == nullptr) return NLOPT_INVALID_ARGS;
    const unsigned n = opt->n;
    const double sgn = opt->maximize ? -1.0 : 1.0;
    kraft::Problem p;
    p.n = (int)n;
    p.meq = (int)opt->eq.size();
    p.mineq = (int)opt->ineq.size();
    p.lb = opt->lb;
    p.ub = opt->ub;
    if (opt->have_xtol) p.xtol_abs = opt->xtol;
    p.maxeval = opt->maxeval;
    for (auto& c : opt->eq) p.ctol.push_back(c.tol);
    for (auto& c : opt->ineq) p.ctol.push_back(c.tol);
    p.f = [&](const double* xx, double* grad) {
        const double v = sgn * opt->f(n, xx, grad, opt->f_data);
        if (grad != nullptr && opt->maximize) for (unsigned i = 0; i < n; ++i) grad[i] = -grad[i];
        return v;
    };
    p.c = [&](int j, const double* xx, double* grad) {
        if (j < p.meq) return opt->eq[j].fn(n, xx, grad, opt->eq[j].data);
        /* NLopt: fc(x) <= 0; Kraft: g(x) >= 0 */
        const auto& c = opt->ineq[j - p.meq];
        const double v = -c.fn(n, xx, grad, c.data);
        if (grad != nullptr) for (unsigned i = 0; i < n; ++i) grad[i] = -grad[i];
        return v;
    };
    kraft::Stats st;
    double minf = 0;
    const kraft::Result r = kraft::slsqp_minimize(p, x, &minf, &st);
    *opt_f = sgn * minf;
    opt->last_iters = st.iters;
    opt->last_evals = st.evals;
    g_calls += 1; g_iters += st.iters; g_evals += st.evals; g_resets += st.resets;
    if (st.iters > g_max_iters) g_max_iters = st.iters;
    if (r == kraft::R_FAILURE) {
        opt->errmsg = "bug: more than iter SQP iterations";  /* NLopt's message for a failed LSQ sub-problem */
        return NLOPT_FAILURE;
    }
    if (r == kraft::R_INVALID_ARGS) { opt->errmsg = "nlopt_kraft: start point outside the bounds"; return NLOPT_INVALID_ARGS; }
    return (nlopt_result)r;
}

/* ---- building blocks for the unit tests (ctypes) ---- */
void kraft_nnls(double* a, int mda, int m, int n, double* b, double* x, double* rnorm, int* mode) {
    std::vector<double> w(n), z(m);
    std::vector<int> index(n);
    kraft::nnls(a, mda, m, n, b, x, rnorm, w.data(), z.data(), index.data(), mode);
}
void kraft_ldp(const double* g, int mg, int m, int n, const double* h, double* x, double* xnorm, double* mult, int* mode) {
    std::vector<double> w((size_t)(n + 1) * (m + 2) + 2 * (size_t)m + 8);
    std::vector<int> index(m + 1);
    kraft::ldp(g, mg, m, n, h, x, xnorm, w.data(), index.data(), mode);
    if (*mode == 1 && mult != nullptr) std::memcpy(mult, w.data(), sizeof(double) * m);
}
void kraft_ldl(int n, double* a, double* z, double sigma) {
    std::vector<double> w(n);
    kraft::ldl_update(n, a, z, sigma, w.data());
}

}  // extern "C"
