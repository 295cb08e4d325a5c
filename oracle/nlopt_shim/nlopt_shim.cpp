/* TEST INFRASTRUCTURE (oracle). Not part of the product path.
 *
 * NLopt provider for the oracle build of the reference (the reference downloads NLopt at build
 * time, reference cmake/libraries.cmake:49-97; it is not available offline).  Implements the 14
 * entry points the reference binds (reference include/optimization.h:137-186,
 * src/model/Consumer.cpp:422).  Only NLOPT_LD_SLSQP -- the reference's default for both the
 * purchasing and the utility optimisation (src/input/ModelInitializer.cpp:1034,1048) -- is
 * provided; every other algorithm id returns NLOPT_INVALID_ARGS with a message.
 *
 * The SLSQP iteration itself is acclimate_b200/csrc/sqp_core.h, run through its HostCtx, which
 * reproduces the device solver's reduction order so that the reference's CPU run and the CUDA
 * run execute the same floating-point sequence (see DESIGN.md "Optimiser").  Its optima are
 * cross-checked against scipy's SLSQP (Kraft's Fortran) in tests/test_sqp_vs_scipy.py.
 * PARITY WITH UPSTREAM NLopt IS UNPINNED.
 */
#include "nlopt.h"

#include <cstdlib>
#include <cstring>
#include <vector>

#include "../../acclimate_b200/csrc/coop.cuh"
#include "../../acclimate_b200/csrc/sqp_core.h"
#include "../../acclimate_b200/csrc/sqp_dense.h"

struct nlopt_opt_s {
    nlopt_algorithm algorithm;
    unsigned n;
    nlopt_func f = nullptr;
    void* f_data = nullptr;
    bool maximize = false;
    nlopt_func con = nullptr;
    void* con_data = nullptr;
    int ctype = accsqp::CON_NONE;
    double ctol = 0.0;
    std::vector<double> lb, ub, xtol;
    std::vector<double> kinks;  // 5 n: scale, target, left slope, right slope, slope at the kink (sqp_core.h "Kinks"); empty = smooth
    int maxeval = 0;
    double maxtime = 0.0;
    const char* errmsg = nullptr;
    int last_iters = 0;
    int last_evals = 0;
};

extern "C" {

nlopt_opt nlopt_create(nlopt_algorithm algorithm, unsigned n) {
    auto* o = new nlopt_opt_s;
    o->algorithm = algorithm;
    o->n = n;
    o->lb.assign(n, -1e300);
    o->ub.assign(n, 1e300);
    o->xtol.assign(n, 0.0);
    return o;
}

void nlopt_destroy(nlopt_opt opt) { delete opt; }

nlopt_result nlopt_set_max_objective(nlopt_opt opt, nlopt_func f, void* f_data) {
    opt->f = f; opt->f_data = f_data; opt->maximize = true;
    return NLOPT_SUCCESS;
}
nlopt_result nlopt_set_min_objective(nlopt_opt opt, nlopt_func f, void* f_data) {
    opt->f = f; opt->f_data = f_data; opt->maximize = false;
    return NLOPT_SUCCESS;
}
nlopt_result nlopt_set_lower_bounds(nlopt_opt opt, const double* lb) {
    opt->lb.assign(lb, lb + opt->n);
    return NLOPT_SUCCESS;
}
nlopt_result nlopt_set_upper_bounds(nlopt_opt opt, const double* ub) {
    opt->ub.assign(ub, ub + opt->n);
    return NLOPT_SUCCESS;
}
nlopt_result nlopt_add_inequality_constraint(nlopt_opt opt, nlopt_func fc, void* fc_data, double tol) {
    if (opt->con != nullptr) { opt->errmsg = "shim supports a single constraint"; return NLOPT_INVALID_ARGS; }
    opt->con = fc; opt->con_data = fc_data; opt->ctype = accsqp::CON_INEQ; opt->ctol = tol;
    return NLOPT_SUCCESS;
}
nlopt_result nlopt_add_equality_constraint(nlopt_opt opt, nlopt_func h, void* h_data, double tol) {
    if (opt->con != nullptr) { opt->errmsg = "shim supports a single constraint"; return NLOPT_INVALID_ARGS; }
    opt->con = h; opt->con_data = h_data; opt->ctype = accsqp::CON_EQ; opt->ctol = tol;
    return NLOPT_SUCCESS;
}
nlopt_result nlopt_set_xtol_abs(nlopt_opt opt, const double* tol) {
    opt->xtol.assign(tol, tol + opt->n);
    return NLOPT_SUCCESS;
}
nlopt_result nlopt_set_maxeval(nlopt_opt opt, int maxeval) { opt->maxeval = maxeval; return NLOPT_SUCCESS; }
nlopt_result nlopt_set_maxtime(nlopt_opt opt, double maxtime) { opt->maxtime = maxtime; return NLOPT_SUCCESS; }
nlopt_result nlopt_set_local_optimizer(nlopt_opt opt, const nlopt_opt) {
    opt->errmsg = "nlopt_shim: subsidiary optimizers are not provided";
    return NLOPT_INVALID_ARGS;
}
nlopt_result nlopt_set_population(nlopt_opt, unsigned) { return NLOPT_SUCCESS; }
const char* nlopt_get_errmsg(nlopt_opt opt) { return opt->errmsg; }
/* Not part of NLopt: describes the per-coordinate kink of the objective (sqp_core.h "Kinks").  The device solver
 * knows the kinks of the purchasing objective from the model; on the oracle side the reference's callbacks cannot tell,
 * so the harness supplies them through nlopt_shim_kinks below (or a test through this setter). */
nlopt_result nlopt_shim_set_kinks(nlopt_opt opt, const double* scale, const double* target, const double* sl, const double* sr, const double* mid) {
    const unsigned n = opt->n;
    opt->kinks.resize(5 * (size_t)n);
    std::memcpy(&opt->kinks[0], scale, n * sizeof(double)); std::memcpy(&opt->kinks[n], target, n * sizeof(double));
    std::memcpy(&opt->kinks[2 * n], sl, n * sizeof(double)); std::memcpy(&opt->kinks[3 * n], sr, n * sizeof(double));
    std::memcpy(&opt->kinks[4 * n], mid, n * sizeof(double));
    return NLOPT_SUCCESS;
}
/* Optional hook (weak): asked once per nlopt_optimize with the objective's user data; returns 1 after filling the five
 * arrays when that objective has kinks.  oracle/ref_harness defines it for PurchasingManager objectives. */
int nlopt_shim_kinks(void* f_data, unsigned n, double* scale, double* target, double* sl, double* sr, double* mid) __attribute__((weak));
int nlopt_shim_last_iterations(nlopt_opt opt) { return opt->last_iters; }
int nlopt_shim_last_evals(nlopt_opt opt) { return opt->last_evals; }

}  // extern "C"

namespace {

struct CallbackEval {
    nlopt_opt o;
    accsqp::Work* w;
    std::vector<double> tmp;
    void operator()(const double* x, bool want_grad, double* F, double* c) {
        const unsigned n = o->n;
        const double sgn = o->maximize ? -1.0 : 1.0;
        if (want_grad) {
            const double fv = o->f(n, x, tmp.data(), o->f_data);
            *F = sgn * fv;
            for (unsigned i = 0; i < n; ++i) w->g[i] = sgn * tmp[i];
        } else {
            *F = sgn * o->f(n, x, nullptr, o->f_data);
        }
        if (o->con != nullptr) {
            const double cs = (o->ctype == accsqp::CON_INEQ) ? -1.0 : 1.0;  // NLopt: fc(x) <= 0
            if (want_grad) {
                const double cv = o->con(n, x, tmp.data(), o->con_data);
                *c = cs * cv;
                for (unsigned i = 0; i < n; ++i) w->a[i] = cs * tmp[i];
            } else {
                *c = cs * o->con(n, x, nullptr, o->con_data);
            }
        } else {
            *c = 0.0;
            if (want_grad) for (unsigned i = 0; i < n; ++i) w->a[i] = 0.0;
        }
    }
};

// the same for sqp_dense.h (gradients go to the arrays the solver names)
struct DenseCallbackEval {
    nlopt_opt o;
    // sqp_dense.h's group convention with the serial executor: F -> scal[0], c -> scal[1]
    void operator()(accsqp::SerialPar&, const double* x, double* g, double* a, double* scal) {
        double* F = &scal[0];
        double* c = &scal[1];
        const unsigned n = o->n;
        const double sgn = o->maximize ? -1.0 : 1.0;
        *F = sgn * o->f(n, x, g, o->f_data);
        for (unsigned i = 0; i < n; ++i) g[i] = sgn * g[i];
        if (o->con != nullptr) {
            const double cs = (o->ctype == accsqp::CON_INEQ) ? -1.0 : 1.0;  // NLopt: fc(x) <= 0
            *c = cs * o->con(n, x, a, o->con_data);
            for (unsigned i = 0; i < n; ++i) a[i] = cs * a[i];
        } else {
            *c = 0.0;
            for (unsigned i = 0; i < n; ++i) a[i] = 0.0;
        }
    }
};

}  // namespace

// thread-group size whose reduction order the diagonal solver replays (accsqp_group_for of the network's largest manager)
static int g_group = ACCSQP_GROUP;
extern "C" void nlopt_provider_set_group(int group) { g_group = group; }
extern "C" const char* nlopt_provider_name(void) { return "shim"; }

// whole-run solver statistics (printed by the harness; serial builds only -- not atomic)
long long g_stat_calls = 0, g_stat_iters = 0, g_stat_evals = 0, g_stat_qp = 0, g_stat_max_iters = 0;
extern "C" void nlopt_provider_stats(long long* out) {
    out[0] = g_stat_calls; out[1] = g_stat_iters; out[2] = g_stat_evals; out[3] = g_stat_qp; out[4] = g_stat_max_iters;
}

extern "C" nlopt_result nlopt_optimize(nlopt_opt opt, double* x, double* opt_f) {
    if (opt->algorithm != NLOPT_LD_SLSQP) {
        opt->errmsg = "nlopt_shim: only NLOPT_LD_SLSQP is provided";
        return NLOPT_INVALID_ARGS;
    }
    if (opt->f == nullptr) return NLOPT_INVALID_ARGS;
    const int n = (int)opt->n;
    std::vector<double> buf((size_t)ACCSQP_NVEC * n);
    double* q = buf.data();
    accsqp::Work w;
    w.x = q; q += n; w.x0 = q; q += n; w.g = q; q += n; w.a = q; q += n; w.d = q; q += n; w.b = q; q += n;
    w.gl = q; q += n; w.xbest = q; q += n;
    w.xl = opt->lb.data();
    w.xu = opt->ub.data();
    w.xtol = opt->xtol.data();
    std::memcpy(w.x, x, sizeof(double) * n);
    std::vector<double> kinks = opt->kinks;
    if (kinks.empty() && nlopt_shim_kinks != nullptr) {
        kinks.resize(5 * (size_t)n);
        if (!nlopt_shim_kinks(opt->f_data, (unsigned)n, &kinks[0], &kinks[n], &kinks[2 * (size_t)n], &kinks[3 * (size_t)n], &kinks[4 * (size_t)n])) kinks.clear();
    }
    if (!kinks.empty()) {
        w.kscale = &kinks[0]; w.ktarget = &kinks[n]; w.ksl_ = &kinks[2 * (size_t)n]; w.ksr_ = &kinks[3 * (size_t)n]; w.kmid_ = &kinks[4 * (size_t)n];
    }

    accsqp::Params p;
    p.n = n; p.ctype = opt->ctype; p.ctol = opt->ctol; p.maxeval = opt->maxeval; p.max_iter = ACCSQP_MAX_ITER;
    accsqp::Stats st;
    int ret;
    if (!kinks.empty()) {
        // separable objective with described kinks (the purchasing managers): diagonal model, sqp_core.h
        acc::HostCtx cx(g_group);  // replays the device group's reduction order
        CallbackEval ev{opt, &w, std::vector<double>((size_t)n)};
        ret = accsqp::sqp_minimize(cx, ev, w, p, st);
        std::memcpy(x, w.xbest, sizeof(double) * n);
    } else {
        // anything else (the consumers' nested-CES utility): dense BFGS model, sqp_dense.h
        std::vector<double> dbuf(accsqp::dense_doubles(n));
        std::vector<int> ibuf(2 * (size_t)n);
        accsqp::DenseWork dw;
        double* r = dbuf.data();
        dw.n = n;
        dw.x = r; r += n; dw.x0 = r; r += n; dw.g = r; r += n; dw.a = r; r += n; dw.d = r; r += n; dw.gl = r; r += n; dw.xbest = r; r += n;
        dw.t1 = r; r += n; dw.t2 = r; r += n; dw.t3 = r; r += n; dw.B = r; r += (size_t)n * n; dw.L = r; r += (size_t)n * n; dw.scal = r;
        dw.state = ibuf.data();
        dw.xl = opt->lb.data(); dw.xu = opt->ub.data(); dw.xtol = opt->xtol.data();
        std::memcpy(dw.x, x, sizeof(double) * n);
        DenseCallbackEval ev{opt};
        accsqp::SerialPar par;
        ret = accsqp::sqp_minimize_dense(par, ev, dw, p, st);
        std::memcpy(x, dw.xbest, sizeof(double) * n);
    }
    *opt_f = opt->maximize ? -st.minf : st.minf;
    opt->last_iters = st.iters;
    opt->last_evals = st.evals;
    g_stat_calls += 1; g_stat_iters += st.iters; g_stat_evals += st.evals; g_stat_qp += st.qp_iters;
    if (st.iters > g_stat_max_iters) g_stat_max_iters = st.iters;
    if (ret == accsqp::SQP_FAILURE) {
        opt->errmsg = "bug: more than iter SQP iterations";  // NLopt's message for Kraft's mode 9
        return NLOPT_FAILURE;
    }
    return (nlopt_result)ret;
}
