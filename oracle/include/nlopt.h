/* TEST INFRASTRUCTURE (oracle). Not part of the product path.
 *
 * Declaration of the subset of NLopt's public C API that the reference binds
 * (reference include/optimization.h:123-190, src/model/Consumer.cpp:422).
 * NLopt itself (github.com/stevengj/nlopt, nominal pin v2.7.0, reference
 * CMakeLists.txt:163) is a build-time download of the reference and is NOT
 * available offline, so the oracle links one of two providers instead:
 *   oracle/nlopt_kraft  an INDEPENDENT from-scratch restatement of Kraft's SLSQP with NLopt's
 *                       documented driver logic (includes nothing of the product) -- the parity anchor;
 *   oracle/nlopt_shim   the product's own solver (csrc/sqp_core.h) behind the same entry points, so
 *                       that everything AROUND the optimiser can be compared bit for bit.
 * Names, argument order and result codes follow NLopt's documented API.
 * No upstream NLopt binary or source is available to compare against.
 */
#ifndef ORACLE_NLOPT_SHIM_H
#define ORACLE_NLOPT_SHIM_H

#ifdef __cplusplus
extern "C" {
#endif

typedef double (*nlopt_func)(unsigned n, const double* x, double* gradient, void* func_data);

/* Enumerator ORDER follows NLopt 2.7's nlopt.h so that integer values stored by the
 * reference in its settings (parameters.h:51-54) keep their meaning. */
typedef enum {
    NLOPT_GN_DIRECT = 0,
    NLOPT_GN_DIRECT_L,
    NLOPT_GN_DIRECT_L_RAND,
    NLOPT_GN_DIRECT_NOSCAL,
    NLOPT_GN_DIRECT_L_NOSCAL,
    NLOPT_GN_DIRECT_L_RAND_NOSCAL,
    NLOPT_GN_ORIG_DIRECT,
    NLOPT_GN_ORIG_DIRECT_L,
    NLOPT_GD_STOGO,
    NLOPT_GD_STOGO_RAND,
    NLOPT_LD_LBFGS_NOCEDAL,
    NLOPT_LD_LBFGS,
    NLOPT_LN_PRAXIS,
    NLOPT_LD_VAR1,
    NLOPT_LD_VAR2,
    NLOPT_LD_TNEWTON,
    NLOPT_LD_TNEWTON_RESTART,
    NLOPT_LD_TNEWTON_PRECOND,
    NLOPT_LD_TNEWTON_PRECOND_RESTART,
    NLOPT_GN_CRS2_LM,
    NLOPT_GN_MLSL,
    NLOPT_GD_MLSL,
    NLOPT_GN_MLSL_LDS,
    NLOPT_GD_MLSL_LDS,
    NLOPT_LD_MMA,
    NLOPT_LN_COBYLA,
    NLOPT_LN_NEWUOA,
    NLOPT_LN_NEWUOA_BOUND,
    NLOPT_LN_NELDERMEAD,
    NLOPT_LN_SBPLX,
    NLOPT_LN_AUGLAG,
    NLOPT_LD_AUGLAG,
    NLOPT_LN_AUGLAG_EQ,
    NLOPT_LD_AUGLAG_EQ,
    NLOPT_LN_BOBYQA,
    NLOPT_GN_ISRES,
    NLOPT_AUGLAG,
    NLOPT_AUGLAG_EQ,
    NLOPT_G_MLSL,
    NLOPT_G_MLSL_LDS,
    NLOPT_LD_SLSQP,
    NLOPT_LD_CCSAQ,
    NLOPT_GN_ESCH,
    NLOPT_GN_AGS,
    NLOPT_NUM_ALGORITHMS
} nlopt_algorithm;

typedef enum {
    NLOPT_FAILURE = -1,
    NLOPT_INVALID_ARGS = -2,
    NLOPT_OUT_OF_MEMORY = -3,
    NLOPT_ROUNDOFF_LIMITED = -4,
    NLOPT_FORCED_STOP = -5,
    NLOPT_SUCCESS = 1,
    NLOPT_STOPVAL_REACHED = 2,
    NLOPT_FTOL_REACHED = 3,
    NLOPT_XTOL_REACHED = 4,
    NLOPT_MAXEVAL_REACHED = 5,
    NLOPT_MAXTIME_REACHED = 6
} nlopt_result;

struct nlopt_opt_s;
typedef struct nlopt_opt_s* nlopt_opt;

nlopt_opt nlopt_create(nlopt_algorithm algorithm, unsigned n);
void nlopt_destroy(nlopt_opt opt);
nlopt_result nlopt_optimize(nlopt_opt opt, double* x, double* opt_f);
nlopt_result nlopt_set_max_objective(nlopt_opt opt, nlopt_func f, void* f_data);
nlopt_result nlopt_set_min_objective(nlopt_opt opt, nlopt_func f, void* f_data);
nlopt_result nlopt_set_lower_bounds(nlopt_opt opt, const double* lb);
nlopt_result nlopt_set_upper_bounds(nlopt_opt opt, const double* ub);
nlopt_result nlopt_add_inequality_constraint(nlopt_opt opt, nlopt_func fc, void* fc_data, double tol);
nlopt_result nlopt_add_equality_constraint(nlopt_opt opt, nlopt_func h, void* h_data, double tol);
nlopt_result nlopt_set_xtol_abs(nlopt_opt opt, const double* tol);
nlopt_result nlopt_set_maxeval(nlopt_opt opt, int maxeval);
nlopt_result nlopt_set_maxtime(nlopt_opt opt, double maxtime);
nlopt_result nlopt_set_local_optimizer(nlopt_opt opt, const nlopt_opt local_opt);
nlopt_result nlopt_set_population(nlopt_opt opt, unsigned pop);
const char* nlopt_get_errmsg(nlopt_opt opt);

/* provider introspection (not part of NLopt): counters of the last nlopt_optimize call, whole-run statistics
 * (calls, iterations, evaluations, QP / reset work, largest iteration count), the provider's name */
int nlopt_shim_last_iterations(nlopt_opt opt);
int nlopt_shim_last_evals(nlopt_opt opt);
void nlopt_provider_stats(long long* out5);
void nlopt_provider_set_group(int group);
const char* nlopt_provider_name(void);

#ifdef __cplusplus
}
#endif

#endif
