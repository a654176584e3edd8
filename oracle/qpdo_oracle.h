/* qpdo_oracle.h -- TEST INFRASTRUCTURE ONLY.
 *
 * Plain-C CPU restatement of QPDO's Newton hot path (reference aldma/qpdo @ 95399ba), used by
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg as the checker.  The product
 * (qpdo_b200/lib/libqpdo_b200.so) never links, imports or calls anything in this directory.
 *
 * Pinned: tests/test_oracle.py checks this restatement against (a) the reference's only
 * known-answer tests, examples/infeasibility_tests.m:30,48,75 (status 1 / -3 / -4), (b) the
 * committed golden vectors in tests/golden/ which were produced by the UNMODIFIED reference
 * compiled from /root/reference (oracle/Makefile -> oracle/_ref/libqpdo_ref.so, generator
 * tests/golden/make_golden.py), and (c) oracle/_ref itself whenever it is present.
 *
 * What is restated how:
 *   - src/qpdo.c, iteration.c, newton.c, linesearch.c, termination.c, scaling.c, lin_alg.c:
 *     operation by operation, same order of floating-point operations (no FMA contraction).
 *   - CHOLMOD 3.0.14 (vendored third party, SuiteSparse 5.8.1): sdmult/scale restated with the
 *     same accumulation order; analyze+factorize+updown+solve restated MATHEMATICALLY as "solve
 *     K dx = rhs with K = Q + sigma I + A' diag(omega) A by a natural-order envelope Cholesky",
 *     where omega tracks the row weights the reference's factor embodies after each
 *     refactorization / rank update (incl. the ldlupdate_mu_changed quirk).  Rounding therefore
 *     differs from CHOLMOD's at the 1e-13 level; iteration counts, status and active sets agree
 *     on every golden case.
 */
#ifndef QPDO_ORACLE_H
#define QPDO_ORACLE_H
#include <stddef.h>

typedef struct {            /* same layout as QPDOSettings (reference include/types.h:96-116, DLONG) */
    double max_time; long max_iter, inner_max_iter;
    double eps_abs, eps_abs_in, eps_prim_inf, eps_dual_inf, rho, theta, delta, mu_min;
    long proximal; double sigma_init, sigma_upd, sigma_min;
    long scaling, verbose, print_interval, reset_newton_iter;
} oracle_settings;

typedef struct {            /* per-iteration trace record */
    long   iter, oter, kind;        /* kind: 0 Newton step, 1 outer update */
    double res_prim_norm, res_dual_norm, res_prim_in_norm, res_dual_in_norm, tau, sigma;
    long   n_active, n_enter, n_leave, factor_mode;   /* factor_mode as in qpdo_dev_scalars */
} oracle_trace;

typedef struct {
    long   iterations, oterations, status_val;
    double res_prim_norm, res_dual_norm, res_prim_in_norm, res_dual_in_norm, objective;
    double solve_seconds;
    long   n_trace;
} oracle_info;

/* CSC with 64-bit indices; Q = lower triangle (sorted rows), A = m x n (sorted rows).
 * x0 / y0 may be NULL (cold start).  x (n), y (m), dx_cert (n), dy_cert (m), active (m) are
 * outputs and may be NULL.  trace may be NULL (cap_trace = 0).  Returns 0, or -1 on bad input. */
int qpdo_oracle_solve(long n, long m, const long *Qp, const long *Qi, const double *Qx,
                      const long *Ap, const long *Ai, const double *Ax,
                      const double *q, double cconst, const double *l, const double *u,
                      const oracle_settings *settings, const double *x0, const double *y0,
                      double *x, double *y, double *dx_cert, double *dy_cert, long *active,
                      oracle_info *info, oracle_trace *trace, long cap_trace);

/* scale_data alone (reference src/scaling.c:24-91): scales Qx, Ax, q, l, u in place */
void qpdo_oracle_scale_data(long n, long m, const long *Qp, const long *Qi, double *Qx,
                            const long *Ap, const long *Ai, double *Ax, double *q, double *l, double *u,
                            long iters, double *D, double *E, double *c);
/* y = A x (which 0), y = A' x (1), y = Q x with Q lower-stored symmetric (2): cholmod_sdmult order */
void qpdo_oracle_spmv(int which, long nrow, long ncol, const long *p, const long *i, const double *v,
                      const double *x, double *y);
/* exact linesearch alone (reference src/linesearch.c:8-158) */
double qpdo_oracle_linesearch(long n, long m, const double *dx, const double *Qdx, const double *df,
                              const double *dy, const double *y, const double *mu, const double *sqrt_mu,
                              const double *Adx, const double *w, const double *l, const double *u);
void oracle_default_settings(oracle_settings *s);
#endif
