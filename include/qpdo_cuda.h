/* qpdo_cuda.h -- thin C ABI between the host C solver (csrc/qpdo_api.c) and the sm_100a kernels.
 *
 * Plain pointers and sizes only.  Each entry point names the reference routine(s) it replaces
 * (paths relative to the reference tree).  All device work is issued on one CUDA stream owned by
 * the context; calls are asynchronous unless they return data to the host.  Every function
 * returns 0 on success or a negative error code; qpdo_dev_last_error() gives the text.  There is
 * no CPU fallback: without a CUDA device qpdo_dev_create fails.
 */
#ifndef QPDO_B200_QPDO_CUDA_H
#define QPDO_B200_QPDO_CUDA_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct qpdo_dev qpdo_dev;

/* problem as it crosses the boundary: 32-bit CSC, sorted rows, no duplicates;
 * Q = lower triangle of the n x n Hessian, A = m x n constraints */
typedef struct {
    int n, m;
    const int    *Qp, *Qi;  const double *Qx;
    const int    *Ap, *Ai;  const double *Ax;
    const double *q, *l, *u;
} qpdo_dev_problem;

/* device-resident scalars mirrored to the host once per iteration */
typedef struct {
    double res_prim_norm, res_dual_norm, res_prim_in_norm, res_dual_in_norm; /* termination.c:35-77 */
    double tau;                 /* last stepsize (linesearch.c:128,144,156)                  */
    double ls_eta, ls_beta;     /* linesearch.c:16-26                                        */
    double objective;
    double inf_a, inf_b, inf_c; /* infeasibility test scratch (termination.c:97-216)        */
    double sigma;               /* proximal weight the kernels read                          */
    int    n_active, n_enter, n_leave;      /* newton.c:96-126                               */
    int    n_mu_changed;                    /* iteration.c:127-168                           */
    int    reset_newton;                    /* QPDOCholmod.reset_newton                      */
    int    factor_mode;   /* last Newton step: 0 refactor omega=act/mu, 1 rank update, 2 Q only */
    int    ls_nL;         /* breakpoints with t > 0                                          */
    int    not_posdef;    /* a non-positive pivot was met by the numeric factorization      */
    int    flag;          /* result of the last infeasibility kernel                        */
    int    n_refactor, n_update, n_qonly;   /* counters since begin_solve                    */
    int    pad;
} qpdo_dev_scalars;

/* cumulative device time per phase (ms), measured with CUDA events when profiling is on */
typedef struct {
    double assemble_ms, factor_ms, solve_ms, spmv_ms, linesearch_ms, residual_ms, other_ms;
    long   n_newton, n_residual, kernel_launches;
} qpdo_dev_profile;

typedef struct {
    int     nsuper, nlevel, max_front, max_nc;
    int64_t lnz, lnz_stored, usize, nnzK, ntriple;
    double  flops;
    double  analyze_seconds;
} qpdo_dev_symbolic_info;

const char *qpdo_dev_last_error(void);
int  qpdo_dev_count(void);

/* qpdo_setup's device half: upload, Ruiz equilibration + cost scaling (scaling.c:24-91 with
 * cholmod_interface.c:162-199 and cholmod_scale.c:137-209), symbolic analysis (replaces
 * cholmod_analyze, cholmod_interface.c:10).  ordering: 0 natural, 1 minimum degree, 2 nested dissection with
 * minimum-degree leaves (the default of qpdo_setup; QPDO_B200_ORDERING=natural|mindeg|nd overrides). */
int  qpdo_dev_create(qpdo_dev **out, const qpdo_dev_problem *p, int scaling_iters, int ordering,
                     int device);
void qpdo_dev_destroy(qpdo_dev *d);
/* D, Dinv (n), E, Einv (m), c, cinv, ||Dinv q||_inf  -> host (QPDOScaling, qpdo.c:150-170) */
int  qpdo_dev_get_scaling(qpdo_dev *d, double *D, double *Dinv, double *E, double *Einv,
                          double *c, double *cinv, double *norm_q);
/* qpdo_warm_start (qpdo.c:217-299) incl. initialize_mu (iteration.c:98-122); x, y are HOST
 * pointers in user (unscaled) units or NULL */
int  qpdo_dev_warm_start(qpdo_dev *d, const double *x, const double *y, double sigma, int proximal,
                         double mu_min, double *objective);
/* qpdo.c:321-325: sigma, reset_newton = TRUE, active_set_old = FALSE */
int  qpdo_dev_begin_solve(qpdo_dev *d, double sigma);
/* compute_outer/inner_residuals (+_norm) (iteration.c:30-93, termination.c:35-77); copies the
 * scalar block to *out (synchronises the stream) */
int  qpdo_dev_residuals(qpdo_dev *d, int proximal, qpdo_dev_scalars *out);
/* update_iterate (iteration.c:11-25) = newton_direction (newton.c:13-91, with the factor /
 * update / solve of cholmod_interface.c:8-102 done as a numeric refactorization with row weights
 * omega) + exact_linesearch (linesearch.c:8-158) + the five axpys.  Asynchronous. */
int  qpdo_dev_newton_step(qpdo_dev *d, int force_reset, int proximal);
/* qpdo.c:370-379 + is_primal_infeasible (termination.c:97-151); *flag = 1 if certified */
int  qpdo_dev_primal_infeasible(qpdo_dev *d, double eps_prim_inf, int *flag);
/* qpdo.c:381-392 + is_dual_infeasible (termination.c:156-216) */
int  qpdo_dev_dual_infeasible(qpdo_dev *d, double eps_dual_inf, int proximal, int *flag);
/* qpdo.c:396-405,426: xbar<-x, ybar<-y, update_mu (iteration.c:127-168, incl. the
 * ldlupdate_mu_changed quirk of cholmod_interface.c:77-93), update_sigma (iteration.c:173-180),
 * res_prim_old <- res_prim.  *sigma is updated in place. */
int  qpdo_dev_outer_update(qpdo_dev *d, int do_mu, int do_sigma, int proximal, double eps_abs,
                           double theta, double delta, double mu_min, double sigma_upd,
                           double sigma_min, double *sigma);
/* compute_objective (iteration.c:185-221) */
int  qpdo_dev_objective(qpdo_dev *d, int proximal, double c_const, double *objective);
/* store_solution (termination.c:82-92) and the host mirrors qpdo_mex.c:237-280 reads; any pointer
 * may be NULL.  active_set is written as 64-bit c_int. */
int  qpdo_dev_store_solution(qpdo_dev *d, double *sol_x, double *sol_y, double *x, double *y,
                             double *dx, double *dy, long *active_set, qpdo_dev_scalars *out);
/* qpdo_update_bounds (qpdo.c:522-544), qpdo_update_q (qpdo.c:549-586); host pointers */
int  qpdo_dev_update_bounds(qpdo_dev *d, const double *l, const double *u);
int  qpdo_dev_update_q(qpdo_dev *d, const double *q, int proximal, double sigma_init,
                       double *sigma, double *c, double *cinv, double *norm_q);
/* qpdo_update_settings with more scaling sweeps (qpdo.c:496-512) */
int  qpdo_dev_rescale(qpdo_dev *d, int extra_iters);

/* diagnostics / tests / bench */
int  qpdo_dev_set_profiling(qpdo_dev *d, int on);
int  qpdo_dev_get_profile(qpdo_dev *d, qpdo_dev_profile *out, int reset);
int  qpdo_dev_get_symbolic_info(qpdo_dev *d, qpdo_dev_symbolic_info *out);
/* copy a named device vector to the host ("x","y","Ax","Qx","Aty","mu","w","dx","dy","omega",
 * "res_prim","res_dual","res_prim_in","res_dual_in","rhs","l","u","q",...): returns length */
long qpdo_dev_get_vector(qpdo_dev *d, const char *name, double *out, long cap);
/* K = Q + sigma I + A' diag(omega) A with the CURRENT omega/sigma: factor, then solve K x = b for
 * a host right-hand side (test hook for the factor/solve kernels; returns the K x - b data the
 * caller needs: x only) */
int  qpdo_dev_test_factor_solve(qpdo_dev *d, const double *omega, double sigma, const double *b,
                                double *x, float *factor_ms, float *solve_ms);
/* y = A x, y = A' x, y = Q x on the device-resident (scaled) matrices; which: 0 A, 1 A', 2 Q */
int  qpdo_dev_test_spmv(qpdo_dev *d, int which, const double *x, double *y);
/* exact_linesearch (linesearch.c:8-158) alone on host vectors of the workspace's sizes (dx, Qdx, df: n;
 * the rest: m): test hook for the breakpoint build / sort / sweep kernels; overwrites the iterate
 * vectors of the workspace, so call it on a workspace that is not solved afterwards */
int  qpdo_dev_test_linesearch(qpdo_dev *d, const double *dx, const double *Qdx, const double *df,
                              const double *dy, const double *y, const double *mu, const double *sqrt_mu,
                              const double *Adx, const double *w, const double *l, const double *u,
                              double *tau);

/* K x = b on the factorization that is in place, WITHOUT allocating: the supernodal triangular solves of
 * qpdo_dev_test_factor_solve / the Newton step (k_fwd, k_bwd) with their scratch (right-hand side, permuted vector,
 * solution, per-front update vectors: 3 n + front rows doubles) in one static device array of
 * qpdo_dev_solve_noalloc_max_doubles() = 524,288 doubles (4 MiB per GPU, held until the library is unloaded).
 * No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  b and x are host or device
 * pointers of n doubles.  Calls no cudaMalloc / cudaFree / allocating routine and launches the kernels directly (no
 * graph instantiation).  Returns 0; -2 with a message naming the limit when the problem needs more scratch than the
 * array holds; -1 on other errors (no factorization yet, CUDA errors).  All calls on one GPU share the array, which
 * keeps the last call's contents (cleared per call where a solve reads before writing): calls that overlap on
 * the GPU race. */
int  qpdo_dev_solve_noalloc(qpdo_dev *d, const double *b, double *x);
long qpdo_dev_solve_noalloc_max_doubles(void);

/* measured FP64 peaks of `device` (< 0: current), TFLOP/s: register-resident mma.sync.m8n8k4.f64 (DMMA) and DFMA
 * loops over all SMs, best of four timed launches.  The roofline denominators of the numeric factorization. */
int  qpdo_dev_measure_fp64(int device, double *dmma_tflops, double *dfma_tflops);

#ifdef __cplusplus
}
#endif
#endif
