/* qpdo_oracle.c -- TEST INFRASTRUCTURE ONLY (see qpdo_oracle.h).
 *
 * Plain-C CPU restatement of QPDO's Newton hot path.  Every function cites the reference file:line
 * it follows (paths relative to the reference tree aldma/qpdo @ 95399ba).  Compiled with
 * -ffp-contract=off so that a + s*b rounds in two steps exactly like the reference's x86-64 build.
 *
 * The only place where the arithmetic is NOT the reference's is the linear solve: CHOLMOD's
 * analyze/factorize/updown/solve (third party, vendored SuiteSparse 5.8.1 / CHOLMOD 3.0.14) is
 * restated mathematically as an envelope LDL' of K = Q + sigma I + A' diag(omega) A in natural
 * order, where omega tracks the row weights the reference's factor embodies.
 */
#include "qpdo_oracle.h"
#include <stdlib.h>
#include <string.h>
#include <math.h>
#include <time.h>

#define ORACLE_INFTY 1e20                    /* include/constants.h:38 */
#define MAX_RANK_UPDATE 100                  /* include/constants.h:69 */
#define MIN_SCALING 1e-9                     /* include/constants.h:62 */

#define c_max(a, b) (((a) > (b)) ? (a) : (b))     /* include/global_opts.h */
#define c_min(a, b) (((a) < (b)) ? (a) : (b))
#define c_absval(x) (((x) < 0) ? -(x) : (x))

void oracle_default_settings(oracle_settings *s)
{   /* include/constants.h:44-67 via src/qpdo.c:24-44 */
    s->max_time = ORACLE_INFTY; s->max_iter = 10000; s->inner_max_iter = 1000;
    s->eps_abs = 1e-6; s->eps_abs_in = 1e0; s->eps_prim_inf = 1e-6; s->eps_dual_inf = 1e-6;
    s->rho = 0.1; s->theta = 0.25; s->delta = 1e-2; s->mu_min = 1e-9;
    s->proximal = 1; s->sigma_init = 1e-3; s->sigma_upd = 1e-1; s->sigma_min = 1e-7;
    s->scaling = 10; s->verbose = 1; s->print_interval = 1; s->reset_newton_iter = 1000;
}

/* ------------------------------------------------------------------ src/lin_alg.c */
static double vec_prod(const double *a, const double *b, long n)
{   /* lin_alg.c:59-71 */
    double prod = 0.0;
    long i = 0;
    if (n >= 4) for (; i <= n - 4; i += 4) prod += (a[i] * b[i] + a[i + 1] * b[i + 1] + a[i + 2] * b[i + 2] + a[i + 3] * b[i + 3]);
    for (; i < n; i++) prod += a[i] * b[i];
    return prod;
}
static double vec_norm_inf(const double *a, long n)
{   /* lin_alg.c:107-140 (a max: order independent) */
    double mx = 0.0;
    for (long i = 0; i < n; i++) { double s = c_absval(a[i]); mx = s > mx ? s : mx; }
    return mx;
}
static void vec_add_scaled(const double *a, const double *b, double *c, double sc, long n)
{   /* lin_alg.c:93-98 */
    for (long i = 0; i < n; i++) c[i] = a[i] + sc * b[i];
}
static void vec_ew_prod(const double *a, const double *b, double *c, long n)
{   /* lin_alg.c:76-81 */
    for (long i = 0; i < n; i++) c[i] = a[i] * b[i];
}
static void vec_self_mult_scalar(double *a, double sc, long n) { for (long i = 0; i < n; i++) a[i] *= sc; }
static void vec_ew_mid_vec(const double *a, const double *bmin, const double *bmax, double *c, long n)
{   /* lin_alg.c:163-168 */
    for (long i = 0; i < n; i++) c[i] = c_max(bmin[i], c_min(a[i], bmax[i]));
}

/* ------------------------------------------------------------------ cholmod_sdmult restated */
void qpdo_oracle_spmv(int which, long nrow, long ncol, const long *Ap, const long *Ai, const double *Ax,
                      const double *x, double *y)
{
    long j, p;
    if (which == 0) {               /* y = A x: t_cholmod_sdmult.c:297-320 (CSC scatter) */
        for (j = 0; j < nrow; j++) y[j] = 0.0;
        for (j = 0; j < ncol; j++) { double xj = x[j]; for (p = Ap[j]; p < Ap[j + 1]; p++) y[Ai[p]] += Ax[p] * xj; }
    } else if (which == 1) {        /* y = A' x: t_cholmod_sdmult.c:141-165 */
        for (j = 0; j < ncol; j++) {
            double yj = 0.0;
            for (p = Ap[j]; p < Ap[j + 1]; p++) yj += Ax[p] * x[Ai[p]];
            y[j] = 0.0 + yj;
        }
    } else {                        /* y = Q x, lower stored: t_cholmod_sdmult.c:440-484 */
        for (j = 0; j < ncol; j++) y[j] = 0.0;
        for (j = 0; j < ncol; j++) {
            double yj = 0.0, xj = x[j];
            for (p = Ap[j]; p < Ap[j + 1]; p++) {
                long i = Ai[p];
                if (i == j) y[i] += Ax[p] * xj;
                else if (i > j) { double aij = Ax[p]; y[i] += aij * xj; yj += aij * x[i]; }
            }
            y[j] += yj;
        }
    }
}

/* ------------------------------------------------------------------ src/scaling.c:24-91 */
void qpdo_oracle_scale_data(long n, long m, const long *Qp, const long *Qi, double *Qx,
                            const long *Ap, const long *Ai, double *Ax, double *q, double *l, double *u,
                            long iters, double *D, double *E, double *c)
{
    double *Dt = (double *)malloc(sizeof(double) * (n ? n : 1)), *Et = (double *)malloc(sizeof(double) * (m ? m : 1));
    long i, j, k, it;
    for (i = 0; i < n; i++) D[i] = 1.0;
    for (i = 0; i < m; i++) E[i] = 1.0;
    for (it = 0; it < iters; it++) {
        for (j = 0; j < n; j++) Dt[j] = 0.0;                       /* cholmod_interface.c:162-175 */
        for (j = 0; j < m; j++) Et[j] = 0.0;                       /* cholmod_interface.c:180-199 */
        for (j = 0; j < n; j++) for (k = Ap[j]; k < Ap[j + 1]; k++) {
            double v = c_absval(Ax[k]);
            Dt[j] = c_max(v, Dt[j]); Et[Ai[k]] = c_max(v, Et[Ai[k]]);
        }
        for (i = 0; i < n; i++) { Dt[i] = Dt[i] < MIN_SCALING ? 1.0 : Dt[i]; Dt[i] = sqrt(Dt[i]); Dt[i] = 1.0 / Dt[i]; }
        for (i = 0; i < m; i++) { Et[i] = Et[i] < MIN_SCALING ? 1.0 : Et[i]; Et[i] = sqrt(Et[i]); Et[i] = 1.0 / Et[i]; }
        for (j = 0; j < n; j++) for (k = Ap[j]; k < Ap[j + 1]; k++) Ax[k] *= Et[Ai[k]];    /* cholmod_scale.c:137-152 */
        for (j = 0; j < n; j++) { double t = Dt[j]; for (k = Ap[j]; k < Ap[j + 1]; k++) Ax[k] *= t; }   /* :154-172 */
        for (i = 0; i < n; i++) D[i] = D[i] * Dt[i];
        for (i = 0; i < m; i++) E[i] = E[i] * Et[i];
    }
    for (j = 0; j < n; j++) { double t = D[j]; for (k = Qp[j]; k < Qp[j + 1]; k++) Qx[k] *= t * D[Qi[k]]; }   /* cholmod_scale.c:174-191 */
    for (i = 0; i < n; i++) q[i] = D[i] * q[i];
    /* cost scaling: work->Qx is all zero at setup (scaling.c:72-73) */
    for (i = 0; i < n; i++) Dt[i] = 0.0 + 1 * q[i];
    *c = 1 / c_max(1.0, vec_norm_inf(Dt, n));
    for (i = 0; i < n; i++) q[i] *= *c;
    for (k = 0; k < Qp[n]; k++) Qx[k] *= *c;
    for (i = 0; i < m; i++) { l[i] = E[i] * l[i]; u[i] = E[i] * u[i]; }
    free(Dt); free(Et);
}

/* ------------------------------------------------------------------ src/linesearch.c */
typedef struct { double x; size_t i; } array_element;       /* include/types.h:14-17 */
static int compare(const void *a, const void *b)
{   /* linesearch.c:204-211 */
    double f = ((const array_element *)a)->x, s = ((const array_element *)b)->x;
    if (f > s) return 1;
    if (f < s) return -1;
    return 0;
}
/* scratch: delta, alpha, t2m (2m doubles each), taus (2m), idxL/P/J (2m longs each), tm (m) */
static double linesearch_core(long n, long m, const double *dx, const double *Qdx, const double *df,
                              const double *dy, const double *y, const double *mu, const double *sqrt_mu,
                              const double *Adx, const double *w, const double *l, const double *u,
                              double *tm, double *delta, double *alpha, double *t2m, array_element *taus,
                              long *idxL, long *idxP, long *idxJ, double *eta_out, double *beta_out)
{
    long i, M2 = 2 * m;
    double eta, beta, a, b;
    /* exact_linesearch, linesearch.c:8-52 */
    vec_ew_prod(dy, mu, tm, m);
    vec_self_mult_scalar(tm, 0.5, m);
    eta = vec_prod(dy, tm, m); eta += vec_prod(dx, Qdx, n); eta *= 0.5;
    beta = vec_prod(y, tm, m); beta += vec_prod(dx, df, n); beta *= 0.5;
    vec_add_scaled(Adx, tm, tm, -1, m);
    vec_ew_prod(tm, sqrt_mu, tm, m);
    memcpy(delta + m, tm, sizeof(double) * m);
    vec_self_mult_scalar(tm, -1, m);
    memcpy(delta, tm, sizeof(double) * m);
    vec_add_scaled(w, l, tm, -1, m); vec_ew_prod(tm, sqrt_mu, tm, m); memcpy(alpha, tm, sizeof(double) * m);
    vec_add_scaled(u, w, tm, -1, m); vec_ew_prod(tm, sqrt_mu, tm, m); memcpy(alpha + m, tm, sizeof(double) * m);
    /* psi'(1) is computed but unused by the reference (linesearch.c:40-49): skipped */
    if (eta_out) *eta_out = eta;
    if (beta_out) *beta_out = beta;
    /* pwa_linesearch, linesearch.c:74-158 */
    size_t nL = 0;
    for (i = 0; i < M2; i++) { t2m[i] = alpha[i] / delta[i]; }
    for (i = 0; i < M2; i++) { if (t2m[i] > 0) { idxL[i] = 1; taus[nL].x = t2m[i]; taus[nL].i = (size_t)i; nL++; } else idxL[i] = 0; }
    for (i = 0; i < M2; i++) idxP[i] = delta[i] > 0 ? 1 : 0;
    for (i = 0; i < M2; i++) idxJ[i] = (idxP[i] + idxL[i]) == 1;
    {
        double p1 = 0.0, p2 = 0.0;             /* vec_prod_ind, linesearch.c:190-199 */
        for (i = 0; i < M2; i++) if (idxJ[i]) p1 += delta[i] * delta[i];
        for (i = 0; i < M2; i++) if (idxJ[i]) p2 += delta[i] * alpha[i];
        a = eta + p1; b = beta - p2;
    }
    if (nL == 0) return -b / a;
    qsort(taus, nL, sizeof(array_element), compare);
    if (b + a * taus[0].x > 0) return -b / a;
    size_t k = 0, iz;
    while (k < nL - 1) {
        iz = taus[k].i;
        if (idxP[iz]) { a = a + delta[iz] * delta[iz]; b = b - delta[iz] * alpha[iz]; }
        else { a = a - delta[iz] * delta[iz]; b = b + delta[iz] * alpha[iz]; }
        k++;
        if (b + a * taus[k].x > 0) return -b / a;
    }
    iz = taus[k].i;
    if (idxP[iz]) { a = a + delta[iz] * delta[iz]; b = b - delta[iz] * alpha[iz]; }
    else { a = a - delta[iz] * delta[iz]; b = b + delta[iz] * alpha[iz]; }
    return -b / a;
}

double qpdo_oracle_linesearch(long n, long m, const double *dx, const double *Qdx, const double *df,
                              const double *dy, const double *y, const double *mu, const double *sqrt_mu,
                              const double *Adx, const double *w, const double *l, const double *u)
{
    long M2 = 2 * m + 1;
    double *buf = (double *)malloc(sizeof(double) * (size_t)(m + 1 + 3 * M2));
    array_element *taus = (array_element *)malloc(sizeof(array_element) * (size_t)M2);
    long *idx = (long *)malloc(sizeof(long) * (size_t)(3 * M2));
    double tau = linesearch_core(n, m, dx, Qdx, df, dy, y, mu, sqrt_mu, Adx, w, l, u, buf, buf + m + 1, buf + m + 1 + M2,
                                 buf + m + 1 + 2 * M2, taus, idx, idx + M2, idx + 2 * M2, NULL, NULL);
    free(buf); free(taus); free(idx);
    return tau;
}

/* ------------------------------------------------------------------ envelope LDL' of K (natural order) */
typedef struct {
    long n, m;
    long *first;         /* first column of row i's envelope */
    long *rowp;          /* offset of row i: entries (i, first[i]..i) */
    double *K;           /* envelope values; after factorization L below the diagonal, D on it */
    double *ytmp;
    long *Rp, *Rj; double *Rx;      /* CSR of (scaled) A */
} envelope;

static void env_build(envelope *E, long n, long m, const long *Qp, const long *Qi, const long *Ap, const long *Ai, const double *Ax)
{
    long i, j, k, nnzA = Ap[n];
    E->n = n; E->m = m;
    E->Rp = (long *)calloc((size_t)m + 1, sizeof(long));
    E->Rj = (long *)malloc(sizeof(long) * (size_t)(nnzA ? nnzA : 1));
    E->Rx = (double *)malloc(sizeof(double) * (size_t)(nnzA ? nnzA : 1));
    for (k = 0; k < nnzA; k++) E->Rp[Ai[k] + 1]++;
    for (i = 0; i < m; i++) E->Rp[i + 1] += E->Rp[i];
    {
        long *nx = (long *)malloc(sizeof(long) * (size_t)(m ? m : 1));
        for (i = 0; i < m; i++) nx[i] = E->Rp[i];
        for (j = 0; j < n; j++) for (k = Ap[j]; k < Ap[j + 1]; k++) { long d = nx[Ai[k]]++; E->Rj[d] = j; E->Rx[d] = Ax[k]; }
        free(nx);
    }
    E->first = (long *)malloc(sizeof(long) * (size_t)n);
    for (i = 0; i < n; i++) E->first[i] = i;
    for (j = 0; j < n; j++) for (k = Qp[j]; k < Qp[j + 1]; k++) { i = Qi[k]; if (i >= j && j < E->first[i]) E->first[i] = j; }
    for (i = 0; i < m; i++) if (E->Rp[i + 1] > E->Rp[i]) {
        long c0 = E->Rj[E->Rp[i]];           /* columns ascending inside a CSR row */
        for (k = E->Rp[i]; k < E->Rp[i + 1]; k++) if (c0 < E->first[E->Rj[k]]) E->first[E->Rj[k]] = c0;
    }
    E->rowp = (long *)malloc(sizeof(long) * (size_t)(n + 1));
    E->rowp[0] = 0;
    for (i = 0; i < n; i++) E->rowp[i + 1] = E->rowp[i] + (i - E->first[i] + 1);
    E->K = (double *)malloc(sizeof(double) * (size_t)E->rowp[n]);
    E->ytmp = (double *)malloc(sizeof(double) * (size_t)n);
}
static void env_free(envelope *E) { free(E->first); free(E->rowp); free(E->K); free(E->ytmp); free(E->Rp); free(E->Rj); free(E->Rx); }
#define ENV(E, i, j) ((E)->K[(E)->rowp[i] + ((j) - (E)->first[i])])

/* K = Q + diag_shift I + A' diag(omega) A, then LDL' in place (replaces cholmod_aat/add/analyze/
 * factorize_p, cholmod_interface.c:8-52, and the cumulative effect of cholmod_updown, :57-93) */
static void env_factor(envelope *E, const long *Qp, const long *Qi, const double *Qx, double shift, const double *omega)
{
    long n = E->n, m = E->m, i, j, k, p, q;
    memset(E->K, 0, sizeof(double) * (size_t)E->rowp[n]);
    for (j = 0; j < n; j++) for (k = Qp[j]; k < Qp[j + 1]; k++) { i = Qi[k]; if (i >= j) ENV(E, i, j) += Qx[k]; }
    for (i = 0; i < n; i++) ENV(E, i, i) += shift;
    for (i = 0; i < m; i++) {
        double om = omega[i];
        if (om == 0.0) continue;
        for (p = E->Rp[i]; p < E->Rp[i + 1]; p++) {
            double t = om * E->Rx[p];
            for (q = E->Rp[i]; q <= p; q++) ENV(E, E->Rj[p], E->Rj[q]) += t * E->Rx[q];
        }
    }
    for (i = 0; i < n; i++) {
        long fi = E->first[i];
        double *y = E->ytmp, d;
        for (j = fi; j < i; j++) {
            long fj = E->first[j], k0 = fi > fj ? fi : fj;
            double s = ENV(E, i, j);
            for (k = k0; k < j; k++) s -= ENV(E, j, k) * y[k];
            y[j] = s;                         /* = L_ij D_j */
        }
        d = ENV(E, i, i);
        for (j = fi; j < i; j++) { double lij = y[j] / ENV(E, j, j); ENV(E, i, j) = lij; d -= lij * y[j]; }
        ENV(E, i, i) = d;
    }
}
static void env_solve(const envelope *E, const double *b, double *x)
{   /* cholmod_solve(CHOLMOD_LDLt), cholmod_interface.c:98-102 */
    long n = E->n, i, j;
    for (i = 0; i < n; i++) { double s = b[i]; for (j = E->first[i]; j < i; j++) s -= ENV(E, i, j) * x[j]; x[i] = s; }
    for (i = 0; i < n; i++) x[i] /= ENV(E, i, i);
    for (i = n - 1; i >= 0; i--) { double xi = x[i]; for (j = E->first[i]; j < i; j++) x[j] -= ENV(E, i, j) * xi; }
}

/* ------------------------------------------------------------------ the solver */
static double now_seconds(void) { struct timespec t; clock_gettime(CLOCK_MONOTONIC, &t); return (double)t.tv_sec + 1e-9 * (double)t.tv_nsec; }

int qpdo_oracle_solve(long n, long m, const long *Qp, const long *Qi, const double *Qx_in,
                      const long *Ap, const long *Ai, const double *Ax_in,
                      const double *q_in, double cconst, const double *l_in, const double *u_in,
                      const oracle_settings *st, const double *x0, const double *y0,
                      double *x_out, double *y_out, double *dx_cert, double *dy_cert, long *active_out,
                      oracle_info *info, oracle_trace *trace, long cap_trace)
{
    long i, k, iter, oter = 0, iter_old = 0, ntr = 0, status = -10;   /* QPDO_UNSOLVED */
    const long nnzQ = Qp[n], nnzA = Ap[n], scaling = st->scaling, proximal = st->proximal;
    for (i = 0; i < m; i++) if (l_in[i] > u_in[i]) return -1;         /* validate.c:9-31 */
#define NV(name, len) double *name = (double *)calloc((size_t)((len) ? (len) : 1), sizeof(double))
    NV(Qv, nnzQ); NV(Av, nnzA); NV(q, n); NV(l, m); NV(u, m);
    NV(x, n); NV(y, m); NV(xbar, n); NV(ybar, m); NV(Ax, m); NV(Qx, n); NV(Aty, n);
    NV(temp_m, m); NV(temp_n, n); NV(mu, m); NV(sqrt_mu, m); NV(w, m); NV(res_prim, m); NV(res_prim_old, m);
    NV(res_dual, n); NV(res_prim_in, m); NV(res_dual_in, n); NV(df, n); NV(dx, n); NV(dy, m); NV(Qdx, n);
    NV(Adx, m); NV(Atdy, n); NV(rhs, n); NV(D, n); NV(Dinv, n); NV(E, m); NV(Einv, m); NV(omega, m);
    NV(ls_delta, 2 * m); NV(ls_alpha, 2 * m); NV(t2m, 2 * m); NV(At_scale, m);
#undef NV
    array_element *taus = (array_element *)malloc(sizeof(array_element) * (size_t)(2 * m + 1));
    long *act = (long *)calloc((size_t)(m ? m : 1), sizeof(long)), *act_old = (long *)calloc((size_t)(m ? m : 1), sizeof(long));
    long *idxs = (long *)malloc(sizeof(long) * (size_t)(6 * m + 3));
    double c = 1.0, cinv = 1.0, sigma, eps_in, tau = 0.0, sqrt_mu_min;
    double rpn = 0, rdn = 0, rpin = 0, rdin = 0, objective = 0;
    long reset_newton, n_active = 0, n_enter = 0, n_leave = 0, n_mu_changed = 0, factor_mode = -1;
    envelope Env;
    double t_start;

    /* ---- qpdo_setup (qpdo.c:49-212) */
    memcpy(Qv, Qx_in, sizeof(double) * (size_t)nnzQ); memcpy(Av, Ax_in, sizeof(double) * (size_t)nnzA);
    memcpy(q, q_in, sizeof(double) * (size_t)n); memcpy(l, l_in, sizeof(double) * (size_t)m); memcpy(u, u_in, sizeof(double) * (size_t)m);
    sigma = st->sigma_init;
    if (scaling) {
        qpdo_oracle_scale_data(n, m, Qp, Qi, Qv, Ap, Ai, Av, q, l, u, scaling, D, E, &c);
        for (i = 0; i < n; i++) Dinv[i] = 1.0 / D[i];
        for (i = 0; i < m; i++) Einv[i] = 1.0 / E[i];
        cinv = 1.0 / c;
    }
    env_build(&Env, n, m, Qp, Qi, Ap, Ai, Av);

    /* ---- qpdo_warm_start (qpdo.c:217-299) */
    if (x0) {
        memcpy(x, x0, sizeof(double) * (size_t)n);
        if (scaling) vec_ew_prod(x, Dinv, x, n);
        memcpy(xbar, x, sizeof(double) * (size_t)n); memcpy(dx, x, sizeof(double) * (size_t)n);
        qpdo_oracle_spmv(2, n, n, Qp, Qi, Qv, dx, Qdx);
        if (proximal) vec_add_scaled(Qdx, x, Qx, sigma, n); else memcpy(Qx, Qdx, sizeof(double) * (size_t)n);
        qpdo_oracle_spmv(0, m, n, Ap, Ai, Av, dx, Adx);
        memcpy(Ax, Adx, sizeof(double) * (size_t)m);
    }
    if (y0) {
        memcpy(y, y0, sizeof(double) * (size_t)m);
        if (scaling) { vec_ew_prod(y, Einv, y, m); vec_self_mult_scalar(y, c, m); }
        memcpy(ybar, y, sizeof(double) * (size_t)m); memcpy(dy, y, sizeof(double) * (size_t)m);
        qpdo_oracle_spmv(1, m, n, Ap, Ai, Av, dy, Atdy);
        memcpy(Aty, Atdy, sizeof(double) * (size_t)n);
    }
    {   /* initialize_mu (iteration.c:98-122) */
        double f = 0.5 * vec_prod(x, Qx, n) + vec_prod(q, x, n);
        vec_ew_mid_vec(Ax, l, u, temp_m, m);
        vec_add_scaled(Ax, temp_m, temp_m, -1, m);
        for (i = 0; i < m; i++)
            mu[i] = c_max(1e-3, c_min(1e3, 0.1 * c_max(1, 0.5 * temp_m[i] * temp_m[i]) / c_max(1, c_absval(f))));
        for (i = 0; i < m; i++) { sqrt_mu[i] = sqrt(mu[i]); sqrt_mu[i] = 1.0 / sqrt_mu[i]; }
        sqrt_mu_min = 1 / sqrt(st->mu_min);
    }

    /* ---- qpdo_solve (qpdo.c:304-476) */
    eps_in = st->eps_abs_in;
    sigma = st->sigma_init;
    reset_newton = 1;
    t_start = now_seconds();
    for (iter = 0; iter < st->max_iter; iter++) {
        /* compute_outer_residuals (iteration.c:30-60) */
        if (scaling) {
            vec_ew_prod(E, y, temp_m, m); vec_self_mult_scalar(temp_m, cinv, m); vec_ew_prod(E, temp_m, temp_m, m);
            vec_add_scaled(Ax, temp_m, temp_m, 1, m);
        } else vec_add_scaled(Ax, y, temp_m, 1, m);
        vec_ew_mid_vec(temp_m, l, u, temp_m, m);
        vec_add_scaled(Ax, temp_m, res_prim, -1, m);
        vec_add_scaled(Qx, q, df, 1, n);
        if (proximal) { vec_add_scaled(df, x, res_dual, -sigma, n); vec_add_scaled(res_dual, Aty, res_dual, 1, n); }
        else vec_add_scaled(df, Aty, res_dual, 1, n);
        /* compute_outer_residuals_norm (termination.c:35-53) */
        if (scaling) {
            vec_ew_prod(Einv, res_prim, temp_m, m); rpn = vec_norm_inf(temp_m, m);
            vec_ew_prod(Dinv, res_dual, temp_n, n); rdn = vec_norm_inf(temp_n, n); rdn *= cinv;
        } else { rpn = vec_norm_inf(res_prim, m); rdn = vec_norm_inf(res_dual, n); }
        /* compute_inner_residuals (iteration.c:65-93) */
        for (i = 0; i < m; i++) w[i] = Ax[i] + mu[i] * (ybar[i] - 0.5 * y[i]);
        vec_ew_mid_vec(w, l, u, temp_m, m);
        for (i = 0; i < m; i++) res_prim_in[i] = Ax[i] + mu[i] * (ybar[i] - y[i]) - temp_m[i];
        if (proximal) vec_add_scaled(df, xbar, df, -sigma, n);
        vec_add_scaled(df, Aty, res_dual_in, 1, n);
        /* compute_inner_residuals_norm (termination.c:58-77) */
        if (scaling) {
            vec_ew_prod(Einv, res_prim_in, temp_m, m); rpin = vec_norm_inf(temp_m, m);
            vec_ew_prod(Dinv, res_dual_in, temp_n, n); rdin = vec_norm_inf(temp_n, n); rdin *= cinv;
        } else { rpin = vec_norm_inf(res_prim_in, m); rdin = vec_norm_inf(res_dual_in, n); }

        /* check_outer_optimality (termination.c:11-23) */
        if (rpn > ORACLE_INFTY || rdn > ORACLE_INFTY) { status = -1; break; }
        if (rpn <= st->eps_abs && rdn <= st->eps_abs) { status = 1; break; }

        if (((iter > iter_old + 1) && (rpin <= eps_in) && (rdin <= eps_in)) || (iter == iter_old + st->inner_max_iter)) {
            if (iter < iter_old + st->inner_max_iter) {
                if (st->eps_prim_inf > 0) {                          /* qpdo.c:370-379 */
                    double en, oob = 0;
                    vec_add_scaled(y, ybar, dy, -1, m);
                    qpdo_oracle_spmv(1, m, n, Ap, Ai, Av, dy, Atdy);
                    /* is_primal_infeasible (termination.c:97-151) */
                    if (scaling) { vec_ew_prod(E, dy, temp_m, m); en = st->eps_prim_inf * vec_norm_inf(temp_m, m); }
                    else en = st->eps_prim_inf * vec_norm_inf(dy, m);
                    if (en != 0) {
                        if (scaling) vec_ew_prod(Dinv, Atdy, Atdy, n);
                        for (i = 0; i < m; i++) {
                            double lim = scaling ? E[i] * ORACLE_INFTY : ORACLE_INFTY;
                            oob += (u[i] < lim) ? u[i] * c_max(dy[i], 0) : 0;
                            oob += (l[i] > -lim) ? l[i] * c_min(dy[i], 0) : 0;
                        }
                        if (vec_norm_inf(Atdy, n) <= en && oob <= -en) {
                            status = -3;
                            if (scaling) { vec_self_mult_scalar(dy, cinv, m); vec_ew_prod(E, dy, dy, m); }
                            break;
                        }
                    }
                }
                if (st->eps_dual_inf > 0) {                          /* qpdo.c:381-392 */
                    double en;
                    int viol = 0;
                    vec_add_scaled(x, xbar, dx, -1, n);
                    qpdo_oracle_spmv(2, n, n, Qp, Qi, Qv, dx, Qdx);
                    qpdo_oracle_spmv(0, m, n, Ap, Ai, Av, dx, Adx);
                    /* is_dual_infeasible (termination.c:156-216) */
                    if (scaling) { vec_ew_prod(D, dx, temp_n, n); en = st->eps_dual_inf * vec_norm_inf(temp_n, n); }
                    else en = st->eps_dual_inf * vec_norm_inf(dx, n);
                    if (en != 0) {
                        if (scaling) vec_ew_prod(Einv, Adx, Adx, m);
                        for (k = 0; k < m && !viol; k++) {
                            double lim = scaling ? E[k] * ORACLE_INFTY : ORACLE_INFTY;
                            if ((u[k] < lim && Adx[k] >= en) || (l[k] > -lim && Adx[k] <= -en)) viol = 1;
                        }
                        if (!viol) {
                            double lim2 = scaling ? c * en : en;
                            if (proximal) vec_add_scaled(Qdx, dx, Qdx, -sigma * tau, n);
                            if (vec_norm_inf(Qdx, n) <= lim2 && vec_prod(q, dx, n) <= -lim2) {
                                status = -4;
                                if (scaling) vec_ew_prod(D, dx, dx, n);
                                break;
                            }
                        }
                    }
                }
            }
            memcpy(xbar, x, sizeof(double) * (size_t)n); memcpy(ybar, y, sizeof(double) * (size_t)m);
            if (oter > 0 && rpn > st->eps_abs) {
                /* update_mu (iteration.c:127-168) */
                double nrm = vec_norm_inf(res_prim, m);
                n_mu_changed = 0;
                for (k = 0; k < m; k++) {
                    At_scale[k] = 1.0;
                    idxs[k] = 0;
                    if (c_absval(res_prim[k]) > c_max(st->eps_abs, st->theta * c_absval(res_prim_old[k]))) {
                        double mu_factor = 1.0 / c_min(1.0, st->delta * nrm / c_absval(res_prim[k]));
                        double mu_new = mu[k] / mu_factor;
                        if (mu_new >= st->mu_min) {
                            if (mu[k] != mu_new) { idxs[k] = 1; n_mu_changed++; }
                            mu[k] = mu_new;
                            mu_factor = sqrt(mu_factor);
                            sqrt_mu[k] = mu_factor * sqrt_mu[k];
                            At_scale[k] = mu_factor;
                        } else {
                            if (mu[k] != st->mu_min) { idxs[k] = 1; n_mu_changed++; }
                            mu[k] = st->mu_min;
                            At_scale[k] = sqrt_mu_min / sqrt_mu[k];
                            sqrt_mu[k] = sqrt_mu_min;
                        }
                    }
                }
                if ((proximal && sigma > st->sigma_min) || (n_mu_changed > 0.25 * MAX_RANK_UPDATE)) reset_newton = 1;
                else if (n_mu_changed != 0) {
                    /* ldlupdate_mu_changed (cholmod_interface.c:77-93): rank update with the columns
                     * sqrt_mu_new * sqrt(1 - 1/s^2) * a_k for EVERY changed row, active or not */
                    for (k = 0; k < m; k++) if (idxs[k]) {
                        double f = sqrt(1 - 1 / (At_scale[k] * At_scale[k]));
                        double col = sqrt_mu[k] * f;
                        omega[k] = omega[k] + col * col;
                    }
                }
            }
            if (proximal && oter > 0 && rdn > st->eps_abs) {
                /* update_sigma (iteration.c:173-180) */
                if (sigma > st->sigma_min) {
                    double sigma_old = sigma;
                    sigma = c_max(sigma * st->sigma_upd, st->sigma_min);
                    reset_newton = 1;
                    vec_add_scaled(Qx, x, Qx, sigma - sigma_old, n);
                }
            }
            if (iter < iter_old + st->inner_max_iter) eps_in = c_max(st->rho * eps_in, 0.1 * st->eps_abs);
            memcpy(res_prim_old, res_prim, sizeof(double) * (size_t)m);
            if (trace && ntr < cap_trace) {
                oracle_trace *T = &trace[ntr++];
                T->iter = iter; T->oter = oter; T->kind = 1; T->res_prim_norm = rpn; T->res_dual_norm = rdn;
                T->res_prim_in_norm = rpin; T->res_dual_in_norm = rdin; T->tau = tau; T->sigma = sigma;
                T->n_active = n_active; T->n_enter = n_mu_changed; T->n_leave = 0; T->factor_mode = -1;
            }
            oter++;
            iter_old = iter;
        } else {
            if (st->reset_newton_iter && iter % st->reset_newton_iter == 0) reset_newton = 1;
            /* ---- update_iterate (iteration.c:11-25) -> newton_direction (newton.c:13-91) */
            n_active = 0;
            for (i = 0; i < m; i++) { act[i] = (w[i] <= l[i]) || (w[i] >= u[i]); n_active += act[i]; }    /* newton.c:96-107 */
            n_enter = n_leave = 0;
            for (i = 0; i < m; i++) { if (act[i] && !act_old[i]) n_enter++; if (!act[i] && act_old[i]) n_leave++; }   /* :112-126 */
            if ((reset_newton && n_active) || (n_enter + n_leave) > MAX_RANK_UPDATE) {
                reset_newton = 0; factor_mode = 0;
                for (i = 0; i < m; i++) omega[i] = act[i] ? 1.0 / mu[i] : 0.0;                /* ldlcholQAtmuA */
            } else if (n_active) {
                factor_mode = 1;
                for (i = 0; i < m; i++) {                                                   /* updown +- */
                    if (act[i] && !act_old[i]) omega[i] = omega[i] + 1.0 / mu[i];
                    if (!act[i] && act_old[i]) omega[i] = omega[i] - 1.0 / mu[i];
                }
            } else {
                factor_mode = 2;
                for (i = 0; i < m; i++) omega[i] = 0.0;                                     /* ldlchol(Q) */
            }
            env_factor(&Env, Qp, Qi, Qv, proximal ? sigma : 0.0, omega);
            for (i = 0; i < m; i++) { dy[i] = res_prim_in[i] / mu[i]; if (!act[i]) dy[i] *= 2; }
            qpdo_oracle_spmv(1, m, n, Ap, Ai, Av, dy, Atdy);
            for (i = 0; i < n; i++) rhs[i] = -res_dual_in[i] - Atdy[i];
            env_solve(&Env, rhs, dx);
            qpdo_oracle_spmv(2, n, n, Qp, Qi, Qv, dx, Qdx);
            if (proximal) vec_add_scaled(Qdx, dx, Qdx, sigma, n);
            qpdo_oracle_spmv(0, m, n, Ap, Ai, Av, dx, Adx);
            for (i = 0; i < m; i++) if (act[i]) dy[i] += (Adx[i] / mu[i]);
            qpdo_oracle_spmv(1, m, n, Ap, Ai, Av, dy, Atdy);
            memcpy(act_old, act, sizeof(long) * (size_t)m);
            /* ---- exact_linesearch (linesearch.c:8-158) */
            tau = linesearch_core(n, m, dx, Qdx, df, dy, y, mu, sqrt_mu, Adx, w, l, u, temp_m, ls_delta, ls_alpha, t2m, taus,
                                  idxs, idxs + 2 * m + 1, idxs + 4 * m + 2, NULL, NULL);
            vec_add_scaled(x, dx, x, tau, n); vec_add_scaled(y, dy, y, tau, m);
            vec_add_scaled(Qx, Qdx, Qx, tau, n); vec_add_scaled(Ax, Adx, Ax, tau, m); vec_add_scaled(Aty, Atdy, Aty, tau, n);
            if (trace && ntr < cap_trace) {
                oracle_trace *T = &trace[ntr++];
                T->iter = iter; T->oter = oter; T->kind = 0; T->res_prim_norm = rpn; T->res_dual_norm = rdn;
                T->res_prim_in_norm = rpin; T->res_dual_in_norm = rdin; T->tau = tau; T->sigma = sigma;
                T->n_active = n_active; T->n_enter = n_enter; T->n_leave = n_leave; T->factor_mode = factor_mode;
            }
        }
        if (now_seconds() - t_start > st->max_time) { status = -6; break; }
    }
    if (status == -10) status = -5;                                  /* QPDO_MAX_ITER_REACHED */
    /* store_solution (termination.c:82-92) + compute_objective (iteration.c:185-221) */
    if (scaling) { vec_self_mult_scalar(y, cinv, m); }
    if (x_out) for (i = 0; i < n; i++) x_out[i] = scaling ? x[i] * D[i] : x[i];
    if (y_out) for (i = 0; i < m; i++) y_out[i] = scaling ? y[i] * E[i] : y[i];
    {
        long j = 0;
        objective = 0;
        if (proximal) {
            if (n >= 4) for (; j <= n - 4; j += 4)
                objective += (0.5 * (Qx[j] - x[j] * sigma) + q[j]) * x[j] + (0.5 * (Qx[j + 1] - x[j + 1] * sigma) + q[j + 1]) * x[j + 1]
                           + (0.5 * (Qx[j + 2] - x[j + 2] * sigma) + q[j + 2]) * x[j + 2] + (0.5 * (Qx[j + 3] - x[j + 3] * sigma) + q[j + 3]) * x[j + 3];
            for (; j < n; j++) objective += (0.5 * (Qx[j] - sigma * x[j]) + q[j]) * x[j];
        } else {
            if (n >= 4) for (; j <= n - 4; j += 4)
                objective += (0.5 * Qx[j] + q[j]) * x[j] + (0.5 * Qx[j + 1] + q[j + 1]) * x[j + 1]
                           + (0.5 * Qx[j + 2] + q[j + 2]) * x[j + 2] + (0.5 * Qx[j + 3] + q[j + 3]) * x[j + 3];
            for (; j < n; j++) objective += (0.5 * Qx[j] + q[j]) * x[j];
        }
        if (scaling) objective *= cinv;
        objective += cconst;
    }
    if (dx_cert) memcpy(dx_cert, dx, sizeof(double) * (size_t)n);
    if (dy_cert) memcpy(dy_cert, dy, sizeof(double) * (size_t)m);
    if (active_out) memcpy(active_out, act, sizeof(long) * (size_t)m);
    if (info) {
        info->iterations = iter; info->oterations = oter; info->status_val = status;
        info->res_prim_norm = rpn; info->res_dual_norm = rdn; info->res_prim_in_norm = rpin; info->res_dual_in_norm = rdin;
        info->objective = objective; info->solve_seconds = now_seconds() - t_start; info->n_trace = ntr;
    }
    env_free(&Env);
    free(Qv); free(Av); free(q); free(l); free(u); free(x); free(y); free(xbar); free(ybar); free(Ax); free(Qx); free(Aty);
    free(temp_m); free(temp_n); free(mu); free(sqrt_mu); free(w); free(res_prim); free(res_prim_old); free(res_dual);
    free(res_prim_in); free(res_dual_in); free(df); free(dx); free(dy); free(Qdx); free(Adx); free(Atdy); free(rhs);
    free(D); free(Dinv); free(E); free(Einv); free(omega); free(ls_delta); free(ls_alpha); free(t2m); free(At_scale);
    free(taus); free(act); free(act_old); free(idxs);
    return 0;
}
