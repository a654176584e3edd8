/* qpdo_api.c -- host C side of the drop-in boundary: the eight qpdo.h entry points.
 *
 * Mirrors the reference's src/qpdo.c (setup 49-212, warm start 217-299, solve loop 304-476,
 * update_* 481-586, cleanup 591-689), src/validate.c and the status strings of src/util.c:50-91.
 * Loop control, validation, ownership and status live here in C; every vector/matrix operation
 * is a call through the C ABI of include/qpdo_cuda.h into the sm_100a kernels.  There is no CPU
 * fallback: if the CUDA layer cannot be brought up, qpdo_setup returns QPDO_NULL.
 */
#define _POSIX_C_SOURCE 200809L
#include "qpdo.h"
#include "qpdo_cuda.h"
#include <stdlib.h>
#include <string.h>
#include <stdio.h>
#include <math.h>
#include <time.h>

struct QPDO_TIMER { struct timespec tic; };

static void tic(QPDOTimer *t) { clock_gettime(CLOCK_MONOTONIC, &t->tic); }
static double toc(QPDOTimer *t)
{
    struct timespec now;
    clock_gettime(CLOCK_MONOTONIC, &now);
    return (double)(now.tv_sec - t->tic.tv_sec) + 1e-9 * (double)(now.tv_nsec - t->tic.tv_nsec);
}
#define eprint(...) do { fprintf(stderr, "ERROR in %s: ", __func__); fprintf(stderr, __VA_ARGS__); fprintf(stderr, "\n"); } while (0)

static qpdo_dev *DEV(const QPDOWorkspace *w) { return (qpdo_dev *)w->chol->LD; }

/* ------------------------------------------------------------------ util.c mirrors -------- */
QPDOSettings *copy_settings(const QPDOSettings *s)
{
    QPDOSettings *c = (QPDOSettings *)malloc(sizeof(QPDOSettings));
    if (c) *c = *s;
    return c;
}

void update_status(QPDOInfo *info, c_int status_val)
{
    const char *s;
    info->status_val = status_val;
    switch (status_val) {                       /* strings: reference src/util.c:56-90 */
    case QPDO_SOLVED:                 s = "solved"; break;
    case QPDO_DUAL_TERMINATED:        s = "dual terminated"; break;
    case QPDO_PRIMAL_INFEASIBLE:      s = "primal infeasible"; break;
    case QPDO_DUAL_INFEASIBLE:        s = "dual infeasible"; break;
    case QPDO_PRIMAL_DUAL_INFEASIBLE: s = "primal-dual infeasible"; break;
    case QPDO_MAX_TIME_REACHED:       s = "max time exceeded"; break;
    case QPDO_MAX_ITER_REACHED:       s = "maximum iterations reached"; break;
    case QPDO_UNSOLVED:               s = "unsolved"; break;
    case QPDO_ERROR:                  s = "error"; break;
    default:                          s = "unrecognised status value"; break;
    }
    strncpy(info->status, s, sizeof info->status - 1);
    info->status[sizeof info->status - 1] = '\0';
}

/* ------------------------------------------------------------------ validate.c mirrors ---- */
static int validate_data(const QPDOData *data)
{
    size_t j;
    if (!data) { eprint("Missing data"); return 0; }
    for (j = 0; j < data->m; j++)
        if (data->l[j] > data->u[j]) {
            eprint("Lower bound at index %d is greater than upper bound: %.4e > %.4e", (int)j, data->l[j], data->u[j]);
            return 0;
        }
    return 1;
}
static int validate_settings(const QPDOSettings *s)
{
    if (!s) { eprint("Missing settings!"); return 0; }
#define REQUIRE(cond, msg) if (!(cond)) { eprint(msg); return 0; }
    REQUIRE(s->max_iter > 0, "max_iter must be positive")
    REQUIRE(s->inner_max_iter > 0, "inner_max_iter must be positive")
    REQUIRE(s->eps_abs > 0, "eps_abs must be positive")
    REQUIRE(s->eps_abs_in > 0, "eps_abs_in must be positive")
    REQUIRE(s->eps_prim_inf >= 0, "eps_prim_inf must be nonnegative")
    REQUIRE(s->eps_dual_inf >= 0, "eps_dual_inf must be nonnegative")
    REQUIRE(s->rho > 0 && s->rho < 1, "rho must be positive and smaller than 1")
    REQUIRE(s->theta > 0 && s->theta <= 1, "theta must be positive and smaller than or equal to 1")
    REQUIRE(s->delta > 0 && s->delta < 1, "delta must be positive and smaller than 1")
    REQUIRE(s->mu_min > 0, "mu_min must be positive")
    REQUIRE(s->proximal == 0 || s->proximal == 1, "proximal must be either 0 or 1")
    REQUIRE(s->sigma_init > 0, "sigma_init must be positive")
    REQUIRE(s->sigma_upd > 0 && s->sigma_upd <= 1, "sigma_upd must be positive and smaller than or equal to 1")
    REQUIRE(!(s->sigma_min > s->sigma_init), "sigma_min must be smaller than or equal to sigma_init")
    REQUIRE(s->scaling >= 0, "scaling must be nonnegative")
    REQUIRE(s->verbose >= 0, "verbose must be either nonnegative")
    REQUIRE(s->print_interval >= 0, "print_interval must be nonnegative")
    REQUIRE(s->reset_newton_iter >= 0, "reset_newton_iter must be nonnegative")
#undef REQUIRE
    return 1;
}

/* ------------------------------------------------------------------ defaults -------------- */
void qpdo_set_default_settings(QPDOSettings *s)
{   /* reference include/constants.h:44-67 */
    s->max_time = QPDO_INFTY;   s->max_iter = 10000;     s->inner_max_iter = 1000;
    s->eps_abs = 1e-6;          s->eps_abs_in = 1e0;     s->eps_prim_inf = 1e-6;   s->eps_dual_inf = 1e-6;
    s->rho = 0.1;               s->theta = 0.25;         s->delta = 1e-2;          s->mu_min = 1e-9;
    s->proximal = 1;            s->sigma_init = 1e-3;    s->sigma_upd = 1e-1;      s->sigma_min = 1e-7;
    s->scaling = 10;            s->verbose = 1;          s->print_interval = 1;    s->reset_newton_iter = 1000;
}

/* ------------------------------------------------------------------ CHOLMOD sparse -> clean CSC */
typedef struct { int *p, *i; double *x; int nnz; } csc32;
static void csc32_free(csc32 *c) { free(c->p); free(c->i); free(c->x); c->p = c->i = NULL; c->x = NULL; }
static long sp_idx(const void *a, size_t k, int itype_long) { return itype_long ? ((const long *)a)[k] : (long)((const int *)a)[k]; }

/* lower_only: keep entries of the triangle selected by stype (transposing an upper-stored Q),
 * mirror CHOLMOD's rule that entries in the other triangle are ignored. */
static int to_csc32(const cholmod_sparse *S, int symmetric, csc32 *out)
{
    const int il = (S->itype == QPDO_CHOLMOD_LONG);
    const size_t ncol = S->ncol;
    const double *Sx = (const double *)S->x;
    size_t j, cnt = 0;
    int *colcount = (int *)calloc(ncol + 1, sizeof(int));
    long p, pend;
    int st = symmetric ? S->stype : 0;
    if (symmetric && st == 0) st = -1;          /* unsymmetric-stored Q: use its lower triangle */
    if (!colcount) return -1;
    for (j = 0; j < ncol; j++) {
        p = sp_idx(S->p, j, il);
        pend = S->packed ? sp_idx(S->p, j + 1, il) : p + sp_idx(S->nz, j, il);
        for (; p < pend; p++) {
            long i = sp_idx(S->i, (size_t)p, il);
            if (st < 0 && i < (long)j) continue;
            if (st > 0 && i > (long)j) continue;
            colcount[(st > 0 ? (size_t)i : j) + 1]++;
            cnt++;
        }
    }
    if (cnt > 2000000000u) { free(colcount); return -1; }
    out->nnz = (int)cnt;
    out->p = (int *)malloc(sizeof(int) * (ncol + 1));
    out->i = (int *)malloc(sizeof(int) * (cnt ? cnt : 1));
    out->x = (double *)malloc(sizeof(double) * (cnt ? cnt : 1));
    if (!out->p || !out->i || !out->x) { free(colcount); csc32_free(out); return -1; }
    out->p[0] = 0;
    for (j = 0; j < ncol; j++) out->p[j + 1] = out->p[j] + colcount[j + 1];
    for (j = 0; j < ncol; j++) colcount[j] = out->p[j];
    for (j = 0; j < ncol; j++) {
        p = sp_idx(S->p, j, il);
        pend = S->packed ? sp_idx(S->p, j + 1, il) : p + sp_idx(S->nz, j, il);
        for (; p < pend; p++) {
            long i = sp_idx(S->i, (size_t)p, il);
            int d;
            if (st < 0 && i < (long)j) continue;
            if (st > 0 && i > (long)j) continue;
            if (st > 0) { d = colcount[i]++; out->i[d] = (int)j; }     /* (i,j) upper -> (j,i) lower */
            else { d = colcount[j]++; out->i[d] = (int)i; }
            out->x[d] = Sx ? Sx[p] : 0.0;
        }
    }
    free(colcount);
    /* sort rows inside each column (insertion sort; inputs are normally already sorted) and sum duplicates */
    {
        int w = 0;
        int *np = (int *)malloc(sizeof(int) * (ncol + 1));
        if (!np) { csc32_free(out); return -1; }
        for (j = 0; j < ncol; j++) {
            int b = out->p[j], e = out->p[j + 1], a, k;
            for (a = b + 1; a < e; a++) {
                int ri = out->i[a]; double rx = out->x[a];
                for (k = a - 1; k >= b && out->i[k] > ri; k--) { out->i[k + 1] = out->i[k]; out->x[k + 1] = out->x[k]; }
                out->i[k + 1] = ri; out->x[k + 1] = rx;
            }
            np[j] = w;
            for (a = b; a < e; a++) {
                if (w > np[j] && out->i[w - 1] == out->i[a]) out->x[w - 1] += out->x[a];
                else { out->i[w] = out->i[a]; out->x[w] = out->x[a]; w++; }
            }
        }
        np[ncol] = w;
        memcpy(out->p, np, sizeof(int) * (ncol + 1));
        out->nnz = w;
        free(np);
    }
    return 0;
}

static cholmod_sparse *copy_sparse_host(const cholmod_sparse *S)
{   /* work->data->Q / A: host copy kept for ownership compatibility (qpdo.c:106-108) */
    const int il = (S->itype == QPDO_CHOLMOD_LONG);
    const size_t isz = il ? sizeof(long) : sizeof(int);
    size_t nz = 0, j;
    cholmod_sparse *C = (cholmod_sparse *)calloc(1, sizeof(cholmod_sparse));
    if (!C) return NULL;
    *C = *S;
    if (S->packed) nz = (size_t)sp_idx(S->p, S->ncol, il);
    else for (j = 0; j < S->ncol; j++) { size_t e = (size_t)(sp_idx(S->p, j, il) + sp_idx(S->nz, j, il)); if (e > nz) nz = e; }
    C->nzmax = nz ? nz : 1;
    C->p = malloc(isz * (S->ncol + 1)); memcpy(C->p, S->p, isz * (S->ncol + 1));
    C->i = malloc(isz * C->nzmax); if (nz) memcpy(C->i, S->i, isz * nz);
    C->x = malloc(sizeof(double) * C->nzmax); if (nz && S->x) memcpy(C->x, S->x, sizeof(double) * nz);
    C->z = NULL;
    if (!S->packed && S->nz) { C->nz = malloc(isz * S->ncol); memcpy(C->nz, S->nz, isz * S->ncol); } else C->nz = NULL;
    return C;
}
static void free_sparse_host(cholmod_sparse *C)
{
    if (!C) return;
    free(C->p); free(C->i); free(C->x); free(C->nz); free(C);
}
static cholmod_dense *dense_shim(size_t n, c_float **xout)
{
    cholmod_dense *D = (cholmod_dense *)calloc(1, sizeof(cholmod_dense));
    D->nrow = n; D->ncol = 1; D->nzmax = n; D->d = n; D->xtype = QPDO_CHOLMOD_REAL; D->dtype = 0;
    D->x = calloc(n ? n : 1, sizeof(c_float));
    *xout = (c_float *)D->x;
    return D;
}
static void dense_free(cholmod_dense *D) { if (D) { free(D->x); free(D); } }
static c_float *vcalloc(size_t n) { return (c_float *)calloc(n ? n : 1, sizeof(c_float)); }
static c_float *vdup(const c_float *a, size_t n)
{
    c_float *b = (c_float *)malloc(sizeof(c_float) * (n ? n : 1));
    if (b && n) memcpy(b, a, sizeof(c_float) * n);
    return b;
}

/* work->data->{q, l, u} mirror the SCALED vectors the device works on, like the reference's work->data after
 * scale_data (src/scaling.c:40-91; qpdo_mex.c reads them); work->data->Q / A keep the caller's values and layout --
 * their scaled copies live in HBM only (32-bit CSC, sorted). */
static void mirror_data_vectors(QPDOWorkspace *work)
{
    qpdo_dev *dev = DEV(work);
    if (work->data->n) qpdo_dev_get_vector(dev, "q", work->data->q, (long)work->data->n);
    if (work->data->m) { qpdo_dev_get_vector(dev, "l", work->data->l, (long)work->data->m); qpdo_dev_get_vector(dev, "u", work->data->u, (long)work->data->m); }
}

/* ------------------------------------------------------------------ qpdo_setup ------------ */
QPDOWorkspace *qpdo_setup(const QPDOData *data, const QPDOSettings *settings)
{
    QPDOWorkspace *work;
    size_t n, m;
    csc32 Q = {0, 0, 0, 0}, A = {0, 0, 0, 0};
    qpdo_dev_problem P;
    qpdo_dev *dev = NULL;
    const char *ord_env;
    int ordering = 2;           /* nested dissection with minimum-degree leaves (falls back to minimum degree on graphs too shallow to cut) */

    if (!validate_data(data)) { eprint("Data validation returned failure"); return QPDO_NULL; }
    if (!validate_settings(settings)) { eprint("Settings validation returned failure"); return QPDO_NULL; }
    n = data->n; m = data->m;
    if (n == 0 || n > 2000000000u || m > 1000000000u) { eprint("problem dimensions out of range"); return QPDO_NULL; }
    if (!data->Q || !data->A || data->Q->nrow != n || data->Q->ncol != n || data->A->ncol != n || data->A->nrow != m) {
        eprint("Q must be n x n and A must be m x n"); return QPDO_NULL;
    }
    work = (QPDOWorkspace *)calloc(1, sizeof(QPDOWorkspace));
    if (!work) { eprint("allocating work failure"); return QPDO_NULL; }
    work->timer = (QPDOTimer *)malloc(sizeof(QPDOTimer));
    tic(work->timer);
    work->settings = copy_settings(settings);
    work->sqrt_delta = sqrt(work->settings->delta);
    work->sigma = work->settings->sigma_init;
    work->chol = (QPDOCholmod *)calloc(1, sizeof(QPDOCholmod));

    /* host copy of the problem (ownership semantics of qpdo.c:100-109) */
    work->data = (QPDOData *)calloc(1, sizeof(QPDOData));
    work->data->n = n; work->data->m = m;
    work->data->l = vdup(data->l, m); work->data->u = vdup(data->u, m); work->data->q = vdup(data->q, n);
    work->data->c = data->c;
    work->data->Q = copy_sparse_host(data->Q);
    work->data->A = copy_sparse_host(data->A);
    work->data->A->stype = 0;

    /* device half */
    if (to_csc32(data->Q, 1, &Q) || to_csc32(data->A, 0, &A)) { eprint("could not convert the sparse inputs"); goto fail; }
    P.n = (int)n; P.m = (int)m;
    P.Qp = Q.p; P.Qi = Q.i; P.Qx = Q.x; P.Ap = A.p; P.Ai = A.i; P.Ax = A.x;
    P.q = data->q; P.l = data->l; P.u = data->u;
    ord_env = getenv("QPDO_B200_ORDERING");
    if (ord_env && !strcmp(ord_env, "natural")) ordering = 0;
    if (ord_env && !strcmp(ord_env, "mindeg")) ordering = 1;
    if (ord_env && !strcmp(ord_env, "nd")) ordering = 2;
    if (qpdo_dev_create(&dev, &P, (int)settings->scaling, ordering, -1)) {
        eprint("CUDA layer: %s", qpdo_dev_last_error());
        goto fail;
    }
    csc32_free(&Q); csc32_free(&A);
    work->chol->LD = (cholmod_factor *)dev;

    /* host mirrors (allocated like qpdo.c:112-191 so that callers and cleanup see the same layout) */
    work->x = vcalloc(n); work->y = vcalloc(m); work->xbar = vcalloc(n); work->ybar = vcalloc(m);
    work->Ax = vcalloc(m); work->Qx = vcalloc(n); work->Aty = vcalloc(n);
    work->initialized = 0;
    work->temp_m = vcalloc(m); work->temp_n = vcalloc(n); work->temp_2m = vcalloc(2 * m);
    work->mu = vcalloc(m); work->sqrt_mu = vcalloc(m); work->n_mu_changed = 0;
    work->z = vcalloc(m); work->w = vcalloc(m); work->res_prim = vcalloc(m); work->res_prim_old = vcalloc(m);
    work->res_dual = vcalloc(n); work->res_prim_in = vcalloc(m); work->res_dual_in = vcalloc(n); work->df = vcalloc(n);
    /* the linesearch scratch lives on the device only; 1-element placeholders keep the pointers valid */
    work->ls_delta = vcalloc(1); work->ls_alpha = vcalloc(1);
    work->ls_taus = (array_element *)calloc(1, sizeof(array_element));
    work->ls_idx_L = (c_int *)calloc(1, sizeof(c_int)); work->ls_idx_P = (c_int *)calloc(1, sizeof(c_int));
    work->ls_idx_J = (c_int *)calloc(1, sizeof(c_int));
    if (settings->scaling) {
        work->scaling = (QPDOScaling *)malloc(sizeof(QPDOScaling));
        work->scaling->D = vcalloc(n); work->scaling->Dinv = vcalloc(n);
        work->scaling->E = vcalloc(m); work->scaling->Einv = vcalloc(m);
        work->chol->E_temp = dense_shim(m, &work->E_temp);
        work->chol->D_temp = dense_shim(n, &work->D_temp);
        if (qpdo_dev_get_scaling(dev, work->scaling->D, work->scaling->Dinv, work->scaling->E, work->scaling->Einv,
                                 &work->scaling->c, &work->scaling->cinv, &work->norm_q)) { eprint("CUDA layer: %s", qpdo_dev_last_error()); goto fail; }
    } else {
        work->scaling = QPDO_NULL;
        if (qpdo_dev_get_scaling(dev, NULL, NULL, NULL, NULL, NULL, NULL, &work->norm_q)) goto fail;
    }
    mirror_data_vectors(work);
    work->chol->linsys_rhs = dense_shim(n, &work->linsys_rhs);
    work->chol->dx = dense_shim(n, &work->dx);
    work->chol->dy = dense_shim(m, &work->dy);
    work->chol->Qdx = dense_shim(n, &work->Qdx);
    work->chol->Adx = dense_shim(m, &work->Adx);
    work->chol->Atdy = dense_shim(n, &work->Atdy);
    work->chol->active_set = (c_int *)calloc(m ? m : 1, sizeof(c_int));
    work->chol->active_set_old = (c_int *)calloc(m ? m : 1, sizeof(c_int));
    work->chol->reset_newton = 1;
    work->chol->enter = (c_int *)calloc(m ? m : 1, sizeof(c_int));
    work->chol->leave = (c_int *)calloc(m ? m : 1, sizeof(c_int));
    work->chol->At_scale = NULL; work->chol->At_sqrt_mu = NULL; work->chol->LD_Q = NULL;
    work->solution = (QPDOSolution *)calloc(1, sizeof(QPDOSolution));
    work->solution->x = vcalloc(n); work->solution->y = vcalloc(m);
    work->info = (QPDOInfo *)calloc(1, sizeof(QPDOInfo));
    update_status(work->info, QPDO_UNSOLVED);
    work->info->solve_time = 0.0; work->info->run_time = 0.0;
    work->info->setup_time = toc(work->timer);
    return work;
fail:
    csc32_free(&Q); csc32_free(&A);
    if (dev && !work->chol->LD) qpdo_dev_destroy(dev);
    if (!work->info) work->info = (QPDOInfo *)calloc(1, sizeof(QPDOInfo));
    qpdo_cleanup(work);
    return QPDO_NULL;
}

/* ------------------------------------------------------------------ qpdo_warm_start ------- */
void qpdo_warm_start(QPDOWorkspace *work, c_float *x_warm_start, c_float *y_warm_start)
{
    double obj = 0.0;
    work->sigma = work->settings->sigma_init;
    if (work->info->status_val != QPDO_UNSOLVED) work->info->setup_time = 0;
    tic(work->timer);
    if (qpdo_dev_warm_start(DEV(work), x_warm_start, y_warm_start, work->sigma, (int)work->settings->proximal,
                            work->settings->mu_min, &obj)) {
        eprint("CUDA layer: %s", qpdo_dev_last_error());
        update_status(work->info, QPDO_ERROR);
        return;
    }
    work->info->objective = x_warm_start ? obj + work->data->c : 0.0;
    work->sqrt_mu_min = 1 / sqrt(work->settings->mu_min);
    work->initialized = 1;
    work->info->setup_time += toc(work->timer);
}

/* ------------------------------------------------------------------ printing (util.c:101-173) */
static void print_header(void)
{
    printf("\n                 QPDO v0.1 on NVIDIA B200 (qpdo_b200)                        \n\n");
    printf("       |---------------------------------------------|--------------------------------|\n");
    printf(" iter  |   objective    res prim   res dual  |  r.p. in  r.d. in   stepsize  |\n");
    printf("       |---------------------------------------------|--------------------------------|\n");
}
static void print_iteration(long iter, const QPDOWorkspace *w, const qpdo_dev_scalars *s)
{
    printf("%6ld | %12.4e %10.2e %10.2e | %10.2e %10.2e %10.2e |\n", iter, w->info->objective, s->res_prim_norm,
           s->res_dual_norm, s->res_prim_in_norm, s->res_dual_in_norm, s->tau);
}

/* ------------------------------------------------------------------ qpdo_solve ------------ */
void qpdo_solve(QPDOWorkspace *work)
{
    const QPDOSettings *st = work->settings;
    qpdo_dev *dev = DEV(work);
    qpdo_dev_scalars sc;
    const int proximal = (int)st->proximal;
    c_int iter, oter = 0, iter_old = 0;
    int failed = 0, flag;

    if (st->verbose) print_header();
    if (!work->initialized) qpdo_warm_start(work, NULL, NULL);
    if (work->info->status_val == QPDO_ERROR && !work->initialized) return;
    work->eps_in = st->eps_abs_in;                                   /* qpdo.c:321-325 */
    work->sigma = st->sigma_init;
    work->chol->reset_newton = 1;
    /* like the reference, the status of a previous solve is NOT reset here (qpdo.c:304-341) */
    tic(work->timer);
    memset(&sc, 0, sizeof sc);
    if (qpdo_dev_begin_solve(dev, work->sigma)) failed = 1;

    for (iter = 0; !failed && iter < st->max_iter; iter++) {
        /* outer + inner residuals and their norms (qpdo.c:346-351) */
        if (qpdo_dev_residuals(dev, proximal, &sc)) { failed = 1; break; }
        work->info->res_prim_norm = sc.res_prim_norm;       work->info->res_dual_norm = sc.res_dual_norm;
        work->info->res_prim_in_norm = sc.res_prim_in_norm; work->info->res_dual_in_norm = sc.res_dual_in_norm;
        work->tau = sc.tau;
        if (st->verbose && st->print_interval && (iter % st->print_interval == 0)) {
            double obj;
            if (!qpdo_dev_objective(dev, proximal, work->data->c, &obj)) work->info->objective = obj;
            print_iteration((long)iter, work, &sc);
        }
        /* check_outer_optimality (termination.c:11-23) */
        if (sc.res_prim_norm > QPDO_INFTY || sc.res_dual_norm > QPDO_INFTY) { update_status(work->info, QPDO_NON_CVX); break; }
        if (sc.res_prim_norm <= st->eps_abs && sc.res_dual_norm <= st->eps_abs) { update_status(work->info, QPDO_SOLVED); break; }

        if (((iter > iter_old + 1) && (sc.res_prim_in_norm <= work->eps_in) && (sc.res_dual_in_norm <= work->eps_in)) ||
            (iter == iter_old + st->inner_max_iter)) {
            /* subproblem solved or stopped (qpdo.c:365-430) */
            if (iter < iter_old + st->inner_max_iter) {
                if (st->eps_prim_inf > 0) {
                    if (qpdo_dev_primal_infeasible(dev, st->eps_prim_inf, &flag)) { failed = 1; break; }
                    if (flag) { update_status(work->info, QPDO_PRIMAL_INFEASIBLE); break; }
                }
                if (st->eps_dual_inf > 0) {
                    if (qpdo_dev_dual_infeasible(dev, st->eps_dual_inf, proximal, &flag)) { failed = 1; break; }
                    if (flag) { update_status(work->info, QPDO_DUAL_INFEASIBLE); break; }
                }
            }
            if (qpdo_dev_outer_update(dev, (oter > 0) && (sc.res_prim_norm > st->eps_abs),
                                      proximal && (oter > 0) && (sc.res_dual_norm > st->eps_abs), proximal, st->eps_abs,
                                      st->theta, st->delta, st->mu_min, st->sigma_upd, st->sigma_min, &work->sigma)) { failed = 1; break; }
            if (iter < iter_old + st->inner_max_iter) {
                double e1 = st->rho * work->eps_in, e2 = 0.1 * st->eps_abs;
                work->eps_in = e1 > e2 ? e1 : e2;
                if (st->verbose && st->print_interval && (iter % st->print_interval == 0))
                    printf("%6ld |-------------------------------------------------------------------|\n", (long)iter);
            } else if (st->verbose && st->print_interval && (iter % st->print_interval == 0))
                printf("%6ld |--  --  --  --  --  --  --  --  --  --  --  --  --  --  --  --  -- |\n", (long)iter);
            oter++;
            iter_old = iter;
        } else {
            /* subproblem iteration (qpdo.c:432-438) */
            int force = (st->reset_newton_iter != 0) && (iter % st->reset_newton_iter == 0);
            if (qpdo_dev_newton_step(dev, force, proximal)) { failed = 1; break; }
        }
        work->info->run_time = work->info->setup_time + toc(work->timer);
        if (work->info->run_time > st->max_time) { update_status(work->info, QPDO_MAX_TIME_REACHED); break; }
    }
    if (failed) {
        eprint("CUDA layer: %s", qpdo_dev_last_error());
        update_status(work->info, QPDO_ERROR);
    } else if (work->info->status_val == QPDO_UNSOLVED) update_status(work->info, QPDO_MAX_ITER_REACHED);
    work->info->iterations = iter;
    work->info->oterations = oter;
    /* store_solution (termination.c:82-92) + host mirrors read by callers (qpdo_mex.c:237-280) */
    if (!failed) {
        double obj;
        if (qpdo_dev_store_solution(dev, work->solution->x, work->solution->y, work->x, work->y, work->dx, work->dy,
                                    work->chol->active_set, &sc) ||
            qpdo_dev_objective(dev, proximal, work->data->c, &obj)) {
            eprint("CUDA layer: %s", qpdo_dev_last_error());
            update_status(work->info, QPDO_ERROR);
        } else {
            work->info->objective = obj;
            work->chol->n_active_set = sc.n_active; work->chol->n_enter = sc.n_enter; work->chol->n_leave = sc.n_leave;
            work->chol->reset_newton = sc.reset_newton; work->n_mu_changed = sc.n_mu_changed;
            work->ls_eta = sc.ls_eta; work->ls_beta = sc.ls_beta; work->tau = sc.tau;
        }
    }
    work->initialized = 0;
    work->info->solve_time = toc(work->timer);
    work->info->run_time = work->info->setup_time + work->info->solve_time;
    if (st->verbose) {
        printf("\n status: %s, iterations: %ld (outer %ld), objective: %.6e\n", work->info->status, (long)work->info->iterations,
               (long)work->info->oterations, work->info->objective);
    }
}

/* ------------------------------------------------------------------ qpdo_update_settings -- */
void qpdo_update_settings(QPDOWorkspace *work, const QPDOSettings *settings)
{
    if (!validate_settings(settings)) {
        eprint("Settings validation returned failure");
        update_status(work->info, QPDO_ERROR);
        return;
    }
    if (work->settings->scaling > settings->scaling) {
        eprint("Decreasing the number of scaling iterations is not allowed");
        update_status(work->info, QPDO_ERROR);
        return;
    } else if (work->settings->scaling < settings->scaling) {
        if (qpdo_dev_rescale(DEV(work), (int)(settings->scaling - work->settings->scaling))) {
            eprint("CUDA layer: %s", qpdo_dev_last_error());
            update_status(work->info, QPDO_ERROR);
            return;
        }
        /* host mirror of the total scaling (qpdo.c:503-511) */
        if (qpdo_dev_get_scaling(DEV(work), work->scaling->D, work->scaling->Dinv, work->scaling->E, work->scaling->Einv,
                                 &work->scaling->c, &work->scaling->cinv, NULL)) {
            eprint("CUDA layer: %s", qpdo_dev_last_error());
            update_status(work->info, QPDO_ERROR);
            return;
        }
        mirror_data_vectors(work);
    }
    free(work->settings);
    work->settings = copy_settings(settings);
    work->sqrt_delta = sqrt(work->settings->delta);
}

/* ------------------------------------------------------------------ qpdo_update_bounds ---- */
void qpdo_update_bounds(QPDOWorkspace *work, const c_float *l, const c_float *u)
{
    size_t j, m = work->data->m;
    if (l != NULL && u != NULL)
        for (j = 0; j < m; j++)
            if (l[j] > u[j]) {
                eprint("Lower bound at index %d is greater than upper bound: %.4e > %.4e", (int)j, l[j], u[j]);
                update_status(work->info, QPDO_ERROR);
                return;
            }
    if (qpdo_dev_update_bounds(DEV(work), l, u)) {
        eprint("CUDA layer: %s", qpdo_dev_last_error());
        update_status(work->info, QPDO_ERROR);
        return;
    }
    /* host mirror of the (scaled) bounds, qpdo.c:538-543 */
    mirror_data_vectors(work);
}

/* ------------------------------------------------------------------ qpdo_update_q --------- */
void qpdo_update_q(QPDOWorkspace *work, const c_float *q)
{
    double c = 1.0, cinv = 1.0, nq = 0.0, sigma = work->sigma;
    if (qpdo_dev_update_q(DEV(work), q, (int)work->settings->proximal, work->settings->sigma_init, &sigma, &c, &cinv, &nq)) {
        eprint("CUDA layer: %s", qpdo_dev_last_error());
        update_status(work->info, QPDO_ERROR);
        return;
    }
    work->sigma = sigma;
    work->norm_q = nq;
    if (work->settings->scaling) { work->scaling->c = c; work->scaling->cinv = cinv; }
    mirror_data_vectors(work);                  /* scaled q, qpdo.c:560-571 */
}

/* ------------------------------------------------------------------ qpdo_cleanup ---------- */
void qpdo_cleanup(QPDOWorkspace *work)
{
    if (!work) return;
    if (work->data) {
        free_sparse_host(work->data->Q); free_sparse_host(work->data->A);
        free(work->data->q); free(work->data->l); free(work->data->u);
        free(work->data);
    }
    if (work->scaling) {
        free(work->scaling->D); free(work->scaling->Dinv); free(work->scaling->E); free(work->scaling->Einv);
        free(work->scaling);
    }
    free(work->x); free(work->y); free(work->Ax); free(work->Qx); free(work->Aty);
    free(work->temp_m); free(work->temp_n); free(work->temp_2m); free(work->mu); free(work->sqrt_mu);
    free(work->z); free(work->w); free(work->res_prim); free(work->res_dual); free(work->res_prim_old);
    free(work->res_prim_in); free(work->res_dual_in); free(work->df); free(work->xbar); free(work->ybar);
    free(work->ls_delta); free(work->ls_alpha); free(work->ls_taus); free(work->ls_idx_L); free(work->ls_idx_P);
    free(work->ls_idx_J);
    free(work->settings);
    if (work->chol) {
        if (work->chol->LD) qpdo_dev_destroy((qpdo_dev *)work->chol->LD);
        dense_free(work->chol->D_temp); dense_free(work->chol->E_temp); dense_free(work->chol->linsys_rhs);
        dense_free(work->chol->dx); dense_free(work->chol->dy); dense_free(work->chol->Qdx);
        dense_free(work->chol->Adx); dense_free(work->chol->Atdy);
        free(work->chol->active_set); free(work->chol->active_set_old); free(work->chol->enter); free(work->chol->leave);
        free(work->chol);
    }
    if (work->solution) { free(work->solution->x); free(work->solution->y); free(work->solution); }
    free(work->timer);
    free(work->info);
    free(work);
}
