/* tests/c_driver/driver.c -- BASELINE.json configs[0] verbatim: "random sparse convex QP n=1000 m=500 density 1 %
 * solved via a C driver calling qpdo_setup/qpdo_solve".
 *
 * This file is compiled ONCE against the REFERENCE's own headers (include/qpdo.h, include/types.h and the
 * SuiteSparse headers they pull in, -DDLONG -DPROFILING: reference include/qpdo.h:14-56, include/types.h:81-224)
 * and the one object is linked twice: against oracle/_ref/libqpdo_ref.so (the unmodified reference) and against
 * qpdo_b200/lib/libqpdo_b200.so (this repo).  Both binaries print the same report; tests/test_c_driver.py diffs
 * them.  That is the compile-time proof of the drop-in boundary: no header of this repo is involved.
 *
 * The problem generator is SURVEY.md 8(d) "C1 inputs": xorshift64 + Box-Muller, fixed in the driver so both
 * libraries get bit-identical inputs.  usage: driver [n m density seed]
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>
#include "qpdo.h"

static unsigned long long rng_state = 88172645463325252ULL;
static unsigned long long xorshift64(void) { unsigned long long x = rng_state; x ^= x << 13; x ^= x >> 7; x ^= x << 17; return rng_state = x; }
static double uniform01(void) { return (double)(xorshift64() >> 11) * (1.0 / 9007199254740992.0); }
static double normal01(void)
{
    double u1 = uniform01(), u2 = uniform01();
    if (u1 < 1e-300) u1 = 1e-300;
    return sqrt(-2.0 * log(u1)) * cos(6.283185307179586 * u2);
}

typedef struct { long r, c; double v; } trip;
static int trip_cmp(const void *a, const void *b)
{
    const trip *x = (const trip *)a, *y = (const trip *)b;
    if (x->c != y->c) return x->c < y->c ? -1 : 1;
    if (x->r != y->r) return x->r < y->r ? -1 : 1;
    return 0;
}
/* triplets (duplicates summed) -> CSC with sorted rows; returns nnz */
static long to_csc(trip *t, long nt, long ncol, c_int *p, c_int *i, c_float *x)
{
    qsort(t, (size_t)nt, sizeof(trip), trip_cmp);
    long nz = 0, k = 0;
    for (long c = 0; c < ncol; c++) {
        p[c] = nz;
        while (k < nt && t[k].c == c) {
            if (nz > p[c] && i[nz - 1] == t[k].r) x[nz - 1] += t[k].v;
            else { i[nz] = t[k].r; x[nz] = t[k].v; nz++; }
            k++;
        }
    }
    p[ncol] = nz;
    return nz;
}
static void fill_sparse(cholmod_sparse *S, size_t nrow, size_t ncol, c_int *p, c_int *i, c_float *x, int stype)
{
    memset(S, 0, sizeof *S);
    S->nrow = nrow; S->ncol = ncol; S->nzmax = (size_t)p[ncol]; S->p = p; S->i = i; S->x = x; S->nz = NULL; S->z = NULL;
    S->stype = stype; S->itype = CHOLMOD_LONG; S->xtype = CHOLMOD_REAL; S->dtype = CHOLMOD_DOUBLE; S->sorted = 1; S->packed = 1;
}

int main(int argc, char **argv)
{
    const long n = argc > 1 ? atol(argv[1]) : 1000, m = argc > 2 ? atol(argv[2]) : 500;
    const double density = argc > 3 ? atof(argv[3]) : 0.01;
    rng_state ^= (unsigned long long)(argc > 4 ? atol(argv[4]) : 1) * 0x9E3779B97F4A7C15ULL;
    /* Q: lower triangle, ~0.5 density n^2 off-diagonals N(0,1), diagonal = sum |row off-diagonals| + U(0,1) */
    const long noff = (long)(0.5 * density * (double)n * (double)n + 0.5);
    trip *tq = (trip *)malloc(sizeof(trip) * (size_t)(noff + n));
    double *rowabs = (double *)calloc((size_t)n, sizeof(double));
    long nt = 0;
    for (long k = 0; k < noff; k++) {
        long r = (long)(uniform01() * (double)n), c = (long)(uniform01() * (double)n);
        double v = normal01();
        if (r == c) continue;
        if (r < c) { long s = r; r = c; c = s; }
        tq[nt].r = r; tq[nt].c = c; tq[nt].v = v; nt++;
        rowabs[r] += fabs(v); rowabs[c] += fabs(v);
    }
    for (long j = 0; j < n; j++) { tq[nt].r = j; tq[nt].c = j; tq[nt].v = rowabs[j] + uniform01(); nt++; }
    c_int *Qp = (c_int *)malloc(sizeof(c_int) * (size_t)(n + 1)), *Qi = (c_int *)malloc(sizeof(c_int) * (size_t)nt);
    c_float *Qx = (c_float *)malloc(sizeof(c_float) * (size_t)nt);
    to_csc(tq, nt, n, Qp, Qi, Qx);
    /* A: density m n triplets N(0,1), duplicates summed */
    const long na = (long)(density * (double)m * (double)n + 0.5);
    trip *ta = (trip *)malloc(sizeof(trip) * (size_t)(na + 1));
    for (long k = 0; k < na; k++) { ta[k].r = (long)(uniform01() * (double)m); ta[k].c = (long)(uniform01() * (double)n); ta[k].v = normal01(); }
    c_int *Ap = (c_int *)malloc(sizeof(c_int) * (size_t)(n + 1)), *Ai = (c_int *)malloc(sizeof(c_int) * (size_t)(na + 1));
    c_float *Ax = (c_float *)malloc(sizeof(c_float) * (size_t)(na + 1));
    to_csc(ta, na, n, Ap, Ai, Ax);
    c_float *q = (c_float *)malloc(sizeof(c_float) * (size_t)n), *l = (c_float *)malloc(sizeof(c_float) * (size_t)m), *u = (c_float *)malloc(sizeof(c_float) * (size_t)m);
    for (long j = 0; j < n; j++) q[j] = normal01();
    for (long i = 0; i < m; i++) { l[i] = -uniform01(); u[i] = uniform01(); }

    cholmod_sparse Q, A;
    fill_sparse(&Q, (size_t)n, (size_t)n, Qp, Qi, Qx, -1);          /* lower triangle stored, like qpdo_mex.c:150 */
    fill_sparse(&A, (size_t)m, (size_t)n, Ap, Ai, Ax, 0);
    QPDOData data;
    data.n = (size_t)n; data.m = (size_t)m; data.Q = &Q; data.A = &A; data.q = q; data.c = 0.0; data.l = l; data.u = u;
    QPDOSettings settings;
    qpdo_set_default_settings(&settings);
    settings.verbose = 0;
    QPDOWorkspace *work = qpdo_setup(&data, &settings);
    if (!work) { printf("qpdo_setup returned NULL\n"); return 2; }
    qpdo_solve(work);
    /* what a C caller reads back: include/types.h:147-224 */
    unsigned long long hash = 1469598103934665603ULL;
    long nact = 0;
    for (long i = 0; i < m; i++) { const int a = work->chol->active_set[i] ? 1 : 0; nact += a; hash = (hash ^ (unsigned long long)(a + 1)) * 1099511628211ULL; }
    double sx = 0, sy = 0, ax = 0, ay = 0;
    for (long j = 0; j < n; j++) { sx += work->solution->x[j]; ax = fmax(ax, fabs(work->solution->x[j])); }
    for (long i = 0; i < m; i++) { sy += work->solution->y[i]; ay = fmax(ay, fabs(work->solution->y[i])); }
    printf("problem n %ld m %ld nnzQ %ld nnzA %ld\n", n, m, (long)Qp[n], (long)Ap[n]);
    printf("status %ld (%s) iterations %ld oterations %ld\n", (long)work->info->status_val, work->info->status, (long)work->info->iterations, (long)work->info->oterations);
    printf("active_set count %ld hash %016llx\n", nact, hash);
    printf("objective %.10e res_prim %.3e res_dual %.3e\n", work->info->objective, work->info->res_prim_norm, work->info->res_dual_norm);
    printf("x sum %.9e max %.9e  y sum %.9e max %.9e\n", sx, ax, sy, ay);
    for (long j = 0; j < n; j += n / 8 > 0 ? n / 8 : 1) printf("x[%ld] %.9e\n", j, work->solution->x[j]);
    for (long i = 0; i < m; i += m / 8 > 0 ? m / 8 : 1) printf("y[%ld] %.9e\n", i, work->solution->y[i]);
    /* second leg of the API: new bounds + warm start on the same workspace (qpdo.c:217-299, 522-544) */
    for (long i = 0; i < m; i++) { l[i] -= 0.05; u[i] += 0.02; }
    c_float *x0 = (c_float *)malloc(sizeof(c_float) * (size_t)n), *y0 = (c_float *)malloc(sizeof(c_float) * (size_t)m);
    memcpy(x0, work->solution->x, sizeof(c_float) * (size_t)n); memcpy(y0, work->solution->y, sizeof(c_float) * (size_t)m);
    qpdo_update_bounds(work, l, u);
    qpdo_warm_start(work, x0, y0);
    qpdo_solve(work);
    printf("resolve status %ld iterations %ld oterations %ld objective %.10e\n", (long)work->info->status_val, (long)work->info->iterations, (long)work->info->oterations, work->info->objective);
    qpdo_cleanup(work);
    return 0;
}
