/* oracle/ref_batch.c -- TEST INFRASTRUCTURE ONLY (CPU baseline of the batch workload).
 *
 * The reference has no batch entry: a user with many QPs loops over qpdo_setup / qpdo_solve /
 * qpdo_cleanup per problem (reference src/qpdo.c:49,304,591).  This driver does exactly that loop
 * over a packed batch (the layout of include/qpdo_batch.h), spread over host threads -- workspaces
 * are independent (SURVEY.md 8(b) "Threading").  The solver library is dlopen'ed by path:
 * oracle/_ref/libqpdo_ref.so (the unmodified reference) for the timed baseline.  include/qpdo.h is
 * only used for its layout-identical public structs.
 */
#define _GNU_SOURCE
#include <dlfcn.h>
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#include "../include/qpdo.h"

typedef struct {
    int    status_val, iterations, oterations, n_active;
    double objective, res_prim_norm, res_dual_norm;
} ref_summary;                              /* same 40 bytes as qpdo_batch_summary */

typedef struct {
    void (*defaults)(QPDOSettings *);
    QPDOWorkspace *(*setup)(const QPDOData *, const QPDOSettings *);
    void (*solve)(QPDOWorkspace *);
    void (*cleanup)(QPDOWorkspace *);
} ref_api;

typedef struct {
    const ref_api *api;
    const QPDOSettings *settings;
    long n, m, nnzQ, nnzA, count;
    const long *Qp, *Qi, *Ap, *Ai;
    const double *Qx, *Ax, *q, *l, *u;
    double *x, *y;
    ref_summary *summary;
    long *next;                             /* shared work-queue head */
    pthread_mutex_t *mu;
} job;

static void fill_sparse(cholmod_sparse *S, long nrow, long ncol, const long *p, const long *i, double *x, int stype)
{
    memset(S, 0, sizeof(*S));
    S->nrow = nrow; S->ncol = ncol; S->nzmax = p[ncol];
    S->p = (void *)p; S->i = (void *)i; S->x = x;
    S->stype = stype; S->itype = 2 /* CHOLMOD_LONG */; S->xtype = 1 /* REAL */; S->dtype = 0;
    S->sorted = 1; S->packed = 1;
}

static void *worker(void *arg)
{
    job *J = (job *)arg;
    double *Qx = malloc(sizeof(double) * J->nnzQ), *Ax = malloc(sizeof(double) * J->nnzA);
    double *q = malloc(sizeof(double) * J->n), *l = malloc(sizeof(double) * J->m), *u = malloc(sizeof(double) * J->m);
    for (;;) {
        pthread_mutex_lock(J->mu);
        long b = (*J->next)++;
        pthread_mutex_unlock(J->mu);
        if (b >= J->count) break;
        memcpy(Qx, J->Qx + b * J->nnzQ, sizeof(double) * J->nnzQ);
        memcpy(Ax, J->Ax + b * J->nnzA, sizeof(double) * J->nnzA);
        memcpy(q, J->q + b * J->n, sizeof(double) * J->n);
        memcpy(l, J->l + b * J->m, sizeof(double) * J->m);
        memcpy(u, J->u + b * J->m, sizeof(double) * J->m);
        cholmod_sparse Q, A;
        fill_sparse(&Q, J->n, J->n, J->Qp, J->Qi, Qx, -1);
        fill_sparse(&A, J->m, J->n, J->Ap, J->Ai, Ax, 0);
        QPDOData d;
        memset(&d, 0, sizeof(d));
        d.n = J->n; d.m = J->m; d.Q = &Q; d.A = &A; d.q = q; d.c = 0.0; d.l = l; d.u = u;
        QPDOWorkspace *w = J->api->setup(&d, J->settings);
        ref_summary *s = J->summary + b;
        memset(s, 0, sizeof(*s));
        if (!w) { s->status_val = -100; continue; }
        J->api->solve(w);
        s->status_val = (int)w->info->status_val; s->iterations = (int)w->info->iterations;
        s->oterations = (int)w->info->oterations;
        s->objective = w->info->objective; s->res_prim_norm = w->info->res_prim_norm; s->res_dual_norm = w->info->res_dual_norm;
        if (J->x) memcpy(J->x + b * J->n, w->solution->x, sizeof(double) * J->n);
        if (J->y) memcpy(J->y + b * J->m, w->solution->y, sizeof(double) * J->m);
        J->api->cleanup(w);
    }
    free(Qx); free(Ax); free(q); free(l); free(u);
    return NULL;
}

/* returns wall seconds of the threaded loop, or a negative value on error */
double ref_batch_run(const char *libpath, long count, long n, long m, const long *Qp, const long *Qi,
                     const long *Ap, const long *Ai, const double *Qx, const double *Ax, const double *q,
                     const double *l, const double *u, long max_iter, int nthreads,
                     double *x, double *y, ref_summary *summary)
{
    void *h = dlopen(libpath, RTLD_NOW | RTLD_LOCAL);
    if (!h) { fprintf(stderr, "ref_batch: %s\n", dlerror()); return -1.0; }
    ref_api api;
    api.defaults = (void (*)(QPDOSettings *))dlsym(h, "qpdo_set_default_settings");
    api.setup = (QPDOWorkspace * (*)(const QPDOData *, const QPDOSettings *)) dlsym(h, "qpdo_setup");
    api.solve = (void (*)(QPDOWorkspace *))dlsym(h, "qpdo_solve");
    api.cleanup = (void (*)(QPDOWorkspace *))dlsym(h, "qpdo_cleanup");
    if (!api.defaults || !api.setup || !api.solve || !api.cleanup) return -2.0;
    QPDOSettings st;
    api.defaults(&st);
    st.verbose = 0;
    if (max_iter > 0) st.max_iter = max_iter;
    long next = 0;
    pthread_mutex_t mu = PTHREAD_MUTEX_INITIALIZER;
    job J = {&api, &st, n, m, Qp[n], Ap[n], count, Qp, Qi, Ap, Ai, Qx, Ax, q, l, u, x, y, summary, &next, &mu};
    if (nthreads < 1) nthreads = 1;
    pthread_t *th = malloc(sizeof(pthread_t) * nthreads);
    struct timespec t0, t1;
    clock_gettime(CLOCK_MONOTONIC, &t0);
    for (int t = 0; t < nthreads; t++) pthread_create(&th[t], NULL, worker, &J);
    for (int t = 0; t < nthreads; t++) pthread_join(th[t], NULL);
    clock_gettime(CLOCK_MONOTONIC, &t1);
    free(th);
    return (double)(t1.tv_sec - t0.tv_sec) + 1e-9 * (double)(t1.tv_nsec - t0.tv_nsec);
}
