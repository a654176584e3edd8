// dev_internal.h -- device context shared by the .cu translation units (not part of the ABI).
#ifndef QPDO_B200_DEV_INTERNAL_H
#define QPDO_B200_DEV_INTERNAL_H

#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <vector>
#include "qpdo_cuda.h"
#include "symbolic.h"

void qpdo_set_error(const char *fmt, ...);

#define CUDA_OK(call)                                                                          \
    do {                                                                                       \
        cudaError_t e_ = (call);                                                               \
        if (e_ != cudaSuccess) {                                                               \
            qpdo_set_error("%s:%d: %s: %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_)); \
            return -1;                                                                         \
        }                                                                                      \
    } while (0)

// one bucket of fronts of one level, launched together
struct FrontLaunch {
    int first, count;      // range in level_sn
    int maxN, maxnc;
    int kind;              // 0 warp-sized smem fronts, 1 smem fronts, 2 big (global, blocked)
};

struct Factor {
    qpdo_symbolic *S = nullptr;
    // device copies of the symbolic structure
    int *sn_start = nullptr, *sn_nr = nullptr, *rows = nullptr, *rel = nullptr, *sn_parent = nullptr;
    int *child_ptr = nullptr, *child_idx = nullptr, *level_sn = nullptr;
    int64_t *sn_rowptr = nullptr, *Lp = nullptr, *Up = nullptr;
    void *fdesc = nullptr;           // FrontDesc per supernode (factor.cu)
    int *perm = nullptr, *pinv = nullptr;
    int *Ki = nullptr, *Kp = nullptr, *Kq = nullptr, *Kt_row = nullptr, *Kt_a1 = nullptr, *Kt_a2 = nullptr;
    int *Kcol = nullptr;             // column of each K entry (diagonal test)
    int64_t *Kdst = nullptr, *Ktp = nullptr;
    double *Kt_val = nullptr;        // a_ir * a_ic per triple (recomputed when A is rescaled)
    double *Lx = nullptr;            // L panels (working values; final for the smem fronts)
    double *Xout = nullptr;          // final L panels of the big (blocked) fronts
    double *Ux = nullptr;            // update matrices
    double *xs = nullptr;            // permuted rhs / solution
    double *uv = nullptr;            // per-front update vectors of the forward solve
    std::vector<std::vector<FrontLaunch>> plan;   // per level
    cudaGraphExec_t g_factor = nullptr, g_solve = nullptr;
    long launches_factor = 0, launches_solve = 0;
    // pattern management (factor.cu): row set P of A that the current symbolic structure covers
    std::vector<int> hQp, hQi, hAp, hAi;          // host copies of the patterns
    std::vector<std::pair<void *, size_t>> pool;  // device arrays that depend on the pattern: slot pool (pointer, capacity in bytes)
    size_t pool_next = 0;                         // next slot of the running analysis
    std::vector<unsigned char> hP;                // host copy of the current pattern row set (empty: superset)
    bool headroom = true;                         // new row sets keep the old rows and take in the rows nearest to a bound
    unsigned char *inP = nullptr;                 // [m] row is in P
    int *cnt = nullptr, *cnt_host = nullptr;      // support / outside counters
    int ordering = 1, npattern = 0, analyses = 0;
    bool dynamic = false, full = false;
    // independent launches of one tree level (size buckets, blocked fronts) run on forked streams inside the captured graphs
    cudaStream_t aux[4] = {nullptr, nullptr, nullptr, nullptr};
    cudaEvent_t ev_fork = nullptr, ev_join[4] = {nullptr, nullptr, nullptr, nullptr};
    int nsing = 0;                                // singleton leaf fronts at the tail of level 0 (element-wise kernels)
    bool no_singletons = false;                   // QPDO_B200_SINGLETONS=0
    bool pdl = true;                              // panel steps chained by programmatic dependent launch (QPDO_B200_PDL=0: off)
    bool solve_cluster = true;                    // big fronts solved by a cluster of CTAs (QPDO_B200_SOLVE=cta: one CTA each)
    bool concurrent = true;                       // QPDO_B200_CONCURRENT=0: one stream (A/B measurements)
    bool panel_v1 = false;                        // QPDO_B200_PANEL=v1: round-1 panel step (A/B measurements)
};

struct qpdo_dev {
    int n = 0, m = 0, device = 0;
    int proximal = 1;                // settings->proximal of the running solve
    cudaStream_t st = nullptr;
    // matrices (scaled in place at setup)
    int nnzQ = 0, nnzA = 0, nnzQf = 0;
    int *Qp = nullptr, *Qi = nullptr;  double *Qx = nullptr;      // lower CSC (values feed K)
    int *Ap = nullptr, *Ai = nullptr;  double *Ax = nullptr;      // CSC (A' y gathers per column)
    int *Rp = nullptr, *Rj = nullptr, *Rsrc = nullptr; double *Rx = nullptr;   // CSR of A
    int *Fp = nullptr, *Fj = nullptr, *Fsrc = nullptr; double *Fx = nullptr;   // full symmetric CSR of Q
    int *Fdiag = nullptr;            // per row: number of entries with column <= row (split point)
    // vectors (n)
    double *q, *x, *xbar, *Qx_, *Aty, *dx, *Qdx, *Atdy, *df, *res_dual, *res_dual_in, *rhs, *tmp_n;
    double *D, *Dinv;
    // vectors (m)
    double *l, *u, *y, *ybar, *Ax_, *dy, *Adx, *w, *res_prim, *res_prim_old, *res_prim_in, *mu;
    double *sqrt_mu, *omega, *tmp_m, *E, *Einv;
    int *act, *act_old;
    // linesearch scratch (2m)
    double *ls_delta = nullptr, *ls_alpha = nullptr;
    unsigned long long *ls_key[2] = {nullptr, nullptr};
    int *ls_idx[2] = {nullptr, nullptr};
    double *ls_part = nullptr;       // per-block partial sums
    double *ls_scan = nullptr;       // per-block prefix sums
    unsigned *ls_hist = nullptr;
    int ls_blocks = 0, sort_blocks = 0;
    // scaling
    int scaling_iters = 0;
    double c = 1.0, cinv = 1.0, norm_q = 0.0;
    // scalars
    qpdo_dev_scalars *sc = nullptr;       // device
    qpdo_dev_scalars *sc_host = nullptr;  // pinned
    double *red = nullptr;                // reduction scratch
    Factor F;
    // profiling
    int profiling = 0;
    qpdo_dev_profile prof = {};
    std::vector<cudaEvent_t> ev_pool;
    std::vector<std::pair<int, int>> ev_used;   // (phase, index of start event)
    size_t ev_next = 0;
    double analyze_seconds = 0;
    std::vector<void *> allocs;
};

// phases for profiling
enum { PH_ASSEMBLE = 0, PH_FACTOR, PH_SOLVE, PH_SPMV, PH_LS, PH_RES, PH_OTHER, PH_COUNT };

struct PhaseTimer {
    qpdo_dev *d; int phase; size_t idx;
    PhaseTimer(qpdo_dev *d_, int phase_);
    ~PhaseTimer();
};
void qpdo_profile_collect(qpdo_dev *d);

template <class T> int dev_alloc(qpdo_dev *d, T **p, size_t count) {
    void *v = nullptr;
    size_t bytes = sizeof(T) * (count ? count : 1);
    cudaError_t e = cudaMalloc(&v, bytes);
    if (e != cudaSuccess) { qpdo_set_error("cudaMalloc(%zu): %s", bytes, cudaGetErrorString(e)); return -1; }
    cudaMemsetAsync(v, 0, bytes, d->st);
    d->allocs.push_back(v);
    *p = (T *)v;
    return 0;
}
template <class T> int dev_upload(qpdo_dev *d, T **p, const T *h, size_t count) {
    if (dev_alloc(d, p, count)) return -1;
    if (count) CUDA_OK(cudaMemcpyAsync(*p, h, sizeof(T) * count, cudaMemcpyHostToDevice, d->st));
    return 0;
}

// factor.cu
int factor_setup(qpdo_dev *d, int ordering);
int factor_refresh_products(qpdo_dev *d);            // after A values change (scaling)
int factor_numeric(qpdo_dev *d);                     // assemble K(omega, sigma) + factorize
int factor_solve(qpdo_dev *d, const double *rhs_unpermuted_neg, double *dx);   // see trisolve
void factor_destroy(qpdo_dev *d);
// linesearch.cu
int linesearch_setup(qpdo_dev *d);
int linesearch_run(qpdo_dev *d);
// spmv (dev_core.cu)
void spmv_A(qpdo_dev *d, const double *x, double *y);      // y = A x      (m)
void spmv_At(qpdo_dev *d, const double *x, double *y);     // y = A' x     (n)
void spmv_Q(qpdo_dev *d, const double *x, double *y);      // y = Q x      (n)

static inline int cdiv(long a, long b) { return (int)((a + b - 1) / b); }

#endif
