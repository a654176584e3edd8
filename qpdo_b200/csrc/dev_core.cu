// dev_core.cu -- device context, Ruiz equilibration, SpMV.  Compiled with -fmad=false so that the
// element-wise arithmetic rounds exactly like the reference's scalar C (x86-64 SSE2, no FMA);
// the accumulation ORDER of the SpMVs also follows cholmod_sdmult, which makes Ax, A'y and Qx
// bit-identical to the reference on identical inputs.
#include "dev_internal.h"
#include <stdarg.h>
#include <string.h>
#include <chrono>
#include <algorithm>

static thread_local char g_err[512] = "";
void qpdo_set_error(const char *fmt, ...)
{
    va_list ap; va_start(ap, fmt); vsnprintf(g_err, sizeof g_err, fmt, ap); va_end(ap);
}
extern "C" const char *qpdo_dev_last_error(void) { return g_err; }
extern "C" int qpdo_dev_count(void)
{
    int c = 0;
    if (cudaGetDeviceCount(&c) != cudaSuccess) return 0;
    return c;
}

// ------------------------------------------------------------------------------- profiling
PhaseTimer::PhaseTimer(qpdo_dev *d_, int phase_) : d(d_), phase(phase_), idx(0)
{
    if (!d->profiling) return;
    if (d->ev_next + 2 > d->ev_pool.size()) {
        size_t old = d->ev_pool.size();
        d->ev_pool.resize(old + 512);
        for (size_t i = old; i < d->ev_pool.size(); i++) cudaEventCreate(&d->ev_pool[i]);
    }
    idx = d->ev_next; d->ev_next += 2;
    cudaEventRecord(d->ev_pool[idx], d->st);
}
PhaseTimer::~PhaseTimer()
{
    if (!d->profiling) return;
    cudaEventRecord(d->ev_pool[idx + 1], d->st);
    d->ev_used.push_back(std::make_pair(phase, (int)idx));
    if (d->ev_used.size() >= 4096) qpdo_profile_collect(d);
}
void qpdo_profile_collect(qpdo_dev *d)
{
    if (d->ev_used.empty()) return;
    cudaStreamSynchronize(d->st);
    double *slot[PH_COUNT] = {&d->prof.assemble_ms, &d->prof.factor_ms, &d->prof.solve_ms, &d->prof.spmv_ms,
                              &d->prof.linesearch_ms, &d->prof.residual_ms, &d->prof.other_ms};
    for (auto &pr : d->ev_used) {
        float ms = 0;
        cudaEventElapsedTime(&ms, d->ev_pool[pr.second], d->ev_pool[pr.second + 1]);
        *slot[pr.first] += ms;
    }
    d->ev_used.clear();
    d->ev_next = 0;
}
extern "C" int qpdo_dev_set_profiling(qpdo_dev *d, int on) { if (d) cudaSetDevice(d->device); qpdo_profile_collect(d); d->profiling = on; return 0; }
extern "C" int qpdo_dev_get_profile(qpdo_dev *d, qpdo_dev_profile *out, int reset)
{
    if (d) cudaSetDevice(d->device);
    qpdo_profile_collect(d);
    *out = d->prof;
    if (reset) d->prof = qpdo_dev_profile();
    return 0;
}

// ------------------------------------------------------------------------------- small kernels
__global__ void k_gather(double *dst, const double *src, const int *idx, int n)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) dst[i] = src[idx[i]];
}
__global__ void k_fill(double *a, double v, int n)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) a[i] = v;
}
__device__ __forceinline__ void atomic_max_nonneg(double *addr, double v)
{   // v >= 0: the IEEE bit pattern is monotone, max is order independent => deterministic
    atomicMax((unsigned long long *)addr, (unsigned long long)__double_as_longlong(v));
}
// column / row infinity norms of A (cholmod_interface.c:162-199)
__global__ void k_colrow_max(const int *Ap, const int *Ai, const double *Ax, double *colmax, double *rowmax, int n)
{
    int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    double mx = 0.0;
    for (int k = Ap[j]; k < Ap[j + 1]; k++) {
        double v = fabs(Ax[k]);
        mx = v > mx ? v : mx;
        atomic_max_nonneg(&rowmax[Ai[k]], v);
    }
    colmax[j] = mx;
}
// limit_scaling, sqrt, reciprocal (scaling.c:13-18,44-52) and the running product (scaling.c:60-61)
__global__ void k_ruiz_vec(double *t, double *acc, int n)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double v = t[i];
    v = v < 1e-9 ? 1.0 : v;
    v = sqrt(v);
    v = 1.0 / v;
    t[i] = v;
    acc[i] = acc[i] * v;
}
// A <- E A (rows) then A <- A D (columns), two separate roundings like two cholmod_scale calls
__global__ void k_ruiz_scaleA(const int *Ap, const int *Ai, double *Ax, const double *Et, const double *Dt, int n)
{
    int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    double dj = Dt[j];
    for (int k = Ap[j]; k < Ap[j + 1]; k++) {
        double v = Ax[k] * Et[Ai[k]];
        Ax[k] = v * dj;
    }
}
// Q <- D Q D (CHOLMOD_SYM: Ax[p] *= t * s[i], cholmod_scale.c:178-190)
__global__ void k_scaleQ_sym(const int *Qp, const int *Qi, double *Qx, const double *D, int n)
{
    int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    double t = D[j];
    for (int k = Qp[j]; k < Qp[j + 1]; k++) { double f = t * D[Qi[k]]; Qx[k] = Qx[k] * f; }
}
__global__ void k_scale_scalar(double *a, double s, int n)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) a[i] = a[i] * s;
}
__global__ void k_ew_prod(const double *a, const double *b, double *c, int n)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) c[i] = a[i] * b[i];
}
__global__ void k_recip(const double *a, double *b, int n)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) b[i] = 1.0 / a[i];
}
// c = a + s*b (vec_add_scaled, lin_alg.c:93-98)
__global__ void k_axpy_out(const double *a, const double *b, double *c, double s, int n)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) c[i] = a[i] + s * b[i];
}
// out[0] = max_i |a_i * (b ? b_i : 1)|  (out zeroed by the caller)
__global__ void k_norm_inf_prod(const double *a, const double *b, double *out, int n)
{
    double mx = 0.0;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        double v = fabs(b ? a[i] * b[i] : a[i]);
        mx = v > mx ? v : mx;
    }
    for (int o = 16; o; o >>= 1) { double t = __shfl_xor_sync(0xffffffffu, mx, o); mx = t > mx ? t : mx; }
    if ((threadIdx.x & 31) == 0 && mx > 0.0) atomic_max_nonneg(out, mx);
}

// ------------------------------------------------------------------------------- SpMV
// y = A x: CSR row gather in ascending column order == the order cholmod_sdmult's CSC scatter
// adds into y[i] (t_cholmod_sdmult.c:297-320)
__global__ void k_spmv_rows(const int *Rp, const int *Rj, const double *Rx, const double *x, double *y, int m)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= m) return;
    double s = 0.0;
    for (int k = Rp[i]; k < Rp[i + 1]; k++) s += Rx[k] * __ldg(&x[Rj[k]]);
    y[i] = s;
}
// y = Q x, Q symmetric stored as full CSR; split sum reproduces t_cholmod_sdmult.c:440-484
__global__ void k_spmv_sym(const int *Fp, const int *Fj, const double *Fx, const int *Fdiag, const double *x,
                           double *y, int n)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int b = Fp[i], mid = b + Fdiag[i], e = Fp[i + 1];
    double s1 = 0.0, s2 = 0.0;
    for (int k = b; k < mid; k++) s1 += Fx[k] * __ldg(&x[Fj[k]]);
    for (int k = mid; k < e; k++) s2 += Fx[k] * __ldg(&x[Fj[k]]);
    y[i] = s1 + s2;
}
void spmv_A(qpdo_dev *d, const double *x, double *y)
{
    if (d->m) k_spmv_rows<<<cdiv(d->m, 128), 128, 0, d->st>>>(d->Rp, d->Rj, d->Rx, x, y, d->m);
}
void spmv_At(qpdo_dev *d, const double *x, double *y)
{   // column j of A = row j of A' (t_cholmod_sdmult.c:141-165)
    k_spmv_rows<<<cdiv(d->n, 128), 128, 0, d->st>>>(d->Ap, d->Ai, d->Ax, x, y, d->n);
}
void spmv_Q(qpdo_dev *d, const double *x, double *y)
{
    k_spmv_sym<<<cdiv(d->n, 128), 128, 0, d->st>>>(d->Fp, d->Fj, d->Fx, d->Fdiag, x, y, d->n);
}

// ------------------------------------------------------------------------------- scale_data
// Ruiz equilibration + cost scaling, scaling.c:24-91.  qx_vec is work->Qx at the time of the call
// (all zero in qpdo_setup, the current Qx when called again from qpdo_update_settings).
static int run_scaling(qpdo_dev *d, int iters)
{
    const int n = d->n, m = d->m, T = 128;
    k_fill<<<cdiv(n, T), T, 0, d->st>>>(d->D, 1.0, n);
    if (m) k_fill<<<cdiv(m, T), T, 0, d->st>>>(d->E, 1.0, m);
    for (int it = 0; it < iters; it++) {
        CUDA_OK(cudaMemsetAsync(d->tmp_m, 0, sizeof(double) * (m ? m : 1), d->st));
        k_colrow_max<<<cdiv(n, T), T, 0, d->st>>>(d->Ap, d->Ai, d->Ax, d->tmp_n, d->tmp_m, n);
        k_ruiz_vec<<<cdiv(n, T), T, 0, d->st>>>(d->tmp_n, d->D, n);
        if (m) k_ruiz_vec<<<cdiv(m, T), T, 0, d->st>>>(d->tmp_m, d->E, m);
        k_ruiz_scaleA<<<cdiv(n, T), T, 0, d->st>>>(d->Ap, d->Ai, d->Ax, d->tmp_m, d->tmp_n, n);
    }
    k_scaleQ_sym<<<cdiv(n, T), T, 0, d->st>>>(d->Qp, d->Qi, d->Qx, d->D, n);
    k_ew_prod<<<cdiv(n, T), T, 0, d->st>>>(d->D, d->q, d->q, n);
    // c = 1 / max(1, || Qx + q ||_inf)
    k_axpy_out<<<cdiv(n, T), T, 0, d->st>>>(d->Qx_, d->q, d->tmp_n, 1.0, n);
    CUDA_OK(cudaMemsetAsync(d->red, 0, sizeof(double), d->st));
    k_norm_inf_prod<<<std::min(cdiv(n, 256), 1024), 256, 0, d->st>>>(d->tmp_n, nullptr, d->red, n);
    double nrm = 0;
    CUDA_OK(cudaMemcpyAsync(&nrm, d->red, sizeof(double), cudaMemcpyDeviceToHost, d->st));
    CUDA_OK(cudaStreamSynchronize(d->st));
    d->c = 1.0 / (nrm > 1.0 ? nrm : 1.0);
    k_scale_scalar<<<cdiv(n, T), T, 0, d->st>>>(d->q, d->c, n);
    if (d->nnzQ) k_scale_scalar<<<cdiv(d->nnzQ, T), T, 0, d->st>>>(d->Qx, d->c, d->nnzQ);
    k_recip<<<cdiv(n, T), T, 0, d->st>>>(d->D, d->Dinv, n);
    if (m) k_recip<<<cdiv(m, T), T, 0, d->st>>>(d->E, d->Einv, m);
    d->cinv = 1.0 / d->c;
    if (m) {
        k_ew_prod<<<cdiv(m, T), T, 0, d->st>>>(d->E, d->l, d->l, m);
        k_ew_prod<<<cdiv(m, T), T, 0, d->st>>>(d->E, d->u, d->u, m);
    }
    return 0;
}
static int refresh_matrix_copies(qpdo_dev *d)
{
    const int T = 128;
    if (d->nnzA) k_gather<<<cdiv(d->nnzA, T), T, 0, d->st>>>(d->Rx, d->Ax, d->Rsrc, d->nnzA);
    if (d->nnzQf) k_gather<<<cdiv(d->nnzQf, T), T, 0, d->st>>>(d->Fx, d->Qx, d->Fsrc, d->nnzQf);
    return factor_refresh_products(d);
}
static int compute_norm_q(qpdo_dev *d)
{
    CUDA_OK(cudaMemsetAsync(d->red, 0, sizeof(double), d->st));
    k_norm_inf_prod<<<std::min(cdiv(d->n, 256), 1024), 256, 0, d->st>>>(d->q, d->scaling_iters ? d->Dinv : nullptr, d->red, d->n);
    CUDA_OK(cudaMemcpyAsync(&d->norm_q, d->red, sizeof(double), cudaMemcpyDeviceToHost, d->st));
    CUDA_OK(cudaStreamSynchronize(d->st));
    return 0;
}

// ------------------------------------------------------------------------------- create / destroy
extern "C" int qpdo_dev_create(qpdo_dev **out, const qpdo_dev_problem *p, int scaling_iters, int ordering, int device)
{
    *out = nullptr;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        qpdo_set_error("no CUDA device available: qpdo_b200 has no CPU fallback");
        return -2;
    }
    if (device < 0) { if (cudaGetDevice(&device) != cudaSuccess) device = 0; }
    CUDA_OK(cudaSetDevice(device));
    qpdo_dev *d = new qpdo_dev();
    d->n = p->n; d->m = p->m; d->device = device; d->scaling_iters = scaling_iters;
    const int n = p->n, m = p->m;
    int rc = -1;
    do {
        if (cudaStreamCreateWithFlags(&d->st, cudaStreamNonBlocking) != cudaSuccess) { qpdo_set_error("cudaStreamCreate failed"); break; }
        d->nnzQ = p->Qp[n]; d->nnzA = p->Ap[n];
        // host-side index structures: CSR of A, full symmetric CSR of Q
        std::vector<int> Rp(m + 1, 0), Rj(d->nnzA), Rsrc(d->nnzA);
        for (int k = 0; k < d->nnzA; k++) Rp[p->Ai[k] + 1]++;
        for (int i = 0; i < m; i++) Rp[i + 1] += Rp[i];
        {
            std::vector<int> nx(Rp.begin(), Rp.end() - 1);
            for (int j = 0; j < n; j++) for (int k = p->Ap[j]; k < p->Ap[j + 1]; k++) { int t = nx[p->Ai[k]]++; Rj[t] = j; Rsrc[t] = k; }
        }
        std::vector<int> Fp(n + 1, 0), Fdiag(n, 0);
        for (int j = 0; j < n; j++) for (int k = p->Qp[j]; k < p->Qp[j + 1]; k++) {
            int i = p->Qi[k];
            Fp[i + 1]++; Fdiag[i]++;                   // (i, j) with j <= i
            if (i != j) Fp[j + 1]++;                   // mirrored (j, i)
        }
        for (int i = 0; i < n; i++) Fp[i + 1] += Fp[i];
        d->nnzQf = Fp[n];
        std::vector<int> Fj(d->nnzQf), Fsrc(d->nnzQf);
        {
            // row i: first the entries (i, j<=i) in ascending j (scan columns ascending), then the
            // mirrored (i, k>i) = column i's sub-diagonal rows in ascending k
            std::vector<int> nx(n);
            for (int i = 0; i < n; i++) nx[i] = Fp[i];
            for (int j = 0; j < n; j++) for (int k = p->Qp[j]; k < p->Qp[j + 1]; k++) { int i = p->Qi[k]; int t = nx[i]++; Fj[t] = j; Fsrc[t] = k; }
            for (int j = 0; j < n; j++) for (int k = p->Qp[j]; k < p->Qp[j + 1]; k++) { int i = p->Qi[k]; if (i != j) { int t = nx[j]++; Fj[t] = i; Fsrc[t] = k; } }
        }
        if (dev_upload(d, &d->Qp, p->Qp, n + 1) || dev_upload(d, &d->Qi, p->Qi, d->nnzQ) || dev_upload(d, &d->Qx, p->Qx, d->nnzQ) ||
            dev_upload(d, &d->Ap, p->Ap, n + 1) || dev_upload(d, &d->Ai, p->Ai, d->nnzA) || dev_upload(d, &d->Ax, p->Ax, d->nnzA) ||
            dev_upload(d, &d->Rp, Rp.data(), m + 1) || dev_upload(d, &d->Rj, Rj.data(), d->nnzA) || dev_upload(d, &d->Rsrc, Rsrc.data(), d->nnzA) ||
            dev_alloc(d, &d->Rx, d->nnzA) ||
            dev_upload(d, &d->Fp, Fp.data(), n + 1) || dev_upload(d, &d->Fj, Fj.data(), d->nnzQf) || dev_upload(d, &d->Fsrc, Fsrc.data(), d->nnzQf) ||
            dev_upload(d, &d->Fdiag, Fdiag.data(), n) || dev_alloc(d, &d->Fx, d->nnzQf)) break;
        double **nv[] = {&d->x, &d->xbar, &d->Qx_, &d->Aty, &d->dx, &d->Qdx, &d->Atdy, &d->df, &d->res_dual,
                         &d->res_dual_in, &d->rhs, &d->tmp_n, &d->D, &d->Dinv};
        double **mv[] = {&d->y, &d->ybar, &d->Ax_, &d->dy, &d->Adx, &d->w, &d->res_prim, &d->res_prim_old,
                         &d->res_prim_in, &d->mu, &d->sqrt_mu, &d->omega, &d->tmp_m, &d->E, &d->Einv};
        bool bad = false;
        for (auto pp : nv) if (dev_alloc(d, pp, n)) bad = true;
        for (auto pp : mv) if (dev_alloc(d, pp, m)) bad = true;
        if (bad) break;
        if (dev_upload(d, &d->q, p->q, n) || dev_upload(d, &d->l, p->l, m) || dev_upload(d, &d->u, p->u, m)) break;
        if (dev_alloc(d, &d->act, m) || dev_alloc(d, &d->act_old, m) || dev_alloc(d, &d->sc, 1) || dev_alloc(d, &d->red, 1024)) break;
        if (cudaMallocHost((void **)&d->sc_host, sizeof(qpdo_dev_scalars)) != cudaSuccess) { qpdo_set_error("cudaMallocHost failed"); break; }
        memset(d->sc_host, 0, sizeof(qpdo_dev_scalars));
        // symbolic analysis on the host (pattern only, independent of the scaling)
        if (factor_setup(d, ordering)) break;            // accumulates d->analyze_seconds
        if (linesearch_setup(d)) break;
        if (scaling_iters > 0) { if (run_scaling(d, scaling_iters)) break; }
        else {
            k_fill<<<cdiv(n, 128), 128, 0, d->st>>>(d->D, 1.0, n); k_fill<<<cdiv(n, 128), 128, 0, d->st>>>(d->Dinv, 1.0, n);
            if (m) { k_fill<<<cdiv(m, 128), 128, 0, d->st>>>(d->E, 1.0, m); k_fill<<<cdiv(m, 128), 128, 0, d->st>>>(d->Einv, 1.0, m); }
        }
        if (refresh_matrix_copies(d)) break;
        if (compute_norm_q(d)) break;
        cudaError_t e = cudaStreamSynchronize(d->st);
        if (e != cudaSuccess) { qpdo_set_error("setup kernels failed: %s", cudaGetErrorString(e)); break; }
        rc = 0;
    } while (0);
    if (rc) { qpdo_dev_destroy(d); return rc; }
    *out = d;
    return 0;
}

extern "C" void qpdo_dev_destroy(qpdo_dev *d)
{
    if (!d) return;
    cudaSetDevice(d->device);
    if (d->st) cudaStreamSynchronize(d->st);
    factor_destroy(d);
    for (void *p : d->allocs) cudaFree(p);
    for (auto &e : d->ev_pool) cudaEventDestroy(e);
    if (d->sc_host) cudaFreeHost(d->sc_host);
    if (d->st) cudaStreamDestroy(d->st);
    delete d;
}

extern "C" int qpdo_dev_get_scaling(qpdo_dev *d, double *D, double *Dinv, double *E, double *Einv, double *c,
                                    double *cinv, double *norm_q)
{
    if (d) cudaSetDevice(d->device);
    const size_t nb = sizeof(double) * d->n, mb = sizeof(double) * d->m;
    if (D) CUDA_OK(cudaMemcpyAsync(D, d->D, nb, cudaMemcpyDeviceToHost, d->st));
    if (Dinv) CUDA_OK(cudaMemcpyAsync(Dinv, d->Dinv, nb, cudaMemcpyDeviceToHost, d->st));
    if (E && d->m) CUDA_OK(cudaMemcpyAsync(E, d->E, mb, cudaMemcpyDeviceToHost, d->st));
    if (Einv && d->m) CUDA_OK(cudaMemcpyAsync(Einv, d->Einv, mb, cudaMemcpyDeviceToHost, d->st));
    CUDA_OK(cudaStreamSynchronize(d->st));
    if (c) *c = d->c;
    if (cinv) *cinv = d->cinv;
    if (norm_q) *norm_q = d->norm_q;
    return 0;
}

// qpdo_update_settings with more scaling sweeps (qpdo.c:496-512): the extra sweeps run on the already scaled data
// exactly as a fresh scale_data would (D, E restart from 1; Q, q, l, u are scaled by the increment; the cost scale is
// recomputed with the CURRENT Qx, as the reference does), then the total D, E, c are the products with the saved ones
// (see the note on D below).
// The iterates are left as they are, like in the reference: callers warm-start afterwards.
extern "C" int qpdo_dev_rescale(qpdo_dev *d, int extra_iters)
{
    if (d) cudaSetDevice(d->device);
    if (extra_iters <= 0) return 0;
    if (d->scaling_iters <= 0) {      // the reference dereferences work->scaling == NULL here (qpdo.c:498): an error instead
        qpdo_set_error("qpdo_update_settings: scaling cannot be switched on after a setup with scaling = 0");
        return -3;
    }
    const int n = d->n, m = d->m, T = 128;
    double *save = nullptr;
    CUDA_OK(cudaMalloc(&save, sizeof(double) * ((size_t)n + m + 1)));
    double *saveD = save, *saveE = save + n;
    const double c_old = d->c;
    int rc = -1;
    do {
        if (cudaMemcpyAsync(saveD, d->D, sizeof(double) * n, cudaMemcpyDeviceToDevice, d->st) != cudaSuccess) break;
        if (m && cudaMemcpyAsync(saveE, d->E, sizeof(double) * m, cudaMemcpyDeviceToDevice, d->st) != cudaSuccess) break;
        if (run_scaling(d, extra_iters)) break;
        // The reference multiplies by work->temp_n, which its own scale_data has meanwhile overwritten with Qx + q
        // (scaling.c:68 after qpdo.c:498): its total D is D_increment o (Qx + q), not D_increment o D_old, and every
        // later un-scaling of x is off.  Default here: the documented intent (product with the saved D).
        // QPDO_B200_COMPAT_RESCALE=1 reproduces the reference bit for bit (d->tmp_n still holds Qx + q).
        const char *compat = getenv("QPDO_B200_COMPAT_RESCALE");
        k_ew_prod<<<cdiv(n, T), T, 0, d->st>>>(d->D, (compat && compat[0] == '1') ? d->tmp_n : saveD, d->D, n);
        if (m) k_ew_prod<<<cdiv(m, T), T, 0, d->st>>>(d->E, saveE, d->E, m);
        d->c = d->c * c_old;
        k_recip<<<cdiv(n, T), T, 0, d->st>>>(d->D, d->Dinv, n);
        if (m) k_recip<<<cdiv(m, T), T, 0, d->st>>>(d->E, d->Einv, m);
        d->cinv = 1.0 / d->c;
        d->scaling_iters += extra_iters;
        if (refresh_matrix_copies(d)) break;       // norm_q is left as it was, like in the reference (nothing reads it)
        rc = 0;
    } while (0);
    cudaStreamSynchronize(d->st);
    cudaFree(save);
    if (rc) qpdo_set_error("qpdo_dev_rescale failed: %s", cudaGetErrorString(cudaGetLastError()));
    return rc;
}

extern "C" int qpdo_dev_get_symbolic_info(qpdo_dev *d, qpdo_dev_symbolic_info *o)
{
    if (d) cudaSetDevice(d->device);
    const qpdo_symbolic *S = d->F.S;
    if (!S) { memset(o, 0, sizeof *o); o->analyze_seconds = d->analyze_seconds; return 0; }      // dynamic pattern, nothing factored yet
    o->nsuper = S->nsuper; o->nlevel = S->nlevel; o->max_front = S->max_front; o->max_nc = S->max_nc;
    o->lnz = S->lnz; o->lnz_stored = S->lnz_stored; o->usize = S->usize; o->nnzK = S->nnzK; o->ntriple = S->ntriple;
    o->flops = S->flops; o->analyze_seconds = d->analyze_seconds;
    return 0;
}
