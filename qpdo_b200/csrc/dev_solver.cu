// dev_solver.cu -- the per-iteration vector work of QPDO's hot loop as fused kernels:
// residuals + norms (iteration.c:30-93, termination.c:35-77), active set and Newton right-hand
// side (newton.c:13-126), iterate update (iteration.c:19-24), penalty updates (iteration.c:98-180),
// infeasibility certificates (termination.c:97-216), objective (iteration.c:185-221).
// Compiled with -fmad=false: every a + s*b rounds like the reference's unfused scalar C.
#include "dev_internal.h"
#include <string.h>
#include <algorithm>

#define TPB 256
#define RED_BLOCKS 120

__device__ __forceinline__ void atomic_max_nonneg(double *addr, double v)
{
    atomicMax((unsigned long long *)addr, (unsigned long long)__double_as_longlong(v));
}
__device__ __forceinline__ double warp_max(double v)
{
    for (int o = 16; o; o >>= 1) { double t = __shfl_xor_sync(0xffffffffu, v, o); v = t > v ? t : v; }
    return v;
}
// fixed-shape block sum (xor tree inside warps, then the warp partials in index order)
__device__ __forceinline__ double block_sum(double v, double *sh)
{
    for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    int w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    __syncthreads();
    if ((threadIdx.x & 31) == 0) sh[w] = v;
    __syncthreads();
    double s = 0.0;
    for (int i = 0; i < nw; i++) s += sh[i];
    return s;
}
__device__ __forceinline__ double clampd(double a, double lo, double hi)
{   // c_max(bmin, c_min(a, bmax))  (lin_alg.c:163-168)
    double t = a < hi ? a : hi;
    return lo > t ? lo : t;
}

// ---------------------------------------------------------------------------- dot products
// part[k*RED_BLOCKS + blk] = partial of pair k; up to 4 pairs; finalised by k_dot_final
struct DotArgs { const double *a[4]; const double *b[4]; int len[4]; int npair; };
__global__ void k_dot_partial(DotArgs A, double *part)
{
    __shared__ double sh[32];
    for (int k = 0; k < A.npair; k++) {
        double s = 0.0;
        const double *a = A.a[k], *b = A.b[k];
        for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < A.len[k]; i += gridDim.x * blockDim.x) s += a[i] * b[i];
        s = block_sum(s, sh);
        if (threadIdx.x == 0) part[k * RED_BLOCKS + blockIdx.x] = s;
    }
}
__global__ void k_dot_final(const double *part, double *out, int npair, int nblk)
{   // one warp per pair, fixed order
    __shared__ double sh[32];
    for (int k = 0; k < npair; k++) {
        double s = 0.0;
        for (int i = threadIdx.x; i < nblk; i += blockDim.x) s += part[k * RED_BLOCKS + i];
        s = block_sum(s, sh);
        if (threadIdx.x == 0) out[k] = s;
    }
}
static void dots(qpdo_dev *d, const DotArgs &A, double *out)
{
    int mx = 1;
    for (int k = 0; k < A.npair; k++) mx = std::max(mx, A.len[k]);
    int nb = std::min(RED_BLOCKS, cdiv(mx, TPB));
    k_dot_partial<<<nb, TPB, 0, d->st>>>(A, d->red + 16);
    k_dot_final<<<1, 128, 0, d->st>>>(d->red + 16, out, A.npair, nb);
}

// ---------------------------------------------------------------------------- generic vector ops
__global__ void k_axpy(const double *a, const double *b, double *c, double s, int n)
{   // vec_add_scaled (lin_alg.c:93-98)
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) c[i] = a[i] + s * b[i];
}
__global__ void k_axpy_dev(const double *a, const double *b, double *c, const double *s, double mult, int n)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) { double sv = (*s) * mult; c[i] = a[i] + sv * b[i]; }
}
__global__ void k_mul(const double *a, const double *b, double *c, int n)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) c[i] = a[i] * b[i];
}
__global__ void k_norm_inf2(const double *a, const double *b, double *out, int n)
{
    double mx = 0.0;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        double v = fabs(b ? a[i] * b[i] : a[i]);
        mx = v > mx ? v : mx;
    }
    mx = warp_max(mx);
    if ((threadIdx.x & 31) == 0 && mx > 0.0) atomic_max_nonneg(out, mx);
}
static void norm_inf(qpdo_dev *d, const double *a, const double *b, double *out, int n)
{
    cudaMemsetAsync(out, 0, sizeof(double), d->st);
    if (n) k_norm_inf2<<<std::min(cdiv(n, TPB), 592), TPB, 0, d->st>>>(a, b, out, n);
}

// ---------------------------------------------------------------------------- warm start
__global__ void k_warm_primal(const double *xin, const double *Dinv, double *x, double *xbar, double *dx, int scaled, int n)
{   // qpdo.c:239-246
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double v = xin[i];
    if (scaled) v = v * Dinv[i];
    x[i] = v; xbar[i] = v; dx[i] = v;
}
__global__ void k_warm_dual(const double *yin, const double *Einv, double c, double *y, double *ybar, double *dy, int scaled, int m)
{   // qpdo.c:269-277
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= m) return;
    double v = yin[i];
    if (scaled) { v = v * Einv[i]; v = v * c; }
    y[i] = v; ybar[i] = v; dy[i] = v;
}
// initialize_mu (iteration.c:98-122); dots[0] = x'Qx, dots[1] = q'x
__global__ void k_init_mu(const double *Ax, const double *l, const double *u, const double *dots, double *mu,
                          double *sqrt_mu, int m)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= m) return;
    double f = 0.5 * dots[0] + dots[1];
    double z = clampd(Ax[i], l[i], u[i]);
    double t = Ax[i] + (-1.0) * z;
    double num = 0.5 * t * t;
    num = num > 1.0 ? num : 1.0;
    double den = fabs(f);
    den = den > 1.0 ? den : 1.0;
    double v = 0.1 * num / den;
    v = v < 1e3 ? v : 1e3;
    v = v > 1e-3 ? v : 1e-3;
    mu[i] = v;
    double s = sqrt(v);
    sqrt_mu[i] = 1.0 / s;
}

// ---------------------------------------------------------------------------- residuals
struct ResArgs {
    int n, m, scaled, proximal;
    double cinv;
    const double *Ax, *y, *ybar, *l, *u, *mu, *E, *Einv;
    const double *Qx, *q, *x, *xbar, *Aty, *Dinv;
    double *res_prim, *res_prim_in, *w, *df, *res_dual, *res_dual_in;
    qpdo_dev_scalars *sc;
};
// one constraint row of the residual block (iteration.c:38-46, 75-81): same statements as the scalar kernel of round 1
__device__ __forceinline__ void res_row(const ResArgs &A, double Ax, double y, double l, double u, double mu, double yb, double E, double Einv,
                                        double &rp, double &w, double &rpi, double &np, double &npi)
{
    double t;
    if (A.scaled) { t = E * y; t = t * A.cinv; t = E * t; t = Ax + t; } else t = Ax + y;
    t = clampd(t, l, u);
    rp = Ax + (-1.0) * t;
    w = Ax + mu * (yb - 0.5 * y);
    const double z = clampd(w, l, u);
    rpi = Ax + mu * (yb - y) - z;
    const double a = fabs(A.scaled ? Einv * rp : rp), b = fabs(A.scaled ? Einv * rpi : rpi);
    np = a > np ? a : np; npi = b > npi ? b : npi;
}
// one variable (iteration.c:52, 89)
__device__ __forceinline__ void res_col(const ResArgs &A, double sigma, double Qx, double q, double x, double xbar, double Aty, double Dinv,
                                        double &rd, double &dfo, double &rdi, double &nd, double &ndi)
{
    double df = Qx + q;
    if (A.proximal) { rd = df + (-sigma) * x; rd = rd + Aty; } else rd = df + Aty;
    if (A.proximal) df = df + (-sigma) * xbar;
    dfo = df;
    rdi = df + Aty;
    const double a = fabs(A.scaled ? Dinv * rd : rd), b = fabs(A.scaled ? Dinv * rdi : rdi);
    nd = a > nd ? a : nd; ndi = b > ndi ? b : ndi;
}
#define LD2(v, p) do { if (two) v = *reinterpret_cast<const double2 *>((p) + i0); else { v.x = (p)[i0]; v.y = 0.0; } } while (0)
#define ST2(p, v) do { if (two) *reinterpret_cast<double2 *>((p) + i0) = v; else (p)[i0] = v.x; } while (0)
__global__ void k_residuals(ResArgs A)
{   // thread t owns elements 2 t, 2 t + 1 of every vector: 128-bit loads and stores
    const int i0 = 2 * (blockIdx.x * blockDim.x + threadIdx.x);
    double np = 0.0, npi = 0.0, nd = 0.0, ndi = 0.0;
    if (i0 < A.m) {
        const bool two = i0 + 1 < A.m;
        double2 Ax, y, l, u, mu, yb, E = make_double2(1.0, 1.0), Ei = make_double2(1.0, 1.0), rp, w, rpi;
        LD2(Ax, A.Ax); LD2(y, A.y); LD2(l, A.l); LD2(u, A.u); LD2(mu, A.mu); LD2(yb, A.ybar);
        if (A.scaled) { LD2(E, A.E); LD2(Ei, A.Einv); }
        res_row(A, Ax.x, y.x, l.x, u.x, mu.x, yb.x, E.x, Ei.x, rp.x, w.x, rpi.x, np, npi);
        if (two) res_row(A, Ax.y, y.y, l.y, u.y, mu.y, yb.y, E.y, Ei.y, rp.y, w.y, rpi.y, np, npi);
        ST2(A.res_prim, rp); ST2(A.w, w); ST2(A.res_prim_in, rpi);
    }
    if (i0 < A.n) {
        const bool two = i0 + 1 < A.n;
        const double sigma = A.sc->sigma;
        double2 Qx, q, x = make_double2(0.0, 0.0), xb = make_double2(0.0, 0.0), Aty, Di = make_double2(1.0, 1.0), rd, df, rdi;
        LD2(Qx, A.Qx); LD2(q, A.q); LD2(Aty, A.Aty);
        if (A.proximal) { LD2(x, A.x); LD2(xb, A.xbar); }
        if (A.scaled) LD2(Di, A.Dinv);
        res_col(A, sigma, Qx.x, q.x, x.x, xb.x, Aty.x, Di.x, rd.x, df.x, rdi.x, nd, ndi);
        if (two) res_col(A, sigma, Qx.y, q.y, x.y, xb.y, Aty.y, Di.y, rd.y, df.y, rdi.y, nd, ndi);
        ST2(A.res_dual, rd); ST2(A.df, df); ST2(A.res_dual_in, rdi);
    }
    np = warp_max(np); npi = warp_max(npi); nd = warp_max(nd); ndi = warp_max(ndi);
    if ((threadIdx.x & 31) == 0) {
        if (np > 0.0) atomic_max_nonneg(&A.sc->res_prim_norm, np);
        if (npi > 0.0) atomic_max_nonneg(&A.sc->res_prim_in_norm, npi);
        if (nd > 0.0) atomic_max_nonneg(&A.sc->res_dual_norm, nd);
        if (ndi > 0.0) atomic_max_nonneg(&A.sc->res_dual_in_norm, ndi);
    }
}
__global__ void k_scale_dual_norms(qpdo_dev_scalars *sc, double cinv)
{   // termination.c:46,72: norm *= cinv
    sc->res_dual_norm = sc->res_dual_norm * cinv;
    sc->res_dual_in_norm = sc->res_dual_in_norm * cinv;
}

// ---------------------------------------------------------------------------- Newton step pieces
__global__ void k_active(const double *w, const double *l, const double *u, int *act, const int *act_old,
                         qpdo_dev_scalars *sc, int m)
{   // active_constraints + enter_leave_constraints (newton.c:96-126)
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    int a = 0, e = 0, lv = 0;
    if (i < m) {
        double wi = w[i];
        a = (wi <= l[i]) || (wi >= u[i]);
        act[i] = a;
        int o = act_old[i];
        e = a && !o; lv = !a && o;
    }
    unsigned ba = __ballot_sync(0xffffffffu, a), be = __ballot_sync(0xffffffffu, e), bl = __ballot_sync(0xffffffffu, lv);
    if ((threadIdx.x & 31) == 0) {
        if (ba) atomicAdd(&sc->n_active, __popc(ba));
        if (be) atomicAdd(&sc->n_enter, __popc(be));
        if (bl) atomicAdd(&sc->n_leave, __popc(bl));
    }
}
__global__ void k_decide(qpdo_dev_scalars *sc, int force_reset)
{   // newton.c:21-33
    int reset = sc->reset_newton | force_reset;
    int mode;
    if ((reset && sc->n_active) || (sc->n_enter + sc->n_leave) > 100) { mode = 0; reset = 0; sc->n_refactor++; }
    else if (sc->n_active) { mode = 1; sc->n_update++; }
    else { mode = 2; sc->n_qonly++; }
    sc->reset_newton = reset;
    sc->factor_mode = mode;
}
// row weights embodied in the factor (DESIGN.md "omega"): refactor -> act/mu, rank update -> +-1/mu
// for entering/leaving rows, Q only -> 0; plus "dy" = (I+P) res_prim_in ./ mu (newton.c:37-40)
__global__ void k_apply_omega(const qpdo_dev_scalars *sc, const int *act, int *act_old, const double *mu,
                              const double *res_prim_in, double *omega, double *dy, int m)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= m) return;
    int mode = sc->factor_mode, a = act[i], o = act_old[i];
    double inv = 1.0 / mu[i];
    double om;
    if (mode == 0) om = a ? inv : 0.0;
    else if (mode == 1) { om = omega[i]; if (a && !o) om = om + inv; if (!a && o) om = om - inv; }
    else om = 0.0;
    omega[i] = om;
    act_old[i] = a;                                         // newton.c:69
    double v = res_prim_in[i] / mu[i];
    if (!a) v = v * 2.0;
    dy[i] = v;
}
__global__ void k_rhs(const double *res_dual_in, const double *Atdy, double *rhs, int n)
{   // newton.c:43-45
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) rhs[i] = -res_dual_in[i] - Atdy[i];
}
__global__ void k_dy_active(const int *act, const double *Adx, const double *mu, double *dy, int m)
{   // newton.c:61-63
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < m && act[i]) dy[i] = dy[i] + Adx[i] / mu[i];
}
struct UpdArgs { int n, m; double *x, *y, *Qx, *Ax, *Aty; const double *dx, *dy, *Qdx, *Adx, *Atdy; const qpdo_dev_scalars *sc; };
// a[i] += tau * b[i] for elements 2 t, 2 t + 1 (128-bit accesses; the arrays are cudaMalloc'ed one by one: 256-byte aligned)
__device__ __forceinline__ void axpy2(double *a, const double *b, double tau, int t, int len)
{
    const int i = 2 * t;
    if (i + 1 < len) {
        double2 x = *reinterpret_cast<double2 *>(a + i);
        const double2 y = *reinterpret_cast<const double2 *>(b + i);
        x.x = x.x + tau * y.x; x.y = x.y + tau * y.y;
        *reinterpret_cast<double2 *>(a + i) = x;
    } else if (i < len) a[i] = a[i] + tau * b[i];
}
__global__ void k_update_iterate(UpdArgs A)
{   // iteration.c:19-24; thread t owns elements 2 t and 2 t + 1 of every vector
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    const double tau = A.sc->tau;
    axpy2(A.x, A.dx, tau, t, A.n);
    axpy2(A.Qx, A.Qdx, tau, t, A.n);
    axpy2(A.Aty, A.Atdy, tau, t, A.n);
    axpy2(A.y, A.dy, tau, t, A.m);
    axpy2(A.Ax, A.Adx, tau, t, A.m);
}

// ---------------------------------------------------------------------------- outer update pieces
// update_mu (iteration.c:127-168).  domega[k] = weight the reference's ldlupdate_mu_changed would
// add for row k (cholmod_interface.c:77-93): (sqrt_mu_new * sqrt(1 - 1/s^2))^2, for EVERY changed
// row, active or not.
__global__ void k_update_mu(const double *res_prim, const double *res_prim_old, const double *nrm, double eps_abs,
                            double theta, double delta, double mu_min, double sqrt_mu_min, double *mu, double *sqrt_mu,
                            double *domega, qpdo_dev_scalars *sc, int m)
{
    int k = blockIdx.x * blockDim.x + threadIdx.x;
    int changed = 0;
    if (k < m) {
        double r = fabs(res_prim[k]), ro = fabs(res_prim_old[k]);
        double thr = theta * ro;
        thr = eps_abs > thr ? eps_abs : thr;
        double dw = 0.0;
        if (r > thr) {
            double t = delta * nrm[0] / r;
            t = t < 1.0 ? t : 1.0;
            double mu_factor = 1.0 / t;
            double mu_old = mu[k];
            double mu_new = mu_old / mu_factor;
            double s;
            if (mu_new >= mu_min) {
                changed = (mu_old != mu_new);
                mu[k] = mu_new;
                s = sqrt(mu_factor);
                sqrt_mu[k] = s * sqrt_mu[k];
            } else {
                changed = (mu_old != mu_min);
                mu[k] = mu_min;
                s = sqrt_mu_min / sqrt_mu[k];
                sqrt_mu[k] = sqrt_mu_min;
            }
            if (changed) {
                double f = sqrt(1.0 - 1.0 / (s * s));
                double col = sqrt_mu[k] * f;
                dw = col * col;
            }
        }
        domega[k] = dw;
    }
    unsigned b = __ballot_sync(0xffffffffu, changed);
    if ((threadIdx.x & 31) == 0 && b) atomicAdd(&sc->n_mu_changed, __popc(b));
}
__global__ void k_mu_apply(qpdo_dev_scalars *sc, int proximal, double sigma_min, const double *domega, double *omega, int m)
{   // iteration.c:161-167
    int k = blockIdx.x * blockDim.x + threadIdx.x;
    int nch = sc->n_mu_changed;
    bool reset = (proximal && sc->sigma > sigma_min) || (nch > 25);
    if (reset) { if (k == 0) sc->reset_newton = 1; return; }
    if (nch == 0) return;
    if (k < m) omega[k] = omega[k] + domega[k];
}
__global__ void k_set_sigma(qpdo_dev_scalars *sc, double sigma, int set_reset)
{
    sc->sigma = sigma;
    if (set_reset) sc->reset_newton = 1;
}
__global__ void k_copy2(const double *a, double *b, int na, const double *c, double *e, int nc)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < na) b[i] = a[i];
    if (i < nc) e[i] = c[i];
}

// ---------------------------------------------------------------------------- infeasibility pieces
// out_of_bounds partial sums (termination.c:126-137); also rescales Atdy by Dinv (termination.c:121-123)
__global__ void k_pinf(const double *u, const double *l, const double *E, const double *dy, int scaled, int m,
                       double *Atdy, const double *Dinv, int n, double *part, double *nrm_out)
{
    __shared__ double sh[32];
    double s = 0.0, mx = 0.0;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < m; i += gridDim.x * blockDim.x) {
        double lim = scaled ? E[i] * 1e20 : 1e20;
        double d = dy[i];
        double dp = d > 0.0 ? d : 0.0, dm = d < 0.0 ? d : 0.0;
        s += (u[i] < lim) ? u[i] * dp : 0.0;
        s += (l[i] > -lim) ? l[i] * dm : 0.0;
    }
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        double v = Atdy[i];
        if (scaled) { v = Dinv[i] * v; Atdy[i] = v; }
        v = fabs(v); mx = v > mx ? v : mx;
    }
    s = block_sum(s, sh);
    if (threadIdx.x == 0) part[blockIdx.x] = s;
    mx = warp_max(mx);
    if ((threadIdx.x & 31) == 0 && mx > 0.0) atomic_max_nonneg(nrm_out, mx);
}
// termination.c:179-195: Adx <- Einv Adx, any violating row clears the certificate
__global__ void k_dinf_rows(const double *u, const double *l, const double *E, const double *Einv, double *Adx,
                            const double *nrm_dx, double eps, int scaled, int m, int *viol)
{
    int k = blockIdx.x * blockDim.x + threadIdx.x;
    double en = eps * nrm_dx[0];
    if (en == 0.0 || k >= m) return;
    double a = Adx[k];
    if (scaled) { a = Einv[k] * a; Adx[k] = a; }
    double lim = scaled ? E[k] * 1e20 : 1e20;
    if ((u[k] < lim && a >= en) || (l[k] > -lim && a <= -en)) *viol = 1;
}
__global__ void k_objective_partial(const double *Qx, const double *x, const double *q, const qpdo_dev_scalars *sc,
                                    int proximal, int n, double *part)
{   // iteration.c:185-221
    __shared__ double sh[32];
    double s = 0.0, sigma = sc->sigma;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        double t = proximal ? 0.5 * (Qx[i] - x[i] * sigma) + q[i] : 0.5 * Qx[i] + q[i];
        s += t * x[i];
    }
    s = block_sum(s, sh);
    if (threadIdx.x == 0) part[blockIdx.x] = s;
}
__global__ void k_sum_final(const double *part, int nblk, double *out)
{
    __shared__ double sh[32];
    double s = 0.0;
    for (int i = threadIdx.x; i < nblk; i += blockDim.x) s += part[i];
    s = block_sum(s, sh);
    if (threadIdx.x == 0) out[0] = s;
}
__global__ void k_store_solution(const double *x, double *y, const double *D, const double *E, double cinv, int scaled,
                                 double *sx, double *sy, int n, int m)
{   // termination.c:82-92 (y is rescaled in place, quirk kept)
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) sx[i] = scaled ? x[i] * D[i] : x[i];
    if (i < m) {
        if (scaled) { double v = y[i] * cinv; y[i] = v; sy[i] = v * E[i]; }
        else sy[i] = y[i];
    }
}
__global__ void k_scale_then_mul(double *a, double s, const double *b, int n)
{   // a <- b .* (a * s)
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) { double v = a[i] * s; a[i] = b[i] * v; }
}
__global__ void k_act_to_long(const int *act, long *out, int m)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < m) out[i] = act[i];
}

// ============================================================================ host entry points
static int read_scalars(qpdo_dev *d, qpdo_dev_scalars *out)
{
    CUDA_OK(cudaMemcpyAsync(d->sc_host, d->sc, sizeof(qpdo_dev_scalars), cudaMemcpyDeviceToHost, d->st));
    CUDA_OK(cudaStreamSynchronize(d->st));
    if (out) *out = *d->sc_host;
    return 0;
}
static int objective_dev(qpdo_dev *d, int proximal, double *out_dev)
{
    int nb = std::min(RED_BLOCKS, std::max(1, cdiv(d->n, TPB)));
    k_objective_partial<<<nb, TPB, 0, d->st>>>(d->Qx_, d->x, d->q, d->sc, proximal, d->n, d->red + 16);
    k_sum_final<<<1, 128, 0, d->st>>>(d->red + 16, nb, out_dev);
    return 0;
}

extern "C" int qpdo_dev_objective(qpdo_dev *d, int proximal, double c_const, double *objective)
{
    if (d) cudaSetDevice(d->device);
    objective_dev(d, proximal, d->red + 8);
    double v = 0;
    CUDA_OK(cudaMemcpyAsync(&v, d->red + 8, sizeof(double), cudaMemcpyDeviceToHost, d->st));
    CUDA_OK(cudaStreamSynchronize(d->st));
    if (d->scaling_iters) v = v * d->cinv;
    *objective = v + c_const;
    return 0;
}

extern "C" int qpdo_dev_warm_start(qpdo_dev *d, const double *x, const double *y, double sigma, int proximal,
                                   double mu_min, double *objective)
{
    if (d) cudaSetDevice(d->device);
    const int n = d->n, m = d->m, scaled = d->scaling_iters > 0;
    (void)mu_min;
    k_set_sigma<<<1, 1, 0, d->st>>>(d->sc, sigma, 0);
    if (x) {
        CUDA_OK(cudaMemcpyAsync(d->tmp_n, x, sizeof(double) * n, cudaMemcpyHostToDevice, d->st));
        k_warm_primal<<<cdiv(n, TPB), TPB, 0, d->st>>>(d->tmp_n, d->Dinv, d->x, d->xbar, d->dx, scaled, n);
        spmv_Q(d, d->dx, d->Qdx);
        if (proximal) k_axpy<<<cdiv(n, TPB), TPB, 0, d->st>>>(d->Qdx, d->x, d->Qx_, sigma, n);
        else CUDA_OK(cudaMemcpyAsync(d->Qx_, d->Qdx, sizeof(double) * n, cudaMemcpyDeviceToDevice, d->st));
        if (m) {
            spmv_A(d, d->dx, d->Adx);
            CUDA_OK(cudaMemcpyAsync(d->Ax_, d->Adx, sizeof(double) * m, cudaMemcpyDeviceToDevice, d->st));
        }
    } else {
        CUDA_OK(cudaMemsetAsync(d->x, 0, sizeof(double) * n, d->st));
        CUDA_OK(cudaMemsetAsync(d->xbar, 0, sizeof(double) * n, d->st));
        CUDA_OK(cudaMemsetAsync(d->Qx_, 0, sizeof(double) * n, d->st));
        if (m) CUDA_OK(cudaMemsetAsync(d->Ax_, 0, sizeof(double) * m, d->st));
    }
    if (y && m) {
        CUDA_OK(cudaMemcpyAsync(d->tmp_m, y, sizeof(double) * m, cudaMemcpyHostToDevice, d->st));
        k_warm_dual<<<cdiv(m, TPB), TPB, 0, d->st>>>(d->tmp_m, d->Einv, d->c, d->y, d->ybar, d->dy, scaled, m);
        spmv_At(d, d->dy, d->Atdy);
        CUDA_OK(cudaMemcpyAsync(d->Aty, d->Atdy, sizeof(double) * n, cudaMemcpyDeviceToDevice, d->st));
    } else {
        if (m) { CUDA_OK(cudaMemsetAsync(d->y, 0, sizeof(double) * m, d->st)); CUDA_OK(cudaMemsetAsync(d->ybar, 0, sizeof(double) * m, d->st)); }
        CUDA_OK(cudaMemsetAsync(d->Aty, 0, sizeof(double) * n, d->st));
    }
    // initialize_mu
    DotArgs A; memset(&A, 0, sizeof A);
    A.npair = 2; A.a[0] = d->x; A.b[0] = d->Qx_; A.len[0] = n; A.a[1] = d->q; A.b[1] = d->x; A.len[1] = n;
    dots(d, A, d->red + 4);
    if (m) k_init_mu<<<cdiv(m, TPB), TPB, 0, d->st>>>(d->Ax_, d->l, d->u, d->red + 4, d->mu, d->sqrt_mu, m);
    if (objective) {
        if (x) return qpdo_dev_objective(d, proximal, 0.0, objective);   // caller adds data->c
        *objective = 0.0;
    }
    CUDA_OK(cudaStreamSynchronize(d->st));
    return 0;
}

extern "C" int qpdo_dev_begin_solve(qpdo_dev *d, double sigma)
{
    if (d) cudaSetDevice(d->device);
    CUDA_OK(cudaMemsetAsync(d->sc, 0, sizeof(qpdo_dev_scalars), d->st));
    k_set_sigma<<<1, 1, 0, d->st>>>(d->sc, sigma, 1);
    if (d->m) CUDA_OK(cudaMemsetAsync(d->act_old, 0, sizeof(int) * d->m, d->st));
    if (d->m) CUDA_OK(cudaMemsetAsync(d->omega, 0, sizeof(double) * d->m, d->st));
    return 0;
}

extern "C" int qpdo_dev_residuals(qpdo_dev *d, int proximal, qpdo_dev_scalars *out)
{
    if (d) cudaSetDevice(d->device);
    {
        PhaseTimer t(d, PH_RES);
        ResArgs A;
        A.n = d->n; A.m = d->m; A.scaled = d->scaling_iters > 0; A.proximal = proximal; A.cinv = d->cinv;
        A.Ax = d->Ax_; A.y = d->y; A.ybar = d->ybar; A.l = d->l; A.u = d->u; A.mu = d->mu; A.E = d->E; A.Einv = d->Einv;
        A.Qx = d->Qx_; A.q = d->q; A.x = d->x; A.xbar = d->xbar; A.Aty = d->Aty; A.Dinv = d->Dinv;
        A.res_prim = d->res_prim; A.res_prim_in = d->res_prim_in; A.w = d->w; A.df = d->df;
        A.res_dual = d->res_dual; A.res_dual_in = d->res_dual_in; A.sc = d->sc;
        CUDA_OK(cudaMemsetAsync(d->sc, 0, 4 * sizeof(double), d->st));
        int N = std::max(d->n, d->m);
        k_residuals<<<cdiv((N + 1) / 2, TPB), TPB, 0, d->st>>>(A);
        if (A.scaled) k_scale_dual_norms<<<1, 1, 0, d->st>>>(d->sc, d->cinv);
        d->prof.n_residual++;
    }
    return read_scalars(d, out);
}

extern "C" int qpdo_dev_newton_step(qpdo_dev *d, int force_reset, int proximal)
{
    if (d) cudaSetDevice(d->device);
    const int n = d->n, m = d->m;
    d->proximal = proximal;
    {
        PhaseTimer t(d, PH_OTHER);
        CUDA_OK(cudaMemsetAsync(&d->sc->n_active, 0, 3 * sizeof(int), d->st));
        if (m) k_active<<<cdiv(m, TPB), TPB, 0, d->st>>>(d->w, d->l, d->u, d->act, d->act_old, d->sc, m);
        k_decide<<<1, 1, 0, d->st>>>(d->sc, force_reset);
        if (m) k_apply_omega<<<cdiv(m, TPB), TPB, 0, d->st>>>(d->sc, d->act, d->act_old, d->mu, d->res_prim_in, d->omega, d->dy, m);
    }
    if (factor_numeric(d)) return -1;
    {
        PhaseTimer t(d, PH_SPMV);
        if (m) spmv_At(d, d->dy, d->Atdy); else CUDA_OK(cudaMemsetAsync(d->Atdy, 0, sizeof(double) * n, d->st));
        k_rhs<<<cdiv(n, TPB), TPB, 0, d->st>>>(d->res_dual_in, d->Atdy, d->rhs, n);
    }
    if (factor_solve(d, d->rhs, d->dx)) return -1;
    {
        PhaseTimer t(d, PH_SPMV);
        spmv_Q(d, d->dx, d->Qdx);
        if (proximal) k_axpy_dev<<<cdiv(n, TPB), TPB, 0, d->st>>>(d->Qdx, d->dx, d->Qdx, &d->sc->sigma, 1.0, n);
        if (m) {
            spmv_A(d, d->dx, d->Adx);
            k_dy_active<<<cdiv(m, TPB), TPB, 0, d->st>>>(d->act, d->Adx, d->mu, d->dy, m);
            spmv_At(d, d->dy, d->Atdy);
        }
    }
    if (linesearch_run(d)) return -1;
    {
        PhaseTimer t(d, PH_OTHER);
        UpdArgs U;
        U.n = n; U.m = m; U.x = d->x; U.y = d->y; U.Qx = d->Qx_; U.Ax = d->Ax_; U.Aty = d->Aty;
        U.dx = d->dx; U.dy = d->dy; U.Qdx = d->Qdx; U.Adx = d->Adx; U.Atdy = d->Atdy; U.sc = d->sc;
        k_update_iterate<<<cdiv((std::max(n, m) + 1) / 2, TPB), TPB, 0, d->st>>>(U);
    }
    d->prof.n_newton++;
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { qpdo_set_error("newton step launch failed: %s", cudaGetErrorString(e)); return -1; }
    return 0;
}

extern "C" int qpdo_dev_outer_update(qpdo_dev *d, int do_mu, int do_sigma, int proximal, double eps_abs, double theta,
                                     double delta, double mu_min, double sigma_upd, double sigma_min, double *sigma)
{
    if (d) cudaSetDevice(d->device);
    const int n = d->n, m = d->m;
    PhaseTimer t(d, PH_OTHER);
    k_copy2<<<cdiv(std::max(n, m), TPB), TPB, 0, d->st>>>(d->x, d->xbar, n, d->y, d->ybar, m);       // qpdo.c:396-397
    if (do_mu && m) {
        norm_inf(d, d->res_prim, nullptr, d->red, m);                                              // iteration.c:130
        CUDA_OK(cudaMemsetAsync(&d->sc->n_mu_changed, 0, sizeof(int), d->st));
        k_update_mu<<<cdiv(m, TPB), TPB, 0, d->st>>>(d->res_prim, d->res_prim_old, d->red, eps_abs, theta, delta, mu_min,
                                                     1.0 / sqrt(mu_min), d->mu, d->sqrt_mu, d->tmp_m, d->sc, m);
        k_mu_apply<<<cdiv(m, TPB), TPB, 0, d->st>>>(d->sc, proximal, sigma_min, d->tmp_m, d->omega, m);
    }
    if (do_sigma && *sigma > sigma_min) {                                                          // iteration.c:173-180
        double s_old = *sigma;
        double s_new = s_old * sigma_upd;
        s_new = s_new > sigma_min ? s_new : sigma_min;
        *sigma = s_new;
        k_set_sigma<<<1, 1, 0, d->st>>>(d->sc, s_new, 1);
        k_axpy<<<cdiv(n, TPB), TPB, 0, d->st>>>(d->Qx_, d->x, d->Qx_, s_new - s_old, n);
    }
    if (m) CUDA_OK(cudaMemcpyAsync(d->res_prim_old, d->res_prim, sizeof(double) * m, cudaMemcpyDeviceToDevice, d->st)); // qpdo.c:426
    return 0;
}

extern "C" int qpdo_dev_primal_infeasible(qpdo_dev *d, double eps, int *flag)
{
    if (d) cudaSetDevice(d->device);
    const int n = d->n, m = d->m, scaled = d->scaling_iters > 0;
    *flag = 0;
    if (!m) return 0;
    PhaseTimer t(d, PH_OTHER);
    k_axpy<<<cdiv(m, TPB), TPB, 0, d->st>>>(d->y, d->ybar, d->dy, -1.0, m);                         // qpdo.c:372
    spmv_At(d, d->dy, d->Atdy);                                                                    // qpdo.c:374
    norm_inf(d, d->dy, scaled ? d->E : nullptr, d->red, m);                                        // termination.c:112-117
    double nrm = 0;
    CUDA_OK(cudaMemcpyAsync(&nrm, d->red, sizeof(double), cudaMemcpyDeviceToHost, d->st));
    CUDA_OK(cudaStreamSynchronize(d->st));
    double en = eps * nrm;
    if (en == 0) return 0;
    int nb = std::min(RED_BLOCKS, std::max(1, cdiv(std::max(n, m), TPB)));
    CUDA_OK(cudaMemsetAsync(d->red + 1, 0, sizeof(double), d->st));
    k_pinf<<<nb, TPB, 0, d->st>>>(d->u, d->l, d->E, d->dy, scaled, m, d->Atdy, d->Dinv, n, d->red + 16, d->red + 1);
    k_sum_final<<<1, 128, 0, d->st>>>(d->red + 16, nb, d->red + 2);
    double h[3];
    CUDA_OK(cudaMemcpyAsync(h, d->red, 3 * sizeof(double), cudaMemcpyDeviceToHost, d->st));
    CUDA_OK(cudaStreamSynchronize(d->st));
    if (h[1] <= en && h[2] <= -en) {                                                               // termination.c:139
        *flag = 1;
        if (scaled) {
            k_scale_then_mul<<<cdiv(m, TPB), TPB, 0, d->st>>>(d->dy, d->cinv, d->E, m);            // termination.c:142-145
        }
    }
    return 0;
}

extern "C" int qpdo_dev_dual_infeasible(qpdo_dev *d, double eps, int proximal, int *flag)
{
    if (d) cudaSetDevice(d->device);
    const int n = d->n, m = d->m, scaled = d->scaling_iters > 0;
    *flag = 0;
    PhaseTimer t(d, PH_OTHER);
    k_axpy<<<cdiv(n, TPB), TPB, 0, d->st>>>(d->x, d->xbar, d->dx, -1.0, n);                         // qpdo.c:383
    spmv_Q(d, d->dx, d->Qdx);                                                                      // qpdo.c:385
    if (m) spmv_A(d, d->dx, d->Adx);                                                               // qpdo.c:387
    norm_inf(d, d->dx, scaled ? d->D : nullptr, d->red, n);                                        // termination.c:172-177
    int *viol = (int *)(d->red + 3);
    CUDA_OK(cudaMemsetAsync(viol, 0, sizeof(double), d->st));
    if (m) k_dinf_rows<<<cdiv(m, TPB), TPB, 0, d->st>>>(d->u, d->l, d->E, d->Einv, d->Adx, d->red, eps, scaled, m, viol);
    double h[4];
    CUDA_OK(cudaMemcpyAsync(h, d->red, 4 * sizeof(double), cudaMemcpyDeviceToHost, d->st));
    CUDA_OK(cudaStreamSynchronize(d->st));
    double en = eps * h[0];
    if (en == 0) return 0;
    int v; memcpy(&v, &h[3], sizeof(int));
    if (v) return 0;
    if (proximal)                                                                                   // termination.c:197-199
        k_axpy_dev<<<cdiv(n, TPB), TPB, 0, d->st>>>(d->Qdx, d->dx, d->Qdx, &d->sc->sigma, -d->sc_host->tau, n);
    norm_inf(d, d->Qdx, nullptr, d->red + 1, n);
    DotArgs A; memset(&A, 0, sizeof A);
    A.npair = 1; A.a[0] = d->q; A.b[0] = d->dx; A.len[0] = n;
    dots(d, A, d->red + 2);
    CUDA_OK(cudaMemcpyAsync(h, d->red, 3 * sizeof(double), cudaMemcpyDeviceToHost, d->st));
    CUDA_OK(cudaStreamSynchronize(d->st));
    double lim = scaled ? d->c * en : en;                                                          // termination.c:200-212
    if (h[1] <= lim && h[2] <= -lim) {
        *flag = 1;
        if (scaled) k_mul<<<cdiv(n, TPB), TPB, 0, d->st>>>(d->D, d->dx, d->dx, n);
    }
    return 0;
}

extern "C" int qpdo_dev_store_solution(qpdo_dev *d, double *sol_x, double *sol_y, double *x, double *y, double *dx,
                                       double *dy, long *active_set, qpdo_dev_scalars *out)
{
    if (d) cudaSetDevice(d->device);
    const int n = d->n, m = d->m;
    const size_t nb = sizeof(double) * n, mb = sizeof(double) * m;
    k_store_solution<<<cdiv(std::max(n, m), TPB), TPB, 0, d->st>>>(d->x, d->y, d->D, d->E, d->cinv, d->scaling_iters > 0,
                                                                    d->tmp_n, d->tmp_m, n, m);
    if (sol_x) CUDA_OK(cudaMemcpyAsync(sol_x, d->tmp_n, nb, cudaMemcpyDeviceToHost, d->st));
    if (sol_y && m) CUDA_OK(cudaMemcpyAsync(sol_y, d->tmp_m, mb, cudaMemcpyDeviceToHost, d->st));
    if (x) CUDA_OK(cudaMemcpyAsync(x, d->x, nb, cudaMemcpyDeviceToHost, d->st));
    if (y && m) CUDA_OK(cudaMemcpyAsync(y, d->y, mb, cudaMemcpyDeviceToHost, d->st));
    if (dx) CUDA_OK(cudaMemcpyAsync(dx, d->dx, nb, cudaMemcpyDeviceToHost, d->st));
    if (dy && m) CUDA_OK(cudaMemcpyAsync(dy, d->dy, mb, cudaMemcpyDeviceToHost, d->st));
    if (active_set && m) {
        long *tmp = (long *)d->ls_key[0];                     // 2m x 8 bytes of scratch, free between linesearches
        k_act_to_long<<<cdiv(m, TPB), TPB, 0, d->st>>>(d->act, tmp, m);
        CUDA_OK(cudaMemcpyAsync(active_set, tmp, sizeof(long) * m, cudaMemcpyDeviceToHost, d->st));
    }
    if (read_scalars(d, out)) return -1;
    qpdo_profile_collect(d);
    return 0;
}

extern "C" int qpdo_dev_update_bounds(qpdo_dev *d, const double *l, const double *u)
{   // qpdo.c:538-543
    if (d) cudaSetDevice(d->device);
    const int m = d->m, scaled = d->scaling_iters > 0;
    if (!m) return 0;
    if (l) {
        CUDA_OK(cudaMemcpyAsync(d->l, l, sizeof(double) * m, cudaMemcpyHostToDevice, d->st));
        if (scaled) k_mul<<<cdiv(m, TPB), TPB, 0, d->st>>>(d->E, d->l, d->l, m);
    }
    if (u) {
        CUDA_OK(cudaMemcpyAsync(d->u, u, sizeof(double) * m, cudaMemcpyHostToDevice, d->st));
        if (scaled) k_mul<<<cdiv(m, TPB), TPB, 0, d->st>>>(d->E, d->u, d->u, m);
    }
    CUDA_OK(cudaStreamSynchronize(d->st));
    return 0;
}

__global__ void k_scale_s(double *a, double s, int n)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) a[i] = a[i] * s;
}
__global__ void k_gather_vals(double *dst, const double *src, const int *idx, int n)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) dst[i] = src[idx[i]];
}

extern "C" int qpdo_dev_update_q(qpdo_dev *d, const double *q, int proximal, double sigma_init, double *sigma,
                                 double *c, double *cinv, double *norm_q)
{   // qpdo.c:549-586
    if (d) cudaSetDevice(d->device);
    const int n = d->n, scaled = d->scaling_iters > 0;
    CUDA_OK(cudaMemcpyAsync(d->q, q, sizeof(double) * n, cudaMemcpyHostToDevice, d->st));
    if (scaled) {
        k_mul<<<cdiv(n, TPB), TPB, 0, d->st>>>(d->D, d->q, d->q, n);
        double c_old = d->c;
        if (proximal) k_axpy<<<cdiv(n, TPB), TPB, 0, d->st>>>(d->Qx_, d->x, d->Qx_, -(*sigma), n);
        k_axpy<<<cdiv(n, TPB), TPB, 0, d->st>>>(d->q, d->Qx_, d->tmp_n, d->cinv, n);
        norm_inf(d, d->tmp_n, nullptr, d->red, n);
        double nrm = 0;
        CUDA_OK(cudaMemcpyAsync(&nrm, d->red, sizeof(double), cudaMemcpyDeviceToHost, d->st));
        CUDA_OK(cudaStreamSynchronize(d->st));
        d->c = 1.0 / (nrm > 1.0 ? nrm : 1.0);
        d->cinv = 1.0 / d->c;
        k_scale_s<<<cdiv(n, TPB), TPB, 0, d->st>>>(d->q, d->c, n);
        double ratio = d->c / c_old;
        if (d->nnzQ) {
            k_scale_s<<<cdiv(d->nnzQ, TPB), TPB, 0, d->st>>>(d->Qx, ratio, d->nnzQ);
            k_gather_vals<<<cdiv(d->nnzQf, TPB), TPB, 0, d->st>>>(d->Fx, d->Qx, d->Fsrc, d->nnzQf);
        }
        k_scale_s<<<cdiv(n, TPB), TPB, 0, d->st>>>(d->Qx_, ratio, n);
        if (proximal) {
            *sigma = sigma_init;
            k_axpy<<<cdiv(n, TPB), TPB, 0, d->st>>>(d->Qx_, d->x, d->Qx_, *sigma, n);
        }
        norm_inf(d, d->q, d->Dinv, d->red, n);
    } else {
        norm_inf(d, d->q, nullptr, d->red, n);
    }
    CUDA_OK(cudaMemcpyAsync(&d->norm_q, d->red, sizeof(double), cudaMemcpyDeviceToHost, d->st));
    CUDA_OK(cudaStreamSynchronize(d->st));
    *c = d->c; *cinv = d->cinv; *norm_q = d->norm_q;
    return 0;
}

extern "C" long qpdo_dev_get_vector(qpdo_dev *d, const char *name, double *out, long cap)
{
    if (d) cudaSetDevice(d->device);
    struct { const char *nm; const double *p; int len; } tab[] = {
        {"x", d->x, d->n}, {"y", d->y, d->m}, {"xbar", d->xbar, d->n}, {"ybar", d->ybar, d->m}, {"Ax", d->Ax_, d->m},
        {"Qx", d->Qx_, d->n}, {"Aty", d->Aty, d->n}, {"mu", d->mu, d->m}, {"sqrt_mu", d->sqrt_mu, d->m}, {"w", d->w, d->m},
        {"dx", d->dx, d->n}, {"dy", d->dy, d->m}, {"Qdx", d->Qdx, d->n}, {"Adx", d->Adx, d->m}, {"Atdy", d->Atdy, d->n},
        {"omega", d->omega, d->m}, {"res_prim", d->res_prim, d->m}, {"res_dual", d->res_dual, d->n},
        {"res_prim_in", d->res_prim_in, d->m}, {"res_dual_in", d->res_dual_in, d->n}, {"rhs", d->rhs, d->n},
        {"l", d->l, d->m}, {"u", d->u, d->m}, {"q", d->q, d->n}, {"df", d->df, d->n}, {"D", d->D, d->n}, {"E", d->E, d->m},
        {"A_values", d->Ax, d->nnzA}, {"Q_values", d->Qx, d->nnzQ}, {"ls_delta", d->ls_delta, 2 * d->m},
        {"ls_alpha", d->ls_alpha, 2 * d->m}};
    for (auto &t : tab) if (!strcmp(t.nm, name)) {
        long len = t.len < cap ? t.len : cap;
        if (len > 0) {
            if (cudaMemcpyAsync(out, t.p, sizeof(double) * len, cudaMemcpyDeviceToHost, d->st) != cudaSuccess) return -1;
            if (cudaStreamSynchronize(d->st) != cudaSuccess) return -1;
        }
        return t.len;
    }
    qpdo_set_error("unknown vector '%s'", name);
    return -1;
}

extern "C" int qpdo_dev_test_spmv(qpdo_dev *d, int which, const double *x, double *y)
{
    if (d) cudaSetDevice(d->device);
    const int nin = which == 1 ? d->m : d->n, nout = which == 0 ? d->m : d->n;
    double *din = which == 1 ? d->tmp_m : d->tmp_n, *dout = which == 0 ? d->Adx : d->Qdx;
    CUDA_OK(cudaMemcpyAsync(din, x, sizeof(double) * nin, cudaMemcpyHostToDevice, d->st));
    if (which == 0) spmv_A(d, din, dout); else if (which == 1) spmv_At(d, din, dout); else spmv_Q(d, din, dout);
    CUDA_OK(cudaMemcpyAsync(y, dout, sizeof(double) * nout, cudaMemcpyDeviceToHost, d->st));
    CUDA_OK(cudaStreamSynchronize(d->st));
    return 0;
}

extern "C" int qpdo_dev_test_linesearch(qpdo_dev *d, const double *dx, const double *Qdx, const double *df, const double *dy,
                                        const double *y, const double *mu, const double *sqrt_mu, const double *Adx,
                                        const double *w, const double *l, const double *u, double *tau)
{
    if (d) cudaSetDevice(d->device);
    struct { double *dst; const double *src; int len; } up[] = {
        {d->dx, dx, d->n}, {d->Qdx, Qdx, d->n}, {d->df, df, d->n}, {d->dy, dy, d->m}, {d->y, y, d->m}, {d->mu, mu, d->m},
        {d->sqrt_mu, sqrt_mu, d->m}, {d->Adx, Adx, d->m}, {d->w, w, d->m}, {d->l, l, d->m}, {d->u, u, d->m}};
    for (auto &v : up)
        if (v.len > 0) CUDA_OK(cudaMemcpyAsync(v.dst, v.src, sizeof(double) * v.len, cudaMemcpyHostToDevice, d->st));
    if (linesearch_run(d)) return -1;
    qpdo_dev_scalars h;
    CUDA_OK(cudaMemcpyAsync(&h, d->sc, sizeof(h), cudaMemcpyDeviceToHost, d->st));
    CUDA_OK(cudaStreamSynchronize(d->st));
    *tau = h.tau;
    return 0;
}

extern "C" int qpdo_dev_test_factor_solve(qpdo_dev *d, const double *omega, double sigma, const double *b, double *x,
                                          float *factor_ms, float *solve_ms)
{
    if (d) cudaSetDevice(d->device);
    cudaEvent_t e0, e1, e2;
    cudaEventCreate(&e0); cudaEventCreate(&e1); cudaEventCreate(&e2);
    if (d->m) CUDA_OK(cudaMemcpyAsync(d->omega, omega, sizeof(double) * d->m, cudaMemcpyHostToDevice, d->st));
    CUDA_OK(cudaMemcpyAsync(d->rhs, b, sizeof(double) * d->n, cudaMemcpyHostToDevice, d->st));
    k_set_sigma<<<1, 1, 0, d->st>>>(d->sc, sigma, 0);
    d->proximal = 1;
    cudaEventRecord(e0, d->st);
    if (factor_numeric(d)) return -1;
    cudaEventRecord(e1, d->st);
    if (factor_solve(d, d->rhs, d->dx)) return -1;
    cudaEventRecord(e2, d->st);
    CUDA_OK(cudaMemcpyAsync(x, d->dx, sizeof(double) * d->n, cudaMemcpyDeviceToHost, d->st));
    CUDA_OK(cudaStreamSynchronize(d->st));
    if (factor_ms) cudaEventElapsedTime(factor_ms, e0, e1);
    if (solve_ms) cudaEventElapsedTime(solve_ms, e1, e2);
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaEventDestroy(e2);
    return 0;
}
