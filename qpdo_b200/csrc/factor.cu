// factor.cu -- numeric multifrontal LDL' factorization of K = Q + sigma I + A' diag(omega) A on the GPU, and
// the supernodal triangular solves.  Replaces, for the Newton hot path, cholmod_aat + cholmod_add
// (K assembly, cholmod_interface.c:45-48), cholmod_factorize_p / super_numeric (numeric
// factorization, cholmod_interface.c:8-30; t_cholmod_super_numeric.c: dsyrk 682, dgemm 708,
// dpotrf 864, dtrsm 997), cholmod_updown (cholmod_interface.c:57-93; here: numeric refactorization
// with changed row weights on the fixed superset pattern) and cholmod_solve (cholmod_interface.c:98-102).
//
// Layout in HBM: one L panel per supernode (N x nc, column major, ld = N = nc + nr) in Lx, one
// update matrix per supernode (nr x nr, lower, ld = nr) in Ux; all fronts resident at once.
// Schedule: levels of the supernodal assembly tree, one launch per (level, size bucket); fronts
// that fit in shared memory are factored by one CTA each, larger ones by a blocked right-looking
// panel loop (LDL' 32x32 -> trsm -> rank-32 SYRK on the FP64 tensor pipe, mma.sync m8n8k4 DMMA).
// The whole sequence is captured once into a CUDA graph and replayed every Newton step.
#include "dev_internal.h"
#include <algorithm>
#include <chrono>
#include <stdlib.h>
#include <string.h>

#define NB 32          // panel width of the blocked path
#define CW 32          // destination-column chunk of the big-front assembly
#define TS 64          // SYRK tile
#define PLD 36         // padded leading dimension of a 64 x 32 panel tile in smem (conflict free)

struct FArgs {
    const int *sn_start, *sn_nr, *rel, *rows, *child_ptr, *child_idx;
    const int64_t *sn_rowptr, *Lp, *Up;
    double *Lx, *Ux;
    double *Xout;              // final L panels of the big (blocked) fronts; Lx keeps their working values
    qpdo_dev_scalars *sc;
};

// ------------------------------------------------------------------------------- K assembly
// Lx[Kdst[e]] = Q_e + sum_t omega[row_t] * (a_ir a_ic)_t (+ sigma on the diagonal): the fused
// replacement of cholmod_aat + cholmod_add + the beta*I of cholmod_factorize_p; Lx was zeroed.
__global__ void k_assemble_K(int64_t nnzK, const int64_t *Kdst, const int *Kq, const int64_t *Ktp, const int *Kt_row,
                             const double *Kt_val, const int *Ki, const int *Kcol, const double *Qx, const double *omega,
                             const qpdo_dev_scalars *sc, int proximal, double *Lx)
{
    int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= nnzK) return;
    double v = 0.0;
    for (int64_t t = Ktp[e]; t < Ktp[e + 1]; t++) v = fma(omega[Kt_row[t]], Kt_val[t], v);
    int q = Kq[e];
    if (q >= 0) v += Qx[q];
    if (proximal && Ki[e] == Kcol[e]) v += sc->sigma;       // beta*I only with factorize_p (cholmod_interface.c:11-16)
    Lx[Kdst[e]] = v;
}
__global__ void k_products(int64_t nt, const int *a1, const int *a2, const double *Ax, double *val)
{
    int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t < nt) val[t] = Ax[a1[t]] * Ax[a2[t]];
}

// u / d for a pivot d whose correctly rounded reciprocal r = __drcp_rn(d) is at hand: q0 = u r, one residual
// correction (Markstein) gives the correctly rounded quotient, i.e. the same bits as u / d, without the
// ~60-instruction slow path that DDIV takes for every zero numerator (structural zeros of a front).
__device__ __forceinline__ double div_by_pivot(double u, double d, double r)
{
    const double q0 = __dmul_rn(u, r);
    return fma(fma(-q0, d, u), r, q0);
}

// ------------------------------------------------------------------------------- smem fronts
// One CTA per front: load the panel, extend-add the children's update matrices in fixed order,
// eliminate the nc pivot columns (right-looking), write the panel and the Schur complement.
template <int NT>
__global__ void __launch_bounds__(NT) k_front_smem(FArgs A, const int *list)
{
    extern __shared__ double F[];
    const int s = list[blockIdx.x];
    const int f = A.sn_start[s], nc = A.sn_start[s + 1] - f, nr = A.sn_nr[s], N = nc + nr, ld = N | 1;
    double *Lp = A.Lx + A.Lp[s];
    const int tid = threadIdx.x;
    for (int idx = tid; idx < N * nc; idx += NT) { int j = idx / N, i = idx - j * N; F[i + j * ld] = Lp[idx]; }
    for (int idx = tid; idx < N * nr; idx += NT) { int j = idx / N, i = idx - j * N; F[i + (nc + j) * ld] = 0.0; }
    __syncthreads();
    for (int t = A.child_ptr[s]; t < A.child_ptr[s + 1]; t++) {
        const int c = A.child_idx[t], rc = A.sn_nr[c];
        const int *rl = A.rel + A.sn_rowptr[c];
        const double *U = A.Ux + A.Up[c];
        for (int idx = tid; idx < rc * rc; idx += NT) {
            int j = idx / rc, i = idx - j * rc;
            if (i >= j) F[rl[i] + rl[j] * ld] += U[idx];
        }
        __syncthreads();
    }
    // right-looking LDL' in blocks of KB pivots.  Panel phase, per pivot: the rows below it compute
    // l_i = a_ik / d_k (unscaled u_i and scaled l_i go to the ub / lb buffers, the L entry is final), then the
    // remaining panel columns are updated.  Trailing phase, once per block: warp per column, lane per row
    // (row values u_i^t stay in registers across the columns), a_ij = fma(-u^3, l^3, fma(... fma(-u^0, l^0, a_ij)))
    // -- the same ascending-k fma sequence per entry as a column-by-column elimination, with a quarter of
    // the shared-memory traffic and barriers.
    constexpr int KB = 4, RMAX = 5;                        // fronts up to 32 * RMAX = 160 rows
    const int lane = tid & 31, wid = tid >> 5, NW = NT / 32;
    double *ub = F + (size_t)ld * N, *lb = ub + KB * N;
    for (int kb = 0; kb < nc; kb += KB) {
        const int nk = min(KB, nc - kb);
        for (int t = 0; t < nk; t++) {
            const int k = kb + t;
            const double dk = F[k + k * ld];
            if (tid == 0 && !(dk > 0.0)) A.sc->not_posdef = 1;
            const double rk = __drcp_rn(dk);
            for (int i = k + 1 + tid; i < N; i += NT) {
                const double u = F[i + k * ld], l = div_by_pivot(u, dk, rk);
                ub[t * N + i] = u; lb[t * N + i] = l; F[i + k * ld] = l;
            }
            __syncthreads();
            if (t + 1 < nk) {
                for (int c = k + 1; c < kb + nk; c++) {
                    const double lc = lb[t * N + c];
                    for (int i = c + tid; i < N; i += NT) F[i + c * ld] = fma(-ub[t * N + i], lc, F[i + c * ld]);
                }
                __syncthreads();
            }
        }
        double ur[RMAX][KB];
#pragma unroll
        for (int r = 0; r < RMAX; r++)
#pragma unroll
            for (int t = 0; t < KB; t++) { const int i = lane + 32 * r; ur[r][t] = (t < nk && i < N) ? ub[t * N + i] : 0.0; }
        for (int j = kb + nk + wid; j < N; j += NW) {
            double lj[KB];
#pragma unroll
            for (int t = 0; t < KB; t++) lj[t] = (t < nk) ? lb[t * N + j] : 0.0;
#pragma unroll
            for (int r = 0; r < RMAX; r++) {
                const int i = lane + 32 * r;
                if (i >= j && i < N) {
                    double v = F[i + j * ld];
#pragma unroll
                    for (int t = 0; t < KB; t++) if (t < nk) v = fma(-ur[r][t], lj[t], v);
                    F[i + j * ld] = v;
                }
            }
        }
        __syncthreads();
    }
    __syncthreads();
    for (int idx = tid; idx < N * nc; idx += NT) { int j = idx / N, i = idx - j * N; Lp[idx] = (i >= j) ? F[i + j * ld] : 0.0; }
    double *U = A.Ux + A.Up[s];
    for (int idx = tid; idx < nr * nr; idx += NT) {
        int j = idx / nr, i = idx - j * nr;
        if (i >= j) U[idx] = F[(nc + i) + (nc + j) * ld];
    }
}

// ------------------------------------------------------------------------------- big fronts
__device__ __forceinline__ double *front_elem(const FArgs &A, int s, int nc, int nr, int N, int i, int j)
{   // element (i, j), i >= j, of the front: L panel for j < nc, update matrix otherwise
    return (j < nc) ? (A.Lx + A.Lp[s] + (int64_t)j * N + i) : (A.Ux + A.Up[s] + (int64_t)(j - nc) * nr + (i - nc));
}
// CTA (front, chunk) owns destination columns [c0, c1): zero their update-matrix part, then add the
// children's contributions that fall into these columns, children in fixed order.
__global__ void __launch_bounds__(256) k_big_assemble(FArgs A, const int *list)
{
    const int s = list[blockIdx.x];
    const int f = A.sn_start[s], nc = A.sn_start[s + 1] - f, nr = A.sn_nr[s], N = nc + nr;
    const int c0 = blockIdx.y * CW;
    if (c0 >= N) return;
    const int c1 = min(N, c0 + CW), tid = threadIdx.x;
    for (int j = max(c0, nc); j < c1; j++) {
        double *col = A.Ux + A.Up[s] + (int64_t)(j - nc) * nr;
        for (int i = j - nc + tid; i < nr; i += 256) col[i] = 0.0;
    }
    __syncthreads();
    for (int t = A.child_ptr[s]; t < A.child_ptr[s + 1]; t++) {
        const int c = A.child_idx[t], rc = A.sn_nr[c];
        const int *rl = A.rel + A.sn_rowptr[c];
        const double *U = A.Ux + A.Up[c];
        int lo = 0, hi = rc;                       // first jj with rl[jj] >= c0
        while (lo < hi) { int mid = (lo + hi) >> 1; if (rl[mid] < c0) lo = mid + 1; else hi = mid; }
        // one warp per source column, 8 independent read-modify-writes in flight per lane (the extend-add
        // is a chain of dependent global accesses: index -> address -> load -> store)
        for (int jj = lo + (tid >> 5); jj < rc; jj += 8) {
            const int dc = rl[jj];
            if (dc >= c1) break;
            const double *ucol = U + (int64_t)jj * rc;
            for (int ii0 = jj + (tid & 31); ii0 < rc; ii0 += 32 * 8) {
                int r[8]; double v[8], cur[8]; double *dst[8];
#pragma unroll
                for (int t = 0; t < 8; t++) { const int ii = ii0 + 32 * t; const bool ok = ii < rc; r[t] = ok ? rl[ii] : -1; v[t] = ok ? ucol[ii] : 0.0; }
#pragma unroll
                for (int t = 0; t < 8; t++) { dst[t] = r[t] >= 0 ? front_elem(A, s, nc, nr, N, r[t], dc) : nullptr; cur[t] = dst[t] ? *dst[t] : 0.0; }
#pragma unroll
                for (int t = 0; t < 8; t++) if (dst[t]) *dst[t] = cur[t] + v[t];
            }
        }
        __syncthreads();
    }
}
// LDL' of the nb x nb diagonal block at (kb, kb): unit lower L11, D on the diagonal.  ONE warp per front,
// lane = block row with its 32 entries in registers; the pivot and the scaled column travel by shuffle, so
// the 32-pivot chain has no barrier and no shared-memory round trip (the chain, not the flops, is what a
// lone CTA at the top of the tree pays for).  Per entry the fma sequence is ascending k, as everywhere.
__global__ void __launch_bounds__(32) k_big_potrf(FArgs A, const int *list, int kb)
{
    const int s = list[blockIdx.x];
    const int f = A.sn_start[s], nc = A.sn_start[s + 1] - f, nr = A.sn_nr[s], N = nc + nr;
    if (kb >= nc) return;
    const int nb = min(NB, nc - kb), lane = threadIdx.x;
    double *Lb = A.Lx + A.Lp[s] + (int64_t)kb * N + kb;
    double a[NB];
#pragma unroll
    for (int j = 0; j < NB; j++) a[j] = (j < nb && lane < nb && lane >= j) ? Lb[(int64_t)j * N + lane] : (lane == j ? 1.0 : 0.0);
    bool bad = false;
#pragma unroll
    for (int k = 0; k < NB; k++) {
        const double dk = __shfl_sync(0xffffffffu, a[k], k);
        bad = bad || (k < nb && !(dk > 0.0));
        const double ak = a[k];
        const double l = div_by_pivot(ak, dk, __drcp_rn(dk));
#pragma unroll
        for (int j = k + 1; j < NB; j++) {
            const double lj = __shfl_sync(0xffffffffu, l, j);
            a[j] = fma(-ak, lj, a[j]);                   // rows above the diagonal carry garbage that is never stored
        }
        if (lane > k) a[k] = l;
    }
    if (bad && lane == 0) A.sc->not_posdef = 1;
#pragma unroll
    for (int j = 0; j < NB; j++)
        if (j < nb && lane < nb) Lb[(int64_t)j * N + lane] = (lane >= j) ? a[j] : 0.0;
}
// rows below the diagonal block: X <- X L11^{-T} D^{-1}, one row per thread
__global__ void __launch_bounds__(128) k_big_trsm(FArgs A, const int *list, int kb)
{
    __shared__ double D[NB][NB + 1];
    const int s = list[blockIdx.x];
    const int f = A.sn_start[s], nc = A.sn_start[s + 1] - f, nr = A.sn_nr[s], N = nc + nr;
    if (kb >= nc) return;
    const int nb = min(NB, nc - kb), tid = threadIdx.x;
    const int r0 = kb + nb + blockIdx.y * 128;
    if (r0 >= N) return;
    double *Lb = A.Lx + A.Lp[s] + (int64_t)kb * N;
    for (int idx = tid; idx < NB * NB; idx += 128) {
        int j = idx / NB, i = idx - j * NB;
        D[i][j] = (i < nb && j < nb) ? Lb[(int64_t)j * N + kb + i] : (i == j ? 1.0 : 0.0);
    }
    __shared__ double Rinv[NB];
    __syncthreads();
    if (tid < NB) Rinv[tid] = __drcp_rn(D[tid][tid]);
    __syncthreads();
    const int i = r0 + tid;
    if (i >= N) return;
    double x[NB];
#pragma unroll
    for (int j = 0; j < NB; j++) x[j] = (j < nb) ? Lb[(int64_t)j * N + i] : 0.0;
    // W = B L11^{-T} (unit lower L11), then X = W D^{-1}
#pragma unroll
    for (int j = 0; j < NB; j++) {
        double v = x[j];
#pragma unroll
        for (int k = 0; k < j; k++) v = fma(-x[k], D[j][k], v);
        x[j] = v;
    }
#pragma unroll
    for (int j = 0; j < NB; j++) x[j] = div_by_pivot(x[j], D[j][j], Rinv[j]);
#pragma unroll
    for (int j = 0; j < NB; j++) if (j < nb) Lb[(int64_t)j * N + i] = x[j];
}
// trailing update C -= (X D) X' for the lower 64 x 64 tiles, X = panel rows x nb; FP64 tensor pipe
__device__ __forceinline__ void dmma884(double &c0, double &c1, double a, double b)
{
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                 : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
__global__ void __launch_bounds__(256) k_big_syrk(FArgs A, const int *list, int kb)
{
    __shared__ double Pa[TS * PLD], Pb[TS * PLD];
    const int s = list[blockIdx.x];
    const int f = A.sn_start[s], nc = A.sn_start[s + 1] - f, nr = A.sn_nr[s], N = nc + nr;
    if (kb >= nc) return;
    const int nb = min(NB, nc - kb);
    const int r0 = kb + nb, R = N - r0;
    if (R <= 0) return;
    const int nt = (R + TS - 1) / TS;
    int ti = (int)((sqrt(8.0 * blockIdx.y + 1.0) - 1.0) * 0.5);
    while ((ti + 1) * (ti + 2) / 2 <= (int)blockIdx.y) ti++;
    while (ti * (ti + 1) / 2 > (int)blockIdx.y) ti--;
    const int tj = blockIdx.y - ti * (ti + 1) / 2;
    if (ti >= nt) return;
    const int tid = threadIdx.x;
    const double *Lb = A.Lx + A.Lp[s] + (int64_t)kb * N;
    const int ra = r0 + ti * TS, rb = r0 + tj * TS;
    for (int idx = tid; idx < TS * NB; idx += 256) {
        int k = idx / TS, r = idx - k * TS;                 // consecutive threads -> consecutive rows
        Pa[r * PLD + k] = (k < nb && ra + r < N) ? Lb[(int64_t)k * N + ra + r] * Lb[(int64_t)k * N + kb + k] : 0.0;
        Pb[r * PLD + k] = (k < nb && rb + r < N) ? Lb[(int64_t)k * N + rb + r] : 0.0;
    }
    __syncthreads();
    const int lane = tid & 31, wid = tid >> 5;
    const int wr = (wid & 1) * 32, wc = (wid >> 1) * 16;    // warp tile: 32 rows x 16 cols
    const int g = lane >> 2, q = lane & 3;
    double acc[4][2][2];
#pragma unroll
    for (int a = 0; a < 4; a++)
#pragma unroll
        for (int b = 0; b < 2; b++) acc[a][b][0] = acc[a][b][1] = 0.0;
#pragma unroll
    for (int ks = 0; ks < NB / 4; ks++) {
        double af[4], bf[2];
#pragma unroll
        for (int a = 0; a < 4; a++) af[a] = Pa[(wr + a * 8 + g) * PLD + ks * 4 + q];
#pragma unroll
        for (int b = 0; b < 2; b++) bf[b] = Pb[(wc + b * 8 + g) * PLD + ks * 4 + q];
#pragma unroll
        for (int a = 0; a < 4; a++)
#pragma unroll
            for (int b = 0; b < 2; b++) dmma884(acc[a][b][0], acc[a][b][1], af[a], bf[b]);
    }
    // read-modify-write of the 16 destination entries: all loads first (one round trip), then the stores
    double *dst[4][2][2];
    double cur[4][2][2];
#pragma unroll
    for (int a = 0; a < 4; a++)
#pragma unroll
        for (int b = 0; b < 2; b++)
#pragma unroll
            for (int e = 0; e < 2; e++) {
                const int gi = ra + wr + a * 8 + g, gj = rb + wc + b * 8 + q * 2 + e;
                dst[a][b][e] = (gi < N && gj <= gi) ? front_elem(A, s, nc, nr, N, gi, gj) : nullptr;
                cur[a][b][e] = dst[a][b][e] ? *dst[a][b][e] : 0.0;
            }
#pragma unroll
    for (int a = 0; a < 4; a++)
#pragma unroll
        for (int b = 0; b < 2; b++)
#pragma unroll
            for (int e = 0; e < 2; e++)
                if (dst[a][b][e]) *dst[a][b][e] = cur[a][b][e] - acc[a][b][e];
}

// One launch per 32-pivot panel step of a big front: every CTA (front, tile pair) redoes the LDL' of the
// diagonal block (one warp, registers) and the triangular solve of ITS two 64-row tiles from the raw panel,
// then applies the rank-32 update of its 64 x 64 tile on the FP64 tensor pipe.  The factored panel goes to
// Xout (tile pairs with tj == 0 write their rows, pair 0 the diagonal block), the raw panel in Lx stays
// intact for the other CTAs -- so potrf -> trsm -> syrk needs no grid-wide ordering, and the serial chain of
// the tree top pays one launch per panel instead of three.  Arithmetic identical to the three-kernel path.
__global__ void __launch_bounds__(256) k_big_panel(FArgs A, const int *list, int kb)
{
    __shared__ double Pa[TS * PLD], Pb[TS * PLD];
    __shared__ double D[NB][NB + 1];
    __shared__ double Rinv[NB];
    const int s = list[blockIdx.x];
    const int f = A.sn_start[s], nc = A.sn_start[s + 1] - f, nr = A.sn_nr[s], N = nc + nr;
    if (kb >= nc) return;
    const int nb = min(NB, nc - kb);
    const int r0 = kb + nb, R = N - r0;
    const int nt = (R + TS - 1) / TS;
    int ti = (int)((sqrt(8.0 * blockIdx.y + 1.0) - 1.0) * 0.5);
    while ((ti + 1) * (ti + 2) / 2 <= (int)blockIdx.y) ti++;
    while (ti * (ti + 1) / 2 > (int)blockIdx.y) ti--;
    const int tj = blockIdx.y - ti * (ti + 1) / 2;
    if (blockIdx.y > 0 && ti >= nt) return;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const double *Lraw = A.Lx + A.Lp[s] + (int64_t)kb * N;          // raw panel, column k at Lraw + k * N
    double *Xo = A.Xout + A.Lp[s] + (int64_t)kb * N;
    // ---- LDL' of the diagonal block (warp 0), result in D: unit lower L11, D on the diagonal
    if (wid == 0) {
        double a[NB];
#pragma unroll
        for (int j = 0; j < NB; j++) a[j] = (j < nb && lane < nb && lane >= j) ? Lraw[(int64_t)j * N + kb + lane] : (lane == j ? 1.0 : 0.0);
        bool bad = false;
#pragma unroll
        for (int k = 0; k < NB; k++) {
            const double dk = __shfl_sync(0xffffffffu, a[k], k);
            bad = bad || (k < nb && !(dk > 0.0));
            const double ak = a[k];
            const double l = div_by_pivot(ak, dk, __drcp_rn(dk));
#pragma unroll
            for (int j = k + 1; j < NB; j++) {
                const double lj = __shfl_sync(0xffffffffu, l, j);
                a[j] = fma(-ak, lj, a[j]);
            }
            if (lane > k) a[k] = l;
        }
        if (bad && lane == 0 && blockIdx.y == 0) A.sc->not_posdef = 1;
#pragma unroll
        for (int j = 0; j < NB; j++) {
            D[lane][j] = (lane >= j) ? a[j] : 0.0;
            if (blockIdx.y == 0 && j < nb && lane < nb) Xo[(int64_t)j * N + kb + lane] = (lane >= j) ? a[j] : 0.0;
        }
        __syncwarp();
        Rinv[lane] = __drcp_rn(D[lane][lane]);
    }
    __syncthreads();
    if (R <= 0) return;
    // ---- triangular solve of the two row tiles: thread t < 64 owns row ra + t, thread 64 + t row rb + t
    const int ra = r0 + ti * TS, rb = r0 + tj * TS;
    if (tid < 2 * TS) {
        const bool second = tid >= TS;
        const int r = second ? tid - TS : tid, i = (second ? rb : ra) + r;
        double x[NB];
        const bool live = i < N && !(second && ti == tj);
#pragma unroll
        for (int j = 0; j < NB; j++) x[j] = (live && j < nb) ? Lraw[(int64_t)j * N + i] : 0.0;
#pragma unroll
        for (int j = 0; j < NB; j++) {
            double v = x[j];
#pragma unroll
            for (int k = 0; k < j; k++) v = fma(-x[k], D[j][k], v);
            x[j] = v;
        }
#pragma unroll
        for (int j = 0; j < NB; j++) x[j] = div_by_pivot(x[j], D[j][j], Rinv[j]);
        double *P = second ? Pb : Pa;
        if (!(second && ti == tj)) {
#pragma unroll
            for (int j = 0; j < NB; j++) {
                const double xv = (j < nb && i < N) ? x[j] : 0.0;
                if (!second) { Pa[r * PLD + j] = xv * D[j][j]; if (ti == tj) Pb[r * PLD + j] = xv; }
                else P[r * PLD + j] = xv;
            }
            if (!second && tj == 0 && i < N) {
#pragma unroll
                for (int j = 0; j < NB; j++) if (j < nb) Xo[(int64_t)j * N + i] = x[j];
            }
        }
    }
    __syncthreads();
    // ---- rank-nb update of the 64 x 64 tile (ti, tj): C -= (X D) X'
    const int wr = (wid & 1) * 32, wc = (wid >> 1) * 16;
    const int g = lane >> 2, q = lane & 3;
    double acc[4][2][2];
#pragma unroll
    for (int a = 0; a < 4; a++)
#pragma unroll
        for (int b = 0; b < 2; b++) acc[a][b][0] = acc[a][b][1] = 0.0;
#pragma unroll
    for (int ks = 0; ks < NB / 4; ks++) {
        double af[4], bf[2];
#pragma unroll
        for (int a = 0; a < 4; a++) af[a] = Pa[(wr + a * 8 + g) * PLD + ks * 4 + q];
#pragma unroll
        for (int b = 0; b < 2; b++) bf[b] = Pb[(wc + b * 8 + g) * PLD + ks * 4 + q];
#pragma unroll
        for (int a = 0; a < 4; a++)
#pragma unroll
            for (int b = 0; b < 2; b++) dmma884(acc[a][b][0], acc[a][b][1], af[a], bf[b]);
    }
    double *dst[4][2][2];
    double cur[4][2][2];
#pragma unroll
    for (int a = 0; a < 4; a++)
#pragma unroll
        for (int b = 0; b < 2; b++)
#pragma unroll
            for (int e = 0; e < 2; e++) {
                const int gi = ra + wr + a * 8 + g, gj = rb + wc + b * 8 + q * 2 + e;
                dst[a][b][e] = (gi < N && gj <= gi) ? front_elem(A, s, nc, nr, N, gi, gj) : nullptr;
                cur[a][b][e] = dst[a][b][e] ? *dst[a][b][e] : 0.0;
            }
#pragma unroll
    for (int a = 0; a < 4; a++)
#pragma unroll
        for (int b = 0; b < 2; b++)
#pragma unroll
            for (int e = 0; e < 2; e++)
                if (dst[a][b][e]) *dst[a][b][e] = cur[a][b][e] - acc[a][b][e];
}

// ------------------------------------------------------------------------------- triangular solves
struct SArgs {
    const int *sn_start, *sn_nr, *rel, *rows, *child_ptr, *child_idx;
    const int64_t *sn_rowptr, *Lp;
    const double *Lx, *Xout;   // fronts of order N > bigN read their panel from Xout
    int bigN;
    double *xs, *uv;
};
// forward: (assembled) t = b_s + children's updates; y_s = L11^{-1} t (unit lower); z_s = D^{-1} y_s is stored;
// u_s = children's pass-through - L21 y_s
template <int NT>
__global__ void __launch_bounds__(NT) k_fwd(SArgs A, const int *list)
{
    extern __shared__ double vec[];                    // N entries + NB for the block solution
    __shared__ double D[NB][NB + 1];
    const int s = list[blockIdx.x];
    const int f = A.sn_start[s], nc = A.sn_start[s + 1] - f, nr = A.sn_nr[s], N = nc + nr;
    const double *L = (N > A.bigN ? A.Xout : A.Lx) + A.Lp[s];
    const int tid = threadIdx.x, lane = tid & 31;
    double *yb = vec + N;
    for (int i = tid; i < N; i += NT) vec[i] = (i < nc) ? A.xs[f + i] : 0.0;
    __syncthreads();
    for (int t = A.child_ptr[s]; t < A.child_ptr[s + 1]; t++) {
        const int c = A.child_idx[t], rc = A.sn_nr[c];
        const int *rl = A.rel + A.sn_rowptr[c];
        const double *u = A.uv + A.sn_rowptr[c];
        for (int i = tid; i < rc; i += NT) vec[rl[i]] += u[i];
        __syncthreads();
    }
    for (int kb = 0; kb < nc; kb += NB) {
        const int nb = min(NB, nc - kb);
        for (int idx = tid; idx < nb * nb; idx += NT) { int j = idx / nb, i = idx - j * nb; D[i][j] = L[(int64_t)(kb + j) * N + kb + i]; }
        __syncthreads();
        if (tid < 32) {
            double v = (lane < nb) ? vec[kb + lane] : 0.0;
            for (int k = 0; k < nb; k++) {
                const double yk = __shfl_sync(0xffffffffu, v, k);
                if (lane > k && lane < nb) v = fma(-D[lane][k], yk, v);
            }
            // y feeds the updates below; z = D^{-1} y is what the backward pass starts from
            if (lane < nb) { yb[lane] = v; A.xs[f + kb + lane] = v / D[lane][lane]; }
        }
        __syncthreads();
        // rows below the block: one row per thread, its NB panel entries fetched as two batches of 16
        // independent loads (the panel streams from HBM/L2 once; latency, not bandwidth, bounds one CTA)
        for (int i = kb + nb + tid; i < N; i += NT) {
            const double *Li = L + (int64_t)kb * N + i;
            double acc = 0.0;
            if (nb == NB) {
#pragma unroll
                for (int h = 0; h < NB; h += 16) {
                    double lv[16];
#pragma unroll
                    for (int k = 0; k < 16; k++) lv[k] = __ldg(Li + (int64_t)(h + k) * N);
#pragma unroll
                    for (int k = 0; k < 16; k++) acc = fma(lv[k], yb[h + k], acc);
                }
            } else {
                for (int k = 0; k < nb; k++) acc = fma(Li[(int64_t)k * N], yb[k], acc);
            }
            vec[i] -= acc;
        }
        __syncthreads();
    }
    double *u = A.uv + A.sn_rowptr[s];
    for (int i = tid; i < nr; i += NT) u[i] = vec[nc + i];
}
// backward: x_s = L11^{-T} (z_s - L21' x_rows), unit diagonal
template <int NT>
__global__ void __launch_bounds__(NT) k_bwd(SArgs A, const int *list)
{
    extern __shared__ double vec[];
    __shared__ double D[NB][NB + 1];
    const int s = list[blockIdx.x];
    const int f = A.sn_start[s], nc = A.sn_start[s + 1] - f, nr = A.sn_nr[s], N = nc + nr;
    const double *L = (N > A.bigN ? A.Xout : A.Lx) + A.Lp[s];
    const int *rows = A.rows + A.sn_rowptr[s];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5, NW = NT / 32;
    for (int i = tid; i < N; i += NT) vec[i] = (i < nc) ? A.xs[f + i] : A.xs[rows[i - nc]];
    __syncthreads();
    const int nblk = (nc + NB - 1) / NB;
    for (int b = nblk - 1; b >= 0; b--) {
        const int kb = b * NB, nb = min(NB, nc - kb);
        for (int idx = tid; idx < nb * nb; idx += NT) { int j = idx / nb, i = idx - j * nb; D[i][j] = L[(int64_t)(kb + j) * N + kb + i]; }
        // t_k = vec[k] - sum_{i >= kb+nb} L[i,k] vec[i], one warp per column
        for (int k = wid; k < nb; k += NW) {
            const double *Lk = L + (int64_t)(kb + k) * N;
            double acc = 0.0;
            int i = kb + nb + lane;
            for (; i + 7 * 32 < N; i += 8 * 32) {            // 8 independent loads in flight per lane
                double lv[8];
#pragma unroll
                for (int r = 0; r < 8; r++) lv[r] = __ldg(Lk + i + r * 32);
#pragma unroll
                for (int r = 0; r < 8; r++) acc = fma(lv[r], vec[i + r * 32], acc);
            }
            for (; i < N; i += 32) acc = fma(Lk[i], vec[i], acc);
            for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
            if (lane == 0) vec[kb + k] -= acc;
        }
        __syncthreads();
        if (tid < 32) {
            double v = (lane < nb) ? vec[kb + lane] : 0.0;
            for (int k = nb - 1; k >= 0; k--) {
                const double xk = __shfl_sync(0xffffffffu, v, k);
                if (lane < k) v = fma(-D[k][lane], xk, v);
            }
            if (lane < nb) { vec[kb + lane] = v; A.xs[f + kb + lane] = v; }
        }
        __syncthreads();
    }
}
__global__ void k_permute_in(const double *rhs, const int *perm, double *xs, int n)
{
    int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k < n) xs[k] = rhs[perm[k]];
}
__global__ void k_permute_out(const double *xs, const int *perm, double *x, int n)
{
    int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k < n) x[perm[k]] = xs[k];
}

// =============================================================================== host side
static const int BUCKET_N[4] = {32, 64, 104, 150};       // smem-front size classes (front order N)
static const int SOLVE_BIG_N = 256;                       // fronts above this order are solved by 1024 threads
static const int SOLVE_WARP_N = 64;                       // fronts up to this order are solved by one warp

static FArgs fargs(qpdo_dev *d)
{
    Factor &F = d->F;
    FArgs A;
    A.sn_start = F.sn_start; A.sn_nr = F.sn_nr; A.rel = F.rel; A.rows = F.rows; A.child_ptr = F.child_ptr;
    A.child_idx = F.child_idx; A.sn_rowptr = F.sn_rowptr; A.Lp = F.Lp; A.Up = F.Up; A.Lx = F.Lx; A.Ux = F.Ux; A.Xout = F.Xout; A.sc = d->sc;
    return A;
}
static SArgs sargs(qpdo_dev *d)
{
    Factor &F = d->F;
    SArgs A;
    A.sn_start = F.sn_start; A.sn_nr = F.sn_nr; A.rel = F.rel; A.rows = F.rows; A.child_ptr = F.child_ptr;
    A.child_idx = F.child_idx; A.sn_rowptr = F.sn_rowptr; A.Lp = F.Lp; A.Lx = F.Lx; A.Xout = F.Xout; A.bigN = BUCKET_N[3]; A.xs = F.xs; A.uv = F.uv;
    return A;
}

// ---- pattern management ---------------------------------------------------------------------------
// K = Q + sigma I + A' diag(omega) A is factored on the pattern of Q plus the rows of A in a ROW SET P that
// covers the support of omega.  P = all rows (the fixed superset, one analysis for the whole solve) is the
// default; for large problems whose support is a small fraction of m (Lasso paths: a handful of active rows
// out of 100,000) the superset factor is orders of magnitude more work than the reference's, which
// re-analyses on the active pattern at every refactor (cholmod_interface.c:10).  There P follows the
// support: a counting kernel checks support(omega) <= P before every factorization, and the host
// re-analyses (60 ms for n = 200k) only when a row outside P gains weight.
static const int DYNAMIC_MIN_ROWS = 20000;                // smaller problems always use the superset
static const int FULL_FRACTION = 8;                       // support > m / FULL_FRACTION: superset from then on

template <class T> static int f_alloc(qpdo_dev *d, T **p, size_t count) { if (dev_alloc(d, p, count)) return -1; d->F.owned.push_back((void *)*p); return 0; }
template <class T> static int f_upload(qpdo_dev *d, T **p, const T *h, size_t count) { if (dev_upload(d, p, h, count)) return -1; d->F.owned.push_back((void *)*p); return 0; }
static void factor_release(qpdo_dev *d)
{
    Factor &F = d->F;
    cudaStreamSynchronize(d->st);
    if (F.g_factor) cudaGraphExecDestroy(F.g_factor);
    if (F.g_solve) cudaGraphExecDestroy(F.g_solve);
    F.g_factor = F.g_solve = nullptr;
    for (void *q : F.owned) {
        for (size_t i = 0; i < d->allocs.size(); i++) if (d->allocs[i] == q) { d->allocs[i] = d->allocs.back(); d->allocs.pop_back(); break; }
        cudaFree(q);
    }
    F.owned.clear();
    if (F.S) qpdo_symbolic_free(F.S);
    F.S = nullptr;
}

__global__ void k_support(const double *omega, const unsigned char *inP, int m, int *cnt)
{   // cnt[0] = rows carrying weight, cnt[1] = those outside the current pattern row set
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    const bool sup = i < m && omega[i] != 0.0, out = sup && !inP[i];
    const unsigned bs = __ballot_sync(0xffffffffu, sup), bo = __ballot_sync(0xffffffffu, out);
    if ((threadIdx.x & 31) == 0) { if (bs) atomicAdd(&cnt[0], __popc(bs)); if (bo) atomicAdd(&cnt[1], __popc(bo)); }
}

static int factor_analyze(qpdo_dev *d, const unsigned char *rowmask);

int factor_setup(qpdo_dev *d, int ordering)
{
    Factor &F = d->F;
    const int n = d->n, m = d->m;
    F.ordering = ordering;
    // host copies of the patterns (re-analysis needs them)
    F.hQp.assign(n + 1, 0); F.hQi.assign(d->nnzQ ? d->nnzQ : 1, 0); F.hAp.assign(n + 1, 0); F.hAi.assign(d->nnzA ? d->nnzA : 1, 0);
    CUDA_OK(cudaMemcpyAsync(F.hQp.data(), d->Qp, sizeof(int) * (n + 1), cudaMemcpyDeviceToHost, d->st));
    CUDA_OK(cudaMemcpyAsync(F.hAp.data(), d->Ap, sizeof(int) * (n + 1), cudaMemcpyDeviceToHost, d->st));
    if (d->nnzQ) CUDA_OK(cudaMemcpyAsync(F.hQi.data(), d->Qi, sizeof(int) * d->nnzQ, cudaMemcpyDeviceToHost, d->st));
    if (d->nnzA) CUDA_OK(cudaMemcpyAsync(F.hAi.data(), d->Ai, sizeof(int) * d->nnzA, cudaMemcpyDeviceToHost, d->st));
    CUDA_OK(cudaStreamSynchronize(d->st));
    if (dev_alloc(d, &F.xs, n) || dev_alloc(d, &F.inP, (size_t)std::max(m, 1)) || dev_alloc(d, &F.cnt, 2)) return -1;
    if (cudaMallocHost((void **)&F.cnt_host, 2 * sizeof(int)) != cudaSuccess) { qpdo_set_error("cudaMallocHost failed"); return -1; }
    cudaFuncSetAttribute(k_front_smem<256>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    cudaFuncSetAttribute(k_front_smem<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
    const char *ev = getenv("QPDO_PATTERN");              // "superset" | "dynamic" override the size rule (tests)
    F.dynamic = (m >= DYNAMIC_MIN_ROWS);
    if (ev && !strcmp(ev, "superset")) F.dynamic = false;
    if (ev && !strcmp(ev, "dynamic") && m > 0) F.dynamic = true;
    if (!F.dynamic) return factor_analyze(d, nullptr);    // superset, once, at setup
    return 0;                                             // dynamic: first analysis at the first factorization
}

// analyse on pattern(Q) + rows of A selected by rowmask (nullptr: all rows) and rebuild the device structures
static int factor_analyze(qpdo_dev *d, const unsigned char *rowmask)
{
    Factor &F = d->F;
    const int n = d->n, m = d->m;
    auto t0 = std::chrono::steady_clock::now();
    factor_release(d);
    std::vector<int> Ap, Ai, rowmap, nzmap;
    const int *pAp = F.hAp.data(), *pAi = F.hAi.data();
    int mr = m;
    std::vector<unsigned char> inP((size_t)std::max(m, 1), rowmask ? 0 : 1);
    if (rowmask) {
        std::vector<int> newid(m, -1);
        mr = 0;
        for (int i = 0; i < m; i++) if (rowmask[i]) { newid[i] = mr++; rowmap.push_back(i); inP[i] = 1; }
        Ap.assign(n + 1, 0);
        for (int j = 0; j < n; j++) {
            for (int t = F.hAp[j]; t < F.hAp[j + 1]; t++) if (newid[F.hAi[t]] >= 0) { Ai.push_back(newid[F.hAi[t]]); nzmap.push_back(t); }
            Ap[j + 1] = (int)Ai.size();
        }
        if (Ai.empty()) Ai.push_back(0);
        pAp = Ap.data(); pAi = Ai.data();
    }
    F.full = (rowmask == nullptr); F.npattern = mr;
    char err[256];
    qpdo_symbolic *S = qpdo_symbolic_analyze(n, mr, F.hQp.data(), F.hQi.data(), pAp, pAi, F.ordering, err);
    if (!S) { qpdo_set_error("symbolic analysis failed: %s", err); return -1; }
    if (rowmask) {                                        // back to the numbering of the full A
        for (int64_t t = 0; t < S->ntriple; t++) { S->Kt_row[t] = rowmap[S->Kt_row[t]]; S->Kt_a1[t] = nzmap[S->Kt_a1[t]]; S->Kt_a2[t] = nzmap[S->Kt_a2[t]]; }
    }
    F.S = S;
    F.analyses++;
    if (m) CUDA_OK(cudaMemcpyAsync(F.inP, inP.data(), (size_t)m, cudaMemcpyHostToDevice, d->st));
    const int ns = S->nsuper;
    const int64_t nrows = S->sn_rowptr[ns];
    std::vector<int> Kcol(S->nnzK);
    for (int k = 0; k < n; k++) for (int e = S->Kp[k]; e < S->Kp[k + 1]; e++) Kcol[e] = k;
    if (f_upload(d, &F.sn_start, (const int *)S->sn_start, ns + 1) || f_upload(d, &F.sn_nr, (const int *)S->sn_nr, ns) ||
        f_upload(d, &F.rows, (const int *)S->rows, nrows) || f_upload(d, &F.rel, (const int *)S->rel, nrows) ||
        f_upload(d, &F.child_ptr, (const int *)S->child_ptr, ns + 1) || f_upload(d, &F.child_idx, (const int *)S->child_idx, ns) ||
        f_upload(d, &F.level_sn, (const int *)S->level_sn, ns) || f_upload(d, &F.sn_rowptr, (const int64_t *)S->sn_rowptr, ns + 1) ||
        f_upload(d, &F.Lp, (const int64_t *)S->Lp, ns + 1) || f_upload(d, &F.Up, (const int64_t *)S->Up, ns + 1) ||
        f_upload(d, &F.perm, (const int *)S->perm, n) || f_upload(d, &F.pinv, (const int *)S->pinv, n) ||
        f_upload(d, &F.Ki, (const int *)S->Ki, S->nnzK) || f_upload(d, &F.Kcol, (const int *)Kcol.data(), S->nnzK) ||
        f_upload(d, &F.Kq, (const int *)S->Kq, S->nnzK) || f_upload(d, &F.Kdst, (const int64_t *)S->Kdst, S->nnzK) ||
        f_upload(d, &F.Ktp, (const int64_t *)S->Ktp, S->nnzK + 1) || f_upload(d, &F.Kt_row, (const int *)S->Kt_row, S->ntriple) ||
        f_upload(d, &F.Kt_a1, (const int *)S->Kt_a1, S->ntriple) || f_upload(d, &F.Kt_a2, (const int *)S->Kt_a2, S->ntriple) ||
        f_alloc(d, &F.Kt_val, S->ntriple) || f_alloc(d, &F.Lx, S->lnz_stored) || f_alloc(d, &F.Xout, S->max_front > BUCKET_N[3] ? S->lnz_stored : 1) || f_alloc(d, &F.Ux, S->usize) ||
        f_alloc(d, &F.uv, nrows))
        return -1;
    // launch plan: per level, fronts are sorted by decreasing order N
    F.plan.assign(S->nlevel, std::vector<FrontLaunch>());
    for (int lv = 0; lv < S->nlevel; lv++) {
        int b = S->level_ptr[lv], e = S->level_ptr[lv + 1];
        auto ordN = [&](int pos) { int s = S->level_sn[pos]; return S->sn_start[s + 1] - S->sn_start[s] + S->sn_nr[s]; };
        auto ncOf = [&](int pos) { int s = S->level_sn[pos]; return S->sn_start[s + 1] - S->sn_start[s]; };
        int pos = b;
        // big fronts first
        int q = pos;
        while (q < e && ordN(q) > BUCKET_N[3]) q++;
        if (q > pos) {
            FrontLaunch L; L.first = pos; L.count = q - pos; L.kind = 2; L.maxN = ordN(pos); L.maxnc = 0;
            for (int t = pos; t < q; t++) L.maxnc = std::max(L.maxnc, ncOf(t));
            F.plan[lv].push_back(L); pos = q;
        }
        for (int bk = 3; bk >= 0 && pos < e; bk--) {
            int lo = bk > 0 ? BUCKET_N[bk - 1] : 0;
            q = pos;
            while (q < e && ordN(q) > lo) q++;
            if (q > pos) {
                FrontLaunch L; L.first = pos; L.count = q - pos; L.kind = (bk == 0) ? 0 : 1; L.maxN = ordN(pos); L.maxnc = bk;
                F.plan[lv].push_back(L); pos = q;
            }
        }
    }
    d->analyze_seconds += std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    return factor_refresh_products(d);
}

// before a factorization: does the pattern row set still cover the support of omega?
static int factor_ensure_pattern(qpdo_dev *d)
{
    Factor &F = d->F;
    const int m = d->m;
    if (F.S && (F.full || !F.dynamic)) return 0;
    CUDA_OK(cudaMemsetAsync(F.cnt, 0, 2 * sizeof(int), d->st));
    if (m) k_support<<<cdiv(m, 256), 256, 0, d->st>>>(d->omega, F.inP, m, F.cnt);
    CUDA_OK(cudaMemcpyAsync(F.cnt_host, F.cnt, 2 * sizeof(int), cudaMemcpyDeviceToHost, d->st));
    CUDA_OK(cudaStreamSynchronize(d->st));
    const int support = F.cnt_host[0], outside = F.cnt_host[1];
    if (F.S && outside == 0 && F.npattern <= 4 * support + 4096) return 0;
    if ((long)support * FULL_FRACTION > (long)m) return factor_analyze(d, nullptr);
    std::vector<double> om((size_t)m);
    CUDA_OK(cudaMemcpyAsync(om.data(), d->omega, sizeof(double) * m, cudaMemcpyDeviceToHost, d->st));
    CUDA_OK(cudaStreamSynchronize(d->st));
    std::vector<unsigned char> mask((size_t)m);
    for (int i = 0; i < m; i++) mask[i] = om[i] != 0.0;
    return factor_analyze(d, mask.data());
}

int factor_refresh_products(qpdo_dev *d)
{
    Factor &F = d->F;
    if (!F.S) return 0;                                   // dynamic pattern: nothing analysed yet
    if (F.S->ntriple) k_products<<<cdiv(F.S->ntriple, 256), 256, 0, d->st>>>(F.S->ntriple, F.Kt_a1, F.Kt_a2, d->Ax, F.Kt_val);
    return 0;
}

static long issue_factor(qpdo_dev *d)
{
    Factor &F = d->F;
    const qpdo_symbolic *S = F.S;
    FArgs A = fargs(d);
    long launches = 0;
    for (int lv = 0; lv < S->nlevel; lv++)
        for (const FrontLaunch &L : F.plan[lv]) {
            const int *list = F.level_sn + L.first;
            if (L.kind == 2) {
                k_big_assemble<<<dim3(L.count, cdiv(L.maxN, CW)), 256, 0, d->st>>>(A, list); launches++;
                for (int kb = 0; kb < L.maxnc; kb += NB) {
                    const int R = L.maxN - kb - 1, nt = R > 0 ? cdiv(R, TS) : 0;
                    k_big_panel<<<dim3(L.count, std::max(1, nt * (nt + 1) / 2)), 256, 0, d->st>>>(A, list, kb); launches++;
                }
            } else {
                size_t smem = sizeof(double) * ((size_t)(L.maxN | 1) * L.maxN + 8 * (size_t)L.maxN);
                if (L.maxnc == 0) k_front_smem<64><<<L.count, 64, smem, d->st>>>(A, list);
                else if (L.maxnc == 1) k_front_smem<128><<<L.count, 128, smem, d->st>>>(A, list);
                else k_front_smem<256><<<L.count, 256, smem, d->st>>>(A, list);
                launches++;
            }
        }
    return launches;
}
static long issue_solve(qpdo_dev *d)
{
    Factor &F = d->F;
    const qpdo_symbolic *S = F.S;
    SArgs A = sargs(d);
    long launches = 0;
    auto ordN = [&](int pos) { int s = S->level_sn[pos]; return S->sn_start[s + 1] - S->sn_start[s] + S->sn_nr[s]; };
    for (int pass = 0; pass < 2; pass++)
        for (int t = 0; t < S->nlevel; t++) {
            int lv = pass == 0 ? t : S->nlevel - 1 - t;
            int b = S->level_ptr[lv], e = S->level_ptr[lv + 1], q = b;
            while (q < e && ordN(q) > SOLVE_WARP_N) q++;
            if (q > b) {
                size_t smem = sizeof(double) * (ordN(b) + NB);
                // fronts wider than SOLVE_BIG_N rows get 1024 threads (one panel row / one panel column per
                // thread / warp), the rest of the CTA-solved fronts 256
                int qb = b;
                while (qb < q && ordN(qb) > SOLVE_BIG_N) qb++;
                if (qb > b) {
                    if (pass == 0) k_fwd<1024><<<qb - b, 1024, smem, d->st>>>(A, F.level_sn + b);
                    else k_bwd<1024><<<qb - b, 1024, smem, d->st>>>(A, F.level_sn + b);
                    launches++;
                }
                if (q > qb) {
                    smem = sizeof(double) * (ordN(qb) + NB);
                    if (pass == 0) k_fwd<256><<<q - qb, 256, smem, d->st>>>(A, F.level_sn + qb);
                    else k_bwd<256><<<q - qb, 256, smem, d->st>>>(A, F.level_sn + qb);
                    launches++;
                }
            }
            if (e > q) {
                size_t smem = sizeof(double) * (ordN(q) + NB);
                if (pass == 0) k_fwd<32><<<e - q, 32, smem, d->st>>>(A, F.level_sn + q);
                else k_bwd<32><<<e - q, 32, smem, d->st>>>(A, F.level_sn + q);
                launches++;
            }
        }
    return launches;
}
static int capture(qpdo_dev *d, long (*issue)(qpdo_dev *), cudaGraphExec_t *out, long *launches)
{
    cudaGraph_t g = nullptr;
    CUDA_OK(cudaStreamBeginCapture(d->st, cudaStreamCaptureModeThreadLocal));
    *launches = issue(d);
    CUDA_OK(cudaStreamEndCapture(d->st, &g));
    CUDA_OK(cudaGraphInstantiate(out, g, 0));
    cudaGraphDestroy(g);
    return 0;
}

int factor_numeric(qpdo_dev *d)
{
    Factor &F = d->F;
    if (factor_ensure_pattern(d)) return -1;
    const qpdo_symbolic *S = F.S;
    {
        PhaseTimer t(d, PH_ASSEMBLE);
        CUDA_OK(cudaMemsetAsync(F.Lx, 0, sizeof(double) * S->lnz_stored, d->st));
        k_assemble_K<<<cdiv(S->nnzK, 256), 256, 0, d->st>>>(S->nnzK, F.Kdst, F.Kq, F.Ktp, F.Kt_row, F.Kt_val, F.Ki, F.Kcol,
                                                            d->Qx, d->omega, d->sc, d->proximal, F.Lx);
        d->prof.kernel_launches += 2;
    }
    {
        PhaseTimer t(d, PH_FACTOR);
        if (!F.g_factor && capture(d, issue_factor, &F.g_factor, &F.launches_factor)) return -1;
        CUDA_OK(cudaGraphLaunch(F.g_factor, d->st));
        d->prof.kernel_launches += F.launches_factor;
    }
    return 0;
}

int factor_solve(qpdo_dev *d, const double *rhs, double *x)
{
    Factor &F = d->F;
    PhaseTimer t(d, PH_SOLVE);
    k_permute_in<<<cdiv(d->n, 256), 256, 0, d->st>>>(rhs, F.perm, F.xs, d->n);
    if (!F.g_solve && capture(d, issue_solve, &F.g_solve, &F.launches_solve)) return -1;
    CUDA_OK(cudaGraphLaunch(F.g_solve, d->st));
    k_permute_out<<<cdiv(d->n, 256), 256, 0, d->st>>>(F.xs, F.perm, x, d->n);
    d->prof.kernel_launches += F.launches_solve + 2;
    return 0;
}

void factor_destroy(qpdo_dev *d)
{
    Factor &F = d->F;
    factor_release(d);
    if (F.cnt_host) cudaFreeHost(F.cnt_host);
    F.cnt_host = nullptr;
}
