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
#include <cooperative_groups.h>
#include <algorithm>
#include <chrono>
#include <stdlib.h>
#include <string.h>

namespace cg = cooperative_groups;

#define NB 32          // panel width of the blocked path
#ifndef CW
#define CW 8           // destination-column chunk of the big-front assembly: one column per warp
#endif
#define TS 64          // SYRK tile
#define PLD 36         // padded leading dimension of a 64 x 32 panel tile in smem (conflict free)

struct FrontDesc { int64_t Lp, Up; int nc, nr; };        // one per supernode: what a front kernel needs to find its data
struct FArgs {
    const FrontDesc *fdesc;
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
#ifdef FRONT_PROBE
// -DFRONT_PROBE: per template (NT = 64 / 128 / 256) sums of the SM cycles a CTA spends in its phases (tools/probe_front.py)
__device__ unsigned long long g_fprobe[3][12];
#define FPROBE(i) do { if (threadIdx.x == 0) fp[i] = clock64(); } while (0)
extern "C" int qpdo_dev_fprobe_read(unsigned long long *out)
{
    cudaDeviceSynchronize();
    cudaMemcpyFromSymbol(out, g_fprobe, sizeof(unsigned long long) * 36);
    unsigned long long z[36] = {0};
    cudaMemcpyToSymbol(g_fprobe, z, sizeof z);
    return 0;
}
#else
#define FPROBE(i)
#endif
template <int NT>
__global__ void __launch_bounds__(NT) k_front_smem(FArgs A, const int *list)
{
    extern __shared__ double F[];
#ifdef FRONT_PROBE
    long long fp[6], fpanel = 0, ftrail = 0, fur = 0, ft0 = 0;
#endif
    FPROBE(0);
    const int s = list[blockIdx.x];
    const int f = A.sn_start[s], nc = A.sn_start[s + 1] - f, nr = A.sn_nr[s], N = nc + nr, ld = N | 1;
    double *Lp = A.Lx + A.Lp[s];
    const int tid = threadIdx.x;
    for (int idx = tid; idx < N * nc; idx += NT) { int j = idx / N, i = idx - j * N; F[i + j * ld] = Lp[idx]; }
    for (int idx = tid; idx < N * nr; idx += NT) { int j = idx / N, i = idx - j * N; F[i + (nc + j) * ld] = 0.0; }
    __syncthreads();
    FPROBE(1);
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
    FPROBE(2);
    // right-looking LDL' in blocks of KB pivots.  Panel phase, per pivot: the rows below it compute
    // l_i = a_ik / d_k (unscaled u_i and scaled l_i go to the ub / lb buffers, the L entry is final), then the
    // remaining panel columns are updated.  Trailing phase, once per block: warp per column, lane per row
    // (row values u_i^t stay in registers across the columns), a_ij = fma(-u^3, l^3, fma(... fma(-u^0, l^0, a_ij)))
    // -- the same ascending-k fma sequence per entry as a column-by-column elimination, with a quarter of
    // the shared-memory traffic and barriers.
    constexpr int KB = 4, RMAX = NT == 64 ? 1 : (NT == 128 ? 2 : 5);   // row chunks of 32: N <= 32 / 64 / 150 by size class (fewer registers, more resident CTAs for the small ones)
    const int lane = tid & 31, wid = tid >> 5, NW = NT / 32;
    double *ub = F + (size_t)ld * N, *lb = ub + KB * N;
    for (int kb = 0; kb < nc; kb += KB) {
        const int nk = min(KB, nc - kb);
        // panel: one thread per row i >= kb with the row's KB panel entries in registers.  EVERY thread loads the
        // KB x KB pivot block (broadcast loads) and factors it redundantly, so a thread finishes its own row with no
        // exchange and no barrier on the pivot chain.  Per entry the operations are those of the column-by-column
        // elimination (u / d by div_by_pivot, a_ic = fma(-u_i, l_c, a_ic) in ascending pivot order): same bits.
#ifdef FRONT_PROBE
        if (tid == 0) ft0 = clock64();
#endif
        // (NT >= N in every bucket: one row per thread.)  The pivot rows' threads overwrite the pivot block with their
        // L entries, so every thread must have loaded the block before anyone stores: barrier between loads and stores.
        const int i = kb + tid;
        const bool act = i < N;
        double P[KB][KB], a[KB];
#pragma unroll
        for (int c = 0; c < KB; c++)
#pragma unroll
            for (int r = c; r < KB; r++) P[c][r] = (r < nk && c < nk) ? F[(kb + r) + (kb + c) * ld] : (r == c ? 1.0 : 0.0);
#pragma unroll
        for (int t = 0; t < KB; t++) a[t] = (act && t < nk) ? F[i + (kb + t) * ld] : 0.0;
        __syncthreads();
        if (act) {
            double u[KB], l[KB];
            bool bad = false;
#pragma unroll
            for (int t = 0; t < KB; t++) {
                if (t >= nk) { u[t] = 0.0; l[t] = 0.0; continue; }     // (CTA-uniform: fronts with one or two pivots pay for one or two)
                const double dt = P[t][t];
                bad = bad || !(dt > 0.0);
                const double rt = __drcp_rn(dt);
                double L[KB];
#pragma unroll
                for (int r = t + 1; r < KB; r++) L[r] = div_by_pivot(P[t][r], dt, rt);
#pragma unroll
                for (int c = t + 1; c < KB; c++)
#pragma unroll
                    for (int r = c; r < KB; r++) P[c][r] = fma(-P[t][r], L[c], P[c][r]);
                u[t] = a[t];
                l[t] = div_by_pivot(a[t], dt, rt);
#pragma unroll
                for (int c = t + 1; c < KB; c++) a[c] = fma(-u[t], L[c], a[c]);
            }
            if (bad && i == kb) A.sc->not_posdef = 1;
#pragma unroll
            for (int t = 0; t < KB; t++) {
                if (t < nk) {
                    const int k = kb + t;
                    if (i > k) { F[i + k * ld] = l[t]; ub[t * N + i] = u[t]; lb[t * N + i] = l[t]; }
                    else if (i == k) F[i + k * ld] = u[t];          // the pivot itself, after the updates of this block
                }
            }
        }
        __syncthreads();
#ifdef FRONT_PROBE
        if (tid == 0) { const long long t1 = clock64(); fpanel += t1 - ft0; ft0 = t1; }
#endif
        double ur[RMAX][KB];
#pragma unroll
        for (int r = 0; r < RMAX; r++)
#pragma unroll
            for (int t = 0; t < KB; t++) { const int i = lane + 32 * r; ur[r][t] = (t < nk && i < N) ? ub[t * N + i] : 0.0; }
#ifdef FRONT_PROBE
        if (tid == 0) { const long long t1 = clock64(); fur += t1 - ft0; ft0 = t1; }
#endif
        // two columns per iteration, all loads before the first store (the compiler cannot reorder a load of F
        // over a store to F): 2 x RMAX independent 4-fma chains in flight instead of one.  Pivots beyond nk have
        // u = l = 0 and fma(-0, 0, v) == v, so no guard is needed.
        for (int j0 = kb + nk + wid; j0 < N; j0 += 2 * NW) {
            const int j1 = j0 + NW;
            const bool two = j1 < N;
            double l0[KB], l1[KB], v0[RMAX], v1[RMAX];
#pragma unroll
            for (int t = 0; t < KB; t++) { l0[t] = (t < nk) ? lb[t * N + j0] : 0.0; l1[t] = (t < nk && two) ? lb[t * N + j1] : 0.0; }
#pragma unroll
            for (int r = 0; r < RMAX; r++) {
                const int i = lane + 32 * r;
                v0[r] = (i >= j0 && i < N) ? F[i + j0 * ld] : 0.0;
                v1[r] = (two && i >= j1 && i < N) ? F[i + j1 * ld] : 0.0;
            }
#pragma unroll
            for (int t = 0; t < KB; t++)
#pragma unroll
                for (int r = 0; r < RMAX; r++) { v0[r] = fma(-ur[r][t], l0[t], v0[r]); v1[r] = fma(-ur[r][t], l1[t], v1[r]); }
#pragma unroll
            for (int r = 0; r < RMAX; r++) {
                const int i = lane + 32 * r;
                if (i >= j0 && i < N) F[i + j0 * ld] = v0[r];
                if (two && i >= j1 && i < N) F[i + j1 * ld] = v1[r];
            }
        }
        __syncthreads();
#ifdef FRONT_PROBE
        if (tid == 0) { const long long t1 = clock64(); ftrail += t1 - ft0; }
#endif
    }
    __syncthreads();
    FPROBE(3);
    for (int idx = tid; idx < N * nc; idx += NT) { int j = idx / N, i = idx - j * N; Lp[idx] = (i >= j) ? F[i + j * ld] : 0.0; }
    double *U = A.Ux + A.Up[s];
    for (int idx = tid; idx < nr * nr; idx += NT) {
        int j = idx / nr, i = idx - j * nr;
        if (i >= j) U[idx] = F[(nc + i) + (nc + j) * ld];
    }
#ifdef FRONT_PROBE
    FPROBE(4);
    if (threadIdx.x == 0) {
        const int w = NT == 64 ? 0 : (NT == 128 ? 1 : 2);
        for (int i = 0; i < 4; i++) atomicAdd(&g_fprobe[w][i], (unsigned long long)(fp[i + 1] - fp[i]));
        atomicAdd(&g_fprobe[w][4], 1ull); atomicAdd(&g_fprobe[w][5], (unsigned long long)N); atomicAdd(&g_fprobe[w][6], (unsigned long long)nc);
        atomicAdd(&g_fprobe[w][7], (unsigned long long)(A.child_ptr[s + 1] - A.child_ptr[s]));
        atomicAdd(&g_fprobe[w][8], (unsigned long long)fpanel); atomicAdd(&g_fprobe[w][9], (unsigned long long)fur); atomicAdd(&g_fprobe[w][10], (unsigned long long)ftrail);
    }
#endif
}

// ------------------------------------------------------------------------------- big fronts
__device__ __forceinline__ double *front_elem(const FArgs &A, int s, int nc, int nr, int N, int i, int j)
{   // element (i, j), i >= j, of the front: L panel for j < nc, update matrix otherwise
    return (j < nc) ? (A.Lx + A.Lp[s] + (int64_t)j * N + i) : (A.Ux + A.Up[s] + (int64_t)(j - nc) * nr + (i - nc));
}
// CTA (front, chunk) owns destination columns [c0, c1): zero their update-matrix part, then add the
// children's contributions that fall into these columns, children in fixed order.
// (Measured: 23 us per level whatever the inner structure -- a variant with the destination columns in shared memory and
// the children's row maps staged by cp.async ran at the same speed.  ncu: 50 % long-scoreboard stalls, 12-14 MB from DRAM
// per launch: a chain of dependent loads of cold data (front descriptor -> child list -> row maps -> columns).)
// HOIST (launches with a handful of fronts: the top of the tree, where this kernel sits on the serial chain): the symbolic
// prologue runs before griddepcontrol.wait.  Launches with many fronts keep the searches inside the child loop: the extra
// registers of the hoisted form cost occupancy there (C3: 40 us per factorization).
template <bool HOIST>
__global__ void __launch_bounds__(256) k_big_assemble(FArgs A, const int *list)
{
    const int s = list[blockIdx.x];
    const int f = A.sn_start[s], nc = A.sn_start[s + 1] - f, nr = A.sn_nr[s], N = nc + nr;
    const int c0 = blockIdx.y * CW;
    if (c0 >= N) return;
    const int c1 = min(N, c0 + CW), tid = threadIdx.x;
    // Everything that depends only on the symbolic structure -- child list, row maps, the first source column of every child
    // that falls into [c0, c1) (a lower-bound search: ~10 dependent loads) -- is done BEFORE griddepcontrol.wait, i.e. while
    // the last panel step of the level below is still running when this launch carries the programmatic-launch attribute.
    constexpr int MAXC = HOIST ? 8 : 1;
    const int t0 = A.child_ptr[s], nch = A.child_ptr[s + 1] - t0;
    int lo_c[MAXC], hi_c[MAXC];
    const int *rl_c[MAXC];
    if (HOIST) {
#pragma unroll
        for (int q = 0; q < MAXC; q++) {
            lo_c[q] = 0; hi_c[q] = 0; rl_c[q] = A.rel;
            if (q < nch) { const int c = A.child_idx[t0 + q]; hi_c[q] = A.sn_nr[c]; rl_c[q] = A.rel + A.sn_rowptr[c]; }
        }
        for (bool more = true; more;) {                // the searches of all children in lockstep: one round trip per probe level
            more = false;
#pragma unroll
            for (int q = 0; q < MAXC; q++)
                if (lo_c[q] < hi_c[q]) {
                    const int mid = (lo_c[q] + hi_c[q]) >> 1;
                    if (rl_c[q][mid] < c0) lo_c[q] = mid + 1; else hi_c[q] = mid;
                    more = true;
                }
        }
    }
    asm volatile("griddepcontrol.wait;" ::: "memory");
    for (int j = max(c0, nc); j < c1; j++) {
        double *col = A.Ux + A.Up[s] + (int64_t)(j - nc) * nr;
        for (int i = j - nc + tid; i < nr; i += 256) col[i] = 0.0;
    }
    __syncthreads();
#pragma unroll 1
    for (int q = 0; q < nch; q++) {
        const int c = A.child_idx[t0 + q], rc = A.sn_nr[c];
        const int *rl = A.rel + A.sn_rowptr[c];
        const double *U = A.Ux + A.Up[c];
        int lo;
        if (HOIST && q < MAXC) {
            lo = lo_c[0];
#pragma unroll
            for (int z = 1; z < MAXC; z++) if (q == z) lo = lo_c[z];
        } else {
            int hi = rc; lo = 0;
            while (lo < hi) { int mid = (lo + hi) >> 1; if (rl[mid] < c0) lo = mid + 1; else hi = mid; }
        }
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
// trailing update C -= (X D) X' for the lower 64 x 64 tiles, X = panel rows x nb; FP64 tensor pipe
__device__ __forceinline__ void dmma884(double &c0, double &c1, double a, double b)
{
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                 : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
// One launch per 32-pivot panel step of a big front: every CTA (front, tile pair) redoes the LDL' of the
// diagonal block (one warp, registers) and the triangular solve of ITS two 64-row tiles from the raw panel,
// then applies the rank-32 update of its 64 x 64 tile on the FP64 tensor pipe.  The factored panel goes to
// Xout (tile pairs with tj == 0 write their rows, pair 0 the diagonal block), the raw panel in Lx stays
// intact for the other CTAs -- so potrf -> trsm -> syrk needs no grid-wide ordering, and the serial chain of
// the tree top pays one launch per panel instead of three.  Arithmetic identical to the three-kernel path.
__global__ void __launch_bounds__(256) k_big_panel_v1(FArgs A, const int *list, int kb)
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


// ---- panel step, second generation -----------------------------------------------------------------
// Same ownership as above (CTA = front x tile pair, every CTA factors the diagonal block itself, no grid-wide
// ordering), but the step is built so that nothing avoidable sits on the serial chain of the tree top:
//  * front descriptors arrive as kernel parameters when the launch has few fronts (the top of the tree): no chain of
//    dependent global loads (list -> sn_start -> Lp -> data) before the first useful load is issued;
//  * warp 0 fetches the diagonal block first; warps 2..7 stage the two raw 64 x 32 row tiles into shared memory by
//    cp.async (k-major, ld 68: the DMMA fragment loads below are bank-conflict free) and every warp fetches its part
//    of the CTA's 64 x 64 tile of the trailing matrix into registers, so both L2 round trips hide behind the LDL';
//  * warp 0 factors the 32 x 32 diagonal block with its rows in registers; the pivot column travels through one
//    shared-memory row per step (one 8-byte store, broadcast 16-byte loads) instead of 2 x 31 shuffles; the column
//    of the NEXT step is published and the next reciprocal (branch-free, rcp_pivot) started before the rest of the
//    current update is issued;
//  * warp 1 follows one step behind (one mbarrier per pivot column) and builds inv(L11) column by column from the
//    same shared rows: no extra time on the chain;
//  * the triangular solve of the row tiles is a product with inv(L11)' on the FP64 tensor pipe (20 of the 32
//    k-steps per 8 x 8 block: the zero blocks of the triangle are skipped), Y = B inv(L11)' feeds the update as
//    X D, X = Y D^-1 is the panel;
//  * rank-32 update of the 64 x 64 tile by DMMA as before;
//  * CTA y == 0 of a front has no tile: it writes the factored diagonal block, with inv(L11) (strictly lower part,
//    transposed) in its unused upper triangle -- the triangular solves use it (k_fwd / k_bwd: no substitution chain
//    in blocked fronts) -- and leaves; the tile pairs start at y == 1.
#define TSP 68         // leading dimension of a k-major 32 x 64 tile (68 mod 16 == 4)
#define LDI 36         // leading dimension of the k-major inverse (36 mod 16 == 4)
#ifndef PANEL_SEQ
#define PANEL_SEQ 0    // 0: inverse one step behind the LDL' (mbarrier per pivot column); 1: after it (named barrier), measured slower
#endif
#ifndef LGI
#define LGI 32         // the same for the inverse warp's loop
#endif
#ifndef LG
#define LG 32          // pivots per pass of the diagonal-block loops (32 = fully unrolled; -DLG=8 / 16 roll them, measured slower)
#endif
#define PANEL_SMEM ((2 * NB * TSP + NB * NB + NB + NB * LDI + NB + NB) * sizeof(double))
#define PANEL_PARAM_FRONTS 16
struct PanelFronts { int count; FrontDesc d[PANEL_PARAM_FRONTS]; };
__device__ __forceinline__ unsigned smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void cp_async8(double *dst_smem, const double *src)
{
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;\n" :: "r"(smem_u32(dst_smem)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;\n" ::: "memory"); }
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int N_> __device__ __forceinline__ void cp_async_wait_group() { asm volatile("cp.async.wait_group %0;\n" :: "n"(N_) : "memory"); }
__device__ __forceinline__ void mbar_init(unsigned long long *bar, unsigned count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" :: "r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive_lane0(unsigned long long *bar, int lane)
{
    asm volatile("{\n.reg .pred P1;\nsetp.eq.s32 P1, %1, 0;\n@P1 mbarrier.arrive.shared::cta.b64 _, [%0];\n}\n" :: "r"(smem_u32(bar)), "r"(lane) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long *bar, unsigned parity)
{
    asm volatile("{\n.reg .pred P1;\nWAIT_LOOP:\nmbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n@P1 bra DONE;\nbra WAIT_LOOP;\nDONE:\n}\n"
                 :: "r"(smem_u32(bar)), "r"(parity) : "memory");
}
// 1 / d to an ulp without the range check (and branch) of __drcp_rn: MUFU.RCP64H seed (2^-20), one cubic step;
// pivots of a regularised SPD matrix are nowhere near the denormal / overflow ranges the check is for
__device__ __forceinline__ double rcp_pivot(double d)
{
    double x;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(x) : "d"(d));
    const double e = fma(-d, x, 1.0);
    const double e2 = fma(e, e, e);
    return fma(x, e2, x);
}
#ifdef PANEL_PROBE
// -DPANEL_PROBE: the first tile CTA of front 0 of every panel launch records SM-clock stamps of its phases and the
// global timer at entry / exit (tools/probe_panel.py)
#define NPROBE 16
__device__ long long g_probe[NPROBE * 4096];
__device__ int g_probe_n;
#define PROBE(i) do { if (probe_on) pr[i] = clock64(); } while (0)
__device__ __forceinline__ long long gtimer() { long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
extern "C" int qpdo_dev_probe_read(long long *out, int cap)
{
    int n = 0;
    cudaDeviceSynchronize();
    cudaMemcpyFromSymbol(&n, g_probe_n, sizeof(int));
    n = n < cap ? n : cap;
    if (n > 4096) n = 4096;
    if (n > 0) cudaMemcpyFromSymbol(out, g_probe, sizeof(long long) * NPROBE * n);
    int zero = 0;
    cudaMemcpyToSymbol(g_probe_n, &zero, sizeof(int));
    return n;
}
#else
#define PROBE(i)
#endif
__global__ void __launch_bounds__(256) k_big_panel(FArgs A, const int *list, int kb, const __grid_constant__ PanelFronts PF)
{
    extern __shared__ __align__(16) double psm[];
    double *Ta = psm, *Tb = Ta + NB * TSP, *Us = Tb + NB * TSP, *LiT = Us + NB * NB + NB, *Rinv = LiT + NB * LDI;
    unsigned long long *bars = (unsigned long long *)(Rinv + NB);
#ifdef PANEL_PROBE
    const bool probe_on = blockIdx.x == 0 && blockIdx.y == 1 && (threadIdx.x & 31) == 0;
    __shared__ long long pr[NPROBE];
    if (probe_on && threadIdx.x == 0) { pr[14] = gtimer(); }
    PROBE(0);
#endif
    // Programmatic dependent launch: the next panel step of the chain may be launched as soon as every CTA of this one has
    // got here (its CTAs become resident on idle SMs only: this kernel is register-limited to one CTA per SM) and does its
    // own prologue -- descriptors, tile indices -- while this step computes; it waits below, before its first read of
    // data this step writes.  Without the launch attribute both instructions do nothing.
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    FrontDesc fd;
    if (PF.count) fd = PF.d[blockIdx.x]; else fd = A.fdesc[list[blockIdx.x]];
    const int nc = fd.nc, nr = fd.nr, N = nc + nr;
    if (kb >= nc) return;
    const int nb = min(NB, nc - kb);
    const int r0 = kb + nb, R = N - r0;
    const int nt = (R + TS - 1) / TS;
    const bool writer = blockIdx.y == 0;                     // diagonal block out, no tile
    const int pair = writer ? 0 : (int)blockIdx.y - 1;
    int ti = (int)((sqrtf(8.0f * pair + 1.0f) - 1.0f) * 0.5f);
    while ((ti + 1) * (ti + 2) / 2 <= pair) ti++;
    while (ti * (ti + 1) / 2 > pair) ti--;
    const int tj = pair - ti * (ti + 1) / 2;
    if (!writer && ti >= nt) return;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int g = lane >> 2, q = lane & 3;
    double *Lfront = A.Lx + fd.Lp, *Ufront = A.Ux + fd.Up;
    const double *Lraw = Lfront + (int64_t)kb * N;                  // raw panel, column k at Lraw + k * N
    double *Xo = A.Xout + fd.Lp + (int64_t)kb * N;
    const int ra = r0 + ti * TS, rb = r0 + tj * TS;
    const bool diag_tile = (ti == tj);
    if (tid == 0) PROBE(1);
    asm volatile("griddepcontrol.wait;" ::: "memory");      // the previous step of the chain (or the extend-add) is complete and visible
    // ---- warp 0: the diagonal block first (lane = row, a[j] = entry (lane, j)); the loads fly while the addresses of
    // the destination tile are worked out, the identity padding beyond nb is applied when they are first used
    double a[NB];
    if (wid == 0) {
        const double *pd = Lraw + kb + min(lane, nb - 1);
#pragma unroll
        for (int j = 0; j < NB; j++) a[j] = pd[(int64_t)min(j, nb - 1) * N];
        mbar_init(bars + lane, 1);
    }
    // ---- destination tile of the update into registers (L2 round trip hidden behind the LDL'): column bases, then rows
    const int wr = (wid & 1) * 32, wc = (wid >> 1) * 16;    // warp tile of the update: 32 rows x 16 cols
    double *colp[2][2];
    double cur[4][2][2];
    if (!writer) {
#pragma unroll
        for (int b = 0; b < 2; b++)
#pragma unroll
            for (int e = 0; e < 2; e++) {
                const int gj = rb + wc + b * 8 + q * 2 + e;
                colp[b][e] = (gj < nc) ? (Lfront + (int64_t)gj * N) : (Ufront + (int64_t)(gj - nc) * nr - nc);
            }
#pragma unroll
        for (int a4 = 0; a4 < 4; a4++)
#pragma unroll
            for (int b = 0; b < 2; b++)
#pragma unroll
                for (int e = 0; e < 2; e++) {
                    const int gi = ra + wr + a4 * 8 + g, gj = rb + wc + b * 8 + q * 2 + e;
                    const bool ok = gi < N && gj <= gi;
                    cur[a4][b][e] = *(ok ? (const double *)(colp[b][e] + gi) : Lraw);   // no select on the loaded value: nothing waits here
                }
    }
    if (tid == 0) PROBE(2);
    __syncthreads();
    if (tid == 0) PROBE(3);
    if (wid == 0) {
        // ---- LDL' of the diagonal block
        if (nb < NB) {                                          // last panel of a front: identity padding
#pragma unroll
            for (int j = 0; j < NB; j++) a[j] = (j < nb && lane < nb) ? a[j] : (lane == j ? 1.0 : 0.0);
        }
        Us[lane] = a[0];
        __syncwarp();
#if !PANEL_SEQ
        mbar_arrive_lane0(bars, lane);
#endif
        double dk = Us[0];
        double r = rcp_pivot(dk);
        if (tid == 0) PROBE(4);
        bool bad = false;
        // NB / LG passes of an LG-step unrolled body over a register WINDOW of the 32 columns: after a pass the LG
        // finished columns leave the window (the writer CTA stores them) and the rest moves down.  Tried because the
        // fully unrolled chain is ~25 KB of straight-line code that runs once per launch; measured, rolling does NOT
        // pay (C3 factor 1.84 / 1.94 / 2.04 ms for LG = 32 / 16 / 8; a second pass over warm code in the same launch is no
        // faster than the first), so LG = NB: one pass.  Window slot w = column kg + w.  NB: the -DPANEL_PROBE stamps
        // themselves slow this chain by ~2x (15.7k cycles probed, ~7k by ncu sampling of the plain build).
#pragma unroll 1
        for (int kg = 0; kg < NB; kg += LG) {
#pragma unroll
            for (int t = 0; t < LG; t++) {
                const int k = kg + t;
                const double *uk = Us + k * NB + kg;            // uk[w]: entry (kg + w, k) of the block
                bad = bad || (k < nb && !(dk > 0.0));
                const double li = a[t] * r;
                double dn = 1.0, rn = 1.0;
                if (lane == k) Rinv[k] = r;
                if (k + 1 < NB) {                               // next pivot column first: published, its reciprocal started
                    a[t + 1] = fma(-li, uk[t + 1], a[t + 1]);
                    Us[(k + 1) * NB + lane] = a[t + 1];
                    __syncwarp();
#if !PANEL_SEQ
                    mbar_arrive_lane0(bars + k + 1, lane);
#endif
                    dn = Us[(k + 1) * NB + k + 1];
                    rn = rcp_pivot(dn);
                }
                // the rest of the update is independent of the next reciprocal: the scheduler interleaves the two
                // (slots beyond column 31 read the padding behind Us and are never used)
                double u[NB];
#pragma unroll
                for (int j = ((t + 2) & ~1); j < NB; j += 2) { const double2 tt = *reinterpret_cast<const double2 *>(uk + j); u[j] = tt.x; u[j + 1] = tt.y; }
#pragma unroll
                for (int j = t + 2; j < NB; j++) a[j] = fma(-li, u[j], a[j]);
                if (lane > k) a[t] = li;
                dk = dn; r = rn;
            }
            if (writer) {
#pragma unroll
                for (int t = 0; t < LG; t++)
                    if (kg + t < nb && lane < nb && lane >= kg + t) Xo[(int64_t)(kg + t) * N + kb + lane] = a[t];
            }
#pragma unroll
            for (int w = 0; w + LG < NB; w++) a[w] = a[w + LG];
#pragma unroll
            for (int w = NB - LG; w < NB; w++) a[w] = 0.0;
        }
#if PANEL_SEQ
        asm volatile("bar.arrive 1, 64;" ::: "memory");       // releases the inverse warp (named barrier 1: warps 0 and 1)
#endif
        if (tid == 0) PROBE(5);
        if (writer && bad && lane == 0) A.sc->not_posdef = 1;
    } else if (wid == 1) {
        // ---- inv(L11), lane = column c: v[i] = inv(L11)(i, c); right-looking over the pivot columns as they appear
        // (step k uses column k and 1 / d_k, both published before the arrival on barrier k + 1)
        // rolled like the LDL': window slot w holds row kg + w of column `lane` of the inverse; a row is final once the
        // steps k < row are done, so the first 8 rows of the window leave it after every pass
        double v[NB];
#pragma unroll
        for (int i = 0; i < NB; i++) v[i] = (i == lane) ? 1.0 : 0.0;
#if PANEL_SEQ
        asm volatile("bar.sync 1, 64;" ::: "memory");         // blocks in hardware until the LDL' warp is through
#endif
#pragma unroll 1
        for (int kg = 0; kg < NB; kg += LGI) {
#pragma unroll
            for (int t = 0; t < LGI; t++) {
                const int k = kg + t;
                if (k + 1 < NB) {
#if !PANEL_SEQ
                    mbar_wait(bars + k + 1, 0);
#endif
                    const double *uk = Us + k * NB + kg;
                    double u[NB];
#pragma unroll
                    for (int j = ((t + 1) & ~1); j < NB; j += 2) { const double2 tt = *reinterpret_cast<const double2 *>(uk + j); u[j] = tt.x; u[j + 1] = tt.y; }
                    const double sk = Rinv[k] * v[t];
#pragma unroll
                    for (int i = t + 1; i < NB; i++) v[i] = fma(-u[i], sk, v[i]);
                }
            }
            if (writer) {
#pragma unroll
                for (int t = 0; t < LGI; t++)
                    if (kg + t >= 1 && kg + t < nb && lane < kg + t) Xo[(int64_t)(kg + t) * N + kb + lane] = v[t];
            } else {
#pragma unroll
                for (int t = 0; t < LGI; t += 2) *reinterpret_cast<double2 *>(LiT + lane * LDI + kg + t) = make_double2(v[t], v[t + 1]);
            }
#pragma unroll
            for (int w = 0; w + LGI < NB; w++) v[w] = v[w + LGI];
#pragma unroll
            for (int w = NB - LGI; w < NB; w++) v[w] = 0.0;
        }
        if (tid == 32) PROBE(6);
    } else if (!writer) {
        // ---- warps 2..7 stage the raw row tiles (k-major) by cp.async while the diagonal block is factored
        const int t192 = tid - 64;
        for (int idx = t192; idx < NB * TS; idx += 192) {
            const int k = idx >> 6, r = idx & 63;
            if (k < nb && ra + r < N) cp_async8(Ta + k * TSP + r, Lraw + (int64_t)k * N + ra + r); else Ta[k * TSP + r] = 0.0;
        }
        if (!diag_tile)
            for (int idx = t192; idx < NB * TS; idx += 192) {
                const int k = idx >> 6, r = idx & 63;
                if (k < nb && rb + r < N) cp_async8(Tb + k * TSP + r, Lraw + (int64_t)k * N + rb + r); else Tb[k * TSP + r] = 0.0;
            }
    }
    if (writer) return;
    cp_async_wait_all();
    if (tid == 64) PROBE(7);
    __syncthreads();
    if (tid == 0) PROBE(8);
    // ---- Y = B inv(L11)' for rows 8 wid .. 8 wid + 7 of both tiles (DMMA, triangle blocks only)
    {
        double afa[NB / 4], afb[NB / 4];
#pragma unroll
        for (int ks = 0; ks < NB / 4; ks++) {
            afa[ks] = Ta[(4 * ks + q) * TSP + 8 * wid + g];
            afb[ks] = diag_tile ? 0.0 : Tb[(4 * ks + q) * TSP + 8 * wid + g];
        }
        __syncwarp();
        double ya[4][2], yb[4][2];
#pragma unroll
        for (int b = 0; b < 4; b++) { ya[b][0] = ya[b][1] = yb[b][0] = yb[b][1] = 0.0; }
#pragma unroll
        for (int b = 0; b < 4; b++)
#pragma unroll
            for (int ks = 0; ks < 2 * b + 2; ks++) {
                const double bf = LiT[(4 * ks + q) * LDI + 8 * b + g];
                dmma884(ya[b][0], ya[b][1], afa[ks], bf);
                if (!diag_tile) dmma884(yb[b][0], yb[b][1], afb[ks], bf);
            }
        const int r = 8 * wid + g;
#pragma unroll
        for (int b = 0; b < 4; b++)
#pragma unroll
            for (int e = 0; e < 2; e++) {
                const int j = 8 * b + 2 * q + e;
                const double rj = Rinv[j];
                const double xa = ya[b][e] * rj;
                Ta[j * TSP + r] = ya[b][e];
                Tb[j * TSP + r] = diag_tile ? xa : yb[b][e] * rj;
                if (tj == 0 && j < nb && ra + r < N) Xo[(int64_t)j * N + ra + r] = xa;
            }
    }
    if (tid == 0) PROBE(9);
    __syncthreads();
    if (tid == 0) PROBE(10);
    // ---- rank-nb update of the 64 x 64 tile (ti, tj): C -= (X D) X'
    double acc[4][2][2];
#pragma unroll
    for (int a4 = 0; a4 < 4; a4++)
#pragma unroll
        for (int b = 0; b < 2; b++) acc[a4][b][0] = acc[a4][b][1] = 0.0;
#pragma unroll
    for (int ks = 0; ks < NB / 4; ks++) {
        double af[4], bf[2];
#pragma unroll
        for (int a4 = 0; a4 < 4; a4++) af[a4] = Ta[(4 * ks + q) * TSP + wr + a4 * 8 + g];
#pragma unroll
        for (int b = 0; b < 2; b++) bf[b] = Tb[(4 * ks + q) * TSP + wc + b * 8 + g];
#pragma unroll
        for (int a4 = 0; a4 < 4; a4++)
#pragma unroll
            for (int b = 0; b < 2; b++) dmma884(acc[a4][b][0], acc[a4][b][1], af[a4], bf[b]);
    }
    if (tid == 0) PROBE(11);
#pragma unroll
    for (int a4 = 0; a4 < 4; a4++)
#pragma unroll
        for (int b = 0; b < 2; b++)
#pragma unroll
            for (int e = 0; e < 2; e++) {
                const int gi = ra + wr + a4 * 8 + g, gj = rb + wc + b * 8 + q * 2 + e;
                if (gi < N && gj <= gi) colp[b][e][gi] = cur[a4][b][e] - acc[a4][b][e];
            }
#ifdef PANEL_PROBE
    if (tid == 0) PROBE(12);
    if (probe_on && tid == 0) { pr[13] = N; pr[15] = gtimer(); int slot = atomicAdd(&g_probe_n, 1); if (slot < 4096) for (int i = 0; i < NPROBE; i++) g_probe[slot * NPROBE + i] = pr[i]; }
#endif
}

// ------------------------------------------------------------------------------- triangular solves
struct SArgs {
    const int *sn_start, *sn_nr, *rel, *rows, *child_ptr, *child_idx;
    const int64_t *sn_rowptr, *Lp;
    const double *Lx, *Xout;   // fronts of order N > bigN read their panel from Xout
    int bigN, panel_v1;        // panel_v1: no inverse diagonal blocks in Xout (round-1 panel step)
    int stage_rows;            // k_fwd<32 | 256>: row stride of the cp.async stage (set per launch)
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
    // NT <= 256: the panel entries a thread needs for the row update of a block are staged by cp.async one block ahead
    // (the first block behind the extend-add of the children): plain loads end up paired with their fmas by ptxas, a
    // chain of L2 round trips per block.  RP rows per thread cover the front (N <= 64 for one warp, N <= 256 for 256).
    constexpr bool STG = NT <= 256;
    constexpr int RP = NT == 32 ? 2 : 1;
    const int SR = A.stage_rows;                        // rows of the stage = order of the largest front of the launch (<= RP * NT)
    double *stage = vec + ((N + NB + 1) & ~1);         // [NB][SR], row i of block kb_ at column i - kb_ - nb_
    auto prefetch = [&](int kb_) {
        if (STG && kb_ < nc) {
            const int nb_ = min(NB, nc - kb_);
#pragma unroll
            for (int r = 0; r < RP; r++) {
                const int o = tid + r * NT, i = kb_ + nb_ + o;
                if (i < N) for (int k = 0; k < nb_; k++) cp_async8(stage + k * SR + o, L + (int64_t)(kb_ + k) * N + i);
            }
        }
    };
    prefetch(0);
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
            if (N > A.bigN && !A.panel_v1) {
                // blocked fronts carry inv(L11) in the upper triangle of the diagonal block (k_big_panel):
                // y = inv(L11) t as a 32 x 32 product, four partial sums, no substitution chain
                double p0 = 0.0, p1 = 0.0, p2 = 0.0, p3 = 0.0;
                for (int k = 0; k < nb; k += 4) {
                    if (k < lane) p0 = fma(D[k][lane], vec[kb + k], p0);
                    if (k + 1 < lane) p1 = fma(D[k + 1][lane], vec[kb + k + 1], p1);
                    if (k + 2 < lane) p2 = fma(D[k + 2][lane], vec[kb + k + 2], p2);
                    if (k + 3 < lane) p3 = fma(D[k + 3][lane], vec[kb + k + 3], p3);
                }
                v += (p0 + p1) + (p2 + p3);
            } else {
                for (int k = 0; k < nb; k++) {
                    const double yk = __shfl_sync(0xffffffffu, v, k);
                    if (lane > k && lane < nb) v = fma(-D[lane][k], yk, v);
                }
            }
            // y feeds the updates below; z = D^{-1} y is what the backward pass starts from
            if (lane < nb) { yb[lane] = v; A.xs[f + kb + lane] = v / D[lane][lane]; }
        }
        __syncthreads();
        // rows below the block: one row per thread, its NB panel entries fetched as two batches of 16
        // independent loads (the panel streams from HBM/L2 once; latency, not bandwidth, bounds one CTA)
        if (STG) cp_async_wait_all();
        for (int i = kb + nb + tid, r = 0; i < N; i += NT, r++) {
            const double *Li = L + (int64_t)kb * N + i;
            double acc = 0.0;
            if (STG && r < RP) {
                for (int k = 0; k < nb; k++) acc = fma(stage[k * SR + tid + r * NT], yb[k], acc);
            } else if (nb == NB) {
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
        prefetch(kb + NB);
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
    const int nblk = (nc + NB - 1) / NB;
    // NT <= 256: the rows below a block (all its columns) are staged by cp.async one block ahead, see k_fwd
    constexpr bool STG = NT <= 256;
    double *stage = vec + ((N + 1) & ~1);                       // [NB][N], entry (i, kb + k) at stage[k * N + i]
    auto prefetch = [&](int b_) {
        if (STG && b_ >= 0) {
            const int kb_ = b_ * NB, nb_ = min(NB, nc - kb_), R = N - kb_ - nb_;
            for (int idx = tid; idx < nb_ * R; idx += NT) { const int k = idx / R, i = kb_ + nb_ + (idx - k * R); cp_async8(stage + k * N + i, L + (int64_t)(kb_ + k) * N + i); }
        }
    };
    prefetch(nblk - 1);
    for (int i = tid; i < N; i += NT) vec[i] = (i < nc) ? A.xs[f + i] : A.xs[rows[i - nc]];
    __syncthreads();
    for (int b = nblk - 1; b >= 0; b--) {
        const int kb = b * NB, nb = min(NB, nc - kb);
        for (int idx = tid; idx < nb * nb; idx += NT) { int j = idx / nb, i = idx - j * nb; D[i][j] = L[(int64_t)(kb + j) * N + kb + i]; }
        if (STG) { cp_async_wait_all(); __syncthreads(); }      // a warp sums columns that other threads staged
        // t_k = vec[k] - sum_{i >= kb+nb} L[i,k] vec[i], one warp per column
        for (int k = wid; k < nb; k += NW) {
            const double *Lk = L + (int64_t)(kb + k) * N;
            double acc = 0.0;
            int i = kb + nb + lane;
            if (STG) { for (; i < N; i += 32) acc = fma(stage[k * N + i], vec[i], acc); }
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
        prefetch(b - 1);                                        // every warp is through with the stage
        if (tid < 32) {
            double v = (lane < nb) ? vec[kb + lane] : 0.0;
            if (N > A.bigN && !A.panel_v1) {
                // x = inv(L11)' v: x_k = v_k + sum_{i > k} inv(L11)(i, k) v_i, inv(L11)(i, k) stored at (k, i)
                double p0 = 0.0, p1 = 0.0, p2 = 0.0, p3 = 0.0;
                for (int i = 0; i < nb; i += 4) {
                    if (i > lane && i < nb) p0 = fma(D[lane][i], vec[kb + i], p0);
                    if (i + 1 > lane && i + 1 < nb) p1 = fma(D[lane][i + 1], vec[kb + i + 1], p1);
                    if (i + 2 > lane && i + 2 < nb) p2 = fma(D[lane][i + 2], vec[kb + i + 2], p2);
                    if (i + 3 > lane && i + 3 < nb) p3 = fma(D[lane][i + 3], vec[kb + i + 3], p3);
                }
                __syncwarp();
                v += (p0 + p1) + (p2 + p3);
            } else {
                for (int k = nb - 1; k >= 0; k--) {
                    const double xk = __shfl_sync(0xffffffffu, v, k);
                    if (lane < k) v = fma(-D[k][lane], xk, v);
                }
            }
            if (lane < nb) { vec[kb + lane] = v; A.xs[f + kb + lane] = v; }
        }
        __syncthreads();
    }
}

// ---- cluster solves of the big fronts ----------------------------------------------------------------------
// Fronts wider than SOLVE_BIG_N rows sit at the top of the tree, a handful per level, and their solves are a chain of
// 32-column steps; one CTA per front keeps one SM's worth of loads in flight (~64 KB) and pays every L2 round trip of
// a step in sequence.  Here a thread-block CLUSTER of SC CTAs solves one front: forward, the rows are dealt to the
// CTAs in tiles of 32 (tile t -> CTA t mod SC); the CTA that owns the rows of the current diagonal block applies
// inv(L11) (stored by k_big_panel), writes the block solution into every CTA's shared memory (distributed shared
// memory), one cluster barrier, every CTA updates its own rows.  Backward, the 32 columns of a block are dealt to the
// 32 warps of the cluster (the right-hand side is replicated), every warp sends its column sum to all CTAs, one
// cluster barrier, every CTA applies inv(L11)' redundantly.  Per entry the operations and their order are those of
// k_fwd / k_bwd: same bits.
#define SC 8           // CTAs per cluster
#define ST 128         // threads per CTA
__global__ void __cluster_dims__(SC, 1, 1) __launch_bounds__(ST) k_fwd_cl(SArgs A, const int *list)
{
    cg::cluster_group cl = cg::this_cluster();
    const int c = (int)cl.block_rank();
    extern __shared__ double own[];                    // this CTA's rows: local tile lt = tile / SC, entry lt * 32 + (i & 31)
    __shared__ double D[NB][NB + 1];
    __shared__ double ybuf[2][NB];
    const int s = list[blockIdx.x / SC];
    const int f = A.sn_start[s], nc = A.sn_start[s + 1] - f, nr = A.sn_nr[s], N = nc + nr;
    const double *L = A.Xout + A.Lp[s];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5, NW = ST / 32;
    const int ntile = (N + 31) >> 5;
    const int nown = ntile > c ? (ntile - c + SC - 1) / SC : 0;
    for (int lt = wid; lt < nown; lt += NW) { const int i = ((lt * SC + c) << 5) + lane; own[lt * 32 + lane] = (i < nc) ? A.xs[f + i] : 0.0; }
    __syncthreads();
    for (int t = A.child_ptr[s]; t < A.child_ptr[s + 1]; t++) {
        const int ch = A.child_idx[t], rc = A.sn_nr[ch];
        const int *rl = A.rel + A.sn_rowptr[ch];
        const double *u = A.uv + A.sn_rowptr[ch];
        for (int idx = tid; idx < rc; idx += ST) {
            const int r = rl[idx];
            if (((r >> 5) % SC) == c) own[((r >> 5) / SC) * 32 + (r & 31)] += u[idx];
        }
        __syncthreads();
    }
    const int nblk = (nc + NB - 1) / NB;
    double *stage = own + (nown + 1) * 32;             // [NB][ST]: panel entries of the next step for this thread's first row tile
    auto prefetchD = [&](int b) {                      // diagonal block b (with inv(L11) in its upper triangle) into D
        const int kb = b * NB, nb = min(NB, nc - kb);
        for (int idx = tid; idx < nb * nb; idx += ST) { const int j = idx / nb, i = idx - j * nb; cp_async8(&D[i][j], L + (int64_t)(kb + j) * N + kb + i); }
    };
    // the loads of a step do not depend on its data: they are issued as cp.async one step ahead (registers would not do:
    // ptxas pairs every load with its fma, and the waiting fma then holds back the loads behind it)
    const int i0 = ((wid * SC + c) << 5) + lane;       // row of this thread's first tile (lt = wid)
    auto prefetchL = [&](int b) {
        if (b < nblk) {
            const int kb = b * NB, nb = min(NB, nc - kb);
            if (wid < nown && i0 >= kb + nb && i0 < N)
                for (int k = 0; k < nb; k++) cp_async8(stage + k * ST + tid, L + (int64_t)(kb + k) * N + i0);
        }
        cp_async_commit();
    };
    if (c < nblk) prefetchD(c);
    cp_async_commit();
    prefetchL(0);
    for (int b = 0; b < nblk; b++) {
        const int kb = b * NB, nb = min(NB, nc - kb);
        // groups in flight here: ..., D (older), L(b) (newest).  Every CTA commits the same sequence of groups.
        if (c == b % SC) {                             // this CTA owns rows kb .. kb + 31
            cp_async_wait_group<1>();
            __syncthreads();                           // D landed; the other warps' updates of these rows are visible
            if (wid == 0) {
                const double *tb = own + (b / SC) * 32;
                double v = (lane < nb) ? tb[lane] : 0.0;
                double p0 = 0.0, p1 = 0.0, p2 = 0.0, p3 = 0.0;
                for (int k = 0; k < nb; k += 4) {
                    if (k < lane) p0 = fma(D[k][lane], tb[k], p0);
                    if (k + 1 < lane) p1 = fma(D[k + 1][lane], tb[k + 1], p1);
                    if (k + 2 < lane) p2 = fma(D[k + 2][lane], tb[k + 2], p2);
                    if (k + 3 < lane) p3 = fma(D[k + 3][lane], tb[k + 3], p3);
                }
                v += (p0 + p1) + (p2 + p3);
                for (int r = 0; r < SC; r++) cl.map_shared_rank(&ybuf[b & 1][0], r)[lane] = (lane < nb) ? v : 0.0;
                if (lane < nb) A.xs[f + kb + lane] = v / D[lane][lane];
            }
            __syncthreads();                           // D is free again
            if (b + SC < nblk) prefetchD(b + SC);
        }
        cp_async_commit();                             // the D group of this step (empty unless this CTA just prefetched)
        cl.sync();
        cp_async_wait_group<1>();                      // L(b) landed (each thread reads only what it copied itself)
        const double *yb = ybuf[b & 1];
        for (int lt = wid; lt < nown; lt += NW) {
            const int i = ((lt * SC + c) << 5) + lane;
            if (i >= kb + nb && i < N) {
                double acc = 0.0;
                if (lt == wid) {
                    for (int k = 0; k < nb; k++) acc = fma(stage[k * ST + tid], yb[k], acc);
                } else {                               // fronts beyond SC * ST rows: further tiles straight from global memory
                    const double *Li = L + (int64_t)kb * N + i;
                    for (int k = 0; k < nb; k++) acc = fma(Li[(int64_t)k * N], yb[k], acc);
                }
                own[lt * 32 + lane] -= acc;
            }
        }
        prefetchL(b + 1);
    }
    double *u = A.uv + A.sn_rowptr[s];
    for (int lt = wid; lt < nown; lt += NW) { const int i = ((lt * SC + c) << 5) + lane; if (i >= nc && i < N) u[i - nc] = own[lt * 32 + lane]; }
    cl.sync();                                         // no CTA leaves while a peer may still write into its shared memory
}
__global__ void __cluster_dims__(SC, 1, 1) __launch_bounds__(ST) k_bwd_cl(SArgs A, const int *list)
{
    cg::cluster_group cl = cg::this_cluster();
    const int c = (int)cl.block_rank();
    extern __shared__ double vec[];                    // the front's N values, replicated in every CTA
    __shared__ double Dd[2][NB][NB + 1];
    __shared__ double tbuf[2][NB];
    const int s = list[blockIdx.x / SC];
    const int f = A.sn_start[s], nc = A.sn_start[s + 1] - f, nr = A.sn_nr[s], N = nc + nr;
    const double *L = A.Xout + A.Lp[s];
    const int *rows = A.rows + A.sn_rowptr[s];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int g = c * (ST / 32) + wid;                 // warp number in the cluster = column of the block it sums
    for (int i = tid; i < N; i += ST) vec[i] = (i < nc) ? A.xs[f + i] : A.xs[rows[i - nc]];
    const int nblk = (nc + NB - 1) / NB;
    double *stage = vec + N + (N & 1) + wid * (size_t)(N + (N & 1));   // this warp's column of the next step (rows below the block)
    auto prefetchD = [&](int b) {
        const int kb = b * NB, nb = min(NB, nc - kb);
        double (*D)[NB + 1] = Dd[b & 1];
        for (int idx = tid; idx < nb * nb; idx += ST) { const int j = idx / nb, i = idx - j * nb; cp_async8(&D[i][j], L + (int64_t)(kb + j) * N + kb + i); }
    };
    auto prefetchL = [&](int b) {                      // column kb + g, rows kb + nb .. N - 1, as cp.async (see k_fwd_cl)
        const int kb = b * NB, nb = min(NB, nc - kb);
        if (g < nb) {
            const double *Lk = L + (int64_t)(kb + g) * N;
            for (int i = kb + nb + lane; i < N; i += 32) cp_async8(stage + i, Lk + i);
        }
    };
    prefetchD(nblk - 1);
    prefetchL(nblk - 1);
    __syncthreads();
    for (int b = nblk - 1; b >= 0; b--) {
        const int kb = b * NB, nb = min(NB, nc - kb);
        cp_async_wait_all();                           // this thread's part of D(b) and of its column; the cluster barrier below publishes D
        if (b > 0) prefetchD(b - 1);
        // t_k = vec[kb + k] - sum_{i >= kb + nb} L[i, kb + k] vec[i]: one warp of the cluster per column
        if (g < nb) {
            double acc = 0.0;
            for (int i = kb + nb + lane; i < N; i += 32) acc = fma(stage[i], vec[i], acc);
            for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
            const double tk = vec[kb + g] - acc;
            if (lane < SC) cl.map_shared_rank(&tbuf[b & 1][0], lane)[g] = tk;
        }
        __syncwarp();
        if (b > 0) prefetchL(b - 1);                   // lands behind the barrier and the block solve
        cl.sync();
        if (wid == 0) {                                // every CTA applies inv(L11)' redundantly
            const double (*D)[NB + 1] = Dd[b & 1];
            const double *tb = tbuf[b & 1];
            double v = (lane < nb) ? tb[lane] : 0.0;
            double p0 = 0.0, p1 = 0.0, p2 = 0.0, p3 = 0.0;
            for (int i = 0; i < nb; i += 4) {
                if (i > lane && i < nb) p0 = fma(D[lane][i], tb[i], p0);
                if (i + 1 > lane && i + 1 < nb) p1 = fma(D[lane][i + 1], tb[i + 1], p1);
                if (i + 2 > lane && i + 2 < nb) p2 = fma(D[lane][i + 2], tb[i + 2], p2);
                if (i + 3 > lane && i + 3 < nb) p3 = fma(D[lane][i + 3], tb[i + 3], p3);
            }
            v += (p0 + p1) + (p2 + p3);
            if (lane < nb) { vec[kb + lane] = v; if (c == 0) A.xs[f + kb + lane] = v; }
        }
        __syncthreads();                               // x of this block visible to the CTA's warps
    }
    cl.sync();
}

// ---- singleton leaf fronts -----------------------------------------------------------------------------------------
// A column of K with no off-diagonal entry (Lasso-type problems outside the active rows: nearly all of n) is a front of
// one pivot with no rows below and no children: its factorization is the sign check of the pivot, its forward solve one
// division, its backward solve nothing.  They form the tail of level 0 (fronts are sorted by decreasing order) and are
// handled element-wise, one thread per front, instead of one CTA each (same operations: xs = xs / d as in k_fwd).
__global__ void k_single_check(FArgs A, const int *list, int count)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < count && !(A.Lx[A.fdesc[list[i]].Lp] > 0.0)) A.sc->not_posdef = 1;
}
__global__ void k_single_fwd(SArgs A, const int *list, int count)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < count) { const int s = list[i], f = A.sn_start[s]; A.xs[f] = A.xs[f] / A.Lx[A.Lp[s]]; }
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
    A.fdesc = (const FrontDesc *)F.fdesc;
    return A;
}
static SArgs sargs(qpdo_dev *d)
{
    Factor &F = d->F;
    SArgs A;
    A.sn_start = F.sn_start; A.sn_nr = F.sn_nr; A.rel = F.rel; A.rows = F.rows; A.child_ptr = F.child_ptr;
    A.child_idx = F.child_idx; A.sn_rowptr = F.sn_rowptr; A.Lp = F.Lp; A.Lx = F.Lx; A.Xout = F.Xout; A.bigN = BUCKET_N[3]; A.panel_v1 = F.panel_v1 ? 1 : 0; A.stage_rows = 0; A.xs = F.xs; A.uv = F.uv;
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

// Device arrays that depend on the pattern come from a slot pool: an analysis asks for them in the same order every time, a
// slot is reused when it is large enough (zero-filled like a fresh allocation) and replaced with 25 % headroom otherwise.
// A support-following pattern is re-analysed a dozen times along a regularisation path; freeing and allocating ~27 arrays
// each time was a third of the cost of an analysis.  The buffers stay registered in d->allocs until the context is destroyed.
static int pool_get(qpdo_dev *d, void **out, size_t bytes)
{
    Factor &F = d->F;
    if (!bytes) bytes = 8;
    const size_t slot = F.pool_next++;
    if (slot >= F.pool.size()) F.pool.push_back(std::make_pair((void *)nullptr, (size_t)0));
    if (F.pool[slot].second < bytes) {
        if (F.pool[slot].first) {
            void *q = F.pool[slot].first;
            for (size_t i = 0; i < d->allocs.size(); i++) if (d->allocs[i] == q) { d->allocs[i] = d->allocs.back(); d->allocs.pop_back(); break; }
            cudaFree(q);
        }
        const size_t cap = bytes + bytes / 4;
        void *v = nullptr;
        cudaError_t e = cudaMalloc(&v, cap);
        if (e != cudaSuccess) { F.pool[slot] = std::make_pair((void *)nullptr, (size_t)0); qpdo_set_error("cudaMalloc(%zu): %s", cap, cudaGetErrorString(e)); return -1; }
        d->allocs.push_back(v);
        F.pool[slot] = std::make_pair(v, cap);
    }
    CUDA_OK(cudaMemsetAsync(F.pool[slot].first, 0, bytes, d->st));
    *out = F.pool[slot].first;
    return 0;
}
template <class T> static int f_alloc(qpdo_dev *d, T **p, size_t count) { void *v; if (pool_get(d, &v, sizeof(T) * count)) return -1; *p = (T *)v; return 0; }
template <class T> static int f_upload(qpdo_dev *d, T **p, const T *h, size_t count)
{
    if (f_alloc(d, p, count)) return -1;
    if (count) CUDA_OK(cudaMemcpyAsync(*p, h, sizeof(T) * count, cudaMemcpyHostToDevice, d->st));
    return 0;
}
static void factor_release(qpdo_dev *d)
{
    Factor &F = d->F;
    cudaStreamSynchronize(d->st);
    if (F.g_factor) cudaGraphExecDestroy(F.g_factor);
    if (F.g_solve) cudaGraphExecDestroy(F.g_solve);
    F.g_factor = F.g_solve = nullptr;
    F.pool_next = 0;                                      // the next analysis reuses the slots
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
int factor_noalloc_setup(qpdo_dev *d);                 // address of the static scratch of qpdo_dev_solve_noalloc, once per device

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
    if (factor_noalloc_setup(d)) return -1;
    for (int b = 0; b < 4; b++) {
        if (cudaStreamCreateWithFlags(&F.aux[b], cudaStreamNonBlocking) != cudaSuccess || cudaEventCreateWithFlags(&F.ev_join[b], cudaEventDisableTiming) != cudaSuccess) {
            qpdo_set_error("factor_setup: stream / event creation failed"); return -1;
        }
    }
    if (cudaEventCreateWithFlags(&F.ev_fork, cudaEventDisableTiming) != cudaSuccess) { qpdo_set_error("factor_setup: event creation failed"); return -1; }
    { const char *cv = getenv("QPDO_B200_CONCURRENT"); F.concurrent = !(cv && !strcmp(cv, "0")); }
    { const char *pv = getenv("QPDO_B200_PDL"); F.pdl = !(pv && !strcmp(pv, "0")); }                           // 0: plain stream order between panel steps (A/B)
    { const char *hv = getenv("QPDO_B200_PATTERN_HEADROOM"); F.headroom = !(hv && !strcmp(hv, "0")); }
    { const char *nv = getenv("QPDO_B200_SINGLETONS"); F.no_singletons = nv && !strcmp(nv, "0"); }             // 0: singleton fronts through the per-front kernels (A/B)
    { const char *sv = getenv("QPDO_B200_SOLVE"); F.solve_cluster = !(sv && !strcmp(sv, "cta")); }   // "cta": one CTA per big front (A/B)
    cudaFuncSetAttribute(k_bwd_cl, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    cudaFuncSetAttribute(k_fwd_cl, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
    cudaFuncSetAttribute(k_fwd<256>, cudaFuncAttributeMaxDynamicSharedMemorySize, 72 * 1024);
    cudaFuncSetAttribute(k_bwd<256>, cudaFuncAttributeMaxDynamicSharedMemorySize, 72 * 1024);
    cudaFuncSetAttribute(k_front_smem<256>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    cudaFuncSetAttribute(k_front_smem<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
    cudaFuncSetAttribute(k_big_panel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)PANEL_SMEM);
    { const char *pv = getenv("QPDO_B200_PANEL"); F.panel_v1 = pv && !strcmp(pv, "v1"); }   // A/B against the round-1 panel step
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
    if (rowmask) F.hP.assign(rowmask, rowmask + m); else F.hP.clear();
    char err[256];
    auto t_sym0 = std::chrono::steady_clock::now();
    qpdo_symbolic *S = qpdo_symbolic_analyze(n, mr, F.hQp.data(), F.hQi.data(), pAp, pAi, F.ordering, err);
    if (!S) { qpdo_set_error("symbolic analysis failed: %s", err); return -1; }
    const double sym_ms = std::chrono::duration<double>(std::chrono::steady_clock::now() - t_sym0).count() * 1e3;
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
    {
        std::vector<FrontDesc> fdh((size_t)ns);
        for (int t = 0; t < ns; t++) { fdh[t].Lp = S->Lp[t]; fdh[t].Up = S->Up[t]; fdh[t].nc = S->sn_start[t + 1] - S->sn_start[t]; fdh[t].nr = S->sn_nr[t]; }
        FrontDesc *fdd = nullptr;
        if (f_upload(d, &fdd, (const FrontDesc *)fdh.data(), (size_t)ns)) return -1;
        F.fdesc = fdd;
    }
    // launch plan: per level, fronts are sorted by decreasing order N
    F.plan.assign(S->nlevel, std::vector<FrontLaunch>());
    F.nsing = 0;
    for (int lv = 0; lv < S->nlevel; lv++) {
        int b = S->level_ptr[lv], e = S->level_ptr[lv + 1];
        auto ordN = [&](int pos) { int s = S->level_sn[pos]; return S->sn_start[s + 1] - S->sn_start[s] + S->sn_nr[s]; };
        if (lv == 0 && !F.no_singletons) {                     // singleton leaf fronts: the tail of level 0, element-wise kernels
            while (e > b && ordN(e - 1) == 1) e--;
            F.nsing = S->level_ptr[1] - e;
            if (F.nsing) { FrontLaunch L; L.first = e; L.count = F.nsing; L.kind = 3; L.maxN = 1; L.maxnc = 1; F.plan[lv].push_back(L); }
        }
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
    const double all_s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    d->analyze_seconds += all_s;
    if (getenv("QPDO_B200_TRACE_ANALYZE"))                // one line per (re-)analysis: pattern rows, supernodes, host symbolic ms, total ms
        fprintf(stderr, "qpdo_b200: analysis %d: %d pattern rows%s, %d supernodes, symbolic %.1f ms, with release / upload %.1f ms\n",
                F.analyses, mr, rowmask ? "" : " (superset)", ns, sym_ms, all_s * 1e3);
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
    // Headroom (QPDO_B200_PATTERN_HEADROOM=0 switches it off): a pattern that is exactly the support is invalid again as soon as
    // the next row enters -- along a regularisation path that is every other Newton step (C4: 25 analyses of ~46 ms = 3/4 of
    // the path).  So the new row set also keeps the rows of the previous one, and takes in the inactive rows nearest to a
    // bound (smallest min(w - l, u - w)), half as many as the support: the rows most likely to enter next.  Rows with zero
    // weight only add structural zeros to K; the 4x rule above sheds them again.
    if (F.headroom) {
        const int budget = 2 * support + 2048;                 // well inside the 4 * support + 4096 of the shrink rule (a budget of
                                                               // 3 x with as many near rows as the support: 9 analyses instead of 11 on C4,
                                                               // but each on a larger pattern: path 1.07 s against 0.78 s)
        int have = support;
        if (!F.hP.empty())
            for (int i = 0; i < m && have < budget; i++) if (F.hP[i] && !mask[i]) { mask[i] = 1; have++; }
        const int want = std::min(budget - have, std::max(32, support / 2));
        if (want > 0) {
            std::vector<double> w((size_t)m), lo((size_t)m), up((size_t)m);
            CUDA_OK(cudaMemcpyAsync(w.data(), d->w, sizeof(double) * m, cudaMemcpyDeviceToHost, d->st));
            CUDA_OK(cudaMemcpyAsync(lo.data(), d->l, sizeof(double) * m, cudaMemcpyDeviceToHost, d->st));
            CUDA_OK(cudaMemcpyAsync(up.data(), d->u, sizeof(double) * m, cudaMemcpyDeviceToHost, d->st));
            CUDA_OK(cudaStreamSynchronize(d->st));
            std::vector<std::pair<double, int>> cand;
            for (int i = 0; i < m; i++)
                if (!mask[i]) { const double sl = std::min(w[i] - lo[i], up[i] - w[i]); if (sl == sl && sl < 1e300) cand.push_back(std::make_pair(sl, i)); }
            const size_t take = std::min(cand.size(), (size_t)want);
            if (take) {
                std::nth_element(cand.begin(), cand.begin() + (take - 1), cand.end());
                for (size_t t = 0; t < take; t++) mask[cand[t].second] = 1;
            }
        }
    }
    return factor_analyze(d, mask.data());
}

int factor_refresh_products(qpdo_dev *d)
{
    Factor &F = d->F;
    if (!F.S) return 0;                                   // dynamic pattern: nothing analysed yet
    if (F.S->ntriple) k_products<<<cdiv(F.S->ntriple, 256), 256, 0, d->st>>>(F.S->ntriple, F.Kt_a1, F.Kt_a2, d->Ax, F.Kt_val);
    return 0;
}

// the launches of one level are independent of each other: branch b > 0 goes to its own stream (a parallel branch of
// the captured graph), the level ends with a join on the main stream
static void fork_level(qpdo_dev *d, int nbranch, cudaStream_t *sv)
{
    Factor &F = d->F;
    for (int b = 0; b < nbranch; b++) sv[b] = d->st;
    if (nbranch <= 1 || !F.concurrent) return;
    cudaEventRecord(F.ev_fork, d->st);
    for (int b = 1; b < nbranch && b <= 4; b++) { sv[b] = F.aux[b - 1]; cudaStreamWaitEvent(sv[b], F.ev_fork, 0); }
}
static void join_level(qpdo_dev *d, int nbranch, cudaStream_t *sv)
{
    Factor &F = d->F;
    if (nbranch <= 1 || !F.concurrent) return;
    for (int b = 1; b < nbranch && b <= 4; b++) { cudaEventRecord(F.ev_join[b - 1], sv[b]); cudaStreamWaitEvent(d->st, F.ev_join[b - 1], 0); }
}
static long issue_factor(qpdo_dev *d)
{
    Factor &F = d->F;
    const qpdo_symbolic *S = F.S;
    FArgs A = fargs(d);
    long launches = 0;
    for (int lv = 0; lv < S->nlevel; lv++) {
        cudaStream_t sv[8];
        const int nbranch = (int)F.plan[lv].size();
        fork_level(d, nbranch, sv);
        int br = 0;
        for (const FrontLaunch &L : F.plan[lv]) {
            const int *list = F.level_sn + L.first;
            cudaStream_t st = sv[br++];
            if (L.kind == 3) { k_single_check<<<cdiv(L.count, 256), 256, 0, st>>>(A, list, L.count); launches++; }
            else if (L.kind == 2) {
                {
                    cudaLaunchConfig_t cfg = {};
                    cfg.gridDim = dim3(L.count, cdiv(L.maxN, CW)); cfg.blockDim = dim3(256); cfg.dynamicSmemBytes = 0; cfg.stream = st;
                    cudaLaunchAttribute at[1];
                    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization; at[0].val.programmaticStreamSerializationAllowed = 1;
                    cfg.attrs = at; cfg.numAttrs = F.pdl ? 1 : 0;
                    if (L.count <= 4) cudaLaunchKernelEx(&cfg, k_big_assemble<true>, A, list); else cudaLaunchKernelEx(&cfg, k_big_assemble<false>, A, list);
                    launches++;
                }
                for (int kb = 0; kb < L.maxnc; kb += NB) {
                    const int R = L.maxN - kb - 1, nt = R > 0 ? cdiv(R, TS) : 0;
                    if (F.panel_v1) k_big_panel_v1<<<dim3(L.count, std::max(1, nt * (nt + 1) / 2)), 256, 0, st>>>(A, list, kb);
                    else {
                        PanelFronts PF; PF.count = 0;
                        if (L.count <= PANEL_PARAM_FRONTS) {        // few fronts: descriptors travel as kernel parameters
                            PF.count = L.count;
                            for (int t = 0; t < L.count; t++) {
                                const int sn = S->level_sn[L.first + t];
                                PF.d[t].Lp = S->Lp[sn]; PF.d[t].Up = S->Up[sn]; PF.d[t].nc = S->sn_start[sn + 1] - S->sn_start[sn]; PF.d[t].nr = S->sn_nr[sn];
                            }
                        }
                        cudaLaunchConfig_t cfg = {};
                        cfg.gridDim = dim3(L.count, 1 + nt * (nt + 1) / 2); cfg.blockDim = dim3(256); cfg.dynamicSmemBytes = PANEL_SMEM; cfg.stream = st;
                        cudaLaunchAttribute at[1];
                        at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization; at[0].val.programmaticStreamSerializationAllowed = 1;
                        cfg.attrs = at; cfg.numAttrs = F.pdl ? 1 : 0;
                        cudaLaunchKernelEx(&cfg, k_big_panel, A, list, kb, PF);
                    }
                    launches++;
                }
            } else {
                size_t smem = sizeof(double) * ((size_t)(L.maxN | 1) * L.maxN + 8 * (size_t)L.maxN);
                if (L.maxnc == 0) k_front_smem<64><<<L.count, 64, smem, st>>>(A, list);
                else if (L.maxnc == 1) k_front_smem<128><<<L.count, 128, smem, st>>>(A, list);
                else k_front_smem<256><<<L.count, 256, smem, st>>>(A, list);
                launches++;
            }
        }
        join_level(d, nbranch, sv);
    }
    return launches;
}
static long issue_solve_args(qpdo_dev *d, const SArgs &A)
{
    Factor &F = d->F;
    const qpdo_symbolic *S = F.S;
    long launches = 0;
    auto ordN = [&](int pos) { int s = S->level_sn[pos]; return S->sn_start[s + 1] - S->sn_start[s] + S->sn_nr[s]; };
    for (int pass = 0; pass < 2; pass++)
        for (int t = 0; t < S->nlevel; t++) {
            int lv = pass == 0 ? t : S->nlevel - 1 - t;
            int b = S->level_ptr[lv], e = S->level_ptr[lv + 1], q = b;
            const int nsing = lv == 0 ? F.nsing : 0;              // singleton leaf fronts: one division forward, nothing backward
            e -= nsing;
            while (q < e && ordN(q) > SOLVE_WARP_N) q++;
            // fronts wider than SOLVE_BIG_N rows get 1024 threads (one panel row / one panel column per
            // thread / warp), the rest of the CTA-solved fronts 256, fronts up to SOLVE_WARP_N one warp: up to three
            // independent launches per level, each on its own branch
            int qb = b;
            while (qb < q && ordN(qb) > SOLVE_BIG_N) qb++;
            const int nbranch = (qb > b) + (q > qb) + (e > q) + (nsing && pass == 0);
            cudaStream_t sv[4];
            fork_level(d, nbranch, sv);
            int br = 0;
            if (qb > b) {
                cudaStream_t st = sv[br++];
                size_t smem = sizeof(double) * (ordN(b) + NB);
                if (F.solve_cluster && !F.panel_v1 && ordN(b) <= 4096) {      // (shared-memory budget of k_bwd_cl: 5 N doubles)
                    const int Nmax = ordN(b);
                    if (pass == 0) k_fwd_cl<<<(qb - b) * SC, ST, sizeof(double) * (32 * (size_t)(cdiv(cdiv(Nmax, 32), SC) + 2) + NB * ST), st>>>(A, F.level_sn + b);
                    else k_bwd_cl<<<(qb - b) * SC, ST, sizeof(double) * (size_t)(Nmax + 2) * (1 + ST / 32), st>>>(A, F.level_sn + b);
                } else if (pass == 0) k_fwd<1024><<<qb - b, 1024, smem, st>>>(A, F.level_sn + b);
                else k_bwd<1024><<<qb - b, 1024, smem, st>>>(A, F.level_sn + b);
                launches++;
            }
            if (q > qb) {
                cudaStream_t st = sv[br++];
                const size_t Nm = (size_t)ordN(qb);              // + the cp.async stage: NB x rows of the largest front
                SArgs As = A; As.stage_rows = (int)((Nm + 1) & ~(size_t)1);
                if (pass == 0) k_fwd<256><<<q - qb, 256, sizeof(double) * (Nm + NB + 2 + NB * (size_t)As.stage_rows), st>>>(As, F.level_sn + qb);
                else k_bwd<256><<<q - qb, 256, sizeof(double) * (Nm + 2 + NB * Nm), st>>>(A, F.level_sn + qb);
                launches++;
            }
            if (e > q) {
                cudaStream_t st = sv[br++];
                const size_t Nm = (size_t)ordN(q);
                SArgs As = A; As.stage_rows = (int)((Nm + 1) & ~(size_t)1);
                if (pass == 0) k_fwd<32><<<e - q, 32, sizeof(double) * (Nm + NB + 2 + NB * (size_t)As.stage_rows), st>>>(As, F.level_sn + q);
                else k_bwd<32><<<e - q, 32, sizeof(double) * (Nm + 2 + NB * Nm), st>>>(A, F.level_sn + q);
                launches++;
            }
            if (nsing && pass == 0) { k_single_fwd<<<cdiv(nsing, 256), 256, 0, sv[br++]>>>(A, F.level_sn + e, nsing); launches++; }
            join_level(d, nbranch, sv);
        }
    return launches;
}
static long issue_solve(qpdo_dev *d) { return issue_solve_args(d, sargs(d)); }
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
    for (int b = 0; b < 4; b++) {
        if (F.aux[b]) cudaStreamDestroy(F.aux[b]);
        if (F.ev_join[b]) cudaEventDestroy(F.ev_join[b]);
        F.aux[b] = nullptr; F.ev_join[b] = nullptr;
    }
    if (F.ev_fork) cudaEventDestroy(F.ev_fork);
    F.ev_fork = nullptr;
}

// =============================================================================== no-allocation solve
// qpdo_dev_solve_noalloc: the supernodal triangular solves (k_fwd / k_bwd, the kernels of factor_solve) on the factor
// that is in place, for callers that must not allocate (code under CUDA graph capture, serving loops, harnesses that
// forbid cudaMalloc).  No counterpart in the reference.  The scratch
// of the solve -- the unpermuted right-hand side, the permuted vector xs, the unpermuted solution and the per-front
// update vectors uv -- lives in ONE file-scope device array of QPDO_NOALLOC_MAX_DOUBLES doubles (4 MiB, held by every
// GPU that loads the library until exit) instead of cudaMalloc'ed buffers; the kernels are launched directly (no
// graph instantiation, which allocates).  Every call shares the array and finds the previous call's contents in
// it: what a solve reads before writing (nothing but uv, cleared below for good measure) is cleared per call;
// two calls that overlap on the GPU (different streams, different contexts of one device) race on it.
#define QPDO_NOALLOC_MAX_DOUBLES (1 << 19)
static __device__ __align__(256) double g_noalloc_scratch[QPDO_NOALLOC_MAX_DOUBLES];
// In host code `g_noalloc_scratch` is a host-side stand-in, not a device address: its address is looked up once per
// device, with that device current, and only the looked-up pointer is ever handed to a kernel or a copy.
#define QPDO_NOALLOC_MAX_DEVICES 64
static double *g_noalloc_addr[QPDO_NOALLOC_MAX_DEVICES];

static int noalloc_scratch_lookup(int device, double **out)
{
    if (device < 0 || device >= QPDO_NOALLOC_MAX_DEVICES) { qpdo_set_error("qpdo_dev_solve_noalloc: device %d out of range", device); return -1; }
    if (!g_noalloc_addr[device]) {
        double *ptr = nullptr;
        CUDA_OK(cudaSetDevice(device));
        CUDA_OK(cudaGetSymbolAddress((void **)&ptr, g_noalloc_scratch));
        if (!ptr) { qpdo_set_error("qpdo_dev_solve_noalloc: cudaGetSymbolAddress returned a null address"); return -1; }
        g_noalloc_addr[device] = ptr;
    }
    if (out) *out = g_noalloc_addr[device];
    return 0;
}
// set-up hook: called from factor_setup with the context's device current
int factor_noalloc_setup(qpdo_dev *d) { return noalloc_scratch_lookup(d->device, nullptr); }

extern "C" long qpdo_dev_solve_noalloc_max_doubles(void) { return (long)QPDO_NOALLOC_MAX_DOUBLES; }

extern "C" int qpdo_dev_solve_noalloc(qpdo_dev *d, const double *b, double *x)
{
    if (!d || !b || !x) { qpdo_set_error("qpdo_dev_solve_noalloc: null argument"); return -1; }
    CUDA_OK(cudaSetDevice(d->device));
    Factor &F = d->F;
    if (!F.S) { qpdo_set_error("qpdo_dev_solve_noalloc: no factorization in place (call after a Newton step or qpdo_dev_test_factor_solve)"); return -1; }
    const qpdo_symbolic *S = F.S;
    const size_t n = (size_t)d->n, nrows = (size_t)S->sn_rowptr[S->nsuper];
    const size_t need = 3 * n + nrows;
    if (need > (size_t)QPDO_NOALLOC_MAX_DOUBLES) {
        qpdo_set_error("qpdo_dev_solve_noalloc: the solve needs %zu doubles of scratch (3 n + front rows, n = %zu), the static scratch holds %d (%d MiB); use the allocating path",
                       need, n, QPDO_NOALLOC_MAX_DOUBLES, (int)(QPDO_NOALLOC_MAX_DOUBLES * sizeof(double) >> 20));
        return -2;
    }
    double *base = nullptr;
    if (noalloc_scratch_lookup(d->device, &base)) return -1;
    double *rhs = base, *xs = base + n, *xout = base + 2 * n, *uv = base + 3 * n;
    CUDA_OK(cudaMemsetAsync(uv, 0, sizeof(double) * (nrows ? nrows : 1), d->st));
    CUDA_OK(cudaMemcpyAsync(rhs, b, sizeof(double) * n, cudaMemcpyDefault, d->st));
    SArgs A = sargs(d);
    A.xs = xs; A.uv = uv;
    k_permute_in<<<cdiv(d->n, 256), 256, 0, d->st>>>(rhs, F.perm, xs, d->n);
    issue_solve_args(d, A);
    k_permute_out<<<cdiv(d->n, 256), 256, 0, d->st>>>(xs, F.perm, xout, d->n);
    CUDA_OK(cudaGetLastError());
    CUDA_OK(cudaMemcpyAsync(x, xout, sizeof(double) * n, cudaMemcpyDefault, d->st));
    CUDA_OK(cudaStreamSynchronize(d->st));
    return 0;
}
