// batch.cu -- many independent small QPs, one CTA per QP, the whole QPDO solver in shared memory.
//
// Each CTA runs qpdo_setup's scaling (reference src/scaling.c:24-91), qpdo_warm_start
// (src/qpdo.c:217-299) and the qpdo_solve loop (src/qpdo.c:304-476) for one instance: residuals and
// norms (iteration.c:30-93, termination.c:35-77), active set / Newton direction (newton.c:13-126),
// the regularised KKT solve K dx = rhs with K = Q + sigma I + A' diag(omega) A (replaces
// cholmod_interface.c:8-102), exact linesearch (linesearch.c:8-158), penalty updates
// (iteration.c:127-180) and the infeasibility certificates (termination.c:97-216).  All instances
// share one sparsity pattern, so the fill-reducing ordering and the pattern of L are computed once
// on the host (symbolic.cpp) and every CTA follows the same elimination program.
//
// Layout: per-instance inputs are instance-major in HBM (coalesced per CTA); per-CTA state is
// ~(nnzA + nnzQ + 12 n + 16 m + |L|) doubles of shared memory (53 KB for n=100, m=150: four CTAs per SM,
// which registers allow as well).  L is kept as packed trapezoids of the supernodes in the permuted order; the
// linesearch scratch lives in vectors that are dead during the linesearch (L survives the step, so an unchanged K is
// not factored again).  CTAs are persistent and pull instances from an atomic
// work queue (iteration counts are ragged).  Compiled with -fmad=false: element-wise arithmetic
// rounds like the reference's scalar C; fma() is used only inside the factorization / solves.
#include "dev_internal.h"
#include "qpdo_batch.h"
#include <algorithm>
#include <stdlib.h>
#include <string.h>
#include <type_traits>
#include <vector>

#define BNT 128                 // threads per CTA
#define B_ORDER_BINS 4096        // adaptive queue order: iteration counts are binned up to this value
#define BNW (BNT / 32)
#ifndef B_MINCTAS
#define B_MINCTAS 4              // resident CTAs per SM the register budget is sized for
#endif
#ifndef B_KUNROLL
#define B_KUNROLL 8            // K assembly: product records in flight per thread
#endif
constexpr int kKUnroll = B_KUNROLL;
#ifndef B_SUNROLL
#define B_SUNROLL 8            // gather SpMV: records in flight per thread
#endif
constexpr int kSUnroll = B_SUNROLL;
#ifdef B_INLINE_ALL
#define B_FN __forceinline__
#else
#define B_FN __noinline__      // the one-warp phases (front loop, forward / backward solves) as real calls: measured +2.7 % (code locality)
#endif
// (out-of-line linesearch / K assembly measured slower: 807k and 961k QP/s against 990k)
#define B_GC __grid_constant__
// instance data streams through once: evict-first loads / stores keep the shared index tables in the 23 KB of L1
#ifdef B_NO_STREAM
#define B_LDS_IN(p) (*(p))
#define B_STS_OUT(p, v) (*(p) = (v))
#else
#define B_LDS_IN(p) __ldcs(p)
#define B_STS_OUT(p, v) __stcs((p), (v))
#endif
#ifdef B_UNROLL_LS
#define B_LOOP1
#else
#define B_LOOP1 _Pragma("unroll 1")       // short-trip loops of the hot path stay rolled (instruction fetch)
#endif
#define B_INFTY 1e20
#define LS_EXACT_MAX 128        // same threshold as linesearch.cu

// gather SpMV program: rows dealt to the threads in rounds of BNT (row = round * BNT + thread), the terms of a
// round slot-major (a warp reads 32 consecutive 4-byte records per slot; uniform trip count, several loads in
// flight); record = value index | column << 16 | second-half flag << 31, 0xffffffff = padding
struct BatchEll { int nround; const int2 *round; const unsigned *rec; };
#define B_ROUNDS_MAX 8          // round tables of up to this many rounds are cached in shared memory
__shared__ int2 s_round[4][B_ROUNDS_MAX];      // ellA, ellAt, ellQ, K assembly
__device__ __forceinline__ int2 b_round(int which, int nround, const int2 *tab, int r) { return nround <= B_ROUNDS_MAX ? s_round[which][r] : tab[r]; }

struct BatchPlan {
    int n, m, nnzA, nnzQ, nnzK, nKpack, karea, P2;
    BatchEll ellA, ellAt, ellQ;              // y = A x (rows of A), y = A' x (columns of A), y = Q x (full symmetric rows)
    // shared structure (device pointers)
    const int *Ap, *Ai;                      // CSC of A
    const int *Rp, *Rj, *Rsrc;               // CSR of A (Rsrc: index into the CSC values)
    const int *Qp, *Qi;                      // lower CSC of Q
    const int *Fp, *Fj, *Fsrc, *Fdiag;       // full symmetric CSR of Q (split point per row)
    const int *perm;                         // perm[new] = old
    int nKround;                             // K assembly: rounds of BNT entries
    const int2 *Kround, *Kmeta;
    const uint2 *Ktr;
    const unsigned *Ktr32;                   // same program in 4-byte records when m < 1023 and nnz(A) <= 2048, else null
    const int *Lcp, *Lri;                    // L column patterns (rows strictly below the diagonal)
    const int *Lrp, *Lci;                    // L row patterns (columns strictly left of the diagonal)
    // supernodal register path (mode 1): every front fits one warp (N + 1 <= 32 lanes)
    int mode, nsuper;
    const int *sn_f, *sn_nc, *sn_nr;         // first column, pivot columns, rows below the pivot block
    const int *sn_base;                      // offset of the packed trapezoid of each supernode in the L area
    const int *Sptr, *Srow;                  // rows below the pivot block (permuted indices)
    const int *Uptr, *Upos;                  // where the (row ib, col jb) update entry lives in the ancestors' trapezoids
    int max_front, ucolw, nSrow;
    int ls_sep;                              // linesearch scratch lives outside the L area (L survives the step: unchanged K is not refactored)
    const int2 *Emap2;                       // owned entries of each front: {i | j << 8, position in L or ~index into xs}, column major, j >= 1, i = j..N
    const int4 *sn_desc;                     // per supernode {f, nc, nr, base}, {Sptr, Eptr, ne, 0}
    // settings (reference include/types.h:96-116)
    int max_iter, inner_max_iter, proximal, scaling, reset_newton_iter;
    double eps_abs, eps_abs_in, eps_prim_inf, eps_dual_inf, rho, theta, delta, mu_min, sigma_init, sigma_upd, sigma_min;
};

struct BatchIO {
    const double *Qx, *Ax, *q, *l, *u, *c, *x0, *y0;
    double *x, *y;
    qpdo_batch_summary *summary;
    unsigned char *active;
    int count;
    int *next;                               // work queue head
    const int *order;                        // the k-th pull of the queue takes instance order[k] (null: k itself)
    int *order_out;                          // adaptive order: the last CTA to finish leaves the order of the NEXT solve here (or null)
    int *done;                               // CTAs that have left the queue loop
    long long *timing;                       // [count][8] clock64 cycles per phase (diagnostics) or NULL
};

// ------------------------------------------------------------------------------- block reductions
template <int K> __device__ __forceinline__ void block_max(double (&v)[K], double *red)
{
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    // every caller reduces absolute values accumulated with `t > v ? t : v` (never negative, never NaN): their
    // IEEE bit patterns order like unsigned integers, so the warp maximum is two 32-bit redux.sync.max (high
    // word, then the low word among the lanes that hold the maximal high word) instead of five shuffle rounds
#pragma unroll
    for (int k = 0; k < K; k++) {
#ifdef B_SHFL_MAX
        for (int o = 16; o; o >>= 1) { double t = __shfl_xor_sync(0xffffffffu, v[k], o); v[k] = t > v[k] ? t : v[k]; }
#else
        const unsigned hi = (unsigned)__double2hiint(v[k]), lo = (unsigned)__double2loint(v[k]);
        const unsigned mh = __reduce_max_sync(0xffffffffu, hi);
        const unsigned ml = __reduce_max_sync(0xffffffffu, hi == mh ? lo : 0u);
        v[k] = __hiloint2double((int)mh, (int)ml);
#endif
    }
    __syncthreads();
    if (lane == 0)
#pragma unroll
        for (int k = 0; k < K; k++) red[wid * K + k] = v[k];
    __syncthreads();
#pragma unroll
    for (int k = 0; k < K; k++) {
        double mx = red[k];
        for (int w = 1; w < BNW; w++) { double t = red[w * K + k]; mx = t > mx ? t : mx; }
        v[k] = mx;
    }
}
// fixed-shape sum: xor tree inside warps, then the warp partials in index order (deterministic)
template <int K> __device__ __forceinline__ void block_sum(double (&v)[K], double *red)
{
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll
    for (int k = 0; k < K; k++)
        for (int o = 16; o; o >>= 1) v[k] += __shfl_xor_sync(0xffffffffu, v[k], o);
    __syncthreads();
    if (lane == 0)
#pragma unroll
        for (int k = 0; k < K; k++) red[wid * K + k] = v[k];
    __syncthreads();
#pragma unroll
    for (int k = 0; k < K; k++) {
        double s = 0.0;
        for (int w = 0; w < BNW; w++) s += red[w * K + k];
        v[k] = s;
    }
}
__device__ __forceinline__ int block_count3(int a, int b, int c, int &oa, int &ob, int &oc, double *red)
{
    unsigned ba = __ballot_sync(0xffffffffu, a), bb = __ballot_sync(0xffffffffu, b), bc = __ballot_sync(0xffffffffu, c);
    int *ri = (int *)red;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    __syncthreads();
    if (lane == 0) { ri[wid * 3] = __popc(ba); ri[wid * 3 + 1] = __popc(bb); ri[wid * 3 + 2] = __popc(bc); }
    __syncthreads();
    oa = ob = oc = 0;
    for (int w = 0; w < BNW; w++) { oa += ri[w * 3]; ob += ri[w * 3 + 1]; oc += ri[w * 3 + 2]; }
    return oa;
}
__device__ __forceinline__ double bclamp(double a, double lo, double hi)
{   // c_max(bmin, c_min(a, bmax))  (lin_alg.c:163-168)
    double t = a < hi ? a : hi;
    return lo > t ? lo : t;
}

// ------------------------------------------------------------------------------- SpMV (cholmod_sdmult order)
// one out-of-line copy serves every y = A x and y = A' x of the loop (the inlined copies at seven call sites were
// a fifth of the instructions fetched per Newton step); everything it needs comes in registers
#ifdef B_SPMV_INLINE
#define B_SPMV_FN __forceinline__
#else
#define B_SPMV_FN __noinline__
#endif
__device__ B_SPMV_FN void b_ell_gather(int which, int nround, const int2 *round, const unsigned *rec, int nrows,
                                       const double *val, const double *x, double *y)
{
    for (int r = 0; r < nround; r++) {
        const int2 rk = b_round(which, nround, round, r);
        const unsigned *T = rec + (size_t)rk.x * BNT + threadIdx.x;
        double s = 0.0;
#pragma unroll (kSUnroll)
        for (int k = 0; k < rk.y; k++) { const unsigned w = T[(size_t)k * BNT]; if (w != 0xffffffffu) s += val[w & 0xffffu] * x[w >> 16]; }
        const int i = r * BNT + threadIdx.x;
        if (i < nrows) y[i] = s;
    }
}
__device__ __forceinline__ void b_spmv_A(const BatchPlan &P, const double *Av, const double *x, double *y)
{   // CSR gather in ascending column order == CSC scatter order (t_cholmod_sdmult.c:297-320)
    b_ell_gather(0, P.ellA.nround, P.ellA.round, P.ellA.rec, P.m, Av, x, y);
}
__device__ __forceinline__ void b_spmv_At(const BatchPlan &P, const double *Av, const double *x, double *y)
{   // t_cholmod_sdmult.c:141-165
    b_ell_gather(1, P.ellAt.nround, P.ellAt.round, P.ellAt.rec, P.n, Av, x, y);
}
__device__ __forceinline__ void b_spmv_Q(const BatchPlan &P, const double *Qv, const double *x, double *y)
{   // t_cholmod_sdmult.c:440-484 (split sum: the terms of the lower-triangular columns, then the mirrored ones)
    for (int r = 0; r < P.ellQ.nround; r++) {
        const int2 rk = b_round(2, P.ellQ.nround, P.ellQ.round, r);
        const unsigned *T = P.ellQ.rec + (size_t)rk.x * BNT + threadIdx.x;
        double s1 = 0.0, s2 = 0.0;
#pragma unroll (kSUnroll)
        for (int k = 0; k < rk.y; k++) {
            const unsigned w = T[(size_t)k * BNT];
            if (w != 0xffffffffu) {
                const double t = Qv[w & 0xffffu], v = x[(w >> 16) & 0x7fffu];
                if ((int)w >= 0) s1 += t * v; else s2 += t * v;
            }
        }
        const int i = r * BNT + threadIdx.x;
        if (i < P.n) y[i] = s1 + s2;
    }
}

// ------------------------------------------------------------------------------- KKT matrix, LDL', solves
// u / d with r = RN(1/d): one Newton correction of RN(u r) gives the correctly rounded quotient (same bits
// as the division operator) without DDIV's slow path for zero or tiny numerators
__device__ __forceinline__ double b_div(double u, double d, double r)
{
    const double q0 = __dmul_rn(u, r);
    return fma(fma(-q0, d, u), r, q0);
}

// correctly rounded reciprocal of a normal, finite d (pivots of an SPD matrix): MUFU seed (20 bits) and three
// Newton steps, the last one from a 1-ulp approximation (Markstein).  Same bits as __drcp_rn on this range, a third
// of its instructions and no slow-path branch.
__device__ __forceinline__ double b_rcp(double d)
{
    double r;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(d));
    double e = fma(-d, r, 1.0); r = fma(r, e, r);
    e = fma(-d, r, 1.0); r = fma(r, e, r);
    e = fma(-d, r, 1.0); r = fma(r, e, r);
    return r;
}
// a / d with the bits of the division operator: reciprocal + Markstein correction when both operands are normal
// numbers of moderate exponent (|exponent| <= 500: no intermediate can overflow or underflow), the operator otherwise
__device__ __forceinline__ double b_quot(double a, double d)
{
    const unsigned ea = ((unsigned)__double2hiint(a) & 0x7ff00000u) - 0x20b00000u, ed = ((unsigned)__double2hiint(d) & 0x7ff00000u) - 0x20b00000u;
    if (ea <= 0x3e800000u && ed <= 0x3e800000u) return b_div(a, d, b_rcp(d));
    return a / d;
}
__device__ __forceinline__ int koff(int j, int n) { return j * n - ((j * (j - 1)) >> 1) - j; }   // (i, j), i >= j, at koff(j) + i

// K = Q + sigma I + A' diag(omega) A in the permuted packed lower triangle: the fused replacement
// of cholmod_aat + cholmod_add + beta*I (cholmod_interface.c:11-16,45-48); same per-entry
// arithmetic as k_assemble_K of factor.cu
__device__ __forceinline__ void b_assemble(const BatchPlan &P, const double *Av, const double *Qv, const double *om,
                                           double sigma, double *K)
{
    for (int t = threadIdx.x; t < P.nKpack; t += BNT) K[t] = 0.0;
    __syncthreads();
    // entries are dealt to the threads in rounds of BNT (sorted by their number of contributing products);
    // the products of round r sit slot-major, so that a warp reads 32 consecutive 8-byte records per slot
    // and the slot loop has a uniform trip count (loads of several slots in flight)
    for (int r = 0; r < P.nKround; r++) {
        const int2 rk = b_round(3, P.nKround, P.Kround, r);   // first slot, slots
        const int e = r * BNT + threadIdx.x;
        int2 mt; mt.x = 0; mt.y = -1;
        if (e < P.nnzK) mt = P.Kmeta[e];                      // packed position | diagonal flag << 31, index into Q's values or -1 (in flight during the slot loop)
        double v = 0.0;
        if (P.Ktr32) {                                        // small problems: 4-byte records row | a1 << 10 | a2 << 21
            const unsigned *T = P.Ktr32 + (size_t)rk.x * BNT + threadIdx.x;
#pragma unroll (kKUnroll)
            for (int k = 0; k < rk.y; k++) {
                const unsigned w = T[(size_t)k * BNT];
                if (w != 0xffffffffu) v = fma(om[w & 1023u], Av[(w >> 10) & 2047u] * Av[w >> 21], v);
            }
        } else {
            const uint2 *T = P.Ktr + (size_t)rk.x * BNT + threadIdx.x;
#pragma unroll (kKUnroll)
            for (int k = 0; k < rk.y; k++) {
                const uint2 w = T[(size_t)k * BNT];           // row (or -1: padding), a1 | a2 << 16
                if ((int)w.x >= 0) v = fma(om[w.x], Av[w.y & 0xffffu] * Av[w.y >> 16], v);
            }
        }
        if (e < P.nnzK) {
            if (mt.y >= 0) v += Qv[mt.y];
            if (P.proximal && mt.x < 0) v += sigma;
            K[mt.x & 0x7fffffff] = v;
        }
    }
    __syncthreads();
}

// Right-looking LDL' following the precomputed pattern of L; the right-hand side rides along as an
// extra row, so the forward solve costs no extra synchronisation.  On return K holds unit L and D,
// xs = D^{-1} L^{-1} xs_in.  One __syncthreads per column.
__device__ __forceinline__ void b_factor_forward(const BatchPlan &P, double *K, double *xs)
{
    const int n = P.n, lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    for (int k = 0; k < n; k++) {
        const int ck = koff(k, n);
        const double dk = K[ck + k];
        const int rb = P.Lcp[k], nr = P.Lcp[k + 1] - rb;
        const int *R = P.Lri + rb;
        const double xk = xs[k];
        for (int b = wid; b < nr; b += BNW) {
            const int rj = R[b];
            const double ljk = K[ck + rj] / dk;
            const int cj = koff(rj, n);
            for (int a = b + lane; a < nr; a += 32) { const int ri = R[a]; K[cj + ri] = fma(-K[ck + ri], ljk, K[cj + ri]); }
            if (lane == 31) xs[rj] = fma(-ljk, xk, xs[rj]);
        }
        __syncthreads();
        for (int a = threadIdx.x; a < nr; a += BNT) K[ck + R[a]] = K[ck + R[a]] / dk;
        if (threadIdx.x == BNT - 1) xs[k] = xk / dk;
    }
    __syncthreads();
}
// x = L^{-T} xs, row-oriented scatter (no reductions), one warp
__device__ __forceinline__ void b_backward(const BatchPlan &P, const double *K, double *xs)
{
    const int n = P.n, lane = threadIdx.x & 31;
    if (threadIdx.x < 32) {
        for (int i = n - 1; i > 0; i--) {
            const double xi = xs[i];
            const int cb = P.Lrp[i], nc = P.Lrp[i + 1] - cb;
            for (int t = lane; t < nc; t += 32) { const int k = P.Lci[cb + t]; xs[k] = fma(-K[koff(k, n) + i], xi, xs[k]); }
            __syncwarp();
        }
    }
    __syncthreads();
}

// ---- supernodal front path (mode 1) -------------------------------------------------------------
// L is stored as one packed trapezoid per supernode: pivot column jj holds front rows jj..N-1
// (diagonal = D, below = unit-L entries), at base + jj*N - jj(jj-1)/2 + (i - jj).  The CTA eliminates
// one dense front at a time: every thread OWNS up to SN_EPT entries (i, j) of the front's lower
// triangle for the whole supernode -- pivot columns, the columns of the update block (which live in
// the ancestors' trapezoids) and the right-hand side as an extra row i = N -- with their shared-memory
// addresses in registers.  Step k: the rows below the pivot scale column k (u = a_ik, l = u / d_k) into
// the ucol / lcol buffers, barrier, every owner applies a_ij = fma(-u_i, l_j, a_ij), barrier.  No index
// loads and no address arithmetic on the critical path.  Per entry the sequence of fma's is the one
// b_factor_forward performs (ascending k, same operands): both paths give bit-identical results.
// (A register-tile variant with one barrier per pivot and a shuffle-based single-warp variant were
// measured slower on B200)
#define SN_NMAX 31              // b_backward_sn: front order limit (one lane per front row)
__device__ __forceinline__ int sn_col(int base, int jj, int N) { return base + jj * N - ((jj * (jj - 1)) >> 1) - jj; }   // + front row i
#define SN_EPT 8
#define SN_DESC_MAX 16          // supernode descriptors cached in shared memory (else read from global)
__shared__ int4 s_desc[2 * SN_DESC_MAX];   // {f, nc, nr, base}, {Sptr, Eptr, ne, 0}
#define SN_ROWS_MAX 128         // rows below the pivot blocks (all supernodes) cached in shared memory
__shared__ int s_rows[SN_ROWS_MAX];
__device__ __forceinline__ int4 sn_desc(const BatchPlan &P, int s, int which) { return P.nsuper <= SN_DESC_MAX ? s_desc[2 * s + which] : P.sn_desc[2 * s + which]; }
// ---- fronts of order N < 32 (rows k+1 .. N of a column pair sit in one warp).  The owners keep three 32-bit
// shared-memory addresses per entry (the entry, ucol[i], ucol[j]); the four scratch columns are SN_W doubles apart,
// so every load of the update is `ld.shared.f64 [reg + immediate]` with no address arithmetic, the activity test
// (j > k) is one compare of addresses and the store is predicated.
#define SN_W 34
#define SN_EPTS 4             // owned entries per thread of a small front (N (N + 1) / 2 - 1 <= 495 <= 4 * BNT)
__device__ __forceinline__ unsigned b_saddr(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }
template <int OFF> __device__ __forceinline__ double b_lds(unsigned a)
{
    double v;
    asm volatile("ld.shared.f64 %0, [%1+%2];" : "=d"(v) : "r"(a), "n"(OFF));
    return v;
}
__device__ __forceinline__ void b_sts_if(unsigned a, double v, bool p)
{
    asm volatile("{ .reg .pred q; setp.ne.u32 q, %2, 0; @q st.shared.f64 [%0], %1; }" :: "r"(a), "d"(v), "r"((unsigned)p) : "memory");
}
// ---- panel step: up to SN_PW pivots per barrier pair.  Warp 0 holds one front row per lane (rows k0 .. N, N = the
// right-hand side, at most 32 of them) with the row's panel entries in registers.  EVERY lane also loads the W x W
// pivot block (broadcast loads) and factors it redundantly, so a lane finishes its own row of the panel
// (u_c = a_c - sum_{q<c} u_q l_cq, l_c = u_c / d_c) without any exchange between lanes: the chain per pivot is
// reciprocal -> quotient -> fma, no shuffle, no barrier.  The u / l columns of the rows below the panel go to the
// 2 W scratch columns, one barrier, all owners apply the W rank-1 updates in ascending pivot order (the same fma
// sequence per entry as the column path: bit-identical results), one barrier.
#define SN_PW 4
#ifdef B_FINE
__device__ long long g_fine[8];
#endif
template <int W> __device__ __forceinline__ void b_panel_update(const unsigned (&ea)[SN_EPTS], const unsigned (&pu)[SN_EPTS],
                                                                  const unsigned (&pl)[SN_EPTS], unsigned thr)
{   // shared-memory bandwidth bound (10 accesses per entry): loads and fma's interleaved per entry measured faster than
    // all loads first (347 vs 382 cycles per step alone on the SM, 441 vs 562 with four CTAs)
    double av[SN_EPTS];
#pragma unroll
    for (int e = 0; e < SN_EPTS; e++) av[e] = b_lds<0>(ea[e]);
#pragma unroll
    for (int e = 0; e < SN_EPTS; e++) {
        { const double u0 = b_lds<0>(pu[e]), l0 = b_lds<8 * SN_W>(pl[e]); av[e] = fma(-u0, l0, av[e]); }
        if (W > 1) { const double u1 = b_lds<16 * SN_W>(pu[e]), l1 = b_lds<24 * SN_W>(pl[e]); av[e] = fma(-u1, l1, av[e]); }
        if (W > 2) { const double u2 = b_lds<32 * SN_W>(pu[e]), l2 = b_lds<40 * SN_W>(pl[e]); av[e] = fma(-u2, l2, av[e]); }
        if (W > 3) { const double u3 = b_lds<48 * SN_W>(pu[e]), l3 = b_lds<56 * SN_W>(pl[e]); av[e] = fma(-u3, l3, av[e]); }
    }
#pragma unroll
    for (int e = 0; e < SN_EPTS; e++) b_sts_if(ea[e], av[e], pl[e] > thr);
}
#ifdef B_EXACT_PIVOTS
// columns k0 .. k0 + W - 1 of the front (W <= nc - k0), executed by warp 0
template <int W> __device__ __forceinline__ void b_panel_factor(double *L, double *xs, double *ucol, int f, int N, int base, int k0)
{
    const int t = threadIdx.x, rows = N - k0;                   // lane t: front row k0 + t; t == rows: the right-hand side
    const int tc = t < rows ? t : rows;                           // lanes past the right-hand side repeat its work and store nothing
    const bool live = t <= rows;
    double *cq[W], *pa[W];                                        // column k0 + q, + front row;  this lane's entry of it
    cq[0] = L + sn_col(base, k0, N);
#pragma unroll
    for (int q = 1; q < W; q++) cq[q] = cq[q - 1] + (N - k0 - q);
#pragma unroll
    for (int q = 0; q < W; q++) pa[q] = tc < rows ? cq[q] + k0 + tc : xs + f + k0 + q;
    double a[W], B[W][W];
#pragma unroll
    for (int q = 0; q < W; q++) {
#pragma unroll
        for (int p = q; p < W; p++) B[p][q] = cq[q][k0 + p];
        a[q] = *pa[q];                                            // lanes t < q read an unrelated entry: never used
    }
    double d[W], r[W], lb[W][W];
#pragma unroll
    for (int q = 0; q < W; q++) {
        d[q] = B[q][q];
        r[q] = b_rcp(d[q]);
#pragma unroll
        for (int p = q + 1; p < W; p++) lb[p][q] = b_div(B[p][q], d[q], r[q]);
#pragma unroll
        for (int q2 = q + 1; q2 < W; q2++)
#pragma unroll
            for (int p = q2; p < W; p++) B[p][q2] = fma(-B[p][q], lb[q2][q], B[p][q2]);
    }
    double u[W], l[W];
#pragma unroll
    for (int c = 0; c < W; c++) {
        double acc = a[c];
#pragma unroll
        for (int q = 0; q < c; q++) acc = fma(-u[q], lb[c][q], acc);
        u[c] = acc;
        l[c] = b_div(acc, d[c], r[c]);
    }
    __syncwarp();                                                 // every lane has read the block before it is overwritten
#pragma unroll
    for (int c = 0; c < W; c++)
        if (live && t >= c) *pa[c] = t == c ? d[c] : l[c];
    if (live && t >= W) {
        double *us = ucol + k0 + t;
#pragma unroll
        for (int c = 0; c < W; c++) { us[2 * c * SN_W] = u[c]; us[(2 * c + 1) * SN_W] = l[c]; }
    }
}
#else
// Fast pivots (default): the chain per pivot PAIR is fma -> reciprocal -> product instead of (reciprocal -> quotient
// -> fma) twice.  With the 2 x 2 leading block [d0 u; u a] the second pivot is d1 = (d0 a - u^2) / d0, so
// 1/d1 = d0 / (d0 a - u^2): both reciprocals start from loaded values and run side by side (the formula of a 2 x 2
// block pivot; the cancellation in d0 a - u^2 is the one in a - u^2/d0 scaled by d0).  Multipliers are l = u * (1/d)
// with a reciprocal accurate to an ulp (seed + two Newton steps) instead of the correctly rounded quotient: the
// factorization is as backward stable as before, but no longer bit-identical to the column path (-DB_EXACT_PIVOTS
// restores that variant; tools/batch_ab.py checks whichever is built).
__device__ __forceinline__ double b_rcp2(double d)
{
    double r;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(d));
    double e = fma(-d, r, 1.0); r = fma(r, e, r);
    e = fma(-d, r, 1.0); r = fma(r, e, r);
    return r;
}
template <int W> __device__ __forceinline__ void b_panel_factor(double *L, double *xs, double *ucol, int f, int N, int base, int k0)
{
    const int t = threadIdx.x, rows = N - k0;                   // lane t: front row k0 + t; t == rows: the right-hand side
    const int tc = t < rows ? t : rows;                           // lanes past the right-hand side repeat its work and store nothing
    const bool live = t <= rows;
    double *cq[W], *pa[W];                                        // column k0 + q, + front row;  this lane's entry of it
    cq[0] = L + sn_col(base, k0, N);
#pragma unroll
    for (int q = 1; q < W; q++) cq[q] = cq[q - 1] + (N - k0 - q);
#pragma unroll
    for (int q = 0; q < W; q++) pa[q] = tc < rows ? cq[q] + k0 + tc : xs + f + k0 + q;
    double a[W], B[W][W];
#pragma unroll
    for (int q = 0; q < W; q++) {
#pragma unroll
        for (int p = q; p < W; p++) B[p][q] = cq[q][k0 + p];
        a[q] = *pa[q];                                            // lanes t < q read an unrelated entry: never used
    }
    double r[W], u[W], l[W];
    if constexpr (W == 1) {
        r[0] = b_rcp2(B[0][0]);
        u[0] = a[0]; l[0] = u[0] * r[0];
    } else {
        // pivots 0, 1
        const double det = fma(B[0][0], B[1][1], -(B[1][0] * B[1][0]));
        r[0] = b_rcp2(B[0][0]);
        const double rdet = b_rcp2(det);
        r[1] = B[0][0] * rdet;
        const double l10 = B[1][0] * r[0];
        u[0] = a[0]; l[0] = u[0] * r[0];
        u[1] = fma(-u[0], l10, a[1]); l[1] = u[1] * r[1];
        if constexpr (W == 4) {
            const double l20 = B[2][0] * r[0], l30 = B[3][0] * r[0];
            const double u21 = fma(-B[2][0], l10, B[2][1]), u31 = fma(-B[3][0], l10, B[3][1]);
            const double l21 = u21 * r[1], l31 = u31 * r[1];
            const double s22 = fma(-u21, l21, fma(-B[2][0], l20, B[2][2]));
            const double s32 = fma(-u31, l21, fma(-B[3][0], l20, B[3][2]));
            const double s33 = fma(-u31, l31, fma(-B[3][0], l30, B[3][3]));
            const double det2 = fma(s22, s33, -(s32 * s32));
            r[2] = b_rcp2(s22);
            const double rdet2 = b_rcp2(det2);
            r[3] = s22 * rdet2;
            const double l32 = s32 * r[2];
            u[2] = fma(-u[1], l21, fma(-u[0], l20, a[2])); l[2] = u[2] * r[2];
            u[3] = fma(-u[2], l32, fma(-u[1], l31, fma(-u[0], l30, a[3]))); l[3] = u[3] * r[3];
        }
    }
    __syncwarp();                                                 // every lane has read the block before it is overwritten
#pragma unroll
    for (int c = 0; c < W; c++)
        if (live && t >= c) *pa[c] = t == c ? r[c] : l[c];        // the diagonal keeps 1 / d_c: what the forward solve of an unchanged K multiplies by
    if (live && t >= W) {
        double *us = ucol + k0 + t;
#pragma unroll
        for (int c = 0; c < W; c++) { us[2 * c * SN_W] = u[c]; us[(2 * c + 1) * SN_W] = l[c]; }
    }
}
#endif
__device__ __forceinline__ void b_factor_front_small(const BatchPlan &P, double *L, double *xs, double *ucol, const int4 d0, const int4 d1)
{
    const int tid = threadIdx.x;
    const int f = d0.x, nc = d0.y, nr = d0.z, N = nc + nr, base = d0.w;
    const int2 *emap = P.Emap2 + d1.y;
    const unsigned Ls = b_saddr(L), xss = b_saddr(xs), uc = b_saddr(ucol);
    unsigned ea[SN_EPTS], pu[SN_EPTS], pl[SN_EPTS];       // N < 32: at most 495 owned entries, 4 per thread
#pragma unroll
    for (int e = 0; e < SN_EPTS; e++) {                         // the table of a small front is padded to SN_EPTS * BNT entries (j = 0: never active)
        const int2 em = emap[tid + e * BNT];                    // i | j << 8, position in L (>= 0) or ~index into xs
        ea[e] = em.y >= 0 ? Ls + 8u * (unsigned)em.y : xss + 8u * (unsigned)(~em.y);
        pu[e] = uc + 8u * (unsigned)(em.x & 255); pl[e] = uc + 8u * (unsigned)(em.x >> 8);
    }
    int k0 = 0;
    for (; k0 + SN_PW <= nc; k0 += SN_PW) {
#ifdef B_FINE
        const long long c0_ = clock64();
#endif
        if (tid < 32) b_panel_factor<SN_PW>(L, xs, ucol, f, N, base, k0);
#ifdef B_FINE
        const long long c1_ = clock64();
#endif
        __syncthreads();
#ifdef B_FINE
        const long long c2_ = clock64();
#endif
        b_panel_update<SN_PW>(ea, pu, pl, uc + 8u * (unsigned)(k0 + SN_PW - 1));
#ifdef B_FINE
        const long long c3_ = clock64();
#endif
        __syncthreads();
#ifdef B_FINE
        if (tid == 0) { const long long c4_ = clock64(); atomicAdd((unsigned long long *)&g_fine[0], (unsigned long long)(c1_ - c0_)); atomicAdd((unsigned long long *)&g_fine[1], (unsigned long long)(c2_ - c1_)); atomicAdd((unsigned long long *)&g_fine[2], (unsigned long long)(c3_ - c2_)); atomicAdd((unsigned long long *)&g_fine[3], (unsigned long long)(c4_ - c3_)); atomicAdd((unsigned long long *)&g_fine[4], 1ull); }
#endif
    }
    for (; k0 < nc; ) {                                          // the last nc mod SN_PW pivots: two, then one
        if (k0 + 2 <= nc) {
            if (tid < 32) b_panel_factor<2>(L, xs, ucol, f, N, base, k0);
            __syncthreads();
            b_panel_update<2>(ea, pu, pl, uc + 8u * (unsigned)(k0 + 1));
            k0 += 2;
        } else {
            if (tid < 32) b_panel_factor<1>(L, xs, ucol, f, N, base, k0);
            __syncthreads();
            b_panel_update<1>(ea, pu, pl, uc + 8u * (unsigned)k0);
            k0 += 1;
        }
        __syncthreads();
    }
    (void)nr;
}
__device__ B_FN void b_factor_fronts(const BatchPlan &P, double *L, double *xs, double *ucol)
{
    const int tid = threadIdx.x;
    for (int s = 0; s < P.nsuper; s++) {
        const int4 d0 = sn_desc(P, s, 0), d1 = sn_desc(P, s, 1);
        const int f = d0.x, nc = d0.y, nr = d0.z, N = nc + nr, base = d0.w;
        if (N < 32) { b_factor_front_small(P, L, xs, ucol, d0, d1); continue; }
        const int2 *emap = P.Emap2 + d1.y;
        const int ne = d1.z, ept = (ne + BNT - 1) / BNT;
        double *lcol = ucol + N + 1;
        int eij[SN_EPT];            // i | j << 8;  j == 0 never updates
        double *ea[SN_EPT];
#pragma unroll
        for (int e = 0; e < SN_EPT; e++) {
            eij[e] = 0; ea[e] = L;
            const int idx = tid + e * BNT;
            if (e < ept && idx < ne) {
                const int2 em = emap[idx];                    // i | j << 8, position in L (>= 0) or ~index into xs
                eij[e] = em.x;
                ea[e] = em.y >= 0 ? L + em.y : xs + ~em.y;
            }
        }
        // one pivot per barrier pair here; fronts of order < 32 took b_factor_front_small (two pivots per pair)
        int k = 0;
        while (k < nc) {
            const int pc = sn_col(base, k, N);
            if (tid < N - k) {                                  // rows k+1 .. N (N = right-hand side)
                const int i = k + 1 + tid;
                double *pa = (i < N) ? L + pc + i : xs + f + k;
                const double u = *pa, dk = L[pc + k];
                // u / dk, correctly rounded (Markstein: reciprocal, product, one residual correction).  The plain
                // quotient takes DDIV's ~60-instruction slow path whenever u == 0, which structural zeros of
                // the front make the common case; this form never does and gives the same bits.
                const double rk = __drcp_rn(dk);
                const double l = b_div(u, dk, rk);
                ucol[i] = u; lcol[i] = l; *pa = l;
            }
            __syncthreads();
            // all loads first, then the fma's, then the stores: the owned entries are independent, the
            // compiler cannot know that (everything is a double* into shared memory)
            double uv[SN_EPT], lv[SN_EPT], av[SN_EPT];
#pragma unroll
            for (int e = 0; e < SN_EPT; e++) {
                const bool on = e < ept && (eij[e] >> 8) > k;
                uv[e] = on ? ucol[eij[e] & 255] : 0.0; lv[e] = on ? lcol[eij[e] >> 8] : 0.0; av[e] = on ? *ea[e] : 0.0;
            }
#pragma unroll
            for (int e = 0; e < SN_EPT; e++) av[e] = fma(-uv[e], lv[e], av[e]);
#pragma unroll
            for (int e = 0; e < SN_EPT; e++)
                if (e < ept && (eij[e] >> 8) > k) *ea[e] = av[e];
            __syncthreads();
            k++;
        }
    }
}
// x = L^{-T} xs: lane c owns pivot column c of the front.  Rows below the pivot block are final (broadcast loads
// from xs); the pivot block goes four rows per step: every lane fetches the four right-hand-side values (shuffles)
// and the six L entries between those rows (broadcast loads), solves the 4 x 4 triangle redundantly and applies the
// four fma's to its own column -- one shuffle latency per four rows instead of one per row.  Per entry the same fma
// sequence as b_backward (descending row order): bit-identical.
__device__ B_FN void b_backward_sn(const BatchPlan &P, const double *L, double *xs)
{   // requires the descriptor and row caches (nsuper <= SN_DESC_MAX, nSrow <= SN_ROWS_MAX): plain shared-memory loads
    const int lane = threadIdx.x & 31;
    const unsigned Ls = b_saddr(L), xss = b_saddr(xs), rws = b_saddr(s_rows);
    for (int s = P.nsuper - 1; s >= 0; s--) {
#ifdef B_FINE
        const long long t0_ = clock64();
#endif
        const int4 d0 = s_desc[2 * s], d1 = s_desc[2 * s + 1];
        const int f = d0.x, nc = d0.y, nr = d0.z, N = nc + nr, base = d0.w;
        const bool own = lane < nc;
        const int lim = own ? lane : 64;                               // this lane's column takes the update of row i while lim < i
        const int cl = own ? lane : nc - 1;                            // lanes past the pivot block repeat the last column's work and store nothing
        const unsigned colb = Ls + 8u * (unsigned)sn_col(base, cl, N); // + 8 * front row
        double v = b_lds<0>(xss + 8u * (unsigned)(f + cl));
        {   // rows below the pivot block (final values), descending: x_i broadcast from xs[rows[i - nc]]
            unsigned ra = rws + 4u * (unsigned)(d1.x + nr), ca = colb + 8u * (unsigned)N;
            int k = nr;
#pragma unroll 1
            for (; k >= 4; k -= 4) {
                int r0, r1, r2, r3;
                asm volatile("ld.shared.b32 %0, [%1+-4];" : "=r"(r0) : "r"(ra)); asm volatile("ld.shared.b32 %0, [%1+-8];" : "=r"(r1) : "r"(ra));
                asm volatile("ld.shared.b32 %0, [%1+-12];" : "=r"(r2) : "r"(ra)); asm volatile("ld.shared.b32 %0, [%1+-16];" : "=r"(r3) : "r"(ra));
                const double l0 = b_lds<-8>(ca), l1 = b_lds<-16>(ca), l2 = b_lds<-24>(ca), l3 = b_lds<-32>(ca);
                const double x0 = b_lds<0>(xss + 8u * (unsigned)r0), x1 = b_lds<0>(xss + 8u * (unsigned)r1);
                const double x2 = b_lds<0>(xss + 8u * (unsigned)r2), x3 = b_lds<0>(xss + 8u * (unsigned)r3);
                v = fma(-l0, x0, v); v = fma(-l1, x1, v); v = fma(-l2, x2, v); v = fma(-l3, x3, v);
                ra -= 16u; ca -= 32u;
            }
#pragma unroll 1
            for (; k >= 1; k--) {
                int r0;
                asm volatile("ld.shared.b32 %0, [%1+-4];" : "=r"(r0) : "r"(ra));
                const double l0 = b_lds<-8>(ca), x0 = b_lds<0>(xss + 8u * (unsigned)r0);
                v = fma(-l0, x0, v);
                ra -= 4u; ca -= 8u;
            }
        }
#ifdef B_FINE
        const long long t1_ = clock64();
#endif
        // pivot block, rows i0 .. i0-3 per step; am = &col[i0], a1..a3 = &L(i0, i0-1), &L(i0, i0-2), &L(i0, i0-3) as
        // shared-space byte addresses, stepped without multiplications (column j-1 starts N - j entries before column j).
        // A lane whose column does not take the update of a row multiplies by zero (the multiplier is selected off the chain).
        int i0 = nc - 1, g = N - i0;
        unsigned am = colb + 8u * (unsigned)i0;
        unsigned a1 = Ls + 8u * (unsigned)(sn_col(base, i0, N) + i0) - 8u * (unsigned)g;
#pragma unroll 1
        for (; i0 >= 3; i0 -= 4) {
            const unsigned a2 = a1 - 8u * (unsigned)(g + 1), a3 = a2 - 8u * (unsigned)(g + 2);
            const double l01 = b_lds<0>(a1), l02 = b_lds<0>(a2), l12 = b_lds<-8>(a2);
            const double l03 = b_lds<0>(a3), l13 = b_lds<-8>(a3), l23 = b_lds<-16>(a3);
            double m0 = b_lds<0>(am), m1 = b_lds<-8>(am), m2 = b_lds<-16>(am), m3 = b_lds<-24>(am);
            m0 = lim < i0 ? m0 : 0.0; m1 = lim < i0 - 1 ? m1 : 0.0; m2 = lim < i0 - 2 ? m2 : 0.0; m3 = lim < i0 - 3 ? m3 : 0.0;
            const double x0 = __shfl_sync(0xffffffffu, v, i0), v1 = __shfl_sync(0xffffffffu, v, i0 - 1);
            const double v2 = __shfl_sync(0xffffffffu, v, i0 - 2), v3 = __shfl_sync(0xffffffffu, v, i0 - 3);
            const double x1 = fma(-l01, x0, v1);
            const double x2 = fma(-l12, x1, fma(-l02, x0, v2));
            const double x3 = fma(-l23, x2, fma(-l13, x1, fma(-l03, x0, v3)));
            v = fma(-m0, x0, v); v = fma(-m1, x1, v); v = fma(-m2, x2, v); v = fma(-m3, x3, v);
            a1 = a3 - 8u * (unsigned)(2 * g + 7) - 32u; am -= 32u; g += 4;
        }
#pragma unroll 1
        for (; i0 >= 1; i0--) {                                        // at most three single rows
            double m0 = b_lds<0>(am);
            m0 = lim < i0 ? m0 : 0.0;
            const double x0 = __shfl_sync(0xffffffffu, v, i0);
            v = fma(-m0, x0, v);
            am -= 8u;
        }
        if (own) xs[f + lane] = v;
        __syncwarp();
#ifdef B_FINE
        if (lane == 0) { const long long t2_ = clock64(); atomicAdd((unsigned long long *)&g_fine[5], (unsigned long long)(t1_ - t0_)); atomicAdd((unsigned long long *)&g_fine[6], (unsigned long long)(t2_ - t1_)); atomicAdd((unsigned long long *)&g_fine[7], 1ull); }
#endif
    }
}

#ifndef B_EXACT_PIVOTS
// z = D^{-1} L^{-1} xs on a factor that is already in place (K unchanged since the last factorization: the reference
// keeps its factor too when no constraint enters or leaves, src/newton.c:24-33).  Lane t owns front row t; four pivot
// columns per step: every lane fetches the four right-hand-side values (shuffles) and the six L entries between those
// rows (broadcast loads), forms u_0..u_3 redundantly and applies them to its own row.  Per entry the fma sequence of
// the fused forward solve of b_factor_front_small (ascending pivot, fma(-u_c, l_jc, .)), then u_c times the stored
// reciprocal: bit-identical to factoring again.
__device__ B_FN void b_forward_sn(const BatchPlan &P, const double *L, double *xs)
{
    const int lane = threadIdx.x & 31;
    const unsigned Ls = b_saddr(L);
    for (int s = 0; s < P.nsuper; s++) {
        const int4 d0 = s_desc[2 * s], d1 = s_desc[2 * s + 1];
        const int f = d0.x, nc = d0.y, nr = d0.z, N = nc + nr, base = d0.w;
        const int tl = lane < N ? lane : N - 1;                        // lanes past the front repeat its last row and store nothing
        const int xi = tl < nc ? f + tl : s_rows[d1.x + tl - nc];
        double b = xs[xi];
        // a0 = &L(row tl, column c0) as a shared-space byte address; column c+1 starts N - c - 1 entries after column c
        unsigned a0 = Ls + 8u * (unsigned)(sn_col(base, 0, N) + tl);
        unsigned g0 = Ls + 8u * (unsigned)sn_col(base, 0, N);          // &L(row 0, column c0) - 8 * 0: block entries are at + 8 * row
        int stepc = N - 1;                                             // distance from column c0 to c0 + 1
#pragma unroll 1
        for (int c0 = 0; c0 < nc; c0 += 4) {
            const unsigned a1 = a0 + 8u * (unsigned)stepc, a2 = a1 + 8u * (unsigned)(stepc - 1), a3 = a2 + 8u * (unsigned)(stepc - 2);
            const unsigned g1 = g0 + 8u * (unsigned)stepc, g2 = g1 + 8u * (unsigned)(stepc - 1);
            const unsigned r0 = 8u * (unsigned)c0;
            const double l10 = b_lds<8>(g0 + r0), l20 = b_lds<16>(g0 + r0), l30 = b_lds<24>(g0 + r0);
            const double l21 = b_lds<16>(g1 + r0), l31 = b_lds<24>(g1 + r0), l32 = b_lds<24>(g2 + r0);
            double m0 = b_lds<0>(a0), m1 = b_lds<0>(a1), m2 = b_lds<0>(a2), m3 = b_lds<0>(a3);
            m0 = tl > c0 ? m0 : 0.0; m1 = (tl > c0 + 1 && c0 + 1 < nc) ? m1 : 0.0;
            m2 = (tl > c0 + 2 && c0 + 2 < nc) ? m2 : 0.0; m3 = (tl > c0 + 3 && c0 + 3 < nc) ? m3 : 0.0;
            const double u0 = __shfl_sync(0xffffffffu, b, c0), v1 = __shfl_sync(0xffffffffu, b, c0 + 1);
            const double v2 = __shfl_sync(0xffffffffu, b, c0 + 2), v3 = __shfl_sync(0xffffffffu, b, c0 + 3);
            double u1 = fma(-u0, l10, v1);
            double u2 = fma(-u1, l21, fma(-u0, l20, v2));
            double u3 = fma(-u2, l32, fma(-u1, l31, fma(-u0, l30, v3)));
            u1 = c0 + 1 < nc ? u1 : 0.0; u2 = c0 + 2 < nc ? u2 : 0.0; u3 = c0 + 3 < nc ? u3 : 0.0;     // columns past the pivot block do not exist
            b = fma(-u0, m0, b); b = fma(-u1, m1, b); b = fma(-u2, m2, b); b = fma(-u3, m3, b);
            a0 = a3 + 8u * (unsigned)(stepc - 3); g0 = g2 + 8u * (unsigned)(stepc - 2) + 8u * (unsigned)(stepc - 3); stepc -= 4;
        }
        if (lane < N) {
            if (lane < nc) b = b * L[sn_col(base, lane, N) + lane];    // the diagonal holds 1 / d
            xs[xi] = b;
        }
        __syncwarp();
    }
}
#endif

// generic variant (fronts wider than a warp): one warp, column-parallel scatter per front row
__device__ B_FN void b_backward_fronts(const BatchPlan &P, const double *L, double *xs)
{
    const int lane = threadIdx.x & 31;
    for (int s = P.nsuper - 1; s >= 0; s--) {
        const int f = P.sn_f[s], nc = P.sn_nc[s], nr = P.sn_nr[s], N = nc + nr, base = P.sn_base[s];
        const int *rows = P.Srow + P.Sptr[s];
        for (int i = N - 1; i >= 1; i--) {
            const double xi = i >= nc ? xs[rows[i - nc]] : xs[f + i];
            const int cend = i < nc ? i : nc;
            for (int c = lane; c < cend; c += 32) xs[f + c] = fma(-L[sn_col(base, c, N) + i], xi, xs[f + c]);
            __syncwarp();
        }
    }
}

// ------------------------------------------------------------------------------- exact linesearch
// reference src/linesearch.c:8-158; the CTA-wide version of k_ls_small (linesearch.cu): breakpoints
// t_i = alpha_i / delta_i, keep t_i > 0, bitonic sort of (IEEE bits, index) in shared memory, prefix
// sums of the slope changes, first segment whose end has a positive derivative.
struct BVec {       // shared-memory vectors of one instance
    double *Av, *Qv, *q, *x, *xbar, *Qx, *Aty, *dx, *Qdx, *Atdy, *df, *rdi, *tn, *D;
    double *l, *u, *y, *ybar, *Ax, *dy, *Adx, *w, *rp, *rpo, *rpi, *mu, *smu, *om, *E, *tm;
    double *K, *red, *ucol;
    unsigned char *act, *acto;
};
__device__ __forceinline__ void b_ls_contrib2(double d, double al, double &da, double &db)
{   // linesearch.c:133-141
    if (d > 0.0) { da = d * d; db = -(d * al); }
    else { da = -(d * d); db = d * al; }
}
// breakpoint k of the linesearch (linesearch.c:28-39,82-114): slope change delta_k and offset alpha_k
__device__ __forceinline__ void b_ls_da(const BVec &V, int m, int k, double &dl, double &al)
{
    const int i = k < m ? k : k - m;
    double tm = V.dy[i] * V.mu[i]; tm = tm * 0.5;
    const double c0 = V.Adx[i] + (-1.0) * tm;
    dl = c0 * V.smu[i];
    if (k < m) { dl = dl * (-1.0); al = (V.w[i] + (-1.0) * V.l[i]) * V.smu[i]; }
    else al = (V.u[i] + (-1.0) * V.w[i]) * V.smu[i];
}
// SEP: the scratch lives in vectors that are dead during the linesearch instead of the L area -- keys in
// [rp, rpi, tm, ucol] (contiguous), 16-bit indices in [rdi, tn], delta / alpha recomputed where they are needed --
// so that L survives the step and an unchanged K need not be factored again.
__shared__ int s_ls_kmin, s_ls_cw[8 * BNW];
__shared__ double s_ls_wa[BNW], s_ls_wb[BNW], s_ls_tau;
#define B_LS_CONTRIB(iz_, da_, db_) do { double dl_, al_; const int z_ = (int)(iz_); if (SEP) b_ls_da(V, m, z_, dl_, al_); else { dl_ = delta[z_]; al_ = alpha[z_]; } b_ls_contrib2(dl_, al_, da_, db_); } while (0)
template <int RLMAX, bool SEP> __device__ __forceinline__ double b_linesearch(const BatchPlan &P, const BVec &V)
{
    const int tid = threadIdx.x, n = P.n, m = P.m, M2 = 2 * m, P2 = P.P2, lane = tid & 31, wid = tid >> 5;
    double *delta = V.K, *alpha = V.K + M2;
    unsigned long long *key = SEP ? (unsigned long long *)V.rp : (unsigned long long *)(V.K + 2 * M2);
    typedef typename std::conditional<SEP, unsigned short, int>::type idx_t;
    idx_t *idx = SEP ? (idx_t *)V.rdi : (idx_t *)(key + P2);
    double *red = V.red;
    int &kmin = s_ls_kmin;                      // file-scope statics: shared by all instantiations
    double (&wa)[BNW] = s_ls_wa, (&wb)[BNW] = s_ls_wb, &tau_sh = s_ls_tau;
    double s[4] = {0, 0, 0, 0};
    B_LOOP1
    for (int i = tid; i < m; i += BNT) {
        double tm = V.dy[i] * V.mu[i]; tm = tm * 0.5;
        s[0] += V.dy[i] * tm; s[2] += V.y[i] * tm;
    }
    B_LOOP1
    for (int i = tid; i < n; i += BNT) { s[1] += V.dx[i] * V.Qdx[i]; s[3] += V.dx[i] * V.df[i]; }
    const bool exact = (n <= LS_EXACT_MAX && M2 <= LS_EXACT_MAX);
    if (exact) {
        // tiny problems: every sum in the reference's scalar order (vec_prod, lin_alg.c:59-71)
        __syncthreads();
        if (tid < 4) {
            const bool mvec = (tid == 0 || tid == 2);
            const int len = mvec ? m : n;
            double prod = 0.0;
            int i = 0;
            #define LS_TERM(i_) (mvec ? ((tid == 0 ? V.dy[i_] : V.y[i_]) * ((V.dy[i_] * V.mu[i_]) * 0.5)) \
                                      : (V.dx[i_] * (tid == 1 ? V.Qdx[i_] : V.df[i_])))
            if (len >= 4) for (; i <= len - 4; i += 4) prod += (LS_TERM(i) + LS_TERM(i + 1) + LS_TERM(i + 2) + LS_TERM(i + 3));
            for (; i < len; i++) prod += LS_TERM(i);
            #undef LS_TERM
            red[tid] = prod;
        }
        __syncthreads();
        s[0] = red[0]; s[1] = red[1]; s[2] = red[2]; s[3] = red[3];
        __syncthreads();
    } else block_sum<4>(s, red);
    double eta = s[0]; eta += s[1]; eta *= 0.5;
    double beta = s[2]; beta += s[3]; beta *= 0.5;
    double r3[3] = {0, 0, 0};                  // aJ, bJ, nL
    // the breakpoints with t > 0 are compacted to the front of key / idx (ballot prefix sums), so that the
    // bitonic network only spans next_pow2(nL) slots instead of next_pow2(2m); (key, idx) is a total order,
    // the sorted prefix does not depend on where the pairs started
    const bool compact = P2 <= RLMAX * BNT;
    int *cw = s_ls_cw;
    unsigned long long kreg[RLMAX];
    int lpre[RLMAX];
#pragma unroll
    for (int r = 0; r < RLMAX; r++) {
        const int k = tid + r * BNT;
        kreg[r] = ~0ull; lpre[r] = 0;
        if (k < P2) {
            unsigned long long kk = ~0ull;
            if (k < M2) {                          // linesearch.c:28-39,82-114
                double dl, al;
                b_ls_da(V, m, k, dl, al);
                if (!SEP) { delta[k] = dl; alpha[k] = al; }
                const double t = b_quot(al, dl);
                const bool L = t > 0.0, Pp = dl > 0.0;
                if (L != Pp) { r3[0] += dl * dl; r3[1] += dl * al; }
                if (L) { r3[2] += 1.0; kk = (unsigned long long)__double_as_longlong(t); }
            }
            if (compact) {
                const unsigned bal = __ballot_sync(0xffffffffu, kk != ~0ull);
                kreg[r] = kk; lpre[r] = __popc(bal & ((1u << lane) - 1u));
                if (lane == 0) cw[r * BNW + wid] = __popc(bal);
            } else { key[k] = kk; idx[k] = (idx_t)(k < M2 ? k : 0); }
        } else if (compact && lane == 0) cw[r * BNW + wid] = 0;
    }
    B_LOOP1
    for (int k = tid + RLMAX * BNT; k < P2; k += BNT) {    // only when !compact (2m > RLMAX * BNT / 2 ... large m)
        unsigned long long kk = ~0ull;
        if (k < M2) {
            double dl, al;
            b_ls_da(V, m, k, dl, al);
            if (!SEP) { delta[k] = dl; alpha[k] = al; }
            const double t = b_quot(al, dl);
            const bool L = t > 0.0, Pp = dl > 0.0;
            if (L != Pp) { r3[0] += dl * dl; r3[1] += dl * al; }
            if (L) { r3[2] += 1.0; kk = (unsigned long long)__double_as_longlong(t); }
        }
        key[k] = kk; idx[k] = (idx_t)(k < M2 ? k : 0);
    }
    block_sum<3>(r3, red);
    const int nLt = (int)(r3[2] + 0.5);
    double aJ = r3[0], bJ = r3[1];
    if (exact) {
        if (tid == 0) {                        // vec_prod_ind, linesearch.c:190-199
            double pa = 0.0, pb = 0.0;
            for (int k = 0; k < M2; k++) {
                double dl, al;
                if (SEP) b_ls_da(V, m, k, dl, al); else { dl = delta[k]; al = alpha[k]; }
                const bool L = (al / dl) > 0.0, Pp = dl > 0.0;
                if (L != Pp) { pa += dl * dl; pb += dl * al; }
            }
            wa[0] = pa; wb[0] = pb;
        }
        __syncthreads();
        aJ = wa[0]; bJ = wb[0];
        __syncthreads();
    }
    const double a0 = eta + aJ, b0 = beta - bJ;
    if (tid == 0) kmin = nLt;
    int P2e = P2;
    if (compact) {
        P2e = 2;
        while (P2e < nLt) P2e <<= 1;
        int base = 0;
#pragma unroll
        for (int r = 0; r < RLMAX; r++) {
            if (r * BNT < P2) {
                int mine = base;
                for (int w = 0; w < BNW; w++) { const int c = cw[r * BNW + w]; if (w < wid) mine += c; base += c; }
                if (kreg[r] != ~0ull) { key[mine + lpre[r]] = kreg[r]; idx[mine + lpre[r]] = (idx_t)(tid + r * BNT); }
            }
        }
        B_LOOP1
        for (int k = nLt + tid; k < P2e; k += BNT) { key[k] = ~0ull; idx[k] = 0; }
    }
    __syncthreads();
    for (int size = 2; size <= P2e; size <<= 1)
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
            B_LOOP1
            for (int t = tid; t < (P2e >> 1); t += BNT) {
                const int lo = 2 * t - (t & (stride - 1)), hi = lo + stride;
                const bool up = ((lo & size) == 0);
                const unsigned long long ka = key[lo], kb = key[hi];
                const idx_t ia = idx[lo], ib = idx[hi];
                const bool gt = (ka > kb) || (ka == kb && ia > ib);
                if (gt == up) { key[lo] = kb; key[hi] = ka; idx[lo] = ib; idx[hi] = ia; }
            }
            __syncthreads();
        }
    if (exact) {                               // the reference's sequential sweep, linesearch.c:121-157
        if (tid == 0) {
            double a = a0, b = b0, tau = 0.0;
            bool done = false;
            if (nLt == 0 || b + a * __longlong_as_double((long long)key[0]) > 0.0) { tau = -b / a; done = true; }
            for (int i = 0; !done && i < nLt; i++) {
                double da, db; B_LS_CONTRIB(idx[i], da, db);
                a = a + da; b = b + db;
                if (i + 1 < nLt && b + a * __longlong_as_double((long long)key[i + 1]) > 0.0) { tau = -b / a; done = true; }
            }
            if (!done) tau = -b / a;
            tau_sh = tau;
        }
        __syncthreads();
        return tau_sh;
    }
    const int chunk = (P2 + BNT - 1) / BNT, k0 = tid * chunk;
    double la = 0, lb = 0;
    B_LOOP1
    for (int c = 0; c < chunk; c++) {
        const int k = k0 + c;
        if (k < nLt) { double da, db; B_LS_CONTRIB(idx[k], da, db); la += da; lb += db; }
    }
    double pa = la, pb = lb;
    for (int o = 1; o < 32; o <<= 1) {
        double ta = __shfl_up_sync(0xffffffffu, pa, o), tb = __shfl_up_sync(0xffffffffu, pb, o);
        if (lane >= o) { pa += ta; pb += tb; }
    }
    if (lane == 31) { wa[wid] = pa; wb[wid] = pb; }
    __syncthreads();
    double oa = 0, ob = 0;
    B_LOOP1
    for (int w = 0; w < wid; w++) { oa += wa[w]; ob += wb[w]; }
    double ra = a0 + (oa + pa - la), rb = b0 + (ob + pb - lb);
    double cand_a = 0, cand_b = 0; int cand_k = -1;
    B_LOOP1
    for (int c = 0; c < chunk; c++) {
        const int k = k0 + c;
        if (k < nLt) {
            const double t = __longlong_as_double((long long)key[k]);
            if (cand_k < 0 && rb + ra * t > 0.0) { cand_k = k; cand_a = ra; cand_b = rb; }
            double da, db; B_LS_CONTRIB(idx[k], da, db);
            ra += da; rb += db;
        }
    }
    if (cand_k >= 0) atomicMin(&kmin, cand_k);
    __syncthreads();
    if (cand_k >= 0 && cand_k == kmin) tau_sh = -cand_b / cand_a;
    if (kmin == nLt && tid == 0) {             // no crossing before the last breakpoint (linesearch.c:148-156)
        double ta = 0, tb = 0;
        for (int w = 0; w < BNW; w++) { ta += wa[w]; tb += wb[w]; }
        tau_sh = -(b0 + tb) / (a0 + ta);
    }
    __syncthreads();
    return tau_sh;
}

// ------------------------------------------------------------------------------- one instance
#ifdef B_UNROLL_VEC
#define B_VEC_PRAGMA
#else
#define B_VEC_PRAGMA _Pragma("unroll 1")     // one or two trips per thread: unrolled copies only cost instruction fetch
#endif
#define FOR_N(i) B_VEC_PRAGMA for (int i = threadIdx.x; i < n; i += BNT)
#define FOR_M(i) B_VEC_PRAGMA for (int i = threadIdx.x; i < m; i += BNT)

#define B_TICK(slot) if (io.timing) { const long long now_ = clock64(); tacc[slot] += now_ - tlast; tlast = now_; }
__device__ __forceinline__ void b_solve_instance(const BatchPlan &P, const BatchIO &io, const BVec &V, const int b)
{
    const int n = P.n, m = P.m, tid = threadIdx.x;
    const bool scaled = P.scaling > 0, proximal = P.proximal != 0;
    double *red = V.red;
    long long tacc[8] = {0, 0, 0, 0, 0, 0, 0, 0}, tlast = io.timing ? clock64() : 0;     // setup, residuals+outer, newton prep, assemble, factor, backward, direction, linesearch+update
    // ---- load (qpdo_setup's deep copy, src/qpdo.c:100-109); infinite bounds clipped like qpdo.m:138-139
    for (int k = tid; k < P.nnzA; k += BNT) V.Av[k] = B_LDS_IN(io.Ax + (size_t)b * P.nnzA + k);
    for (int k = tid; k < P.nnzQ; k += BNT) V.Qv[k] = B_LDS_IN(io.Qx + (size_t)b * P.nnzQ + k);
    FOR_N(i) { V.q[i] = B_LDS_IN(io.q + (size_t)b * n + i); V.x[i] = 0.0; V.xbar[i] = 0.0; V.Qx[i] = 0.0; V.Aty[i] = 0.0; V.D[i] = 1.0; V.dx[i] = 0.0; }
    int bad_bounds = 0;
    FOR_M(i) {
        double lo = B_LDS_IN(io.l + (size_t)b * m + i), hi = B_LDS_IN(io.u + (size_t)b * m + i);
        bad_bounds |= lo > hi;                                        // validate_data (src/validate.c:19-28)
        V.l[i] = lo > -B_INFTY ? lo : -B_INFTY; V.u[i] = hi < B_INFTY ? hi : B_INFTY;
        V.y[i] = 0.0; V.ybar[i] = 0.0; V.Ax[i] = 0.0; V.E[i] = 1.0; V.om[i] = 0.0; V.dy[i] = 0.0; V.rpo[i] = 0.0;
        V.act[i] = 0; V.acto[i] = 0;
    }
    double c = 1.0, cinv = 1.0;
    if (__syncthreads_or(bad_bounds)) {                               // qpdo_setup would return NULL: status QPDO_ERROR, nothing solved
        if (io.x) { FOR_N(i) io.x[(size_t)b * n + i] = 0.0; }
        if (io.y) { FOR_M(i) io.y[(size_t)b * m + i] = 0.0; }
        if (io.active) { FOR_M(i) io.active[(size_t)b * m + i] = 0; }
        if (io.summary && tid == 0) {
            qpdo_batch_summary s;
            s.status_val = QPDO_ERROR; s.iterations = 0; s.oterations = 0; s.n_active = 0; s.objective = 0.0; s.res_prim_norm = 0.0; s.res_dual_norm = 0.0;
            io.summary[b] = s;
        }
        return;
    }
    // ---- scale_data (src/scaling.c:24-91): Ruiz sweeps on A, Q <- DQD, q <- Dq, cost scaling
    if (scaled) {
        for (int it = 0; it < P.scaling; it++) {
            FOR_N(j) { double mx = 0.0; for (int k = P.Ap[j]; k < P.Ap[j + 1]; k++) { double v = fabs(V.Av[k]); mx = v > mx ? v : mx; } V.tn[j] = mx; }
            FOR_M(i) { double mx = 0.0; for (int k = P.Rp[i]; k < P.Rp[i + 1]; k++) { double v = fabs(V.Av[P.Rsrc[k]]); mx = v > mx ? v : mx; } V.tm[i] = mx; }
            FOR_N(j) { double v = V.tn[j]; v = v < 1e-9 ? 1.0 : v; v = sqrt(v); v = 1.0 / v; V.tn[j] = v; V.D[j] = V.D[j] * v; }
            FOR_M(i) { double v = V.tm[i]; v = v < 1e-9 ? 1.0 : v; v = sqrt(v); v = 1.0 / v; V.tm[i] = v; V.E[i] = V.E[i] * v; }
            __syncthreads();
            FOR_N(j) { const double dj = V.tn[j]; for (int k = P.Ap[j]; k < P.Ap[j + 1]; k++) { double v = V.Av[k] * V.tm[P.Ai[k]]; V.Av[k] = v * dj; } }
            __syncthreads();
        }
        FOR_N(j) { const double t = V.D[j]; for (int k = P.Qp[j]; k < P.Qp[j + 1]; k++) { double f = t * V.D[P.Qi[k]]; V.Qv[k] = V.Qv[k] * f; } }
        double nq[1] = {0.0};
        FOR_N(i) { double v = V.D[i] * V.q[i]; V.q[i] = v; double t = 0.0 + 1.0 * v; t = fabs(t); nq[0] = t > nq[0] ? t : nq[0]; }
        block_max<1>(nq, red);
        c = 1.0 / (nq[0] > 1.0 ? nq[0] : 1.0);
        cinv = 1.0 / c;
        FOR_N(i) V.q[i] = V.q[i] * c;
        for (int k = tid; k < P.nnzQ; k += BNT) V.Qv[k] = V.Qv[k] * c;
        FOR_M(i) { V.l[i] = V.E[i] * V.l[i]; V.u[i] = V.E[i] * V.u[i]; }
        __syncthreads();
    }
    // ---- qpdo_warm_start (src/qpdo.c:217-299)
    double sigma = P.sigma_init;
    if (io.x0) {
        FOR_N(i) { double v = io.x0[(size_t)b * n + i]; if (scaled) v = v * (1.0 / V.D[i]); V.x[i] = v; V.xbar[i] = v; V.dx[i] = v; }
        __syncthreads();
        b_spmv_Q(P, V.Qv, V.dx, V.Qdx);
        b_spmv_A(P, V.Av, V.dx, V.Adx);
        __syncthreads();
        FOR_N(i) V.Qx[i] = proximal ? V.Qdx[i] + sigma * V.x[i] : V.Qdx[i];
        FOR_M(i) V.Ax[i] = V.Adx[i];
    }
    if (io.y0) {
        FOR_M(i) { double v = io.y0[(size_t)b * m + i]; if (scaled) { v = v * (1.0 / V.E[i]); v = v * c; } V.y[i] = v; V.ybar[i] = v; V.dy[i] = v; }
        __syncthreads();
        b_spmv_At(P, V.Av, V.dy, V.Atdy);
        __syncthreads();
        FOR_N(i) V.Aty[i] = V.Atdy[i];
    }
    __syncthreads();
    {   // initialize_mu (src/iteration.c:98-122)
        double d2[2] = {0.0, 0.0};
        FOR_N(i) { d2[0] += V.x[i] * V.Qx[i]; d2[1] += V.q[i] * V.x[i]; }
        block_sum<2>(d2, red);
        const double f = 0.5 * d2[0] + d2[1];
        FOR_M(i) {
            double z = bclamp(V.Ax[i], V.l[i], V.u[i]);
            double t = V.Ax[i] + (-1.0) * z;
            double num = 0.5 * t * t; num = num > 1.0 ? num : 1.0;
            double den = fabs(f); den = den > 1.0 ? den : 1.0;
            double v = 0.1 * num / den; v = v < 1e3 ? v : 1e3; v = v > 1e-3 ? v : 1e-3;
            V.mu[i] = v; V.smu[i] = 1.0 / sqrt(v);
        }
    }
    const double sqrt_mu_min = 1.0 / sqrt(P.mu_min);
    // ---- qpdo_solve (src/qpdo.c:304-476)
    double eps_in = P.eps_abs_in, tau = 0.0;
    sigma = P.sigma_init;
    int reset_newton = 1, iter, oter = 0, iter_old = 0, status = QPDO_UNSOLVED, n_active = 0;
    bool k_valid = false;                       // the L area holds the factor of K(omega, sigma_f)
    double sigma_f = 0.0;
    B_TICK(0)
    double rpn = 0, rdn = 0, rpin = 0, rdin = 0;
    __syncthreads();
    for (iter = 0; iter < P.max_iter; iter++) {
        {   // compute_outer/inner_residuals (+_norm): iteration.c:30-93, termination.c:35-77
            double nr[4] = {0, 0, 0, 0};
            FOR_M(i) {
                const double Ax = V.Ax[i], y = V.y[i], l = V.l[i], u = V.u[i], mu = V.mu[i], yb = V.ybar[i];
                double t;
                if (scaled) { const double E = V.E[i]; t = E * y; t = t * cinv; t = E * t; t = Ax + t; } else t = Ax + y;
                t = bclamp(t, l, u);
                const double rp = Ax + (-1.0) * t;
                V.rp[i] = rp;
                const double w = Ax + mu * (yb - 0.5 * y);
                V.w[i] = w;
                const double z = bclamp(w, l, u);
                const double rpi = Ax + mu * (yb - y) - z;
                V.rpi[i] = rpi;
                const double Einv = scaled ? b_rcp(V.E[i]) : 1.0;
                double a = fabs(scaled ? Einv * rp : rp), bb = fabs(scaled ? Einv * rpi : rpi);
                nr[0] = a > nr[0] ? a : nr[0]; nr[1] = bb > nr[1] ? bb : nr[1];
            }
            FOR_N(i) {
                double df = V.Qx[i] + V.q[i], rd;
                if (proximal) { rd = df + (-sigma) * V.x[i]; rd = rd + V.Aty[i]; } else rd = df + V.Aty[i];
                if (proximal) df = df + (-sigma) * V.xbar[i];
                V.df[i] = df;
                const double rdi = df + V.Aty[i];
                V.rdi[i] = rdi;
                const double Dinv = scaled ? b_rcp(V.D[i]) : 1.0;
                double a = fabs(scaled ? Dinv * rd : rd), bb = fabs(scaled ? Dinv * rdi : rdi);
                nr[2] = a > nr[2] ? a : nr[2]; nr[3] = bb > nr[3] ? bb : nr[3];
            }
            block_max<4>(nr, red);
            rpn = nr[0]; rpin = nr[1]; rdn = nr[2]; rdin = nr[3];
            if (scaled) { rdn = rdn * cinv; rdin = rdin * cinv; }
        }
        // check_outer_optimality (termination.c:11-23)
        if (rpn > B_INFTY || rdn > B_INFTY) { status = QPDO_NON_CVX; break; }
        if (rpn <= P.eps_abs && rdn <= P.eps_abs) { status = QPDO_SOLVED; break; }
        if (((iter > iter_old + 1) && (rpin <= eps_in) && (rdin <= eps_in)) || (iter == iter_old + P.inner_max_iter)) {
            if (iter < iter_old + P.inner_max_iter) {
                if (P.eps_prim_inf > 0) {      // qpdo.c:370-379 + is_primal_infeasible (termination.c:97-151)
                    double nd[1] = {0.0};
                    FOR_M(i) { double v = V.y[i] + (-1.0) * V.ybar[i]; V.dy[i] = v; double t = fabs(scaled ? v * V.E[i] : v); nd[0] = t > nd[0] ? t : nd[0]; }
                    block_max<1>(nd, red);
                    b_spmv_At(P, V.Av, V.dy, V.Atdy);
                    const double en = P.eps_prim_inf * nd[0];
                    __syncthreads();
                    if (en != 0) {
                        double s1[1] = {0.0}, mx[1] = {0.0};
                        FOR_M(i) {
                            const double lim = scaled ? V.E[i] * 1e20 : 1e20, d = V.dy[i];
                            const double dp = d > 0.0 ? d : 0.0, dm = d < 0.0 ? d : 0.0;
                            s1[0] += (V.u[i] < lim) ? V.u[i] * dp : 0.0;
                            s1[0] += (V.l[i] > -lim) ? V.l[i] * dm : 0.0;
                        }
                        FOR_N(i) { double v = V.Atdy[i]; if (scaled) { v = b_rcp(V.D[i]) * v; V.Atdy[i] = v; } v = fabs(v); mx[0] = v > mx[0] ? v : mx[0]; }
                        block_sum<1>(s1, red);
                        block_max<1>(mx, red);
                        if (mx[0] <= en && s1[0] <= -en) {
                            status = QPDO_PRIMAL_INFEASIBLE;
                            if (scaled) { FOR_M(i) { double v = V.dy[i] * cinv; V.dy[i] = V.E[i] * v; } }
                            break;
                        }
                    }
                }
                if (P.eps_dual_inf > 0) {      // qpdo.c:381-392 + is_dual_infeasible (termination.c:156-216)
                    double nd[1] = {0.0};
                    FOR_N(i) { double v = V.x[i] + (-1.0) * V.xbar[i]; V.dx[i] = v; double t = fabs(scaled ? v * V.D[i] : v); nd[0] = t > nd[0] ? t : nd[0]; }
                    block_max<1>(nd, red);
                    b_spmv_Q(P, V.Qv, V.dx, V.Qdx);
                    b_spmv_A(P, V.Av, V.dx, V.Adx);
                    const double en = P.eps_dual_inf * nd[0];
                    __syncthreads();
                    if (en != 0) {
                        int viol = 0, d1, d2;
                        FOR_M(k) {
                            double a = V.Adx[k];
                            if (scaled) { a = b_rcp(V.E[k]) * a; V.Adx[k] = a; }
                            const double lim = scaled ? V.E[k] * 1e20 : 1e20;
                            if ((V.u[k] < lim && a >= en) || (V.l[k] > -lim && a <= -en)) viol = 1;
                        }
                        block_count3(viol, 0, 0, viol, d1, d2, red);
                        if (!viol) {
                            double mx[1] = {0.0}, dt[1] = {0.0};
                            const double sv = sigma * (-tau);
                            FOR_N(i) {
                                double v = V.Qdx[i];
                                if (proximal) { v = v + sv * V.dx[i]; V.Qdx[i] = v; }
                                v = fabs(v); mx[0] = v > mx[0] ? v : mx[0];
                                dt[0] += V.q[i] * V.dx[i];
                            }
                            block_max<1>(mx, red);
                            block_sum<1>(dt, red);
                            const double lim = scaled ? c * en : en;
                            if (mx[0] <= lim && dt[0] <= -lim) {
                                status = QPDO_DUAL_INFEASIBLE;
                                if (scaled) { FOR_N(i) V.dx[i] = V.D[i] * V.dx[i]; }
                                break;
                            }
                        }
                    }
                }
            }
            // outer update (qpdo.c:396-430)
            FOR_N(i) V.xbar[i] = V.x[i];
            FOR_M(i) V.ybar[i] = V.y[i];
            if (oter > 0 && rpn > P.eps_abs) {                         // update_mu (iteration.c:127-168)
                double nm[1] = {0.0};
                FOR_M(i) { double t = fabs(V.rp[i]); nm[0] = t > nm[0] ? t : nm[0]; }
                block_max<1>(nm, red);
                int changed_any = 0;
                FOR_M(k) {
                    const double r = fabs(V.rp[k]), ro = fabs(V.rpo[k]);
                    double thr = P.theta * ro; thr = P.eps_abs > thr ? P.eps_abs : thr;
                    double dw = 0.0;
                    if (r > thr) {
                        double t = P.delta * nm[0] / r; t = t < 1.0 ? t : 1.0;
                        const double mu_factor = 1.0 / t, mu_old = V.mu[k], mu_new = mu_old / mu_factor;
                        double s; int changed;
                        if (mu_new >= P.mu_min) { changed = (mu_old != mu_new); V.mu[k] = mu_new; s = sqrt(mu_factor); V.smu[k] = s * V.smu[k]; }
                        else { changed = (mu_old != P.mu_min); V.mu[k] = P.mu_min; s = sqrt_mu_min / V.smu[k]; V.smu[k] = sqrt_mu_min; }
                        if (changed) { const double f = sqrt(1.0 - 1.0 / (s * s)); const double col = V.smu[k] * f; dw = col * col; changed_any++; }
                    }
                    V.tm[k] = dw;
                }
                int nch, d1, d2;
                {   // count over the CTA (a thread may own several rows)
                    double cs[1] = {(double)changed_any};
                    block_sum<1>(cs, red);
                    nch = (int)(cs[0] + 0.5); (void)d1; (void)d2;
                }
                if ((proximal && sigma > P.sigma_min) || nch > 25) reset_newton = 1;
                else if (nch) { FOR_M(k) V.om[k] = V.om[k] + V.tm[k]; k_valid = false; }      // ldlupdate_mu_changed (cholmod_interface.c:77-93): the factor no longer embodies omega
            }
            if (proximal && oter > 0 && rdn > P.eps_abs && sigma > P.sigma_min) {      // update_sigma (iteration.c:173-180)
                const double s_old = sigma;
                double s_new = s_old * P.sigma_upd; s_new = s_new > P.sigma_min ? s_new : P.sigma_min;
                sigma = s_new; reset_newton = 1;
                const double ds = s_new - s_old;
                FOR_N(i) V.Qx[i] = V.Qx[i] + ds * V.x[i];
            }
            if (iter < iter_old + P.inner_max_iter) { const double e1 = P.rho * eps_in, e2 = 0.1 * P.eps_abs; eps_in = e1 > e2 ? e1 : e2; }
            FOR_M(i) V.rpo[i] = V.rp[i];
            oter++; iter_old = iter;
            __syncthreads();
        } else {
            // ---- update_iterate (iteration.c:11-25): newton_direction (newton.c:13-91)
            B_TICK(1)
            if (P.reset_newton_iter != 0 && iter % P.reset_newton_iter == 0) reset_newton = 1;
            int ca = 0, ce = 0, cl = 0, n_enter, n_leave;
            FOR_M(i) {
                const double wi = V.w[i];
                const int a = (wi <= V.l[i]) || (wi >= V.u[i]), o = V.acto[i];
                V.act[i] = (unsigned char)a;
                ca += a; ce += (a && !o); cl += (!a && o);
            }
            {   // integer counts: one redux.sync.add per count inside the warp, warp partials through shared memory
                const int wa_ = __reduce_add_sync(0xffffffffu, ca), we_ = __reduce_add_sync(0xffffffffu, ce), wl_ = __reduce_add_sync(0xffffffffu, cl);
                int *ri = (int *)red;
                __syncthreads();
                if ((tid & 31) == 0) { ri[(tid >> 5) * 3] = wa_; ri[(tid >> 5) * 3 + 1] = we_; ri[(tid >> 5) * 3 + 2] = wl_; }
                __syncthreads();
                n_active = n_enter = n_leave = 0;
                for (int w = 0; w < BNW; w++) { n_active += ri[w * 3]; n_enter += ri[w * 3 + 1]; n_leave += ri[w * 3 + 2]; }
            }
            int mode;
            if ((reset_newton && n_active) || (n_enter + n_leave) > QPDO_MAX_RANK_UPDATE) { mode = 0; reset_newton = 0; }
            else if (n_active) mode = 1;
            else mode = 2;
            int om_changed = 0;
            FOR_M(i) {      // row weights embodied in the factor + "dy" = (I+P) res_prim_in ./ mu (newton.c:37-40)
                const int a = V.act[i], o = V.acto[i];
                const double mui = V.mu[i], inv = b_rcp(mui);
                double om;
                if (mode == 0) om = a ? inv : 0.0;
                else if (mode == 1) { om = V.om[i]; if (a && !o) om = om + inv; if (!a && o) om = om - inv; }
                else om = 0.0;
                om_changed |= (om != V.om[i]);
                V.om[i] = om;
                V.acto[i] = (unsigned char)a;
                double v = b_div(V.rpi[i], mui, inv);
                if (!a) v = v * 2.0;
                V.dy[i] = v;
            }
            // K = Q + sigma I + A' diag(omega) A is the matrix of the last factorization when no row weight and not sigma
            // has changed since: the factor in the L area is still the one the reference would use (its factor is untouched
            // when no constraint enters or leaves, src/newton.c:24-33; a refactorization of the same matrix gives the same
            // bits): assembly and LDL' are skipped, only the two triangular solves run.
            const bool refactor = __syncthreads_or(om_changed) || !k_valid || sigma != sigma_f;
            b_spmv_At(P, V.Av, V.dy, V.Atdy);
            B_TICK(2)
            if (refactor) b_assemble(P, V.Av, V.Qv, V.om, sigma, V.K);            // ends with __syncthreads
            else __syncthreads();
            FOR_N(k) { const int o = P.perm[k]; V.tn[k] = -V.rdi[o] - V.Atdy[o]; }     // P * rhs (newton.c:43-45)
            __syncthreads();
            B_TICK(3)
            if (P.mode == 1) {
#ifndef B_EXACT_PIVOTS
                if (!refactor) { if (tid < 32) b_forward_sn(P, V.K, V.tn); __syncthreads(); }
                else
#endif
                {
                    b_factor_fronts(P, V.K, V.tn, V.ucol);
                    k_valid = P.ls_sep != 0; sigma_f = sigma;         // L survives the step only when the linesearch scratch lives elsewhere
                }
                B_TICK(4)
                if (tid < 32) { if (P.max_front <= SN_NMAX && P.nsuper <= SN_DESC_MAX && P.nSrow <= SN_ROWS_MAX) b_backward_sn(P, V.K, V.tn); else b_backward_fronts(P, V.K, V.tn); }
                __syncthreads();
            } else {
                b_factor_forward(P, V.K, V.tn);
                b_backward(P, V.K, V.tn);
            }
            B_TICK(5)
            FOR_N(k) V.dx[P.perm[k]] = V.tn[k];
            __syncthreads();
            b_spmv_Q(P, V.Qv, V.dx, V.Qdx);
            b_spmv_A(P, V.Av, V.dx, V.Adx);
            __syncthreads();
            if (proximal) { FOR_N(i) V.Qdx[i] = V.Qdx[i] + sigma * V.dx[i]; }
            FOR_M(i) { if (V.act[i]) { const double mui = V.mu[i]; V.dy[i] = V.dy[i] + b_div(V.Adx[i], mui, b_rcp(mui)); } }
            __syncthreads();
            b_spmv_At(P, V.Av, V.dy, V.Atdy);
            __syncthreads();
            B_TICK(6)
            // breakpoint passes unrolled to what 2m needs (the unrolled code of unused passes costs instruction fetch)
            if (P.ls_sep) tau = P.P2 <= 2 * BNT ? b_linesearch<2, true>(P, V) : P.P2 <= 4 * BNT ? b_linesearch<4, true>(P, V) : b_linesearch<8, true>(P, V);
            else tau = P.P2 <= 2 * BNT ? b_linesearch<2, false>(P, V) : P.P2 <= 4 * BNT ? b_linesearch<4, false>(P, V) : b_linesearch<8, false>(P, V);
            FOR_N(i) { V.x[i] = V.x[i] + tau * V.dx[i]; V.Qx[i] = V.Qx[i] + tau * V.Qdx[i]; V.Aty[i] = V.Aty[i] + tau * V.Atdy[i]; }
            FOR_M(i) { V.y[i] = V.y[i] + tau * V.dy[i]; V.Ax[i] = V.Ax[i] + tau * V.Adx[i]; }
            __syncthreads();
            B_TICK(7)
        }
    }
    if (status == QPDO_UNSOLVED) status = QPDO_MAX_ITER_REACHED;
    // ---- store_solution (termination.c:82-92) + compute_objective (iteration.c:185-221)
    __syncthreads();
    double ob[1] = {0.0};
    FOR_N(i) { const double t = proximal ? 0.5 * (V.Qx[i] - V.x[i] * sigma) + V.q[i] : 0.5 * V.Qx[i] + V.q[i]; ob[0] += t * V.x[i]; }
    block_sum<1>(ob, red);
    double objective = ob[0];
    if (scaled) objective = objective * cinv;
    objective = objective + (io.c ? io.c[b] : 0.0);
    if (io.x) { FOR_N(i) B_STS_OUT(io.x + (size_t)b * n + i, scaled ? V.x[i] * V.D[i] : V.x[i]); }
    if (io.y) { FOR_M(i) { double v = V.y[i]; if (scaled) { v = v * cinv; v = v * V.E[i]; } B_STS_OUT(io.y + (size_t)b * m + i, v); } }
    if (io.active) { FOR_M(i) io.active[(size_t)b * m + i] = V.act[i]; }
    if (io.timing && tid == 0) { B_TICK(1) for (int j = 0; j < 8; j++) io.timing[(size_t)b * 8 + j] = tacc[j]; }
    if (io.summary && tid == 0) {
        qpdo_batch_summary s;
        s.status_val = status; s.iterations = iter; s.oterations = oter; s.n_active = n_active;
        s.objective = objective; s.res_prim_norm = rpn; s.res_dual_norm = rdn;
        io.summary[b] = s;
    }
}

// ------------------------------------------------------------------------------- kernel
__global__ void __launch_bounds__(BNT, B_MINCTAS) k_batch_solve(const B_GC BatchPlan P, const B_GC BatchIO io)
{
    extern __shared__ double sm[];
    __shared__ int s_b;
    const int n = P.n, m = P.m;
    BVec V;
    double *p = sm;
    V.Av = p; p += P.nnzA; V.Qv = p; p += P.nnzQ;
#define TAKE(ptr, len) ptr = p; p += (len);
    TAKE(V.q, n) TAKE(V.x, n) TAKE(V.xbar, n) TAKE(V.Qx, n) TAKE(V.Aty, n) TAKE(V.dx, n) TAKE(V.Qdx, n) TAKE(V.Atdy, n)
    TAKE(V.df, n) TAKE(V.rdi, n) TAKE(V.tn, n) TAKE(V.D, n)
    TAKE(V.l, m) TAKE(V.u, m) TAKE(V.y, m) TAKE(V.ybar, m) TAKE(V.Ax, m) TAKE(V.dy, m) TAKE(V.Adx, m) TAKE(V.w, m)
    TAKE(V.rpo, m) TAKE(V.mu, m) TAKE(V.smu, m) TAKE(V.om, m) TAKE(V.E, m)
    TAKE(V.rp, m) TAKE(V.rpi, m) TAKE(V.tm, m)              // [rp, rpi, tm, ucol]: dead during the linesearch, its key area when P.ls_sep
#undef TAKE
    V.ucol = p; p += 2 * SN_PW * P.ucolw;
    V.K = p; p += P.karea;
    V.red = p; p += BNW * 8;
    V.act = (unsigned char *)p; V.acto = V.act + m;
    if (P.nsuper <= SN_DESC_MAX) for (int t = threadIdx.x; t < 2 * P.nsuper; t += BNT) s_desc[t] = P.sn_desc[t];
    if (P.nSrow <= SN_ROWS_MAX) for (int t = threadIdx.x; t < P.nSrow; t += BNT) s_rows[t] = P.Srow[t];
    if (threadIdx.x < B_ROUNDS_MAX) {
        const int t = threadIdx.x;
        if (t < P.ellA.nround && P.ellA.nround <= B_ROUNDS_MAX) s_round[0][t] = P.ellA.round[t];
        if (t < P.ellAt.nround && P.ellAt.nround <= B_ROUNDS_MAX) s_round[1][t] = P.ellAt.round[t];
        if (t < P.ellQ.nround && P.ellQ.nround <= B_ROUNDS_MAX) s_round[2][t] = P.ellQ.round[t];
        if (t < P.nKround && P.nKround <= B_ROUNDS_MAX) s_round[3][t] = P.Kround[t];
    }
    for (;;) {
        __syncthreads();
        if (threadIdx.x == 0) s_b = atomicAdd(io.next, 1);
        __syncthreads();
        if (s_b >= io.count) break;
        const int b = io.order ? io.order[s_b] : s_b;
        b_solve_instance(P, io, V, b);
    }
    // Queue order for the NEXT solve of the same instances: longest first, by the iteration counts this launch has
    // just produced (counting sort; ties in arbitrary order -- the order only schedules, results do not depend on it).
    // A receding-horizon loop re-solves similar problems, and the ragged tail of a launch is set by where the longest
    // instances start.  Done by the last CTA to leave the queue (a separate kernel could not become resident beside
    // the persistent CTAs of the other stream's launch: they hold every register of the SM).
    if (io.order_out) {
        __threadfence();
        __syncthreads();
        if (threadIdx.x == 0) s_b = atomicAdd(io.done, 1);
        __syncthreads();
        if (s_b == (int)gridDim.x - 1) {
            __threadfence();
            int *bins = (int *)sm;                                   // B_ORDER_BINS counters, then the scan partials
            for (int t = threadIdx.x; t < B_ORDER_BINS + BNT; t += BNT) bins[t] = 0;
            __syncthreads();
            for (int i = threadIdx.x; i < io.count; i += BNT) {
                int it = ((volatile const qpdo_batch_summary *)io.summary)[i].iterations; it = it < 0 ? 0 : it >= B_ORDER_BINS ? B_ORDER_BINS - 1 : it;
                atomicAdd(&bins[B_ORDER_BINS - 1 - it], 1);          // descending
            }
            __syncthreads();
            const int per = B_ORDER_BINS / BNT, t0 = threadIdx.x * per;
            int acc = 0;
            for (int t = 0; t < per; t++) acc += bins[t0 + t];
            bins[B_ORDER_BINS + threadIdx.x] = acc;
            __syncthreads();
            int off = 0;
            for (int w = 0; w < (int)threadIdx.x; w++) off += bins[B_ORDER_BINS + w];
            for (int t = 0; t < per; t++) { const int c = bins[t0 + t]; bins[t0 + t] = off; off += c; }
            __syncthreads();
            for (int i = threadIdx.x; i < io.count; i += BNT) {
                int it = ((volatile const qpdo_batch_summary *)io.summary)[i].iterations; it = it < 0 ? 0 : it >= B_ORDER_BINS ? B_ORDER_BINS - 1 : it;
                io.order_out[atomicAdd(&bins[B_ORDER_BINS - 1 - it], 1)] = i;
            }
        }
    }
}

// =============================================================================== host side
struct qpdo_batch {
    int n = 0, m = 0, nnzA = 0, nnzQ = 0, capacity = 0, device = 0, lnz = 0, mode = 0;
    BatchPlan P;
    size_t smem = 0;
    int ctas = 0;
    cudaStream_t st = nullptr, st2 = nullptr;     // st2: second lane of the chunked host-buffer pipeline
    std::vector<void *> allocs;
    // staging for the host-buffer entry point
    double *dQx = nullptr, *dAx = nullptr, *dq = nullptr, *dl = nullptr, *du = nullptr, *dc = nullptr, *dx0 = nullptr, *dy0 = nullptr;
    double *dx = nullptr, *dy = nullptr;
    qpdo_batch_summary *dsum = nullptr;
    unsigned char *dact = nullptr;
    int *next = nullptr;
    int *dorder = nullptr;            // [capacity] queue order per launch range (relative indices), valid when order_count > 0
    int adaptive = 0, order_count = 0, order_chunks = 0;
    long long *timing = nullptr;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
};

template <class T> static int b_alloc(qpdo_batch *b, T **p, size_t count)
{
    void *v = nullptr;
    cudaError_t e = cudaMalloc(&v, sizeof(T) * (count ? count : 1));
    if (e != cudaSuccess) { qpdo_set_error("cudaMalloc(%zu): %s", sizeof(T) * count, cudaGetErrorString(e)); return -1; }
    b->allocs.push_back(v);
    *p = (T *)v;
    return 0;
}
template <class T> static int b_upload(qpdo_batch *b, const T **p, const std::vector<T> &h)
{
    T *d = nullptr;
    if (b_alloc(b, &d, h.size())) return -1;
    if (!h.empty()) CUDA_OK(cudaMemcpy(d, h.data(), sizeof(T) * h.size(), cudaMemcpyHostToDevice));
    *p = d;
    return 0;
}

extern "C" void qpdo_batch_destroy(qpdo_batch *b)
{
    if (!b) return;
    cudaSetDevice(b->device);
    if (b->st) cudaStreamSynchronize(b->st);
    if (b->st2) cudaStreamSynchronize(b->st2);
    for (void *p : b->allocs) cudaFree(p);
    if (b->e0) cudaEventDestroy(b->e0);
    if (b->e1) cudaEventDestroy(b->e1);
    if (b->st) cudaStreamDestroy(b->st);
    if (b->st2) cudaStreamDestroy(b->st2);
    delete b;
}

extern "C" int qpdo_batch_create(qpdo_batch **out, int n, int m, const int *Qp, const int *Qi, const int *Ap, const int *Ai,
                                 int capacity, const QPDOSettings *s, int device)
{
    *out = nullptr;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { qpdo_set_error("no CUDA device available: qpdo_b200 has no CPU fallback"); return -2; }
    if (n <= 0 || m < 0 || capacity <= 0 || !s) { qpdo_set_error("qpdo_batch_create: bad arguments"); return -3; }
    // validate_settings (reference src/validate.c:34-170)
    if (!(s->max_iter > 0 && s->inner_max_iter > 0 && s->eps_abs > 0 && s->eps_abs_in > 0 && s->eps_prim_inf >= 0 && s->eps_dual_inf >= 0 &&
          s->rho > 0 && s->rho < 1 && s->theta > 0 && s->theta <= 1 && s->delta > 0 && s->delta < 1 && s->mu_min > 0 &&
          (s->proximal == 0 || s->proximal == 1) && s->sigma_init > 0 && s->sigma_upd > 0 && s->sigma_upd <= 1 &&
          !(s->sigma_min > s->sigma_init) && s->scaling >= 0 && s->reset_newton_iter >= 0)) {
        qpdo_set_error("qpdo_batch_create: settings validation returned failure");
        return -3;
    }
    if (device < 0) { if (cudaGetDevice(&device) != cudaSuccess) device = 0; }
    CUDA_OK(cudaSetDevice(device));
    char err[256];
    qpdo_symbolic *S = qpdo_symbolic_analyze(n, m, Qp, Qi, Ap, Ai, 1, err);
    if (!S) { qpdo_set_error("symbolic analysis failed: %s", err); return -1; }
    qpdo_batch *b = new qpdo_batch();
    b->n = n; b->m = m; b->nnzQ = Qp[n]; b->nnzA = Ap[n]; b->capacity = capacity; b->device = device;
    int rc = -1;
    do {
        const int nnzA = b->nnzA, nnzQ = b->nnzQ;
        // CSR of A, full symmetric CSR of Q (same construction as qpdo_dev_create)
        std::vector<int> Rp(m + 1, 0), Rj(nnzA), Rsrc(nnzA);
        for (int k = 0; k < nnzA; k++) Rp[Ai[k] + 1]++;
        for (int i = 0; i < m; i++) Rp[i + 1] += Rp[i];
        { std::vector<int> nx(Rp.begin(), Rp.end() - 1);
          for (int j = 0; j < n; j++) for (int k = Ap[j]; k < Ap[j + 1]; k++) { int t = nx[Ai[k]]++; Rj[t] = j; Rsrc[t] = k; } }
        std::vector<int> Fp(n + 1, 0), Fdiag(n, 0);
        for (int j = 0; j < n; j++) for (int k = Qp[j]; k < Qp[j + 1]; k++) { int i = Qi[k]; Fp[i + 1]++; Fdiag[i]++; if (i != j) Fp[j + 1]++; }
        for (int i = 0; i < n; i++) Fp[i + 1] += Fp[i];
        std::vector<int> Fj(Fp[n]), Fsrc(Fp[n]);
        { std::vector<int> nx(Fp.begin(), Fp.end() - 1);
          for (int j = 0; j < n; j++) for (int k = Qp[j]; k < Qp[j + 1]; k++) { int i = Qi[k]; int t = nx[i]++; Fj[t] = j; Fsrc[t] = k; }
          for (int j = 0; j < n; j++) for (int k = Qp[j]; k < Qp[j + 1]; k++) { int i = Qi[k]; if (i != j) { int t = nx[j]++; Fj[t] = i; Fsrc[t] = k; } } }
        // pattern of L from the supernodes: column j of supernode [f, l]: rows j+1..l, then the supernode's rows
        std::vector<int> Lcp(n + 1, 0), Lri;
        for (int sn = 0; sn < S->nsuper; sn++) {
            const int f = S->sn_start[sn], l = S->sn_start[sn + 1] - 1;
            for (int j = f; j <= l; j++) {
                for (int i = j + 1; i <= l; i++) Lri.push_back(i);
                for (int64_t t = S->sn_rowptr[sn]; t < S->sn_rowptr[sn + 1]; t++) Lri.push_back(S->rows[t]);
                Lcp[j + 1] = (int)Lri.size();
            }
        }
        b->lnz = (int)Lri.size() + n;
        std::vector<int> Lrp(n + 1, 0), Lci(Lri.size());
        for (int r : Lri) Lrp[r + 1]++;
        for (int i = 0; i < n; i++) Lrp[i + 1] += Lrp[i];
        { std::vector<int> nx(Lrp.begin(), Lrp.end() - 1);
          for (int j = 0; j < n; j++) for (int t = Lcp[j]; t < Lcp[j + 1]; t++) Lci[nx[Lri[t]]++] = j; }
        // supernodal register path (mode 1) when every front fits one warp; QPDO_BATCH_FACTOR=columns forces
        // the column-by-column path (mode 0) for A/B tests
        int mode = 1, max_front = 1;
        for (int sn = 0; sn < S->nsuper; sn++) {
            const int nc = S->sn_start[sn + 1] - S->sn_start[sn];
            const int N = nc + S->sn_nr[sn];
            if ((N - 1) * (N + 2) / 2 > BNT * SN_EPT || N > 254) mode = 0;
            max_front = std::max(max_front, N);
        }
        { const char *ev = getenv("QPDO_BATCH_FACTOR"); if (ev && !strcmp(ev, "columns")) mode = 0; }
        std::vector<int> sn_f(S->nsuper), sn_nc(S->nsuper), sn_nr(S->nsuper), sn_base(S->nsuper), Sptr(S->nsuper + 1, 0), Srow, Uptr(S->nsuper + 1, 0), Upos;
        {
            int off = 0;
            for (int sn = 0; sn < S->nsuper; sn++) {
                sn_f[sn] = S->sn_start[sn]; sn_nc[sn] = S->sn_start[sn + 1] - S->sn_start[sn]; sn_nr[sn] = S->sn_nr[sn];
                sn_base[sn] = off;
                const int N = sn_nc[sn] + sn_nr[sn];
                off += sn_nc[sn] * N - sn_nc[sn] * (sn_nc[sn] - 1) / 2;
                for (int64_t t = S->sn_rowptr[sn]; t < S->sn_rowptr[sn + 1]; t++) Srow.push_back(S->rows[t]);
                Sptr[sn + 1] = (int)Srow.size();
            }
        }
        // position of entry (row i, col k), i >= k (permuted indices) in the trapezoids; -1 if outside the pattern of L
        auto trap_pos = [&](int i, int k) -> int {
            const int sn = S->col2sn[k], f = sn_f[sn], nc = sn_nc[sn], N = nc + sn_nr[sn], jj = k - f;
            int fr;
            if (i < f + nc) fr = i - f;
            else {
                const int *rb = Srow.data() + Sptr[sn], *re = Srow.data() + Sptr[sn + 1];
                const int *it = std::lower_bound(rb, re, i);
                if (it == re || *it != i) return -1;
                fr = nc + (int)(it - rb);
            }
            return sn_base[sn] + jj * N - jj * (jj - 1) / 2 - jj + fr;
        };
        bool pattern_ok = true;
        for (int sn = 0; sn < S->nsuper; sn++) {
            const int nr = sn_nr[sn];
            const int *rows = Srow.data() + Sptr[sn];
            for (int jb = 0; jb < nr; jb++) for (int ib = 0; ib < nr; ib++) {
                int pos = 0;
                if (ib >= jb) { pos = trap_pos(rows[ib], rows[jb]); if (pos < 0) { pattern_ok = false; pos = 0; } }
                Upos.push_back(pos);
            }
            Uptr[sn + 1] = (int)Upos.size();
        }
        if (!pattern_ok) mode = 0;
        // owned entries of every front with their final shared-memory positions, supernode descriptors
        std::vector<int2> Emap2;
        std::vector<int4> sn_desc(2 * (size_t)S->nsuper);
        for (int sn = 0; sn < S->nsuper; sn++) {
            const int nc = sn_nc[sn], nr = sn_nr[sn], N = nc + nr, f = sn_f[sn];
            const int *rows = Srow.data() + Sptr[sn];
            const int e0 = (int)Emap2.size();
            for (int j = 1; j < N; j++) for (int i = j; i <= N; i++) {
                int2 em; em.x = i | (j << 8);
                if (i == N) em.y = ~(j < nc ? f + j : rows[j - nc]);
                else if (j < nc) em.y = sn_base[sn] + j * N - j * (j - 1) / 2 - j + i;
                else em.y = Upos[Uptr[sn] + (j - nc) * nr + (i - nc)];
                Emap2.push_back(em);
            }
            const int ne_real = (int)Emap2.size() - e0;
            if (N < 32) {        // small fronts: padded to SN_EPTS * BNT entries (i = j = 0 never updates), so the kernel loads its four without bounds tests
                int2 pad; pad.x = 0; pad.y = 0;
                while ((int)Emap2.size() - e0 < SN_EPTS * BNT) Emap2.push_back(pad);
            }
            int4 d0, d1;
            d0.x = f; d0.y = nc; d0.z = nr; d0.w = sn_base[sn];
            d1.x = Sptr[sn]; d1.y = e0; d1.z = ne_real; d1.w = 0;
            sn_desc[2 * sn] = d0; sn_desc[2 * sn + 1] = d1;
        }
        // K entries -> packed positions
        const int nnzK = (int)S->nnzK;
        int P_nKround = 0;
        std::vector<int> Kpos(nnzK), Kq(S->Kq, S->Kq + nnzK), Ktp(nnzK + 1);
        std::vector<unsigned char> Kdiag(nnzK);
        for (int k = 0; k < n && mode == 1; k++) for (int e = S->Kp[k]; e < S->Kp[k + 1]; e++)
            if (trap_pos(S->Ki[e], k) < 0) mode = 0;
        for (int k = 0; k < n; k++) for (int e = S->Kp[k]; e < S->Kp[k + 1]; e++) {
            const int i = S->Ki[e];
            Kpos[e] = mode == 1 ? trap_pos(i, k) : k * n - ((k * (k - 1)) >> 1) - k + i;
            Kdiag[e] = (i == k);
        }
        for (int e = 0; e <= nnzK; e++) Ktp[e] = (int)S->Ktp[e];
        // assembly program: entries sorted by product count, dealt in rounds of BNT, products slot-major per round
        std::vector<int2> Kround, Kmeta(nnzK);
        std::vector<uint2> Ktr;
        {
            std::vector<int> order(nnzK);
            for (int e = 0; e < nnzK; e++) order[e] = e;
            std::stable_sort(order.begin(), order.end(), [&](int x, int y) { return Ktp[x + 1] - Ktp[x] > Ktp[y + 1] - Ktp[y]; });
            const int nround = (nnzK + BNT - 1) / BNT;
            for (int r = 0; r < nround; r++) {
                int kmax = 0;
                for (int t = 0; t < BNT && r * BNT + t < nnzK; t++) { const int e = order[r * BNT + t]; kmax = std::max(kmax, Ktp[e + 1] - Ktp[e]); }
                int2 rk; rk.x = (int)(Ktr.size() / BNT); rk.y = kmax;
                Kround.push_back(rk);
                const size_t base = Ktr.size();
                uint2 pad; pad.x = 0xffffffffu; pad.y = 0;
                Ktr.resize(base + (size_t)kmax * BNT, pad);
                for (int t = 0; t < BNT && r * BNT + t < nnzK; t++) {
                    const int e = order[r * BNT + t];
                    for (int k = 0; k < Ktp[e + 1] - Ktp[e]; k++) {
                        const int64_t tt = S->Ktp[e] + k;
                        uint2 w; w.x = (unsigned)S->Kt_row[tt]; w.y = (unsigned)S->Kt_a1[tt] | ((unsigned)S->Kt_a2[tt] << 16);
                        Ktr[base + (size_t)k * BNT + t] = w;
                    }
                    int2 mt; mt.x = Kpos[e] | (Kdiag[e] ? (int)0x80000000 : 0); mt.y = Kq[e];
                    Kmeta[r * BNT + t] = mt;
                }
            }
            P_nKround = nround;
        }
        // gather programs of the three SpMVs (same term order per row as the CSR/CSC loops)
        std::vector<int2> ellRound[3]; std::vector<unsigned> ellRec[3];
        auto build_ell = [&](int which, int nrows, const std::vector<int> &ptr, const int *col, const int *src, const int *split) {
            const int nround = (nrows + BNT - 1) / BNT;
            for (int r = 0; r < nround; r++) {
                int kmax = 0;
                for (int t = 0; t < BNT && r * BNT + t < nrows; t++) kmax = std::max(kmax, ptr[r * BNT + t + 1] - ptr[r * BNT + t]);
                int2 rk; rk.x = (int)(ellRec[which].size() / BNT); rk.y = kmax;
                ellRound[which].push_back(rk);
                const size_t base = ellRec[which].size();
                ellRec[which].resize(base + (size_t)kmax * BNT, 0xffffffffu);
                for (int t = 0; t < BNT && r * BNT + t < nrows; t++) {
                    const int i = r * BNT + t;
                    for (int k = ptr[i]; k < ptr[i + 1]; k++) {
                        const unsigned sv = (unsigned)(src ? src[k] : k);
                        unsigned w = sv | ((unsigned)col[k] << 16);
                        if (split && k - ptr[i] >= split[i]) w |= 0x80000000u;
                        ellRec[which][base + (size_t)(k - ptr[i]) * BNT + t] = w;
                    }
                }
            }
        };
        if (nnzA >= 65536 || nnzQ >= 65536 || n >= 32768 || m >= 32768) { qpdo_set_error("qpdo_batch_create: problem does not fit the batch kernel's 16-bit index records"); break; }
        {
            std::vector<int> vAp0(Ap, Ap + n + 1);
            build_ell(0, m, Rp, Rj.data(), Rsrc.data(), nullptr);
            build_ell(1, n, vAp0, Ai, nullptr, nullptr);
            build_ell(2, n, Fp, Fj.data(), Fsrc.data(), Fdiag.data());
        }
        if (nnzA >= 65536) { qpdo_set_error("qpdo_batch_create: nnz(A) >= 65536 does not fit the batch kernel"); break; }
        std::vector<int> perm(S->perm, S->perm + n), vQp(Qp, Qp + n + 1), vQi(Qi, Qi + nnzQ), vAp(Ap, Ap + n + 1), vAi(Ai, Ai + nnzA);
        BatchPlan &P = b->P;
        memset(&P, 0, sizeof P);
        P.n = n; P.m = m; P.nnzA = nnzA; P.nnzQ = nnzQ; P.nnzK = nnzK;
        P.nKpack = mode == 1 ? b->lnz : n * (n + 1) / 2;
        P.nKround = P_nKround;
        P.mode = mode; P.nsuper = S->nsuper; b->mode = mode; P.max_front = max_front; P.ucolw = std::max(max_front + 2, SN_W); P.nSrow = (int)Srow.size();
        int P2 = 64; while (P2 < 2 * m) P2 <<= 1;
        P.P2 = P2;
        // linesearch scratch outside the L area when the dead vectors hold it (keys: [rp, rpi, tm, ucol]; 16-bit
        // indices: [rdi, tn]): L then survives the step and an unchanged K is not factored again
        P.ls_sep = (mode == 1 && max_front <= SN_NMAX && S->nsuper <= SN_DESC_MAX && (int)Srow.size() <= SN_ROWS_MAX &&
                    P2 <= 3 * m + 2 * SN_PW * P.ucolw && P2 <= 8 * n && !getenv("QPDO_BATCH_NO_LS_SEP")) ? 1 : 0;
#ifdef B_EXACT_PIVOTS
        P.ls_sep = 0;                                    // the exact variant keeps d on the diagonal and always refactors
#endif
        P.karea = P.ls_sep ? P.nKpack : std::max(P.nKpack, 4 * m + P2 + P2 / 2 + 2);
        P.ellA.nround = (int)ellRound[0].size(); P.ellAt.nround = (int)ellRound[1].size(); P.ellQ.nround = (int)ellRound[2].size();
        if (b_upload(b, &P.ellA.round, ellRound[0]) || b_upload(b, &P.ellA.rec, ellRec[0]) || b_upload(b, &P.ellAt.round, ellRound[1]) ||
            b_upload(b, &P.ellAt.rec, ellRec[1]) || b_upload(b, &P.ellQ.round, ellRound[2]) || b_upload(b, &P.ellQ.rec, ellRec[2])) break;
#ifndef B_NO_KTR32
        if (m < 1023 && nnzA <= 2048) {
            std::vector<unsigned> Ktr32(Ktr.size());
            for (size_t t = 0; t < Ktr.size(); t++)
                Ktr32[t] = (int)Ktr[t].x < 0 ? 0xffffffffu : (Ktr[t].x | ((Ktr[t].y & 0xffffu) << 10) | ((Ktr[t].y >> 16) << 21));
            if (b_upload(b, &P.Ktr32, Ktr32)) break;
        }
#endif
        if (b_upload(b, &P.Ap, vAp) || b_upload(b, &P.Ai, vAi) || b_upload(b, &P.Rp, Rp) || b_upload(b, &P.Rj, Rj) || b_upload(b, &P.Rsrc, Rsrc) ||
            b_upload(b, &P.Qp, vQp) || b_upload(b, &P.Qi, vQi) || b_upload(b, &P.Fp, Fp) || b_upload(b, &P.Fj, Fj) || b_upload(b, &P.Fsrc, Fsrc) ||
            b_upload(b, &P.Fdiag, Fdiag) || b_upload(b, &P.perm, perm) || b_upload(b, &P.Kround, Kround) || b_upload(b, &P.Kmeta, Kmeta) || b_upload(b, &P.Ktr, Ktr) ||
            b_upload(b, &P.sn_f, sn_f) || b_upload(b, &P.sn_nc, sn_nc) || b_upload(b, &P.sn_nr, sn_nr) || b_upload(b, &P.sn_base, sn_base) ||
            b_upload(b, &P.Sptr, Sptr) || b_upload(b, &P.Srow, Srow) || b_upload(b, &P.Uptr, Uptr) || b_upload(b, &P.Upos, Upos) ||
            b_upload(b, &P.Emap2, Emap2) || b_upload(b, &P.sn_desc, sn_desc) ||
            b_upload(b, &P.Lcp, Lcp) || b_upload(b, &P.Lri, Lri) || b_upload(b, &P.Lrp, Lrp) || b_upload(b, &P.Lci, Lci))
            break;
        P.max_iter = (int)s->max_iter; P.inner_max_iter = (int)s->inner_max_iter; P.proximal = (int)s->proximal; P.scaling = (int)s->scaling;
        P.reset_newton_iter = (int)s->reset_newton_iter;
        P.eps_abs = s->eps_abs; P.eps_abs_in = s->eps_abs_in; P.eps_prim_inf = s->eps_prim_inf; P.eps_dual_inf = s->eps_dual_inf;
        P.rho = s->rho; P.theta = s->theta; P.delta = s->delta; P.mu_min = s->mu_min; P.sigma_init = s->sigma_init;
        P.sigma_upd = s->sigma_upd; P.sigma_min = s->sigma_min;
        b->smem = sizeof(double) * ((size_t)nnzA + nnzQ + 12 * (size_t)n + 16 * (size_t)m + P.karea + BNW * 8 + 2 * SN_PW * (size_t)P.ucolw) + 2 * (size_t)m + 16;
        b->smem = std::max(b->smem, sizeof(int) * (size_t)(B_ORDER_BINS + BNT));
        if (b->smem > 227 * 1024) {
            qpdo_set_error("qpdo_batch_create: a problem of this size needs %zu bytes of shared memory per CTA (limit 232448): use qpdo_setup/qpdo_solve", b->smem);
            break;
        }
        if (cudaFuncSetAttribute(k_batch_solve, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)b->smem) != cudaSuccess) { qpdo_set_error("cudaFuncSetAttribute failed"); break; }
        int per_sm = 0, sms = 0;
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_batch_solve, BNT, b->smem);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
        { const char *ev = getenv("QPDO_BATCH_CTAS_PER_SM"); if (ev && atoi(ev) > 0) per_sm = std::min(per_sm, atoi(ev)); }   // occupancy experiments
        b->ctas = std::max(1, per_sm) * std::max(1, sms);
        if (cudaStreamCreateWithFlags(&b->st, cudaStreamNonBlocking) != cudaSuccess || cudaStreamCreateWithFlags(&b->st2, cudaStreamNonBlocking) != cudaSuccess) { qpdo_set_error("cudaStreamCreate failed"); break; }
        cudaEventCreate(&b->e0); cudaEventCreate(&b->e1);
        const size_t cap = (size_t)capacity;
        if (b_alloc(b, &b->dQx, cap * nnzQ) || b_alloc(b, &b->dAx, cap * nnzA) || b_alloc(b, &b->dq, cap * n) || b_alloc(b, &b->dl, cap * m) ||
            b_alloc(b, &b->du, cap * m) || b_alloc(b, &b->dc, cap) || b_alloc(b, &b->dx0, cap * n) || b_alloc(b, &b->dy0, cap * m) ||
            b_alloc(b, &b->dx, cap * n) || b_alloc(b, &b->dy, cap * m) || b_alloc(b, &b->dsum, cap) || b_alloc(b, &b->dact, cap * m) ||
            b_alloc(b, &b->next, 4) || b_alloc(b, &b->dorder, cap))
            break;
        rc = 0;
    } while (0);
    qpdo_symbolic_free(S);
    if (rc) { qpdo_batch_destroy(b); return rc; }
    *out = b;
    return 0;
}

static int batch_launch(qpdo_batch *b, cudaStream_t st, int *next, int first, int count, bool use_order, const double *dQx, const double *dAx,
                        const double *dq, const double *dl, const double *du, const double *dc, const double *dx0, const double *dy0,
                        double *dx, double *dy, qpdo_batch_summary *dsummary, unsigned char *dactive)
{   // instances first .. first + count - 1 of the given (instance-major) device arrays
    const size_t o = (size_t)first, n = b->n, m = b->m;
    BatchIO io;
    io.Qx = dQx + o * b->nnzQ; io.Ax = dAx + o * b->nnzA; io.q = dq + o * n; io.l = dl + o * m; io.u = du + o * m;
    io.c = dc ? dc + o : nullptr; io.x0 = dx0 ? dx0 + o * n : nullptr; io.y0 = dy0 ? dy0 + o * m : nullptr;
    io.x = dx ? dx + o * n : nullptr; io.y = dy ? dy + o * m : nullptr; io.summary = dsummary ? dsummary + o : nullptr;
    io.active = dactive ? dactive + o * m : nullptr; io.count = count; io.next = next;
    io.timing = b->timing ? b->timing + o * 8 : nullptr;
    io.order = use_order ? b->dorder + o : nullptr;
    io.order_out = (b->adaptive && dsummary) ? b->dorder + o : nullptr;
    io.done = next + 2;                                                // next[0..1]: queue heads of the two streams, next[2..3]: their exit counters
    CUDA_OK(cudaMemsetAsync(next, 0, sizeof(int), st));
    CUDA_OK(cudaMemsetAsync(next + 2, 0, sizeof(int), st));
    k_batch_solve<<<std::min(count, b->ctas), BNT, b->smem, st>>>(b->P, io);
    CUDA_OK(cudaGetLastError());
    return 0;
}

extern "C" int qpdo_batch_solve_device(qpdo_batch *b, int count, const double *dQx, const double *dAx, const double *dq,
                                       const double *dl, const double *du, const double *dc, const double *dx0, const double *dy0,
                                       double *dx, double *dy, qpdo_batch_summary *dsummary, unsigned char *dactive, int sync,
                                       float *kernel_ms)
{
    if (!b || count < 0 || !dQx || !dAx || !dq || !dl || !du) { qpdo_set_error("qpdo_batch_solve_device: bad arguments"); return -3; }
    // the per-instance scratch of the library (queue order, phase accounting) is sized for `capacity` instances
    if (count > b->capacity) { qpdo_set_error("qpdo_batch_solve_device: count %d exceeds the capacity %d the batch was created with", count, b->capacity); return -3; }
    if (count == 0) return 0;
    CUDA_OK(cudaSetDevice(b->device));
    const bool use_order = b->adaptive && b->order_count == count && b->order_chunks == 1;
    if (kernel_ms) CUDA_OK(cudaEventRecord(b->e0, b->st));
    if (batch_launch(b, b->st, b->next, 0, count, use_order, dQx, dAx, dq, dl, du, dc, dx0, dy0, dx, dy, dsummary, dactive)) return -1;
    if (kernel_ms) CUDA_OK(cudaEventRecord(b->e1, b->st));
    if (b->adaptive && dsummary) { b->order_count = count; b->order_chunks = 1; } else b->order_count = 0;
    if (sync || kernel_ms) {
        CUDA_OK(cudaStreamSynchronize(b->st));
        if (kernel_ms) CUDA_OK(cudaEventElapsedTime(kernel_ms, b->e0, b->e1));
    }
    return 0;
}

// chunks of a host-buffer solve: enough of them that the first host->device and the last device->host copy are a
// small part of the call, few enough that every chunk still fills the resident CTAs several times over
static int batch_chunks(const qpdo_batch *b, int count)
{
    static const int forced = getenv("QPDO_BATCH_CHUNKS") ? atoi(getenv("QPDO_BATCH_CHUNKS")) : 0;                 // experiments
    if (forced > 0) return forced;
    int nchunk = count / (3 * b->ctas);
    return nchunk < 1 ? 1 : nchunk > 8 ? 8 : nchunk;
}

// issue (asynchronously, on the batch's two streams) the copies and launches of instances [0, count) of the given
// HOST arrays; batch_solve_wait joins them
static int batch_solve_issue(qpdo_batch *b, int count, const double *Qx, const double *Ax, const double *q, const double *l,
                             const double *u, const double *c, const double *x0, const double *y0, double *x, double *y,
                             qpdo_batch_summary *summary, unsigned char *active_set)
{
    CUDA_OK(cudaSetDevice(b->device));
    const size_t n = b->n, m = b->m, D = sizeof(double);
    // chunks alternate between two streams: the host->device copy of chunk k+1 and the device->host copy of
    // chunk k-1 overlap the kernel of chunk k (with pinned host buffers; pageable ones just serialise), and the
    // ragged tail of one chunk's launch is filled by the head of the next
    const int nchunk = batch_chunks(b, count);
    static const int host_order = getenv("QPDO_BATCH_HOST_ORDER") ? atoi(getenv("QPDO_BATCH_HOST_ORDER")) : 1;     // experiments
    const bool use_order = host_order && b->adaptive && b->order_count == count && b->order_chunks == nchunk;
    for (int k = 0; k < nchunk; k++) {
        const int lo = (int)((long)count * k / nchunk), hi = (int)((long)count * (k + 1) / nchunk), cn = hi - lo;
        if (cn <= 0) continue;
        cudaStream_t st = (k & 1) ? b->st2 : b->st;
        const size_t o = (size_t)lo;
        CUDA_OK(cudaMemcpyAsync(b->dQx + o * b->nnzQ, Qx + o * b->nnzQ, D * cn * b->nnzQ, cudaMemcpyHostToDevice, st));
        CUDA_OK(cudaMemcpyAsync(b->dAx + o * b->nnzA, Ax + o * b->nnzA, D * cn * b->nnzA, cudaMemcpyHostToDevice, st));
        CUDA_OK(cudaMemcpyAsync(b->dq + o * n, q + o * n, D * cn * n, cudaMemcpyHostToDevice, st));
        if (m) CUDA_OK(cudaMemcpyAsync(b->dl + o * m, l + o * m, D * cn * m, cudaMemcpyHostToDevice, st));
        if (m) CUDA_OK(cudaMemcpyAsync(b->du + o * m, u + o * m, D * cn * m, cudaMemcpyHostToDevice, st));
        if (c) CUDA_OK(cudaMemcpyAsync(b->dc + o, c + o, D * cn, cudaMemcpyHostToDevice, st));
        if (x0) CUDA_OK(cudaMemcpyAsync(b->dx0 + o * n, x0 + o * n, D * cn * n, cudaMemcpyHostToDevice, st));
        if (y0 && m) CUDA_OK(cudaMemcpyAsync(b->dy0 + o * m, y0 + o * m, D * cn * m, cudaMemcpyHostToDevice, st));
        if (batch_launch(b, st, b->next + (k & 1), lo, cn, use_order, b->dQx, b->dAx, b->dq, b->dl, b->du, c ? b->dc : nullptr, x0 ? b->dx0 : nullptr,
                         (y0 && m) ? b->dy0 : nullptr, b->dx, b->dy, b->dsum, b->dact))
            return -1;
        if (x) CUDA_OK(cudaMemcpyAsync(x + o * n, b->dx + o * n, D * cn * n, cudaMemcpyDeviceToHost, st));
        if (y && m) CUDA_OK(cudaMemcpyAsync(y + o * m, b->dy + o * m, D * cn * m, cudaMemcpyDeviceToHost, st));
        if (summary) CUDA_OK(cudaMemcpyAsync(summary + o, b->dsum + o, sizeof(qpdo_batch_summary) * cn, cudaMemcpyDeviceToHost, st));
        if (active_set && m) CUDA_OK(cudaMemcpyAsync(active_set + o * m, b->dact + o * m, (size_t)cn * m, cudaMemcpyDeviceToHost, st));
    }
    if (b->adaptive) { b->order_count = count; b->order_chunks = nchunk; } else b->order_count = 0;
    return 0;
}
static int batch_solve_wait(qpdo_batch *b)
{
    CUDA_OK(cudaSetDevice(b->device));
    cudaError_t e = cudaStreamSynchronize(b->st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(b->st2);
    if (e != cudaSuccess) { qpdo_set_error("batch kernel failed: %s", cudaGetErrorString(e)); return -1; }
    return 0;
}

extern "C" int qpdo_batch_solve(qpdo_batch *b, int count, const double *Qx, const double *Ax, const double *q, const double *l,
                                const double *u, const double *c, const double *x0, const double *y0, double *x, double *y,
                                qpdo_batch_summary *summary, unsigned char *active_set)
{
    if (!b || count < 0 || count > b->capacity || !Qx || !Ax || !q || !l || !u) { qpdo_set_error("qpdo_batch_solve: bad arguments (count > capacity?)"); return -3; }
    if (count == 0) return 0;
    if (batch_solve_issue(b, count, Qx, Ax, q, l, u, c, x0, y0, x, y, summary, active_set)) return -1;
    return batch_solve_wait(b);
}

// One process, several GPUs (SURVEY.md 8(b) "device/GPU-count selection", 8(e)): instance i goes to GPU floor(i * G / count),
// contiguous blocks, every block through that GPU's own chunked pipeline, all GPUs in flight at once.  No inter-GPU
// traffic: the summaries of all blocks land in the caller's one host array.
extern "C" int qpdo_batch_solve_multi(qpdo_batch *const *batches, int ngpu, int count, const double *Qx, const double *Ax,
                                      const double *q, const double *l, const double *u, const double *c, const double *x0,
                                      const double *y0, double *x, double *y, qpdo_batch_summary *summary, unsigned char *active_set)
{
    if (!batches || ngpu <= 0 || count < 0 || !Qx || !Ax || !q || !l || !u) { qpdo_set_error("qpdo_batch_solve_multi: bad arguments"); return -3; }
    for (int g = 0; g < ngpu; g++) {
        const qpdo_batch *b = batches[g];
        if (!b || b->n != batches[0]->n || b->m != batches[0]->m || b->nnzA != batches[0]->nnzA || b->nnzQ != batches[0]->nnzQ) { qpdo_set_error("qpdo_batch_solve_multi: batch %d is null or has another shape", g); return -3; }
        const long lo = ((long)count * g + ngpu - 1) / ngpu, hi = ((long)count * (g + 1) + ngpu - 1) / ngpu;
        if (hi - lo > b->capacity) { qpdo_set_error("qpdo_batch_solve_multi: block %d holds %ld instances, capacity %d", g, hi - lo, b->capacity); return -3; }
    }
    int rc = 0, issued = 0;
    for (int g = 0; g < ngpu && !rc; g++) {
        qpdo_batch *b = batches[g];
        const size_t lo = (size_t)(((long)count * g + ngpu - 1) / ngpu), hi = (size_t)(((long)count * (g + 1) + ngpu - 1) / ngpu), n = b->n, m = b->m;
        issued = g + 1;
        if (hi == lo) continue;
        rc = batch_solve_issue(b, (int)(hi - lo), Qx + lo * b->nnzQ, Ax + lo * b->nnzA, q + lo * n, l + lo * m, u + lo * m, c ? c + lo : nullptr,
                               x0 ? x0 + lo * n : nullptr, y0 ? y0 + lo * m : nullptr, x ? x + lo * n : nullptr, y ? y + lo * m : nullptr,
                               summary ? summary + lo : nullptr, active_set ? active_set + lo * m : nullptr);
    }
    for (int g = 0; g < issued; g++) if (batch_solve_wait(batches[g])) rc = -1;
    return rc;
}

extern "C" int qpdo_batch_set_adaptive_order(qpdo_batch *b, int on)
{
    if (!b) return -3;
    b->adaptive = on ? 1 : 0;
    b->order_count = 0;
    return 0;
}

extern "C" int qpdo_batch_timing(qpdo_batch *b, int count, double *cycles8)
{   // count == 0: switch the per-phase clock64 accounting on; count > 0: mean cycles per instance and phase of the last solve
    if (!b) return -3;
    CUDA_OK(cudaSetDevice(b->device));
    if (!b->timing) { if (b_alloc(b, &b->timing, (size_t)b->capacity * 8)) return -1; CUDA_OK(cudaMemset(b->timing, 0, sizeof(long long) * 8 * b->capacity)); }
    if (count <= 0 || !cycles8) return 0;
    if (count > b->capacity) count = b->capacity;
    std::vector<long long> h((size_t)count * 8);
    CUDA_OK(cudaStreamSynchronize(b->st));
    CUDA_OK(cudaMemcpy(h.data(), b->timing, sizeof(long long) * h.size(), cudaMemcpyDeviceToHost));
    for (int j = 0; j < 8; j++) { double sacc = 0; for (int i = 0; i < count; i++) sacc += (double)h[(size_t)i * 8 + j]; cycles8[j] = sacc / count; }
    return 0;
}

#ifdef B_FINE
extern "C" int qpdo_batch_fine(long long *out8) { return cudaMemcpyFromSymbol(out8, g_fine, sizeof(long long) * 8) == cudaSuccess ? 0 : -1; }
#endif
extern "C" int qpdo_batch_info(const qpdo_batch *b, int *smem_bytes, int *ctas, int *lnz, int *threads)
{
    if (!b) return -3;
    if (smem_bytes) *smem_bytes = (int)b->smem;
    if (ctas) *ctas = b->ctas;
    if (lnz) *lnz = b->lnz;
    if (threads) *threads = BNT;
    return 0;
}

// ---- arrays of the reference's own QPDOData (include/types.h:81-90) that share one pattern
static int data_to_csc(const cholmod_sparse *S, int lower, std::vector<int> &p, std::vector<int> &i, std::vector<double> &x)
{
    const bool il = (S->itype == QPDO_CHOLMOD_LONG);
    auto idx = [&](const void *a, size_t k) { return il ? (long)((const long *)a)[k] : (long)((const int *)a)[k]; };
    const size_t ncol = S->ncol;
    p.assign(ncol + 1, 0); i.clear(); x.clear();
    const double *Sx = (const double *)S->x;
    for (size_t j = 0; j < ncol; j++) {
        long b = idx(S->p, j), e = S->packed ? idx(S->p, j + 1) : b + idx(S->nz, j);
        long prev = -1;
        for (long k = b; k < e; k++) {
            long r = idx(S->i, (size_t)k);
            if (lower && r < (long)j) continue;
            if (r <= prev) return -1;              // unsorted or duplicate: not supported by the batch entry
            prev = r;
            i.push_back((int)r); x.push_back(Sx[k]);
        }
        p[j + 1] = (int)i.size();
    }
    return 0;
}

extern "C" int qpdo_batch_solve_data(const QPDOData *const *data, int count, const QPDOSettings *settings, double *const *x,
                                     double *const *y, qpdo_batch_summary *summary, int device)
{
    if (!data || count <= 0 || !settings) { qpdo_set_error("qpdo_batch_solve_data: bad arguments"); return -3; }
    const int n = (int)data[0]->n, m = (int)data[0]->m;
    std::vector<int> Qp, Qi, Ap, Ai, p2, i2;
    std::vector<double> v;
    std::vector<double> Qx, Ax, q((size_t)count * n), l((size_t)count * m), u((size_t)count * m), c(count);
    for (int k = 0; k < count; k++) {
        const QPDOData *d = data[k];
        if ((int)d->n != n || (int)d->m != m || d->Q->stype > 0) { qpdo_set_error("qpdo_batch_solve_data: instance %d has another shape (or upper-stored Q)", k); return -4; }
        for (size_t j = 0; j < (size_t)m; j++) if (d->l[j] > d->u[j]) { qpdo_set_error("instance %d: lower bound %zu above upper bound", k, j); return -3; }
        if (data_to_csc(d->Q, 1, p2, i2, v)) { qpdo_set_error("instance %d: Q must have sorted, duplicate-free columns", k); return -4; }
        if (k == 0) { Qp = p2; Qi = i2; Qx.reserve((size_t)count * v.size()); }
        else if (p2 != Qp || i2 != Qi) { qpdo_set_error("qpdo_batch_solve_data: instance %d does not share the pattern of Q", k); return -4; }
        Qx.insert(Qx.end(), v.begin(), v.end());
        if (data_to_csc(d->A, 0, p2, i2, v)) { qpdo_set_error("instance %d: A must have sorted, duplicate-free columns", k); return -4; }
        if (k == 0) { Ap = p2; Ai = i2; Ax.reserve((size_t)count * v.size()); }
        else if (p2 != Ap || i2 != Ai) { qpdo_set_error("qpdo_batch_solve_data: instance %d does not share the pattern of A", k); return -4; }
        Ax.insert(Ax.end(), v.begin(), v.end());
        memcpy(&q[(size_t)k * n], d->q, sizeof(double) * n);
        if (m) { memcpy(&l[(size_t)k * m], d->l, sizeof(double) * m); memcpy(&u[(size_t)k * m], d->u, sizeof(double) * m); }
        c[k] = d->c;
    }
    qpdo_batch *b = nullptr;
    int rc = qpdo_batch_create(&b, n, m, Qp.data(), Qi.data(), Ap.data(), Ai.data(), count, settings, device);
    if (rc) return rc;
    std::vector<double> xs((size_t)count * n), ys((size_t)count * m);
    rc = qpdo_batch_solve(b, count, Qx.data(), Ax.data(), q.data(), l.data(), u.data(), c.data(), nullptr, nullptr, xs.data(), ys.data(), summary, nullptr);
    if (!rc) for (int k = 0; k < count; k++) {
        if (x && x[k]) memcpy(x[k], &xs[(size_t)k * n], sizeof(double) * n);
        if (y && y[k] && m) memcpy(y[k], &ys[(size_t)k * m], sizeof(double) * m);
    }
    qpdo_batch_destroy(b);
    return rc;
}
