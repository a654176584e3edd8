// linesearch.cu -- exact linesearch on the convex piecewise-quadratic merit function
// (reference src/linesearch.c:8-158): build eta, beta and the 2m breakpoints t_i = alpha_i/delta_i,
// keep t_i > 0, sort them (the reference's qsort -> device radix sort of the IEEE bit patterns, or
// an in-smem bitonic sort when 2m <= 2048), prefix-sum the slope changes and take the first
// segment whose end has a positive derivative: tau = -b/a.
// Compiled with -fmad=false (element-wise arithmetic rounds like the reference).
#include "dev_internal.h"
#include <algorithm>

#define LS_SMALL_MAX 2048
#define LS_EXACT_MAX 128     // up to this many variables / breakpoints the linesearch sums run in reference order
#define SORT_TILE 2048
#define SORT_THREADS 256

__device__ __forceinline__ double bsum(double v, double *sh)
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

struct LsArgs {
    int n, m;
    const double *dx, *Qdx, *df, *dy, *y, *mu, *sqrt_mu, *Adx, *w, *l, *u;
    double *delta, *alpha;
    unsigned long long *key;
    int *idx;
    qpdo_dev_scalars *sc;
};
// breakpoint k of 2m (linesearch.c:28-39,82-114): returns key (all ones when t <= 0 or NaN) and
// adds this breakpoint's share of the initial slope sums over J = L xor P
__device__ __forceinline__ unsigned long long ls_breakpoint(const LsArgs &A, int k, double &aJ, double &bJ, int &nL)
{
    const int m = A.m, i = k < m ? k : k - m;
    double tm = A.dy[i] * A.mu[i]; tm = tm * 0.5;
    double c0 = A.Adx[i] + (-1.0) * tm;
    double dl = c0 * A.sqrt_mu[i];
    double al;
    if (k < m) { dl = dl * (-1.0); al = (A.w[i] + (-1.0) * A.l[i]) * A.sqrt_mu[i]; }
    else al = (A.u[i] + (-1.0) * A.w[i]) * A.sqrt_mu[i];
    A.delta[k] = dl; A.alpha[k] = al;
    const double t = al / dl;
    const bool L = t > 0.0, P = dl > 0.0;
    if (L != P) { aJ += dl * dl; bJ += dl * al; }
    if (L) { nL++; return (unsigned long long)__double_as_longlong(t); }
    return ~0ull;
}
__device__ __forceinline__ void ls_contrib(const double *delta, const double *alpha, int iz, double &da, double &db)
{   // linesearch.c:133-141
    const double d = delta[iz], al = alpha[iz];
    if (d > 0.0) { da = d * d; db = -(d * al); }
    else { da = -(d * d); db = d * al; }
}

// ------------------------------------------------------------------------------- small path
__global__ void __launch_bounds__(1024) k_ls_small(LsArgs A)
{
    __shared__ unsigned long long key[LS_SMALL_MAX];
    __shared__ int idx[LS_SMALL_MAX];
    __shared__ double sh[32];
    __shared__ double wa[32], wb[32];
    __shared__ int kmin;
    const int tid = threadIdx.x, n = A.n, m = A.m, M2 = 2 * m;
    double s1 = 0, s2 = 0, s3 = 0, s4 = 0;
    for (int i = tid; i < m; i += 1024) {
        double tm = A.dy[i] * A.mu[i]; tm = tm * 0.5;
        s1 += A.dy[i] * tm; s3 += A.y[i] * tm;
    }
    for (int i = tid; i < n; i += 1024) { s2 += A.dx[i] * A.Qdx[i]; s4 += A.dx[i] * A.df[i]; }
    // tiny problems: every sum in exactly the reference's scalar order (vec_prod's 4-way unrolling,
    // lin_alg.c:59-71; vec_prod_ind, linesearch.c:190-199), so tau is bit-identical whenever dx, dy are
    const bool exact = (n <= LS_EXACT_MAX && M2 <= LS_EXACT_MAX);
    if (exact) {
        if (tid < 4) {
            const bool mvec = (tid == 0 || tid == 2);
            const int len = mvec ? m : n;
            double prod = 0.0;
            int i = 0;
            #define LS_TERM(i_) (mvec ? ((tid == 0 ? A.dy[i_] : A.y[i_]) * ((A.dy[i_] * A.mu[i_]) * 0.5)) \
                                      : (A.dx[i_] * (tid == 1 ? A.Qdx[i_] : A.df[i_])))
            if (len >= 4) for (; i <= len - 4; i += 4) prod += (LS_TERM(i) + LS_TERM(i + 1) + LS_TERM(i + 2) + LS_TERM(i + 3));
            for (; i < len; i++) prod += LS_TERM(i);
            #undef LS_TERM
            wa[tid] = prod;
        }
        __syncthreads();
        s1 = wa[0]; s2 = wa[1]; s3 = wa[2]; s4 = wa[3];
        __syncthreads();
    } else {
        s1 = bsum(s1, sh); s2 = bsum(s2, sh); s3 = bsum(s3, sh); s4 = bsum(s4, sh);
    }
    double eta = s1; eta += s2; eta *= 0.5;
    double beta = s3; beta += s4; beta *= 0.5;
    int P2 = 64; while (P2 < M2) P2 <<= 1;
    double aJ = 0, bJ = 0; int nL = 0;
    for (int k = tid; k < P2; k += 1024) {
        if (k < M2) { key[k] = ls_breakpoint(A, k, aJ, bJ, nL); idx[k] = k; }
        else { key[k] = ~0ull; idx[k] = 0; }
    }
    aJ = bsum(aJ, sh); bJ = bsum(bJ, sh);
    const int nLt = (int)(bsum((double)nL, sh) + 0.5);
    if (exact) {
        __syncthreads();
        if (tid == 0) {
            double pa = 0.0, pb = 0.0;
            for (int k = 0; k < M2; k++) {
                const double dl = A.delta[k], al = A.alpha[k];
                const bool L = (al / dl) > 0.0, P = dl > 0.0;
                if (L != P) { pa += dl * dl; pb += dl * al; }
            }
            wa[0] = pa; wb[0] = pb;
        }
        __syncthreads();
        aJ = wa[0]; bJ = wb[0];
        __syncthreads();
    }
    const double a0 = eta + aJ, b0 = beta - bJ;
    if (tid == 0) kmin = nLt;
    __syncthreads();
    // bitonic sort of (key, idx), ties broken by index so the result is deterministic
    for (int size = 2; size <= P2; size <<= 1)
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
            for (int t = tid; t < (P2 >> 1); t += 1024) {
                int lo = 2 * t - (t & (stride - 1)), hi = lo + stride;
                bool up = ((lo & size) == 0);
                unsigned long long ka = key[lo], kb = key[hi];
                int ia = idx[lo], ib = idx[hi];
                bool gt = (ka > kb) || (ka == kb && ia > ib);
                if (gt == up) { key[lo] = kb; key[hi] = ka; idx[lo] = ib; idx[hi] = ia; }
            }
            __syncthreads();
        }
    if (exact) {
        // the reference's sequential sweep, linesearch.c:121-157
        if (tid == 0) {
            double a = a0, b = b0, tau;
            bool done = false;
            if (nLt == 0 || b + a * __longlong_as_double((long long)key[0]) > 0.0) { tau = -b / a; done = true; }
            for (int i = 0; !done && i < nLt; i++) {
                double da, db; ls_contrib(A.delta, A.alpha, idx[i], da, db);
                a = a + da; b = b + db;
                if (i + 1 < nLt && b + a * __longlong_as_double((long long)key[i + 1]) > 0.0) { tau = -b / a; done = true; }
            }
            if (!done) tau = -b / a;
            A.sc->tau = tau; A.sc->ls_eta = eta; A.sc->ls_beta = beta; A.sc->ls_nL = nLt;
        }
        return;
    }
    // prefix sums of the slope changes in sorted order; thread owns a contiguous chunk
    const int chunk = (P2 + 1023) / 1024;             // 1 or 2
    const int k0 = tid * chunk;
    double la = 0, lb = 0;
    for (int c = 0; c < chunk; c++) {
        int k = k0 + c;
        if (k < nLt) { double da, db; ls_contrib(A.delta, A.alpha, idx[k], da, db); la += da; lb += db; }
    }
    // block exclusive scan of (la, lb)
    double pa = la, pb = lb;
    const int lane = tid & 31, wid = tid >> 5;
    for (int o = 1; o < 32; o <<= 1) {
        double ta = __shfl_up_sync(0xffffffffu, pa, o), tb = __shfl_up_sync(0xffffffffu, pb, o);
        if (lane >= o) { pa += ta; pb += tb; }
    }
    if (lane == 31) { wa[wid] = pa; wb[wid] = pb; }
    __syncthreads();
    double oa = 0, ob = 0;
    for (int w = 0; w < wid; w++) { oa += wa[w]; ob += wb[w]; }
    double ea = oa + pa - la, eb = ob + pb - lb;       // exclusive prefix at k0
    double ra = a0 + ea, rb = b0 + eb;
    double cand_a = 0, cand_b = 0; int cand_k = -1;
    for (int c = 0; c < chunk; c++) {
        int k = k0 + c;
        if (k < nLt) {
            double t = __longlong_as_double((long long)key[k]);
            if (cand_k < 0 && rb + ra * t > 0.0) { cand_k = k; cand_a = ra; cand_b = rb; }
            double da, db; ls_contrib(A.delta, A.alpha, idx[k], da, db);
            ra += da; rb += db;
        }
    }
    if (cand_k >= 0) atomicMin(&kmin, cand_k);
    __syncthreads();
    if (cand_k >= 0 && cand_k == kmin) A.sc->tau = -cand_b / cand_a;
    if (kmin == nLt) {
        // no crossing before the last breakpoint: all slope changes applied (linesearch.c:148-156)
        double ta = 0, tb = 0;
        for (int w = 0; w < 32; w++) { ta += wa[w]; tb += wb[w]; }
        if (tid == 0) A.sc->tau = -(b0 + tb) / (a0 + ta);
    }
    if (tid == 0) { A.sc->ls_eta = eta; A.sc->ls_beta = beta; A.sc->ls_nL = nLt; }
}

// ------------------------------------------------------------------------------- large path
// part layout per block: [0] dy.tm [1] dx.Qdx [2] y.tm [3] dx.df [4] aJ [5] bJ [6] nL
#define LS_PART 8
__global__ void __launch_bounds__(256) k_ls_build(LsArgs A, double *part)
{
    __shared__ double sh[32];
    const int n = A.n, m = A.m, M2 = 2 * m;
    double s1 = 0, s2 = 0, s3 = 0, s4 = 0, aJ = 0, bJ = 0; int nL = 0;
    const int stride = gridDim.x * blockDim.x, g = blockIdx.x * blockDim.x + threadIdx.x;
    for (int i = g; i < m; i += stride) {
        double tm = A.dy[i] * A.mu[i]; tm = tm * 0.5;
        s1 += A.dy[i] * tm; s3 += A.y[i] * tm;
    }
    for (int i = g; i < n; i += stride) { s2 += A.dx[i] * A.Qdx[i]; s4 += A.dx[i] * A.df[i]; }
    for (int k = g; k < M2; k += stride) { A.key[k] = ls_breakpoint(A, k, aJ, bJ, nL); A.idx[k] = k; }
    s1 = bsum(s1, sh); s2 = bsum(s2, sh); s3 = bsum(s3, sh); s4 = bsum(s4, sh);
    aJ = bsum(aJ, sh); bJ = bsum(bJ, sh);
    double cnt = bsum((double)nL, sh);
    if (threadIdx.x == 0) {
        double *p = part + (size_t)blockIdx.x * LS_PART;
        p[0] = s1; p[1] = s2; p[2] = s3; p[3] = s4; p[4] = aJ; p[5] = bJ; p[6] = cnt;
    }
}
// ab[0] = a0, ab[1] = b0
__global__ void k_ls_reduce(const double *part, int nblk, qpdo_dev_scalars *sc, double *ab)
{
    __shared__ double sh[32];
    double v[7];
    for (int j = 0; j < 7; j++) {
        double s = 0;
        for (int b = threadIdx.x; b < nblk; b += blockDim.x) s += part[(size_t)b * LS_PART + j];
        v[j] = bsum(s, sh);
    }
    if (threadIdx.x == 0) {
        double eta = v[0]; eta += v[1]; eta *= 0.5;
        double beta = v[2]; beta += v[3]; beta *= 0.5;
        sc->ls_eta = eta; sc->ls_beta = beta; sc->ls_nL = (int)(v[6] + 0.5);
        ab[0] = eta + v[4]; ab[1] = beta - v[5];
    }
}

// LSD radix sort, 8-bit digits, stable.  Tile = 2048 keys per CTA, 8 warps x 8 rows of 32.
// Per-warp digit counts come from __match_any_sync; the scatter pass recomputes them so that the
// rank of every key inside its tile follows the original order.
__global__ void __launch_bounds__(SORT_THREADS) k_sort_hist(const unsigned long long *key, int N, int shift, unsigned *hist, int nblk, unsigned *tot)
{
    __shared__ unsigned cnt[256];
    const int tid = threadIdx.x;
    cnt[tid] = 0;
    __syncthreads();
    const int base = blockIdx.x * SORT_TILE;
    for (int k = base + tid; k < min(N, base + SORT_TILE); k += SORT_THREADS)
        atomicAdd(&cnt[(unsigned)(key[k] >> shift) & 255u], 1u);
    __syncthreads();
    hist[(size_t)tid * nblk + blockIdx.x] = cnt[tid];
    if (cnt[tid]) atomicAdd(&tot[tid], cnt[tid]);              // digit totals of the pass (integers: order does not matter)
}
// Exclusive scan of the digit-major table hist[digit][block], in place: one CTA per digit.  The start of a digit's row is
// the sum of the totals of the smaller digits (256 values, accumulated by k_sort_hist), the row itself is scanned in
// chunks of 256 with a carry.  (Round 1 scanned the whole table in ONE CTA: 17 us per pass at 2m = 360,600, the
// largest single item of the linesearch.)
__global__ void __launch_bounds__(256) k_sort_scan_rows(unsigned *hist, const unsigned *tot, int nblk)
{
    __shared__ unsigned ws[8];
    __shared__ unsigned carry_s;
    const int d = blockIdx.x, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    unsigned v = (tid < d) ? tot[tid] : 0u;
    for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) ws[wid] = v;
    __syncthreads();
    if (tid == 0) { unsigned b = 0; for (int w = 0; w < 8; w++) b += ws[w]; carry_s = b; }
    __syncthreads();
    unsigned *row = hist + (size_t)d * nblk;
    for (int c0 = 0; c0 < nblk; c0 += 256) {
        const int i = c0 + tid;
        const unsigned x = (i < nblk) ? row[i] : 0u;
        unsigned p = x;
        for (int o = 1; o < 32; o <<= 1) { const unsigned t = __shfl_up_sync(0xffffffffu, p, o); if (lane >= o) p += t; }
        __syncthreads();                                         // ws / carry_s of the previous chunk consumed
        if (lane == 31) ws[wid] = p;
        __syncthreads();
        unsigned off = carry_s;
        for (int w = 0; w < wid; w++) off += ws[w];
        if (i < nblk) row[i] = off + p - x;
        __syncthreads();
        if (tid == 255) carry_s = off + p;
    }
}
__global__ void __launch_bounds__(SORT_THREADS) k_sort_scatter(const unsigned long long *key, const int *idx, int N, int shift,
                                                              const unsigned *hist, int nblk, unsigned long long *okey, int *oidx)
{
    __shared__ unsigned cnt[8][256];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    for (int i = tid; i < 8 * 256; i += SORT_THREADS) (&cnt[0][0])[i] = 0;
    __syncthreads();
    const int base = blockIdx.x * SORT_TILE + wid * 256;       // this warp's 256 keys
    for (int r = 0; r < 8; r++) {
        const int k = base + r * 32 + lane;
        const bool ok = k < N;
        const unsigned dgt = ok ? ((unsigned)(key[k] >> shift) & 255u) : 256u + lane;   // inactive lanes match nobody
        const unsigned peers = __match_any_sync(0xffffffffu, dgt);
        if (ok && lane == __ffs(peers) - 1) cnt[wid][dgt] += __popc(peers);
        __syncwarp();
    }
    __syncthreads();
    {   // thread tid = digit: turn per-warp counts into per-warp start offsets (global)
        unsigned run = hist[(size_t)tid * nblk + blockIdx.x];
        for (int w = 0; w < 8; w++) { unsigned c = cnt[w][tid]; cnt[w][tid] = run; run += c; }
    }
    __syncthreads();
    for (int r = 0; r < 8; r++) {
        const int k = base + r * 32 + lane;
        const bool ok = k < N;
        unsigned long long kv = ok ? key[k] : 0ull;
        const unsigned dgt = ok ? ((unsigned)(kv >> shift) & 255u) : 256u + lane;
        const unsigned peers = __match_any_sync(0xffffffffu, dgt);
        const unsigned rank = __popc(peers & ((1u << lane) - 1u));
        unsigned pos = 0;
        if (ok) pos = cnt[wid][dgt] + rank;
        __syncwarp();
        if (ok && lane == __ffs(peers) - 1) cnt[wid][dgt] += __popc(peers);
        __syncwarp();
        if (ok) { okey[pos] = kv; oidx[pos] = idx[k]; }
    }
}

// sweep over the sorted breakpoints
__global__ void __launch_bounds__(256) k_sweep_sums(const int *idx, const double *delta, const double *alpha,
                                                    const qpdo_dev_scalars *sc, double *bs)
{
    __shared__ double sh[32];
    const int nL = sc->ls_nL, base = blockIdx.x * SORT_TILE;
    double la = 0, lb = 0;
    for (int k = base + threadIdx.x; k < min(nL, base + SORT_TILE); k += 256) {
        double da, db; ls_contrib(delta, alpha, idx[k], da, db); la += da; lb += db;
    }
    la = bsum(la, sh); lb = bsum(lb, sh);
    if (threadIdx.x == 0) { bs[2 * blockIdx.x] = la; bs[2 * blockIdx.x + 1] = lb; }
}
__global__ void k_sweep_scan(double *bs, int nblk)
{   // exclusive scan of the per-tile sums, sequential in tile order (nblk is small); total at the end
    if (threadIdx.x == 0) {
        double ra = 0, rb = 0;
        for (int b = 0; b < nblk; b++) { double a = bs[2 * b], c = bs[2 * b + 1]; bs[2 * b] = ra; bs[2 * b + 1] = rb; ra += a; rb += c; }
        bs[2 * nblk] = ra; bs[2 * nblk + 1] = rb;
    }
}
// candidate per tile: cand[3*blk] = k (or -1), a_k, b_k
__global__ void __launch_bounds__(256) k_sweep_find(const unsigned long long *key, const int *idx, const double *delta,
                                                    const double *alpha, const qpdo_dev_scalars *sc, const double *bs,
                                                    const double *ab, double *cand)
{
    __shared__ double wa[8], wb[8];
    __shared__ int kmin;
    const int nL = sc->ls_nL, base = blockIdx.x * SORT_TILE, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    if (tid == 0) kmin = 0x7fffffff;
    const int k0 = base + tid * 8;                      // 8 consecutive breakpoints per thread
    double la = 0, lb = 0;
    for (int c = 0; c < 8; c++) {
        int k = k0 + c;
        if (k < nL) { double da, db; ls_contrib(delta, alpha, idx[k], da, db); la += da; lb += db; }
    }
    double pa = la, pb = lb;
    for (int o = 1; o < 32; o <<= 1) {
        double ta = __shfl_up_sync(0xffffffffu, pa, o), tb = __shfl_up_sync(0xffffffffu, pb, o);
        if (lane >= o) { pa += ta; pb += tb; }
    }
    if (lane == 31) { wa[wid] = pa; wb[wid] = pb; }
    __syncthreads();
    double oa = 0, ob = 0;
    for (int w = 0; w < wid; w++) { oa += wa[w]; ob += wb[w]; }
    double ra = ab[0] + bs[2 * blockIdx.x] + (oa + pa - la);
    double rb = ab[1] + bs[2 * blockIdx.x + 1] + (ob + pb - lb);
    double ca = 0, cb = 0; int ck = -1;
    for (int c = 0; c < 8; c++) {
        int k = k0 + c;
        if (k < nL) {
            double t = __longlong_as_double((long long)key[k]);
            if (ck < 0 && rb + ra * t > 0.0) { ck = k; ca = ra; cb = rb; }
            double da, db; ls_contrib(delta, alpha, idx[k], da, db); ra += da; rb += db;
        }
    }
    if (ck >= 0) atomicMin(&kmin, ck);
    __syncthreads();
    if (tid == 0 && kmin == 0x7fffffff) cand[3 * blockIdx.x] = -1.0;
    if (ck >= 0 && ck == kmin) { cand[3 * blockIdx.x] = (double)ck; cand[3 * blockIdx.x + 1] = ca; cand[3 * blockIdx.x + 2] = cb; }
}
__global__ void k_sweep_final(const double *cand, int nblk, const double *bs, const double *ab, qpdo_dev_scalars *sc)
{
    if (threadIdx.x == 0) {
        for (int b = 0; b < nblk; b++)
            if (cand[3 * b] >= 0.0) { sc->tau = -cand[3 * b + 2] / cand[3 * b + 1]; return; }
        sc->tau = -(ab[1] + bs[2 * nblk + 1]) / (ab[0] + bs[2 * nblk]);
    }
}

// =============================================================================== host side
int linesearch_setup(qpdo_dev *d)
{
    const size_t M2 = 2 * (size_t)d->m;
    d->sort_blocks = std::max(1, cdiv((long)M2, SORT_TILE));
    d->ls_blocks = std::max(1, std::min(592, cdiv(std::max(d->n, 2 * d->m), 256)));
    if (dev_alloc(d, &d->ls_delta, M2) || dev_alloc(d, &d->ls_alpha, M2) || dev_alloc(d, &d->ls_key[0], M2) ||
        dev_alloc(d, &d->ls_key[1], M2) || dev_alloc(d, &d->ls_idx[0], M2) || dev_alloc(d, &d->ls_idx[1], M2) ||
        dev_alloc(d, &d->ls_part, (size_t)d->ls_blocks * LS_PART + 8) ||
        dev_alloc(d, &d->ls_scan, (size_t)d->sort_blocks * 5 + 8) ||
        dev_alloc(d, &d->ls_hist, (size_t)d->sort_blocks * 256 + 8 * 256))
        return -1;
    return 0;
}

int linesearch_run(qpdo_dev *d)
{
    PhaseTimer t(d, PH_LS);
    LsArgs A;
    A.n = d->n; A.m = d->m; A.dx = d->dx; A.Qdx = d->Qdx; A.df = d->df; A.dy = d->dy; A.y = d->y; A.mu = d->mu;
    A.sqrt_mu = d->sqrt_mu; A.Adx = d->Adx; A.w = d->w; A.l = d->l; A.u = d->u; A.delta = d->ls_delta; A.alpha = d->ls_alpha;
    A.key = d->ls_key[0]; A.idx = d->ls_idx[0]; A.sc = d->sc;
    const int M2 = 2 * d->m;
    if (M2 <= LS_SMALL_MAX) {
        k_ls_small<<<1, 1024, 0, d->st>>>(A);
        d->prof.kernel_launches += 1;
        return 0;
    }
    const int nsb = d->sort_blocks;
    double *ab = d->ls_part + (size_t)d->ls_blocks * LS_PART;
    k_ls_build<<<d->ls_blocks, 256, 0, d->st>>>(A, d->ls_part);
    k_ls_reduce<<<1, 256, 0, d->st>>>(d->ls_part, d->ls_blocks, d->sc, ab);
    int cur = 0;
    unsigned *tot = d->ls_hist + (size_t)nsb * 256;            // [8][256] digit totals, one row per pass
    CUDA_OK(cudaMemsetAsync(tot, 0, sizeof(unsigned) * 8 * 256, d->st));
    for (int pass = 0; pass < 8; pass++) {
        const int shift = 8 * pass;
        k_sort_hist<<<nsb, SORT_THREADS, 0, d->st>>>(d->ls_key[cur], M2, shift, d->ls_hist, nsb, tot + 256 * pass);
        k_sort_scan_rows<<<256, 256, 0, d->st>>>(d->ls_hist, tot + 256 * pass, nsb);
        k_sort_scatter<<<nsb, SORT_THREADS, 0, d->st>>>(d->ls_key[cur], d->ls_idx[cur], M2, shift, d->ls_hist, nsb,
                                                        d->ls_key[cur ^ 1], d->ls_idx[cur ^ 1]);
        cur ^= 1;
    }
    double *bs = d->ls_scan, *cand = d->ls_scan + 2 * (size_t)nsb + 2;
    k_sweep_sums<<<nsb, 256, 0, d->st>>>(d->ls_idx[cur], d->ls_delta, d->ls_alpha, d->sc, bs);
    k_sweep_scan<<<1, 32, 0, d->st>>>(bs, nsb);
    k_sweep_find<<<nsb, 256, 0, d->st>>>(d->ls_key[cur], d->ls_idx[cur], d->ls_delta, d->ls_alpha, d->sc, bs, ab, cand);
    k_sweep_final<<<1, 32, 0, d->st>>>(cand, nsb, bs, ab, d->sc);
    d->prof.kernel_launches += 2 + 1 + 24 + 4;
    return 0;
}
