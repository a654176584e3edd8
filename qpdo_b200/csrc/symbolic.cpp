// symbolic.cpp -- see symbolic.h.  Host-side, runs once per qpdo_setup.
#include "symbolic.h"
#include <thread>
#include <functional>
#include <atomic>
#include <vector>
#include <algorithm>
#include <numeric>
#include <cstdio>
#include <cstdlib>
#include <cstring>

namespace {

template <class T> T *dup(const std::vector<T> &v) {
    T *p = (T *)malloc(sizeof(T) * (v.size() ? v.size() : 1));
    if (p && !v.empty()) memcpy(p, v.data(), sizeof(T) * v.size());
    return p;
}

// relaxed amalgamation thresholds: merge a child into its parent when the number of explicit
// zeros this introduces into L stays small (same idea as cholmod_super_symbolic.c:362-600, own
// parameters tuned for GPU fronts: wider pivot blocks amortise launch/sync latency)
const int    RELAX_SMALL_NC = 8;      // always merge while the merged pivot block is this narrow
const double RELAX_ZERO_FRAC_1 = 0.35, RELAX_ZERO_FRAC_2 = 0.15, RELAX_ZERO_FRAC_3 = 0.05;
const int    RELAX_NC_1 = 32, RELAX_NC_2 = 96;

} // namespace

extern "C" void qpdo_symbolic_free(qpdo_symbolic *S)
{
    if (!S) return;
    free(S->perm); free(S->pinv); free(S->sn_start); free(S->sn_nr); free(S->sn_rowptr);
    free(S->rows); free(S->rel); free(S->sn_parent); free(S->child_ptr); free(S->child_idx);
    free(S->Lp); free(S->Up); free(S->sn_level); free(S->level_ptr); free(S->level_sn);
    free(S->col2sn); free(S->Kp); free(S->Ki); free(S->Kdst); free(S->Kq); free(S->Ktp);
    free(S->Kt_row); free(S->Kt_a1); free(S->Kt_a2);
    free(S);
}

static qpdo_symbolic *analyze_full(int n, int m, const int *Qp, const int *Qi,
                                   const int *Ap, const int *Ai, int ordering,
                                   char *err)
{
    if (err) err[0] = 0;
    // ------------------------------------------------------------------ CSR pattern of A
    std::vector<int64_t> Rp(m + 1, 0);
    const int64_t nnzA = Ap[n];
    for (int64_t k = 0; k < nnzA; k++) Rp[Ai[k] + 1]++;
    for (int i = 0; i < m; i++) Rp[i + 1] += Rp[i];
    std::vector<int> Rj(nnzA), Rk(nnzA);            // column, and index into CSC values
    {
        std::vector<int64_t> nxt(Rp.begin(), Rp.end() - 1);
        for (int j = 0; j < n; j++)
            for (int k = Ap[j]; k < Ap[j + 1]; k++) { int64_t d = nxt[Ai[k]]++; Rj[d] = j; Rk[d] = k; }
    }
    double pairs = 0;
    for (int i = 0; i < m; i++) { double r = (double)(Rp[i + 1] - Rp[i]); pairs += r * (r + 1) / 2; }
    if (pairs > 4.0e8) {
        if (err) snprintf(err, 256, "A'A pattern too dense for the superset-pattern factorization (%.3g row-pair products)", pairs);
        return NULL;
    }
    // ------------------------------------------------------------------ adjacency graph of K
    std::vector<int64_t> gp(n + 1, 0);
    std::vector<int> gi;
    {
        // upper neighbours of Q come from a transpose pass
        std::vector<std::vector<int>> qup(n);
        for (int j = 0; j < n; j++)
            for (int k = Qp[j]; k < Qp[j + 1]; k++) if (Qi[k] != j) qup[Qi[k]].push_back(j);
        std::vector<int> mark(n, -1);
        gi.reserve((size_t)(2 * pairs < 1e9 ? 2 * pairs : 1e9) / 2 + 2 * (size_t)Qp[n]);
        for (int j = 0; j < n; j++) {
            mark[j] = j;
            size_t start = gi.size();
            for (int k = Qp[j]; k < Qp[j + 1]; k++) { int i = Qi[k]; if (mark[i] != j) { mark[i] = j; gi.push_back(i); } }
            for (int i : qup[j]) if (mark[i] != j) { mark[i] = j; gi.push_back(i); }
            for (int k = Ap[j]; k < Ap[j + 1]; k++) {
                int r = Ai[k];
                for (int64_t t = Rp[r]; t < Rp[r + 1]; t++) { int c = Rj[t]; if (mark[c] != j) { mark[c] = j; gi.push_back(c); } }
            }
            std::sort(gi.begin() + start, gi.end());
            gp[j + 1] = (int64_t)gi.size();
        }
    }
    // ------------------------------------------------------------------ ordering
    std::vector<int> perm(n), pinv(n);
    if (ordering == 0) std::iota(perm.begin(), perm.end(), 0);
    else if (ordering == 2) {
        const char *lf = getenv("QPDO_ND_LEAF");
        if (qpdo_order_nd(n, gp.data(), gi.data(), perm.data(), lf ? atoi(lf) : 256) != 0) {
            if (err) snprintf(err, 256, "nested-dissection ordering failed");
            return NULL;
        }
    }
    else if (qpdo_order_mindeg(n, gp.data(), gi.data(), perm.data()) != 0) {
        if (err) snprintf(err, 256, "minimum-degree ordering failed");
        return NULL;
    }
    for (int k = 0; k < n; k++) pinv[perm[k]] = k;
    // ------------------------------------------------------------------ etree of P K P'
    // Liu's algorithm with path compression over the rows of the lower triangle.
    auto etree_of = [&](const std::vector<int> &pv, const std::vector<int> &pi, std::vector<int> &parent) {
        std::vector<int> anc(n, -1);
        parent.assign(n, -1);
        for (int k = 0; k < n; k++) {               // k = permuted column; neighbours with smaller index
            int old = pv[k];
            for (int64_t t = gp[old]; t < gp[old + 1]; t++) {
                int i = pi[gi[t]];
                while (i != -1 && i < k) {
                    int nx = anc[i];
                    anc[i] = k;
                    if (nx == -1) parent[i] = k;
                    i = nx;
                }
            }
        }
    };
    std::vector<int> parent;
    etree_of(perm, pinv, parent);
    // ------------------------------------------------------------------ postorder, relabel
    {
        std::vector<int> head(n, -1), nxt(n, -1), post; post.reserve(n);
        for (int j = n - 1; j >= 0; j--) if (parent[j] >= 0) { nxt[j] = head[parent[j]]; head[parent[j]] = j; }
        std::vector<int> stack;
        for (int r = 0; r < n; r++) {
            if (parent[r] != -1) continue;
            stack.push_back(r);
            while (!stack.empty()) {
                int v = stack.back();
                int c = head[v];
                if (c == -1) { post.push_back(v); stack.pop_back(); }
                else { head[v] = nxt[c]; stack.push_back(c); }
            }
        }
        std::vector<int> np(n);
        for (int k = 0; k < n; k++) np[k] = perm[post[k]];
        perm.swap(np);
        for (int k = 0; k < n; k++) pinv[perm[k]] = k;
        etree_of(perm, pinv, parent);
    }
    // ------------------------------------------------------------------ lower pattern of P K P' (CSC)
    std::vector<int> Kp(n + 1, 0);
    for (int k = 0; k < n; k++) {
        int old = perm[k], cnt = 1;
        for (int64_t t = gp[old]; t < gp[old + 1]; t++) if (pinv[gi[t]] > k) cnt++;
        Kp[k + 1] = cnt;
    }
    for (int k = 0; k < n; k++) Kp[k + 1] += Kp[k];
    const int64_t nnzK = Kp[n];
    std::vector<int> Ki(nnzK);
    for (int k = 0; k < n; k++) {
        int old = perm[k], d = Kp[k];
        Ki[d++] = k;
        for (int64_t t = gp[old]; t < gp[old + 1]; t++) { int i = pinv[gi[t]]; if (i > k) Ki[d++] = i; }
        std::sort(Ki.begin() + Kp[k] + 1, Ki.begin() + d);
    }
    // ------------------------------------------------------------------ column counts of L
    // row-subtree traversal: for row i, every column on the etree paths from the nonzeros of
    // K(i, 0:i-1) up towards i gets one entry.  O(|L|) time, O(n) memory.
    std::vector<int> cc(n, 1);
    {
        std::vector<int> mk(n, -1);
        // need row access: build the strict upper pattern by transposing the lower CSC
        std::vector<int> Tp(n + 1, 0);
        for (int64_t e = 0; e < nnzK; e++) Tp[Ki[e] + 1]++;
        for (int k = 0; k < n; k++) Tp[k + 1] += Tp[k];
        std::vector<int> Tj(nnzK), nx(Tp.begin(), Tp.end() - 1);
        for (int k = 0; k < n; k++) for (int e = Kp[k]; e < Kp[k + 1]; e++) Tj[nx[Ki[e]]++] = k;
        for (int i = 0; i < n; i++) {
            mk[i] = i;
            for (int t = Tp[i]; t < Tp[i + 1]; t++) {
                int j = Tj[t];
                while (j != -1 && j < i && mk[j] != i) { mk[j] = i; cc[j]++; j = parent[j]; }
            }
        }
    }
    // ------------------------------------------------------------------ fundamental supernodes
    std::vector<int> nchild(n, 0);
    for (int j = 0; j < n; j++) if (parent[j] >= 0) nchild[parent[j]]++;
    std::vector<int> sn_start;               // first column of each fundamental supernode
    for (int j = 0; j < n; j++) {
        bool join = j > 0 && parent[j - 1] == j && cc[j - 1] == cc[j] + 1 && nchild[j] == 1;
        if (!join) sn_start.push_back(j);
    }
    int ns = (int)sn_start.size();
    sn_start.push_back(n);
    // ------------------------------------------------------------------ relaxed amalgamation
    // The tree is postordered, so the supernode that ends right before s starts is either the
    // last remaining child of s or outside its subtree.  Walking s upwards and repeatedly trying
    // to absorb that predecessor keeps every merged supernode a contiguous column range.  A merged
    // node keeps the row structure of its top piece: nr = cc[last column] - 1.
    {
        std::vector<int> col2sn0(n);
        for (int s = 0; s < ns; s++) for (int j = sn_start[s]; j < sn_start[s + 1]; j++) col2sn0[j] = s;
        std::vector<int> spar(ns, -1), rep(ns);
        for (int s = 0; s < ns; s++) { int l = sn_start[s + 1] - 1; if (parent[l] >= 0) spar[s] = col2sn0[parent[l]]; rep[s] = s; }
        auto find = [&](int s) { while (rep[s] != s) { rep[s] = rep[rep[s]]; s = rep[s]; } return s; };
        std::vector<int> first(sn_start.begin(), sn_start.end() - 1);
        std::vector<int64_t> truenz(ns, 0);
        for (int s = 0; s < ns; s++) for (int j = sn_start[s]; j < sn_start[s + 1]; j++) truenz[s] += cc[j];
        std::vector<char> dead(ns, 0);
        for (int s = 0; s < ns; s++) {
            const int64_t nr = cc[sn_start[s + 1] - 1] - 1;
            while (first[s] > 0) {
                int c = col2sn0[first[s] - 1];                 // alive: nothing above it merged yet
                if (spar[c] < 0 || find(spar[c]) != s) break;  // not a child of s
                int64_t ncm = sn_start[s + 1] - first[c];
                int64_t stored = ncm * (ncm + 1) / 2 + ncm * nr;
                int64_t tnz = truenz[s] + truenz[c];
                double frac = (double)(stored - tnz) / (double)stored;
                bool ok = (ncm <= RELAX_SMALL_NC) || (ncm <= RELAX_NC_1 && frac <= RELAX_ZERO_FRAC_1) ||
                          (ncm <= RELAX_NC_2 && frac <= RELAX_ZERO_FRAC_2) || (frac <= RELAX_ZERO_FRAC_3);
                if (!ok) break;
                first[s] = first[c]; truenz[s] = tnz; dead[c] = 1; rep[c] = s;
            }
        }
        std::vector<int> ns_start;
        for (int s = 0; s < ns; s++) if (!dead[s]) ns_start.push_back(first[s]);
        ns_start.push_back(n);
        sn_start.swap(ns_start);
        ns = (int)sn_start.size() - 1;
    }
    // ------------------------------------------------------------------ supernodal tree + row structure
    std::vector<int> col2sn(n);
    for (int s = 0; s < ns; s++) for (int j = sn_start[s]; j < sn_start[s + 1]; j++) col2sn[j] = s;
    std::vector<int> sn_parent(ns, -1);
    for (int s = 0; s < ns; s++) { int l = sn_start[s + 1] - 1; if (parent[l] >= 0) sn_parent[s] = col2sn[parent[l]]; }
    std::vector<int> child_ptr(ns + 1, 0), child_idx(ns > 0 ? ns : 1);
    for (int s = 0; s < ns; s++) if (sn_parent[s] >= 0) child_ptr[sn_parent[s] + 1]++;
    for (int s = 0; s < ns; s++) child_ptr[s + 1] += child_ptr[s];
    {
        std::vector<int> nx(child_ptr.begin(), child_ptr.end() - 1);
        for (int s = 0; s < ns; s++) if (sn_parent[s] >= 0) child_idx[nx[sn_parent[s]]++] = s;   // ascending
    }
    std::vector<int64_t> rowptr(ns + 1, 0);
    std::vector<int> rows;
    {
        int64_t tot = 0;
        for (int s = 0; s < ns; s++) tot += cc[sn_start[s + 1] - 1] - 1;
        rows.reserve((size_t)tot);
        std::vector<int> mk(n, -1);
        for (int s = 0; s < ns; s++) {
            int f = sn_start[s], l = sn_start[s + 1] - 1;
            size_t b = rows.size();
            for (int j = f; j <= l; j++)
                for (int e = Kp[j]; e < Kp[j + 1]; e++) { int i = Ki[e]; if (i > l && mk[i] != s) { mk[i] = s; rows.push_back(i); } }
            for (int t = child_ptr[s]; t < child_ptr[s + 1]; t++) {
                int c = child_idx[t];
                for (int64_t q = rowptr[c]; q < rowptr[c + 1]; q++) { int i = rows[q]; if (i > l && mk[i] != s) { mk[i] = s; rows.push_back(i); } }
            }
            std::sort(rows.begin() + b, rows.end());
            rowptr[s + 1] = (int64_t)rows.size();
            if ((int64_t)rows.size() - (int64_t)b != cc[l] - 1) {
                if (err) snprintf(err, 256, "symbolic: row count mismatch at supernode %d (%lld vs %d)", s,
                                  (long long)(rows.size() - b), cc[l] - 1);
                return NULL;
            }
        }
    }
    // relative maps into the parent front, levels, storage offsets
    std::vector<int> rel(rows.size() ? rows.size() : 1), sn_nr(ns), level(ns, 0);
    std::vector<int64_t> Lp(ns + 1, 0), Up(ns + 1, 0);
    int max_front = 0, max_nc = 0, nlevel = 0;
    for (int s = 0; s < ns; s++) {
        int nc = sn_start[s + 1] - sn_start[s];
        int nr = (int)(rowptr[s + 1] - rowptr[s]);
        sn_nr[s] = nr;
        int N = nc + nr;
        Lp[s + 1] = Lp[s] + (int64_t)N * nc;
        Up[s + 1] = Up[s] + (int64_t)nr * nr;
        max_front = std::max(max_front, N); max_nc = std::max(max_nc, nc);
        int p = sn_parent[s];
        if (p >= 0) {
            int fp = sn_start[p], lp = sn_start[p + 1] - 1, ncp = lp - fp + 1;
            int64_t q = rowptr[p];
            for (int64_t t = rowptr[s]; t < rowptr[s + 1]; t++) {
                int i = rows[t];
                if (i <= lp) rel[t] = i - fp;
                else {
                    while (q < rowptr[p + 1] && rows[q] < i) q++;
                    if (q >= rowptr[p + 1] || rows[q] != i) { if (err) snprintf(err, 256, "symbolic: child row not in parent front"); return NULL; }
                    rel[t] = ncp + (int)(q - rowptr[p]);
                }
            }
        }
    }
    for (int s = 0; s < ns; s++) { int p = sn_parent[s]; if (p >= 0) level[p] = std::max(level[p], level[s] + 1); }
    for (int s = 0; s < ns; s++) nlevel = std::max(nlevel, level[s] + 1);
    std::vector<int> level_ptr(nlevel + 1, 0), level_sn(ns > 0 ? ns : 1);
    for (int s = 0; s < ns; s++) level_ptr[level[s] + 1]++;
    for (int k = 0; k < nlevel; k++) level_ptr[k + 1] += level_ptr[k];
    {
        std::vector<int> nx(level_ptr.begin(), level_ptr.end() - 1);
        for (int s = 0; s < ns; s++) level_sn[nx[level[s]]++] = s;
        auto order_of = [&](int s) { return sn_start[s + 1] - sn_start[s] + sn_nr[s]; };
        for (int k = 0; k < nlevel; k++)
            std::stable_sort(level_sn.begin() + level_ptr[k], level_sn.begin() + level_ptr[k + 1],
                             [&](int a, int b) { return order_of(a) > order_of(b); });
    }
    // ------------------------------------------------------------------ where K's entries land in Lx
    std::vector<int64_t> Kdst(nnzK);
    for (int k = 0; k < n; k++) {
        int s = col2sn[k], f = sn_start[s], l = sn_start[s + 1] - 1, nc = l - f + 1, N = nc + sn_nr[s];
        int64_t base = Lp[s] + (int64_t)(k - f) * N;
        const int *rb = rows.data() + rowptr[s], *re = rows.data() + rowptr[s + 1];
        for (int e = Kp[k]; e < Kp[k + 1]; e++) {
            int i = Ki[e];
            int loc = (i <= l) ? i - f : nc + (int)(std::lower_bound(rb, re, i) - rb);
            Kdst[e] = base + loc;
        }
    }
    auto find_entry = [&](int r, int c) -> int {      // permuted lower entry (r >= c)
        const int *b = Ki.data() + Kp[c], *e = Ki.data() + Kp[c + 1];
        const int *p = std::lower_bound(b, e, r);
        return (p != e && *p == r) ? (int)(p - Ki.data()) : -1;
    };
    std::vector<int> Kq(nnzK, -1);
    for (int j = 0; j < n; j++)
        for (int k = Qp[j]; k < Qp[j + 1]; k++) {
            int a = pinv[Qi[k]], b = pinv[j];
            int e = find_entry(std::max(a, b), std::min(a, b));
            if (e < 0) { if (err) snprintf(err, 256, "symbolic: Q entry missing from K pattern"); return NULL; }
            Kq[e] = k;
        }
    // products a_ir * a_ic grouped by K entry, constraint rows in ascending order inside a group
    std::vector<int64_t> Ktp(nnzK + 1, 0);
    // pass 1 (the binary searches): rows split over threads, every row writes its own stretch of pair_e
    std::vector<int64_t> prow((size_t)m + 1, 0);
    for (int r = 0; r < m; r++) { const int64_t d = Rp[r + 1] - Rp[r]; prow[r + 1] = prow[r] + d * (d + 1) / 2; }
    std::vector<int> pair_e((size_t)prow[m]);
    {
        int nth = (int)std::min<unsigned>(8u, std::max(1u, std::thread::hardware_concurrency()));
        if (prow[m] < 100000) nth = 1;
        { const char *ev = getenv("QPDO_SYM_THREADS"); if (ev && atoi(ev) >= 1) nth = std::min(nth, atoi(ev)); }
        std::atomic<int> missing(0);
        auto work = [&](int r0, int r1) {
            for (int r = r0; r < r1; r++) {
                int64_t g = prow[r];
                for (int64_t t1 = Rp[r]; t1 < Rp[r + 1]; t1++)
                    for (int64_t t2 = t1; t2 < Rp[r + 1]; t2++) {
                        int a = pinv[Rj[t1]], b = pinv[Rj[t2]];
                        int e = find_entry(std::max(a, b), std::min(a, b));
                        if (e < 0) { missing = 1; e = 0; }
                        pair_e[g++] = e;
                        __atomic_fetch_add(&Ktp[e + 1], (int64_t)1, __ATOMIC_RELAXED);
                    }
            }
        };
        std::vector<std::thread> th;
        for (int t = 1; t < nth; t++) {                     // equal shares of the pairs, not of the rows
            const int r0 = (int)(std::lower_bound(prow.begin(), prow.end(), prow[m] * t / nth) - prow.begin());
            const int r1 = (int)(std::lower_bound(prow.begin(), prow.end(), prow[m] * (t + 1) / nth) - prow.begin());
            th.emplace_back(work, std::min(r0, m), t + 1 == nth ? m : std::min(r1, m));
        }
        work(0, nth > 1 ? std::min((int)(std::lower_bound(prow.begin(), prow.end(), prow[m] / nth) - prow.begin()), m) : m);
        for (auto &x : th) x.join();
        if (missing) { if (err) snprintf(err, 256, "symbolic: A'A entry missing from K pattern"); return NULL; }
    }
    for (int64_t e = 0; e < nnzK; e++) Ktp[e + 1] += Ktp[e];
    const int64_t ntriple = Ktp[nnzK];
    std::vector<int> Kt_row(ntriple ? ntriple : 1), Kt_a1(ntriple ? ntriple : 1), Kt_a2(ntriple ? ntriple : 1);
    {
        std::vector<int64_t> nx(Ktp.begin(), Ktp.end() - 1);
        size_t g = 0;
        for (int r = 0; r < m; r++)
            for (int64_t t1 = Rp[r]; t1 < Rp[r + 1]; t1++)
                for (int64_t t2 = t1; t2 < Rp[r + 1]; t2++) {
                    int64_t d = nx[pair_e[g++]]++;
                    Kt_row[d] = r; Kt_a1[d] = Rk[t1]; Kt_a2[d] = Rk[t2];
                }
    }
    // ------------------------------------------------------------------ pack the result
    qpdo_symbolic *S = (qpdo_symbolic *)calloc(1, sizeof(qpdo_symbolic));
    if (!S) return NULL;
    S->n = n; S->m = m; S->nsuper = ns; S->nlevel = nlevel;
    S->nnzK = nnzK; S->ntriple = ntriple;
    {
        // the copies are independent and mostly first-touch page faults of fresh allocations: the large ones (entry maps, product
        // tables, row structure) each on a thread of their own when the problem is large, the rest here
        const bool par = (size_t)nnzK + (size_t)ntriple > 400000 && std::thread::hardware_concurrency() >= 4;
        std::vector<std::thread> th;
        auto go = [&](std::function<void()> f) { if (par) th.emplace_back(f); else f(); };
        go([&] { S->Kt_row = dup(Kt_row); }); go([&] { S->Kt_a1 = dup(Kt_a1); }); go([&] { S->Kt_a2 = dup(Kt_a2); });
        go([&] { S->Kdst = dup(Kdst); S->Kq = dup(Kq); }); go([&] { S->Ktp = dup(Ktp); S->Ki = dup(Ki); });
        go([&] { S->rows = dup(rows); S->rel = dup(rel); });
        S->perm = dup(perm); S->pinv = dup(pinv); S->sn_start = dup(sn_start); S->sn_nr = dup(sn_nr);
        S->sn_rowptr = dup(rowptr); S->sn_parent = dup(sn_parent);
        S->child_ptr = dup(child_ptr); S->child_idx = dup(child_idx); S->Lp = dup(Lp); S->Up = dup(Up);
        S->sn_level = dup(level); S->level_ptr = dup(level_ptr); S->level_sn = dup(level_sn);
        S->col2sn = dup(col2sn); S->Kp = dup(Kp);
        for (auto &x : th) x.join();
    }
    S->lnz = 0; S->flops = 0;
    for (int j = 0; j < n; j++) { S->lnz += cc[j]; S->flops += (double)cc[j] * (double)cc[j]; }
    S->lnz_stored = Lp[ns]; S->usize = Up[ns]; S->max_front = max_front; S->max_nc = max_nc;
    return S;
}

// ------------------------------------------------------------------------------------------------ singleton columns
// A column of K without off-diagonal entries (no off-diagonal of Q, and every row of A that touches it touches nothing
// else) is a front of its own: one pivot, no rows below, no children, no parent.  With a support-following pattern
// (factor.cu, Lasso-type problems) these are nearly all of n, and pushing them through ordering, tree, counts and
// supernodes is most of an analysis (33 ms at n = 200k, 11 times along the C4 path).  So: analyse the subproblem on the
// other columns with the code above (order-preserving relabelling: every tie-break falls the same way, the elimination
// order of every connected component is the one the full analysis produces), then append the singletons -- permuted
// columns nsub .. n-1, supernodes of their own at the tail of level 0, one K entry each (the diagonal, with the products
// a_rj^2 of its column of A).
namespace {
template <class T> bool grow(T *&p, size_t newn)
{
    T *q = (T *)realloc(p, sizeof(T) * (newn ? newn : 1));
    if (!q) return false;
    p = q;
    return true;
}
}

extern "C" qpdo_symbolic *qpdo_symbolic_analyze(int n, int m, const int *Qp, const int *Qi,
                                                const int *Ap, const int *Ai, int ordering,
                                                char *err)
{
    if (err) err[0] = 0;
    { const char *ev = getenv("QPDO_SYM_COMPRESS"); if ((ev && !strcmp(ev, "0")) || n < 4096 || ordering == 0) return analyze_full(n, m, Qp, Qi, Ap, Ai, ordering, err); }
    // ---- which columns have an off-diagonal entry in K?
    std::vector<unsigned char> coupled((size_t)n, 0);
    for (int c = 0; c < n; c++)
        for (int k = Qp[c]; k < Qp[c + 1]; k++) if (Qi[k] != c) { coupled[c] = 1; coupled[Qi[k]] = 1; }
    {
        std::vector<int> rowcnt((size_t)std::max(m, 1), 0);
        for (int k = 0; k < Ap[n]; k++) rowcnt[Ai[k]]++;
        for (int c = 0; c < n; c++)
            for (int k = Ap[c]; k < Ap[c + 1]; k++) if (rowcnt[Ai[k]] >= 2) { coupled[c] = 1; break; }
    }
    int nsub = 0;
    for (int c = 0; c < n; c++) nsub += coupled[c];
    const int nsing = n - nsub;
    if (nsing < n / 2 || nsub == 0) return analyze_full(n, m, Qp, Qi, Ap, Ai, ordering, err);
    // ---- the subproblem: columns (and Q rows) relabelled in ascending order, positions of the kept entries remembered
    std::vector<int> subid((size_t)n, -1), subv((size_t)nsub), singv((size_t)nsing);
    for (int c = 0, a = 0, b = 0; c < n; c++) { if (coupled[c]) { subid[c] = a; subv[a++] = c; } else singv[b++] = c; }
    std::vector<int> sQp((size_t)nsub + 1, 0), sQi, qmap, sAp((size_t)nsub + 1, 0), sAi, amap;
    for (int a = 0; a < nsub; a++) {
        const int c = subv[a];
        for (int k = Qp[c]; k < Qp[c + 1]; k++) { sQi.push_back(subid[Qi[k]]); qmap.push_back(k); }     // rows of a coupled column are coupled or the diagonal
        sQp[a + 1] = (int)sQi.size();
        for (int k = Ap[c]; k < Ap[c + 1]; k++) { sAi.push_back(Ai[k]); amap.push_back(k); }
        sAp[a + 1] = (int)sAi.size();
    }
    if (sQi.empty()) sQi.push_back(0);
    if (sAi.empty()) sAi.push_back(0);
    qpdo_symbolic *S = analyze_full(nsub, m, sQp.data(), sQi.data(), sAp.data(), sAi.data(), ordering, err);
    if (!S) return NULL;
    // ---- back to the caller's numbering: columns, positions in Q and A
    const int ns0 = S->nsuper, ns = ns0 + nsing;
    const int64_t nnzK0 = S->nnzK, nt0 = S->ntriple, nrows0 = S->sn_rowptr[ns0];
    for (int64_t e = 0; e < nnzK0; e++) if (S->Kq[e] >= 0) S->Kq[e] = qmap[S->Kq[e]];
    for (int64_t t = 0; t < nt0; t++) { S->Kt_a1[t] = amap[S->Kt_a1[t]]; S->Kt_a2[t] = amap[S->Kt_a2[t]]; }
    int64_t nt_sing = 0;
    for (int t = 0; t < nsing; t++) nt_sing += Ap[singv[t] + 1] - Ap[singv[t]];
    const int64_t nnzK = nnzK0 + nsing, ntriple = nt0 + nt_sing;
    bool ok = grow(S->perm, n) && grow(S->pinv, n) && grow(S->col2sn, n) && grow(S->Kp, (size_t)n + 1) &&
              grow(S->sn_start, (size_t)ns + 1) && grow(S->sn_nr, ns) && grow(S->sn_rowptr, (size_t)ns + 1) && grow(S->sn_parent, ns) &&
              grow(S->child_ptr, (size_t)ns + 1) && grow(S->child_idx, ns) && grow(S->Lp, (size_t)ns + 1) && grow(S->Up, (size_t)ns + 1) &&
              grow(S->sn_level, ns) && grow(S->level_sn, ns) && grow(S->Ki, nnzK) && grow(S->Kdst, nnzK) && grow(S->Kq, nnzK) &&
              grow(S->Ktp, (size_t)nnzK + 1) && grow(S->Kt_row, ntriple) && grow(S->Kt_a1, ntriple) && grow(S->Kt_a2, ntriple);
    if (!ok) { qpdo_symbolic_free(S); if (err) snprintf(err, 256, "symbolic: out of memory"); return NULL; }
    // permutation: perm[new] = old; the subproblem's columns first, then the singletons in ascending order
    for (int a = 0; a < nsub; a++) S->perm[a] = subv[S->perm[a]];
    for (int t = 0; t < nsing; t++) S->perm[nsub + t] = singv[t];
    for (int k = 0; k < n; k++) S->pinv[S->perm[k]] = k;
    // level lists: the singletons go to the tail of level 0 (fronts of a level are sorted by decreasing order)
    {
        const int l0e = S->level_ptr[1];
        memmove(S->level_sn + l0e + nsing, S->level_sn + l0e, sizeof(int) * (size_t)(ns0 - l0e));
        for (int t = 0; t < nsing; t++) S->level_sn[l0e + t] = ns0 + t;
        for (int lv = 1; lv <= S->nlevel; lv++) S->level_ptr[lv] += nsing;
    }
    int64_t tpos = nt0;
    for (int t = 0; t < nsing; t++) {
        const int s = ns0 + t, col = nsub + t, c = singv[t];
        S->sn_start[s + 1] = col + 1;                      // (sn_start[ns0] == nsub already)
        S->sn_nr[s] = 0; S->sn_rowptr[s + 1] = nrows0; S->sn_parent[s] = -1;
        S->child_ptr[s + 1] = S->child_ptr[ns0];
        S->Lp[s + 1] = S->Lp[s] + 1; S->Up[s + 1] = S->Up[s];
        S->sn_level[s] = 0; S->col2sn[col] = s;
        const int64_t e = nnzK0 + t;
        S->Kp[col + 1] = (int)(e + 1);
        S->Ki[e] = col; S->Kdst[e] = S->Lp[s];
        S->Kq[e] = (Qp[c + 1] > Qp[c]) ? Qp[c] : -1;       // an uncoupled column of Q holds its diagonal entry or nothing
        for (int k = Ap[c]; k < Ap[c + 1]; k++) { S->Kt_row[tpos] = Ai[k]; S->Kt_a1[tpos] = k; S->Kt_a2[tpos] = k; tpos++; }
        S->Ktp[e + 1] = tpos;
    }
    S->n = n; S->nsuper = ns; S->nnzK = nnzK; S->ntriple = ntriple;
    S->lnz += nsing; S->flops += (double)nsing; S->lnz_stored = S->Lp[ns]; S->usize = S->Up[ns];
    if (S->max_front < 1) S->max_front = 1;
    if (S->max_nc < 1) S->max_nc = 1;
    return S;
}
