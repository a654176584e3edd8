// ordering.cpp -- fill-reducing minimum-degree ordering on a quotient graph (host, setup-time).
//
// Replaces the ordering step of cholmod_analyze in the reference (which QPDO hard-wires to
// NATURAL, src/cholmod_interface.c:120; BASELINE north_star asks for AMD on the host instead).
// Own implementation of the approximate-minimum-degree idea (Amestoy/Davis/Duff 1996): quotient
// graph with element absorption, approximate external degrees via the |Le \ Lp| trick, mass
// elimination and hash-based supervariable detection.  Adjacency is kept in per-node vectors.
#include "symbolic.h"
#include <vector>
#include <algorithm>
#include <cstdint>
#include <cstring>

namespace {

struct MinDeg {
    int n;
    std::vector<std::vector<int>> adjv;   // variable -> adjacent variables (pruned lazily)
    std::vector<std::vector<int>> adje;   // variable -> adjacent elements
    std::vector<std::vector<int>> elv;    // element  -> its variables (principal supervariables)
    std::vector<int> nv;                  // supervariable weight; 0 = absorbed into another
    std::vector<int> state;               // 0 live variable, 1 element, 2 dead (absorbed/eliminated)
    std::vector<int> deg;                 // approximate external degree (weighted)
    std::vector<int> head, next, prev;    // degree buckets
    std::vector<int> w;                   // work flags for |Le \ Lp|
    std::vector<int> mark;                // generic marker
    std::vector<int> merged_into;         // for absorbed supervariables: principal
    int wflg = 1, mflg = 0, mindeg = 0;

    void bucket_remove(int i) {
        int d = deg[i];
        if (prev[i] >= 0) next[prev[i]] = next[i]; else head[d] = next[i];
        if (next[i] >= 0) prev[next[i]] = prev[i];
        next[i] = prev[i] = -1;
    }
    void bucket_insert(int i) {
        int d = deg[i];
        prev[i] = -1; next[i] = head[d];
        if (head[d] >= 0) prev[head[d]] = i;
        head[d] = i;
        if (d < mindeg) mindeg = d;
    }
};

} // namespace

extern "C" int qpdo_order_mindeg(int n, const int64_t *gp, const int *gi, int *perm_out)
{
    if (n <= 0) return 0;
    MinDeg G;
    G.n = n;
    G.adjv.resize(n); G.adje.resize(n); G.elv.resize(n);
    G.nv.assign(n, 1); G.state.assign(n, 0); G.deg.assign(n, 0);
    G.head.assign(n + 1, -1); G.next.assign(n, -1); G.prev.assign(n, -1);
    G.w.assign(n, 0); G.mark.assign(n, 0); G.merged_into.assign(n, -1);
    for (int j = 0; j < n; j++) {
        G.adjv[j].assign(gi + gp[j], gi + gp[j + 1]);
        G.deg[j] = (int)G.adjv[j].size();
    }
    G.mindeg = n;
    for (int j = 0; j < n; j++) G.bucket_insert(j);

    std::vector<int> order;            // principal variables in elimination order
    order.reserve(n);
    std::vector<int> Lp, hashv(n, 0), cand;
    std::vector<int64_t> hkey(n, 0);
    int nleft = n;                      // weighted count of uneliminated variables

    auto new_mark = [&]() { if (++G.mflg == INT32_MAX) { std::fill(G.mark.begin(), G.mark.end(), 0); G.mflg = 1; } return G.mflg; };

    while (nleft > 0) {
        while (G.mindeg <= n && G.head[G.mindeg] < 0) G.mindeg++;
        int p = G.head[G.mindeg];
        G.bucket_remove(p);
        // ---- form the new element p: Lp = adjv[p] U (U_{e in adje[p]} elv[e]) \ {p}
        int tag = new_mark();
        G.mark[p] = tag;
        Lp.clear();
        int lpw = 0;
        for (int v : G.adjv[p]) if (G.state[v] == 0 && G.nv[v] > 0 && G.mark[v] != tag) { G.mark[v] = tag; Lp.push_back(v); lpw += G.nv[v]; }
        for (int e : G.adje[p]) {
            if (G.state[e] != 1) continue;
            for (int v : G.elv[e]) if (G.state[v] == 0 && G.nv[v] > 0 && G.mark[v] != tag) { G.mark[v] = tag; Lp.push_back(v); lpw += G.nv[v]; }
            G.state[e] = 2; std::vector<int>().swap(G.elv[e]);       // absorbed into p
        }
        std::vector<int>().swap(G.adjv[p]); std::vector<int>().swap(G.adje[p]);
        G.state[p] = 1;
        order.push_back(p);
        int pw = G.nv[p];
        nleft -= pw;
        // ---- remove the members of Lp from the degree lists, prune their lists
        for (int i : Lp) G.bucket_remove(i);
        // |Le \ Lp| for every element adjacent to a member of Lp
        if (G.wflg > INT32_MAX - 2 * n - 2) { std::fill(G.w.begin(), G.w.end(), 0); G.wflg = 1; }
        int wf = G.wflg;
        for (int i : Lp) {
            std::vector<int> &ae = G.adje[i];
            size_t k = 0;
            for (int e : ae) {
                if (G.state[e] != 1) continue;          // absorbed element: drop
                ae[k++] = e;
                if (G.w[e] < wf) {                     // first touch: weighted size of Le
                    int sz = 0;
                    for (int v : G.elv[e]) if (G.state[v] == 0) sz += G.nv[v];
                    G.w[e] = sz + wf;
                }
                G.w[e] -= G.nv[i];
            }
            ae.resize(k);
        }
        // ---- degree update, aggressive absorption, mass elimination, hashing
        cand.clear();
        for (int i : Lp) {
            std::vector<int> &ae = G.adje[i];
            std::vector<int> &av = G.adjv[i];
            int64_t d = 0, h = 0;
            size_t k = 0;
            for (int e : ae) {
                int ext = G.w[e] - wf;
                if (ext > 0) { d += ext; ae[k++] = e; h += e; }
                else { G.state[e] = 2; std::vector<int>().swap(G.elv[e]); }   // Le subset of Lp
            }
            ae.resize(k);
            k = 0;
            for (int v : av) {
                if (G.state[v] != 0 || G.nv[v] == 0 || G.mark[v] == tag) continue;  // dead or covered by p
                av[k++] = v; d += G.nv[v]; h += v;
            }
            av.resize(k);
            if (ae.empty() && av.empty()) {
                // mass elimination: i is indistinguishable from p from now on
                G.state[i] = 2; G.merged_into[i] = p; G.nv[p] += G.nv[i]; nleft -= G.nv[i];
                lpw -= G.nv[i]; G.nv[i] = 0;
                continue;
            }
            ae.push_back(p); h += p;
            int64_t dnew = d + (lpw - G.nv[i]);
            int64_t dold = (int64_t)G.deg[i] + (lpw - G.nv[i]);
            int64_t cap = nleft - G.nv[i];
            int64_t dd = std::min(std::min(dnew, dold), cap);
            if (dd < 0) dd = 0;
            G.deg[i] = (int)dd;
            hkey[i] = h;
            cand.push_back(i);
        }
        G.wflg = wf + 2 * n + 2 > 0 ? wf + n + 1 : 1;
        // ---- supervariable detection among the survivors of Lp (same hash -> compare lists)
        if (cand.size() > 1) {
            std::sort(cand.begin(), cand.end(), [&](int a, int b) { return hkey[a] != hkey[b] ? hkey[a] < hkey[b] : a < b; });
            for (size_t a = 0; a < cand.size(); ) {
                size_t b = a + 1;
                while (b < cand.size() && hkey[cand[b]] == hkey[cand[a]]) b++;
                for (size_t x = a; x < b; x++) {
                    int i = cand[x];
                    if (G.nv[i] == 0) continue;
                    int t2 = new_mark();
                    for (int e : G.adje[i]) G.mark[e] = t2;
                    for (int v : G.adjv[i]) G.mark[v] = t2;
                    for (size_t y = x + 1; y < b; y++) {
                        int j = cand[y];
                        if (G.nv[j] == 0) continue;
                        if (G.adje[j].size() != G.adje[i].size() || G.adjv[j].size() != G.adjv[i].size()) continue;
                        bool same = true;
                        for (int e : G.adje[j]) if (G.mark[e] != t2) { same = false; break; }
                        if (same) for (int v : G.adjv[j]) if (G.mark[v] != t2) { same = false; break; }
                        if (!same) continue;
                        // j joins supervariable i
                        G.nv[i] += G.nv[j]; G.deg[i] = std::max(0, G.deg[i] - G.nv[j]);
                        G.nv[j] = 0; G.state[j] = 2; G.merged_into[j] = i;
                        std::vector<int>().swap(G.adje[j]); std::vector<int>().swap(G.adjv[j]);
                    }
                }
                a = b;
            }
        }
        // ---- finalise element p and put the survivors back into the degree lists
        std::vector<int> &ev = G.elv[p];
        ev.clear();
        for (int i : Lp) if (G.state[i] == 0 && G.nv[i] > 0) { ev.push_back(i); G.bucket_insert(i); }
        (void)pw;
    }
    // ---- expand supervariables: each principal followed by everything merged into it
    std::vector<std::vector<int>> kids(n);
    for (int j = 0; j < n; j++) if (G.merged_into[j] >= 0) kids[G.merged_into[j]].push_back(j);
    int k = 0;
    std::vector<int> stack;
    for (int p : order) {
        stack.push_back(p);
        while (!stack.empty()) {
            int v = stack.back(); stack.pop_back();
            perm_out[k++] = v;
            for (int c : kids[v]) stack.push_back(c);
        }
    }
    return k == n ? 0 : -1;
}
