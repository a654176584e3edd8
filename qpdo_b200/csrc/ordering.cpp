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
#include <cstdlib>
#include <atomic>
#include <thread>

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

// ------------------------------------------------------------------------------------------------
// Nested dissection by level structures (George's automatic nested dissection) with minimum-degree
// leaves.  A PDE-type graph (CONT-300: a 13-point stencil on a 300 x 300 grid) under minimum degree
// alone ends in a tall, thin top of the assembly tree -- one big front after another, i.e. one serial
// chain of 32-pivot panel steps on the GPU.  Dissection gives the top a logarithmic height and several
// fronts per level.  Separator of a connected piece = one BFS level from a pseudo-peripheral vertex
// (edges only join equal or adjacent levels), chosen among the levels that split the piece 35/65 or
// better for the fewest vertices, then thinned (vertices without a neighbour in the next level move to
// the first part).  Pieces of at most `leaf` vertices, and pieces whose level structure is too shallow
// to cut (random graphs), are ordered by minimum degree on the induced subgraph.
namespace {

// State shared by all runs of the dissection.  The arrays are indexed by vertex; concurrent runs work on disjoint
// vertex sets, BFS stamps come from one atomic counter (a vertex may carry a stale stamp of an ancestor's sweep).
struct NDWork {
    int n;
    const int64_t *gp;
    const int *gi;
    std::vector<int> piece;      // piece id of every vertex (only vertices of the current piece carry its id)
    std::vector<int> level;      // BFS level, valid when lstamp[v] == the sweep's stamp
    std::vector<int> lstamp;
    std::vector<int> local;      // global -> local index inside a leaf
    std::atomic<int> stamp{0};
    std::atomic<int> idgen{1};   // piece labels (0 = the whole graph); labels are only ever compared
    int leaf = 256;
};

// BFS over the vertices with piece[v] == id from `root`; returns the vertices in visiting order, `ends` = level starts
static void nd_bfs(NDWork &W, int id, int root, std::vector<int> &out, std::vector<int> &ends)
{
    out.clear(); ends.clear();
    const int stamp = ++W.stamp;
    W.lstamp[root] = stamp; W.level[root] = 0;
    out.push_back(root);
    size_t head = 0;
    int cur = 0;
    ends.push_back(0);
    while (head < out.size()) {
        const int v = out[head++];
        if (W.level[v] != cur) { cur = W.level[v]; ends.push_back((int)head - 1); }
        for (int64_t t = W.gp[v]; t < W.gp[v + 1]; t++) {
            const int u = W.gi[t];
            if (W.piece[u] != id || W.lstamp[u] == stamp) continue;
            W.lstamp[u] = stamp; W.level[u] = W.level[v] + 1;
            out.push_back(u);
        }
    }
    ends.push_back((int)out.size());
}

static void nd_leaf(NDWork &W, const std::vector<int> &verts, int id, std::vector<int> &order)
{
    const int k = (int)verts.size();
    if (k <= 2) { for (int v : verts) order.push_back(v); return; }
    for (int i = 0; i < k; i++) W.local[verts[i]] = i;
    std::vector<int64_t> lp(k + 1, 0);
    std::vector<int> li;
    for (int i = 0; i < k; i++) {
        const int v = verts[i];
        for (int64_t t = W.gp[v]; t < W.gp[v + 1]; t++) if (W.piece[W.gi[t]] == id) li.push_back(W.local[W.gi[t]]);
        lp[i + 1] = (int64_t)li.size();
    }
    std::vector<int> lperm(k);
    if (li.empty() || qpdo_order_mindeg(k, lp.data(), li.data(), lperm.data()) != 0) { for (int v : verts) order.push_back(v); return; }
    for (int i = 0; i < k; i++) order.push_back(verts[lperm[i]]);
}

// Dissect the piece `V0` (its vertices carry the label `id0`) and append its elimination order to `order`.  While
// `par` > 0 the two parts of a cut of a large piece are dissected by two threads, each with its own output, concatenated
// in the order the one-thread run produces (first part, second part, separator): the result does not depend on `par`.
static void nd_run(NDWork &W, std::vector<int> V0, int id0, std::vector<int> &order, int par)
{
    const int64_t *gp = W.gp; const int *gi = W.gi; const int leaf = W.leaf;
    // explicit stack of tasks: {table slot, kind}; kind 0 = dissect the piece, kind 1 = emit a stored separator
    struct Task { int slot; int kind; };
    std::vector<Task> stack;
    std::vector<std::vector<int>> verts(1), seps(1);        // per slot: the piece's vertices / its separator
    std::vector<int> ids(1, id0);                           // per slot: the piece's label
    verts[0].swap(V0);
    stack.push_back({0, 0});
    std::vector<int> bfs, ends, bfs2, ends2;
    auto new_piece = [&]() { verts.emplace_back(); seps.emplace_back(); ids.push_back(W.idgen++); return (int)verts.size() - 1; };
    while (!stack.empty()) {
        const Task tk = stack.back(); stack.pop_back();
        const int slot = tk.slot, id = ids[slot];
        if (tk.kind == 1) { nd_leaf(W, seps[slot], -1 - id, order); std::vector<int>().swap(seps[slot]); continue; }
        std::vector<int> V; V.swap(verts[slot]);
        const int k = (int)V.size();
        if (k == 0) continue;
        if (k <= leaf) { nd_leaf(W, V, id, order); continue; }
        // connected components of the piece in one sweep (a pattern with few active rows is mostly isolated
        // vertices): small components go straight to the order, the others become pieces of their own
        nd_bfs(W, id, V[0], bfs, ends);
        if ((int)bfs.size() < k) {
            std::vector<int> todo;                          // new slots, dissected in discovery order
            size_t vi = 0;
            std::vector<int> comp; comp.swap(bfs);
            for (;;) {
                if ((int)comp.size() <= leaf && comp.size() <= 2) { for (int v : comp) { W.piece[v] = -2000000000; order.push_back(v); } }
                else {
                    const int sc = new_piece(), idc = ids[sc];
                    verts[sc] = comp;
                    for (int v : comp) W.piece[v] = idc;
                    todo.push_back(sc);
                }
                while (vi < V.size() && W.piece[V[vi]] != id) vi++;
                if (vi >= V.size()) break;
                nd_bfs(W, id, V[vi], comp, ends);
            }
            for (size_t t = todo.size(); t-- > 0;) stack.push_back({todo[t], 0});
            continue;
        }
        // pseudo-peripheral root: restart from a minimum-degree vertex of the last level while the depth grows
        int root = V[0], depth = (int)ends.size() - 1;
        for (int it = 0; it < 4; it++) {
            const int lb = ends[ends.size() - 2], le = ends[ends.size() - 1];
            int best = bfs[lb]; int64_t bd = gp[best + 1] - gp[best];
            for (int t = lb; t < le; t++) { const int64_t dgr = gp[bfs[t] + 1] - gp[bfs[t]]; if (dgr < bd) { bd = dgr; best = bfs[t]; } }
            nd_bfs(W, id, best, bfs2, ends2);
            const int d2 = (int)ends2.size() - 1;
            if (d2 > depth) { depth = d2; root = best; bfs.swap(bfs2); ends.swap(ends2); } else break;
        }
        (void)root;
        const int nlev = (int)ends.size() - 1;
        if (nlev < 5) { nd_leaf(W, V, id, order); continue; }    // too shallow to cut (dense / random graph): minimum degree
        // separator level: among levels leaving >= 35 % on both sides the one with the fewest vertices
        int bestl = -1, bests = k + 1;
        for (int l = 1; l + 1 < nlev; l++) {
            const int before = ends[l], after = k - ends[l + 1], sz = ends[l + 1] - ends[l];
            if (before * 100 >= 35 * k && after * 100 >= 35 * k && sz < bests) { bests = sz; bestl = l; }
        }
        if (bestl < 0) {                                   // no balanced level: the one nearest the middle
            int bestd = k + 1;
            for (int l = 1; l + 1 < nlev; l++) { const int dlt = abs(ends[l] - (k - ends[l + 1])); if (dlt < bestd) { bestd = dlt; bestl = l; } }
        }
        if (bestl < 0) { nd_leaf(W, V, id, order); continue; }
        const int sa = new_piece(), sb = new_piece(), ida = ids[sa], idb = ids[sb];
        std::vector<int> &Va = verts[sa], &Vb = verts[sb], &Sp = seps[slot];
        // levels carry the stamp of the BFS that is in `bfs` now: re-run it if the last probe did not win
        nd_bfs(W, id, bfs[0], bfs, ends);
        for (int t = 0; t < k; t++) {
            const int v = bfs[t], l = W.level[v];
            if (l < bestl) { W.piece[v] = ida; Va.push_back(v); }
            else if (l > bestl) { W.piece[v] = idb; Vb.push_back(v); }
        }
        for (int t = ends[bestl]; t < ends[bestl + 1]; t++) {
            const int v = bfs[t];
            bool touches = false;
            for (int64_t e = gp[v]; e < gp[v + 1] && !touches; e++) touches = (W.piece[gi[e]] == idb);
            if (touches) { W.piece[v] = -1 - id; Sp.push_back(v); }
            else { W.piece[v] = ida; Va.push_back(v); }
        }
        if (par > 0 && (int)Va.size() > 4096 && (int)Vb.size() > 4096) {
            std::vector<int> oa, ob, A, B;
            A.swap(Va); B.swap(Vb);
            std::thread th([&W, &A, ida, &oa, par]() { nd_run(W, std::move(A), ida, oa, par - 1); });
            nd_run(W, std::move(B), idb, ob, par - 1);
            th.join();
            order.insert(order.end(), oa.begin(), oa.end());
            order.insert(order.end(), ob.begin(), ob.end());
            nd_leaf(W, seps[slot], -1 - id, order);
            std::vector<int>().swap(seps[slot]);
            continue;
        }
        stack.push_back({slot, 1});        // emitted after both parts
        stack.push_back({sb, 0});
        stack.push_back({sa, 0});
    }
}

} // namespace

extern "C" int qpdo_order_nd(int n, const int64_t *gp, const int *gi, int *perm_out, int leaf)
{
    if (n <= 0) return 0;
    if (leaf < 8) leaf = 8;
    NDWork W;
    W.n = n; W.gp = gp; W.gi = gi; W.leaf = leaf;
    W.piece.assign(n, 0); W.level.assign(n, 0); W.lstamp.assign(n, 0); W.local.assign(n, 0);
    std::vector<int> order; order.reserve(n);
    std::vector<int> all(n);
    for (int v = 0; v < n; v++) all[v] = v;
    // QPDO_ND_THREADS: 1 = one thread; default: up to 8 (three levels of two-way forks).  Same order either way.
    int par = 3;
    { const char *ev = getenv("QPDO_ND_THREADS"); if (ev) { const int t = atoi(ev); par = t <= 1 ? 0 : (t < 4 ? 1 : (t < 8 ? 2 : 3)); } }
    if (std::thread::hardware_concurrency() < 2) par = 0;
    nd_run(W, std::move(all), 0, order, par);
    if ((int)order.size() != n) return -1;
    for (int i = 0; i < n; i++) perm_out[i] = order[i];
    return 0;
}
