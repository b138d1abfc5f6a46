// Host symbolic analysis (see symbolic.h).  Own implementation of the standard algorithms the
// reference uses (Liu etree, cs_post ordering, Gilbert-Ng-Peyton counts) plus a supernodal
// row-structure builder that is O(sum_s m_s) instead of the reference's O(n^2) schol.
#include "symbolic.h"

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <mutex>
#include <numeric>
#include <stdexcept>

namespace spchol {

// SPCHOL_TIME_HOST=1: print the wall time of every host phase to stderr
struct PhaseTimer {
    bool on = std::getenv("SPCHOL_TIME_HOST") != nullptr;
    std::chrono::steady_clock::time_point t0 = std::chrono::steady_clock::now();
    void lap(const char* what) {
        if (!on) return;
        const auto t1 = std::chrono::steady_clock::now();
        std::fprintf(stderr, "[spchol host] %-28s %8.1f ms\n", what, std::chrono::duration<double, std::milli>(t1 - t0).count());
        t0 = t1;
    }
};

// ---------------------------------------------------------------------------------------------
// Elimination tree (semantics of reference include/chol.hpp:378-410).
// ---------------------------------------------------------------------------------------------
void etree(int n, const int* Ap, const int* Ai, int* parent) {
    std::vector<int> up(n, -1);  // compressed "virtual root" pointers
    for (int col = 0; col < n; ++col) {
        parent[col] = -1;
        for (int p = Ap[col]; p < Ap[col + 1]; ++p) {
            int node = Ai[p];
            // climb from `node` towards the current root, re-pointing everything at `col`
            while (node >= 0 && node < col) {
                const int nxt = up[node];
                up[node] = col;
                if (nxt < 0) { parent[node] = col; break; }
                node = nxt;
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------
// Postorder of the forest, children in increasing index (reference :466-499 / :445-460).
// ---------------------------------------------------------------------------------------------
void post_order(int n, const int* parent, int* post) {
    std::vector<int> first_child(n, -1), sibling(n, -1), stk(n > 0 ? n : 1);
    for (int v = n - 1; v >= 0; --v) {
        const int p = parent[v];
        if (p < 0) continue;
        sibling[v] = first_child[p];
        first_child[p] = v;
    }
    int out = 0;
    for (int root = 0; root < n; ++root) {
        if (parent[root] >= 0) continue;
        int sp = 0;
        stk[0] = root;
        while (sp >= 0) {
            const int v = stk[sp];
            const int c = first_child[v];
            if (c < 0) { post[out++] = v; --sp; }
            else { first_child[v] = sibling[c]; stk[++sp] = c; }
        }
    }
}

// ---------------------------------------------------------------------------------------------
// Column counts of L (reference :567-622; Gilbert, Ng & Peyton skeleton-graph algorithm).
// A is upper CSC, so "row j of the lower triangle" is read through the transposed pattern.
// ---------------------------------------------------------------------------------------------
void col_count(int n, const int* Ap, const int* Ai, const int* parent, const int* post, int* cc) {
    const int nz = Ap[n];
    std::vector<int> Tp(n + 1, 0), Ti(nz > 0 ? nz : 1);
    for (int p = 0; p < nz; ++p) ++Tp[Ai[p] + 1];
    for (int j = 0; j < n; ++j) Tp[j + 1] += Tp[j];
    {
        std::vector<int> cur(Tp.begin(), Tp.end() - 1);
        for (int j = 0; j < n; ++j)
            for (int p = Ap[j]; p < Ap[j + 1]; ++p) Ti[cur[Ai[p]]++] = j;
    }
    std::vector<int> first(n, -1), maxfirst(n, -1), prevleaf(n, -1), setp(n), delta(n, 0);
    std::iota(setp.begin(), setp.end(), 0);
    for (int k = 0; k < n; ++k) {
        int j = post[k];
        delta[j] = (first[j] < 0) ? 1 : 0;
        while (j >= 0 && first[j] < 0) { first[j] = k; j = parent[j]; }
    }
    for (int k = 0; k < n; ++k) {
        const int j = post[k];
        if (parent[j] >= 0) --delta[parent[j]];
        for (int p = Tp[j]; p < Tp[j + 1]; ++p) {
            const int i = Ti[p];
            if (i <= j || first[j] <= maxfirst[i]) continue;   // not a skeleton leaf of row i
            maxfirst[i] = first[j];
            const int prev = prevleaf[i];
            ++delta[j];
            if (prev >= 0) {
                int q = prev;
                while (q != setp[q]) q = setp[q];
                for (int s = prev; s != q;) { const int nx = setp[s]; setp[s] = q; s = nx; }
                --delta[q];
            }
            prevleaf[i] = j;
        }
        if (parent[j] >= 0) setp[j] = parent[j];
    }
    for (int j = 0; j < n; ++j) cc[j] = delta[j];
    for (int j = 0; j < n; ++j) if (parent[j] >= 0) cc[parent[j]] += cc[j];
}

// ---------------------------------------------------------------------------------------------
// ereach with the reference's exact contract (:689-739): the caller owns s/w/x, `w[i] == k`
// means visited, output is s[top..n) in topological order.
// ---------------------------------------------------------------------------------------------
int ereach(int n, const int* Ap, const int* Ai, const double* Ax, int k, const int* parent,
           int* s, int* w, double* x, int top) {
    std::vector<int> chain;
    (void)n;
    for (int p = Ap[k]; p < Ap[k + 1]; ++p) {
        int i = Ai[p];
        if (i > k) continue;
        if (x && Ax) x[i] = Ax[p];
        chain.clear();
        for (; i != -1 && w[i] != k; i = parent[i]) { chain.push_back(i); w[i] = k; }
        for (auto it = chain.rbegin(); it != chain.rend(); ++it) s[--top] = *it;
    }
    return top;
}

// ---------------------------------------------------------------------------------------------
// compute_levels (reference src/chol.cpp:7-40): depth from root, deepest bucket first.
// ---------------------------------------------------------------------------------------------
int compute_levels(int n, const int* parent, int* level_ptr, int* level_nodes) {
    if (n <= 0) return 0;
    std::vector<int> depth(n, -1), chain;
    int maxd = 0;
    for (int j = 0; j < n; ++j) {
        chain.clear();
        int v = j;
        for (; v >= 0 && depth[v] < 0; v = parent[v]) chain.push_back(v);
        int d = (v < 0) ? 0 : depth[v] + 1;
        for (auto it = chain.rbegin(); it != chain.rend(); ++it) depth[*it] = d++;
    }
    for (int j = 0; j < n; ++j) maxd = std::max(maxd, depth[j]);
    const int nlev = maxd + 1;
    std::fill(level_ptr, level_ptr + nlev + 1, 0);
    for (int j = 0; j < n; ++j) ++level_ptr[maxd - depth[j] + 1];
    for (int l = 0; l < nlev; ++l) level_ptr[l + 1] += level_ptr[l];
    std::vector<int> cur(level_ptr, level_ptr + nlev);
    for (int j = 0; j < n; ++j) level_nodes[cur[maxd - depth[j]]++] = j;
    return nlev;
}

// ---------------------------------------------------------------------------------------------
// Fundamental supernodes by the reference's rule (src/chol.cpp:75-85).
// ---------------------------------------------------------------------------------------------
int compute_supernodes(int n, const int* parent, const int64_t* Lp, int* sn_id, int* sn_start) {
    if (n <= 0) return 0;
    int ns = 0;
    sn_start[0] = 0;
    sn_id[0] = 0;
    for (int j = 1; j < n; ++j) {
        const bool chain = parent[j - 1] == j &&
                           (Lp[j + 1] - Lp[j]) == (Lp[j] - Lp[j - 1]) - 1;
        if (!chain) sn_start[++ns] = j;
        sn_id[j] = ns;
    }
    sn_start[++ns] = n;
    return ns;
}

// Assembly tree (reference src/chol.cpp:102-135 scans every row of every column; the minimum
// supernode id among the below-rows is the supernode of parent[last column], which is what we use).
void atree_from_etree(int n, const int* parent, const int* sn_id, int ns, const int* sn_start,
                      int* super_parent) {
    (void)n;
    for (int s = 0; s < ns; ++s) {
        const int pj = parent[sn_start[s + 1] - 1];
        super_parent[s] = pj < 0 ? -1 : sn_id[pj];
    }
}

// ---------------------------------------------------------------------------------------------
// COO -> upper CSC with duplicate summation (reference :309-369). Two-pass bucket sort:
// by row then (stable) by column, so duplicates add up in input order.
// ---------------------------------------------------------------------------------------------
int64_t triplet_to_csc(int64_t nz, const int* ti, const int* tj, const double* tx, int n,
                       int* Ap, int* Ai, double* Ax) {
    std::vector<int64_t> cnt(static_cast<size_t>(n) + 1, 0);
    std::vector<int> r(nz), c(nz);
    for (int64_t k = 0; k < nz; ++k) {
        int i = ti[k], j = tj[k];
        if (i < 0 || i >= n || j < 0 || j >= n) return -1;
        if (j < i) std::swap(i, j);
        r[k] = i; c[k] = j;
    }
    // pass 1: order by row
    std::vector<int64_t> ord1(nz), ord2(nz);
    for (int64_t k = 0; k < nz; ++k) ++cnt[r[k] + 1];
    for (int i = 0; i < n; ++i) cnt[i + 1] += cnt[i];
    for (int64_t k = 0; k < nz; ++k) ord1[cnt[r[k]]++] = k;
    // pass 2: stable by column
    std::fill(cnt.begin(), cnt.end(), 0);
    for (int64_t k = 0; k < nz; ++k) ++cnt[c[k] + 1];
    for (int j = 0; j < n; ++j) cnt[j + 1] += cnt[j];
    for (int64_t q = 0; q < nz; ++q) { const int64_t k = ord1[q]; ord2[cnt[c[k]]++] = k; }
    std::fill(Ap, Ap + n + 1, 0);
    int64_t m = 0;
    int lr = -1, lc = -1;
    for (int64_t q = 0; q < nz; ++q) {
        const int64_t k = ord2[q];
        if (r[k] == lr && c[k] == lc) { Ax[m - 1] += tx[k]; continue; }
        lr = r[k]; lc = c[k];
        Ai[m] = lr; Ax[m] = tx[k]; ++Ap[lc + 1]; ++m;
    }
    for (int j = 0; j < n; ++j) Ap[j + 1] += Ap[j];
    return m;
}

// ---------------------------------------------------------------------------------------------
// Row structure of a supernode partition: for each row i (ascending) and each A(k,i), k < i, walk
// the assembly tree from the supernode of k towards the supernode of i, appending i to every
// supernode not yet marked for row i.  This is the reference's ereach-based schol (:910-943)
// lifted to supernodes: rows come out ascending per supernode, total work = sum_s (m_s - w_s).
// ---------------------------------------------------------------------------------------------
static void walk_rows(int n, const int* Ap, const int* Ai, int ns, const int* sn_id, const int* sn_parent,
                      const int64_t* ptr, int* rows) {
    std::vector<int64_t> cur(ptr, ptr + ns);
    std::vector<int> mark(ns, -1);
    for (int i = 0; i < n; ++i) {
        const int si = sn_id[i];
        for (int p = Ap[i]; p < Ap[i + 1]; ++p) {
            const int k = Ai[p];
            if (k >= i) continue;
            for (int s = sn_id[k]; s != si && s >= 0 && mark[s] != i; s = sn_parent[s]) {
                mark[s] = i;
                if (cur[s] >= ptr[s + 1]) throw std::runtime_error("spchol: internal row-structure overflow");
                rows[cur[s]++] = i;
            }
        }
    }
    for (int s = 0; s < ns; ++s)
        if (cur[s] != ptr[s + 1]) throw std::runtime_error("spchol: internal row-structure count mismatch");
}

// ---------------------------------------------------------------------------------------------
// analyze(): everything schol() produces, at supernode granularity.
// The amalgamation needs the column counts only, so it runs before the row structure, and the row
// structure is built for the fronts alone (walk_rows on the relaxed assembly tree).
// ---------------------------------------------------------------------------------------------
void analyze(int n, const int* Ap, const int* Ai, const RelaxParams& rp, Symbolic& S) {
    if (n <= 0) throw std::runtime_error("spchol: n must be positive");
    S = Symbolic();
    S.n = n;
    S.nnzA = Ap[n];
    S.Ap.assign(Ap, Ap + n + 1);
    S.Ai.assign(Ai, Ai + S.nnzA);
    for (int j = 0; j < n; ++j) {
        if (Ap[j + 1] < Ap[j]) throw std::runtime_error("spchol: column pointers not monotone");
        for (int p = Ap[j]; p < Ap[j + 1]; ++p) {
            if (Ai[p] < 0 || Ai[p] > j) throw std::runtime_error("spchol: A must hold the upper triangle (row <= col)");
            if (p > Ap[j] && Ai[p] <= Ai[p - 1]) throw std::runtime_error("spchol: rows must be strictly ascending in a column");
        }
    }
    PhaseTimer tm;
    S.parent.resize(n); S.post.resize(n); S.colcount.resize(n); S.Lp.resize(n + 1);
    tm.lap("validate + copy A");
    etree(n, Ap, Ai, S.parent.data());
    tm.lap("etree");
    post_order(n, S.parent.data(), S.post.data());
    tm.lap("post_order");
    col_count(n, Ap, Ai, S.parent.data(), S.post.data(), S.colcount.data());
    tm.lap("col_count");
    int64_t nnz = 0; double fl = 0;
    for (int j = 0; j < n; ++j) {
        S.Lp[j] = nnz; nnz += S.colcount[j];
        fl += static_cast<double>(S.colcount[j]) * S.colcount[j];
    }
    S.Lp[n] = nnz; S.nnzL = nnz; S.flops = fl;

    // fundamental supernodes + assembly tree
    S.fsn_id.resize(n); S.fsn_start.resize(n + 1);
    S.nsf = compute_supernodes(n, S.parent.data(), S.Lp.data(), S.fsn_id.data(), S.fsn_start.data());
    S.fsn_start.resize(S.nsf + 1);
    S.fsn_parent.resize(S.nsf);
    atree_from_etree(n, S.parent.data(), S.fsn_id.data(), S.nsf, S.fsn_start.data(), S.fsn_parent.data());

    tm.lap("fundamental supernodes");
    const int nsf = S.nsf;
    S.fbelow_ptr.assign(nsf + 1, 0);
    for (int f = 0; f < nsf; ++f) {
        const int wf = S.fsn_start[f + 1] - S.fsn_start[f];
        S.fbelow_ptr[f + 1] = S.fbelow_ptr[f] + (S.colcount[S.fsn_start[f]] - wf);
    }

    // relaxed amalgamation of contiguous parent/child chains (needs the column counts only)
    S.sn_start.clear(); S.sn_top.clear(); S.sn_padded.clear();
    S.sn_start.reserve(nsf + 1);
    auto mf = [&](int f) { return S.colcount[S.fsn_start[f]]; };
    int g_first = 0;                 // first fundamental supernode of the open group
    int64_t g_w = S.fsn_start[1] - S.fsn_start[0];
    int64_t g_zeros = 0;
    auto lower_entries = [](int64_t w, int64_t m) { return w * m - w * (w - 1) / 2; };
    std::vector<int> nchild(rp.par_min_w > 0 ? nsf : 0, 0);
    if (rp.par_min_w > 0)
        for (int f = 0; f < nsf; ++f) if (S.fsn_parent[f] >= 0) ++nchild[S.fsn_parent[f]];
    auto close_group = [&](int f_end) {  // group = [g_first, f_end)
        S.sn_start.push_back(S.fsn_start[g_first]);
        S.sn_top.push_back(f_end - 1);
        S.sn_padded.push_back(g_zeros > 0 ? 1 : 0);
        S.padded_zeros += g_zeros;
    };
    for (int f = 1; f < nsf; ++f) {
        const int wf = S.fsn_start[f + 1] - S.fsn_start[f];
        bool merge = false;
        int64_t nz_new = 0;
        if (rp.enable && S.fsn_parent[f - 1] == f) {
            // merged rows = columns of the group ++ rows(f): every column of the group is padded
            // up to (its remaining group columns) + m_f
            const int64_t m_group_old = g_w + (mf(f - 1) - (S.fsn_start[f] - S.fsn_start[f - 1]));
            const int64_t m_new = g_w + mf(f);
            nz_new = g_zeros + g_w * (m_new - m_group_old);
            const int64_t w_new = g_w + wf;
            const double frac = static_cast<double>(nz_new) / static_cast<double>(lower_entries(w_new, m_new));
            if (rp.max_m > 0 && m_new > rp.max_m) merge = false;
            else if (rp.par_min_w > 0 && g_w >= rp.par_min_w && nchild[f] > 1) merge = false;
            else if (nz_new == 0) merge = true;   // nothing introduced: always merge
            else if (w_new <= rp.nrelax[0]) merge = true;
            else if (w_new <= rp.nrelax[1]) merge = frac <= rp.zrelax[0];
            else if (w_new <= rp.nrelax[2]) merge = frac <= rp.zrelax[1];
            else merge = frac <= rp.zrelax[2];
        }
        if (merge) { g_w += wf; g_zeros = nz_new; }
        else { close_group(f); g_first = f; g_w = wf; g_zeros = 0; }
    }
    close_group(nsf);
    S.ns = static_cast<int>(S.sn_start.size());
    S.sn_start.push_back(n);
    S.sn_id.resize(n);
    for (int s = 0; s < S.ns; ++s)
        for (int j = S.sn_start[s]; j < S.sn_start[s + 1]; ++j) S.sn_id[j] = s;
    S.sn_parent.resize(S.ns);
    for (int s = 0; s < S.ns; ++s) {
        const int pj = S.parent[S.sn_start[s + 1] - 1];
        S.sn_parent[s] = pj < 0 ? -1 : S.sn_id[pj];
    }
    tm.lap("amalgamation");

    // below-row lists of the fronts: a front's rows below its own columns are those of its top
    // fundamental supernode, so the walk runs on the (much smaller) relaxed assembly tree
    const int ns = S.ns;
    S.below_ptr.assign(ns + 1, 0);
    for (int s = 0; s < ns; ++s) {
        const int f = S.sn_top[s];
        S.below_ptr[s + 1] = S.below_ptr[s] + (S.fbelow_ptr[f + 1] - S.fbelow_ptr[f]);
    }
    S.below_rows.resize(S.below_ptr[ns] > 0 ? S.below_ptr[ns] : 0);
    walk_rows(n, Ap, Ai, ns, S.sn_id.data(), S.sn_parent.data(), S.below_ptr.data(), S.below_rows.data());
    S.panel_entries = 0; S.sum_m = 0;
    for (int s = 0; s < S.ns; ++s) {
        S.panel_entries += static_cast<int64_t>(S.m(s)) * S.w(s);
        S.sum_m += S.m(s);
    }
    S.lazy = std::make_shared<Symbolic::Lazy>();
    tm.lap("below-row lists");
}

// the true (unpadded) pattern below every fundamental supernode: only the 32-bit CSC views of L
// need it (S.i(), the CSC export map), so it is built on first use
const std::vector<int>& Symbolic::fundamental_below() const {
    std::call_once(lazy->once, [&] {
        PhaseTimer tm;
        lazy->rows.resize(fbelow_ptr[nsf] > 0 ? fbelow_ptr[nsf] : 0);
        walk_rows(n, Ap.data(), Ai.data(), nsf, fsn_id.data(), fsn_parent.data(), fbelow_ptr.data(), lazy->rows.data());
        tm.lap("fundamental below-rows");
    });
    return lazy->rows;
}

void Symbolic::fill_Li(int* Li) const {
    const std::vector<int>& fbelow = fundamental_below();
    for (int f = 0; f < nsf; ++f) {
        const int last = fsn_start[f + 1];
        const int* b = fbelow.data() + fbelow_ptr[f];
        const int64_t nb = fbelow_ptr[f + 1] - fbelow_ptr[f];
        for (int j = fsn_start[f]; j < last; ++j) {
            int* out = Li + Lp[j];
            for (int r = j; r < last; ++r) *out++ = r;
            std::copy(b, b + nb, out);
        }
    }
}

}  // namespace spchol
