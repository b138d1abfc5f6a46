#include "ordering.h"

#include <algorithm>
#include <cmath>
#include <stdexcept>

namespace spchol {

// ---------------------------------------------------------------------------------------------
// Geometric nested dissection of an nx x ny x nz grid: split the longest side
// (ties: x, then y), order left box, right box, then the separator plane; boxes are emitted in
// natural order (z outer, y, x inner).
// ---------------------------------------------------------------------------------------------
namespace {
struct Box { int x0, x1, y0, y1, z0, z1; bool leaf; };
}

void nd_perm_grid(int nx, int ny, int nz, int64_t* perm) {
    int64_t pos = 0;
    std::vector<Box> stack;
    stack.push_back({0, nx, 0, ny, 0, nz, false});
    while (!stack.empty()) {
        const Box b = stack.back();
        stack.pop_back();
        const int dx = b.x1 - b.x0, dy = b.y1 - b.y0, dz = b.z1 - b.z0;
        if (dx <= 0 || dy <= 0 || dz <= 0) continue;
        if (b.leaf || static_cast<int64_t>(dx) * dy * dz <= 8 || std::max(dx, std::max(dy, dz)) < 3) {
            for (int z = b.z0; z < b.z1; ++z)
                for (int y = b.y0; y < b.y1; ++y) {
                    const int64_t base = static_cast<int64_t>(nx) * (y + static_cast<int64_t>(ny) * z);
                    for (int x = b.x0; x < b.x1; ++x) perm[pos++] = base + x;
                }
            continue;
        }
        // pushed in reverse so they pop as: left, right, separator
        if (dx >= dy && dx >= dz) {
            const int m = b.x0 + dx / 2;
            stack.push_back({m, m + 1, b.y0, b.y1, b.z0, b.z1, true});
            stack.push_back({m + 1, b.x1, b.y0, b.y1, b.z0, b.z1, false});
            stack.push_back({b.x0, m, b.y0, b.y1, b.z0, b.z1, false});
        } else if (dy >= dz) {
            const int m = b.y0 + dy / 2;
            stack.push_back({b.x0, b.x1, m, m + 1, b.z0, b.z1, true});
            stack.push_back({b.x0, b.x1, m + 1, b.y1, b.z0, b.z1, false});
            stack.push_back({b.x0, b.x1, b.y0, m, b.z0, b.z1, false});
        } else {
            const int m = b.z0 + dz / 2;
            stack.push_back({b.x0, b.x1, b.y0, b.y1, m, m + 1, true});
            stack.push_back({b.x0, b.x1, b.y0, b.y1, m + 1, b.z1, false});
            stack.push_back({b.x0, b.x1, b.y0, b.y1, b.z0, m, false});
        }
    }
    if (pos != static_cast<int64_t>(nx) * ny * nz) throw std::runtime_error("nd_perm_grid: internal count mismatch");
}

// ---------------------------------------------------------------------------------------------
// Approximate minimum degree (Amestoy, Davis & Duff, SIAM J. Matrix Anal. Appl. 17(4), 1996).
// Quotient graph: a live variable i keeps the variables it is still directly adjacent to (var[i])
// and the elements (eliminated pivots) it belongs to (elm[i]); an element e keeps its variables
// (mem[e]).  Indistinguishable variables are merged into supervariables of weight nv[i]; degrees
// are the paper's upper bound d_i = min(n_left - nv_i, d_i + |Lp \ i|, |A_i| + |Lp \ i| +
// sum_e |Le \ Lp|) with the |Le \ Lp| obtained by the one-pass counter trick (w[]), and elements
// with Le \ Lp empty are absorbed into the new one.  Rows with more than 10 sqrt(n) entries are
// ordered last.  The result is composed with a postorder of the elimination tree so that the
// columns of a supernode are contiguous.
// ---------------------------------------------------------------------------------------------
void etree(int n, const int* Ap, const int* Ai, int* parent);      // symbolic.cpp
void post_order(int n, const int* parent, int* post);

namespace {

struct DegreeLists {
    std::vector<int> head, next, prev;
    int mindeg;
    explicit DegreeLists(int n) : head(n + 1, -1), next(n, -1), prev(n, -1), mindeg(n) {}
    void insert(int i, int d) {
        next[i] = head[d]; prev[i] = -1;
        if (head[d] >= 0) prev[head[d]] = i;
        head[d] = i;
        if (d < mindeg) mindeg = d;
    }
    void remove(int i, int d) {
        if (prev[i] >= 0) next[prev[i]] = next[i]; else head[d] = next[i];
        if (next[i] >= 0) prev[next[i]] = prev[i];
    }
};

}  // namespace

void order_amd(int n, const int* Ap, const int* Ai, int* perm) {
    if (n <= 0) return;
    // symmetric adjacency without the diagonal
    std::vector<std::vector<int>> var(n), elm(n), mem(n);
    {
        std::vector<int> cnt(n, 0);
        for (int j = 0; j < n; ++j)
            for (int p = Ap[j]; p < Ap[j + 1]; ++p) {
                const int i = Ai[p];
                if (i < 0 || i >= n) throw std::runtime_error("order_amd: row index out of range");
                if (i != j) { ++cnt[i]; ++cnt[j]; }
            }
        for (int i = 0; i < n; ++i) var[i].reserve(cnt[i]);
        for (int j = 0; j < n; ++j)
            for (int p = Ap[j]; p < Ap[j + 1]; ++p) {
                const int i = Ai[p];
                if (i != j) { var[i].push_back(j); var[j].push_back(i); }
            }
        for (int i = 0; i < n; ++i) {          // duplicates in the input would break the counts
            std::sort(var[i].begin(), var[i].end());
            var[i].erase(std::unique(var[i].begin(), var[i].end()), var[i].end());
        }
    }
    enum : char { LIVE = 0, ELEMENT = 1, MERGED = 2, DENSE = 3, DEAD_ELEMENT = 4 };
    std::vector<char> state(n, LIVE);
    std::vector<int> nv(n, 1), deg(n), esize(n, 0), w(n, 0), mark(n, -1), merged_into(n, -1);
    std::vector<int> order;                    // principal variables in elimination order
    order.reserve(n);
    int n_left = n;

    // dense rows last
    std::vector<int> dense;
    {
        double lim = 10.0 * std::sqrt(static_cast<double>(n));
        if (lim < 16) lim = 16;
        for (int i = 0; i < n; ++i)
            if (static_cast<double>(var[i].size()) > lim) { state[i] = DENSE; dense.push_back(i); }
        if (!dense.empty()) {
            for (int i = 0; i < n; ++i) {
                if (state[i] == DENSE) continue;
                auto& v = var[i];
                v.erase(std::remove_if(v.begin(), v.end(), [&](int j) { return state[j] == DENSE; }), v.end());
            }
            n_left -= static_cast<int>(dense.size());
        }
    }
    DegreeLists dl(n);
    for (int i = 0; i < n; ++i)
        if (state[i] == LIVE) { deg[i] = static_cast<int>(var[i].size()); dl.insert(i, deg[i]); }

    int wflg = 1;                              // w[e] < wflg: counter of e not initialised in this step
    auto bump_wflg = [&](int by) {
        if (wflg > 0x7fffffff - by - n - 2) { std::fill(w.begin(), w.end(), 0); wflg = 1; }
        else wflg += by;
    };
    std::vector<int> Lp, hash_of(n, 0), bucket_head(n, -1), bucket_next(n, -1), touched;
    int tag = 0;                               // mark[] stamps

    while (n_left > 0) {
        while (dl.mindeg <= n && dl.head[dl.mindeg] < 0) ++dl.mindeg;
        const int p = dl.head[dl.mindeg];
        dl.remove(p, deg[p]);

        // ---- Lp = (A_p u U_{e in E_p} L_e) \ p over live supervariables -------------------
        Lp.clear();
        ++tag;
        mark[p] = tag;
        int deg_p = 0;
        auto take = [&](int i) {
            if (state[i] != LIVE || mark[i] == tag) return;
            mark[i] = tag; Lp.push_back(i); deg_p += nv[i];
        };
        for (int i : var[p]) take(i);
        for (int e : elm[p]) {
            if (state[e] != ELEMENT) continue;
            for (int i : mem[e]) take(i);
            state[e] = DEAD_ELEMENT;           // absorbed into p
            std::vector<int>().swap(mem[e]);
        }
        std::vector<int>().swap(var[p]);
        std::vector<int>().swap(elm[p]);
        state[p] = ELEMENT;
        order.push_back(p);
        n_left -= nv[p];
        for (int i : Lp) dl.remove(i, deg[i]);

        // ---- pass 1: w[e] - wflg = |L_e \ Lp| for every element adjacent to Lp -------------
        for (int i : Lp)
            for (int e : elm[i]) {
                if (state[e] != ELEMENT) continue;
                if (w[e] >= wflg) w[e] -= nv[i];
                else w[e] = esize[e] + wflg - nv[i];
            }

        // ---- pass 2: prune lists, new degrees, mass elimination ----------------------------
        touched.clear();
        int n_mass = 0;
        for (int i : Lp) {
            auto& E = elm[i];
            auto& V = var[i];
            int64_t d = 0;
            int ke = 0;
            unsigned h = 0;
            for (int e : E) {
                if (state[e] != ELEMENT) continue;
                const int ext = w[e] - wflg;
                if (ext > 0) { d += ext; E[ke++] = e; h += static_cast<unsigned>(e); }
                else { state[e] = DEAD_ELEMENT; std::vector<int>().swap(mem[e]); }     // aggressive absorption
            }
            E.resize(ke);
            int kv = 0;
            for (int j : V) {
                if (state[j] != LIVE || mark[j] == tag) continue;                      // eliminated, merged or now reached through p
                d += nv[j]; V[kv++] = j; h += static_cast<unsigned>(j);
            }
            V.resize(kv);
            if (ke == 0 && kv == 0) {
                // mass elimination: i is adjacent to nothing outside Lp, it can follow p right away
                state[i] = MERGED; merged_into[i] = p;
                deg_p -= nv[i]; nv[p] += nv[i]; n_left -= nv[i]; nv[i] = 0;
                ++n_mass;
                continue;
            }
            E.push_back(p);
            h += static_cast<unsigned>(p);
            // without |Lp \ i|, which is added once Lp and the supervariable weights are final
            if (d < deg[i]) deg[i] = static_cast<int>(d);
            hash_of[i] = static_cast<int>(h % static_cast<unsigned>(n));
        }
        bump_wflg(n + 1);                      // all counters of this step are stale now

        // ---- supervariable detection among Lp (same var and elm lists) ---------------------
        if (n_mass) {
            int k = 0;
            for (int i : Lp) if (state[i] == LIVE) Lp[k++] = i;
            Lp.resize(k);
        }
        for (int i : Lp) {
            const int h = hash_of[i];
            if (bucket_head[h] < 0) touched.push_back(h);
            bucket_next[i] = bucket_head[h]; bucket_head[h] = i;
        }
        for (int h : touched) {
            for (int i = bucket_head[h]; i >= 0; i = bucket_next[i]) {
                if (state[i] != LIVE || bucket_next[i] < 0) continue;
                ++tag;
                for (int e : elm[i]) mark[e] = tag;      // element and variable ids share mark[]: an id is one or the other
                for (int j : var[i]) mark[j] = tag;
                int prev = i;
                for (int j = bucket_next[i]; j >= 0; j = bucket_next[prev]) {
                    bool same = state[j] == LIVE && elm[j].size() == elm[i].size() && var[j].size() == var[i].size();
                    if (same) for (int e : elm[j]) if (mark[e] != tag) { same = false; break; }
                    if (same) for (int v : var[j]) if (mark[v] != tag) { same = false; break; }
                    if (same) {
                        state[j] = MERGED; merged_into[j] = i;
                        nv[i] += nv[j]; nv[j] = 0;
                        std::vector<int>().swap(var[j]);
                        std::vector<int>().swap(elm[j]);
                        bucket_next[prev] = bucket_next[j];
                    } else prev = j;
                }
            }
            bucket_head[h] = -1;
        }

        // ---- element p ----------------------------------------------------------------------
        {
            int k = 0;
            deg_p = 0;
            for (int i : Lp) if (state[i] == LIVE) { Lp[k++] = i; deg_p += nv[i]; }
            Lp.resize(k);
        }
        for (int i : Lp) {
            int d = deg[i] + deg_p - nv[i];
            const int cap = n_left - nv[i];
            if (d > cap) d = cap;
            if (d < 0) d = 0;
            deg[i] = d;
            dl.insert(i, d);
        }
        esize[p] = deg_p;
        if (Lp.empty()) state[p] = DEAD_ELEMENT; else mem[p] = Lp;
    }

    // ---- expand supervariables: every merged variable follows its principal ----------------
    std::vector<int> child_head(n, -1), child_next(n, -1);
    for (int i = n - 1; i >= 0; --i)
        if (merged_into[i] >= 0) { child_next[i] = child_head[merged_into[i]]; child_head[merged_into[i]] = i; }
    std::vector<int> amd;
    amd.reserve(n);
    std::vector<int> stk;
    auto emit_tree = [&](int root) {
        stk.clear(); stk.push_back(root);
        while (!stk.empty()) {
            const int v = stk.back(); stk.pop_back();
            amd.push_back(v);
            for (int c = child_head[v]; c >= 0; c = child_next[c]) stk.push_back(c);
        }
    };
    for (int p : order) emit_tree(p);
    for (int d : dense) emit_tree(d);
    if (static_cast<int>(amd.size()) != n) throw std::runtime_error("order_amd: internal count mismatch");

    // ---- compose with an elimination-tree postorder of the permuted matrix -----------------
    std::vector<int> Bp(n + 1), Bi(Ap[n] > 0 ? Ap[n] : 1);
    permute_upper(n, Ap, Ai, nullptr, amd.data(), Bp.data(), Bi.data(), nullptr, nullptr);
    std::vector<int> parent(n), post(n);
    etree(n, Bp.data(), Bi.data(), parent.data());
    post_order(n, parent.data(), post.data());
    for (int k = 0; k < n; ++k) perm[k] = amd[post[k]];
}

// ---------------------------------------------------------------------------------------------
// triu(P A P^T): entry (i, j) of A goes to (min, max) of (inv[i], inv[j]); a counting sort by row
// then by column leaves the rows of every column ascending.
// ---------------------------------------------------------------------------------------------
void permute_upper(int n, const int* Ap, const int* Ai, const double* Ax, const int* perm,
                   int* Bp, int* Bi, double* Bx, int64_t* map) {
    std::vector<int> inv(n, -1);
    for (int k = 0; k < n; ++k) {
        if (perm[k] < 0 || perm[k] >= n || inv[perm[k]] >= 0) throw std::runtime_error("permute_upper: not a permutation");
        inv[perm[k]] = k;
    }
    const int nz = Ap[n];
    std::vector<int> r(nz), c(nz);
    for (int j = 0; j < n; ++j)
        for (int p = Ap[j]; p < Ap[j + 1]; ++p) {
            const int a = inv[Ai[p]], b = inv[j];
            r[p] = a < b ? a : b; c[p] = a < b ? b : a;
        }
    // stable counting sort by row, then by column
    std::vector<int> cnt(n + 1, 0), by_row(nz);
    for (int p = 0; p < nz; ++p) ++cnt[r[p] + 1];
    for (int i = 0; i < n; ++i) cnt[i + 1] += cnt[i];
    for (int p = 0; p < nz; ++p) by_row[cnt[r[p]]++] = p;
    std::fill(Bp, Bp + n + 1, 0);
    for (int p = 0; p < nz; ++p) ++Bp[c[p] + 1];
    for (int j = 0; j < n; ++j) Bp[j + 1] += Bp[j];
    std::vector<int> cur(Bp, Bp + n);
    for (int q = 0; q < nz; ++q) {
        const int p = by_row[q];
        const int dst = cur[c[p]]++;
        Bi[dst] = r[p];
        if (Bx && Ax) Bx[dst] = Ax[p];
        if (map) map[p] = dst;
    }
}

}  // namespace spchol
