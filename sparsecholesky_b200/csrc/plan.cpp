// Host planner (see plan.h): fronts, relative-index maps, workspace layout, launch program.
#include "plan.h"

#include <algorithm>
#include <map>
#include <stdexcept>

namespace spchol {

namespace {

inline int even_up(int v) { return (v + 1) & ~1; }
inline int64_t even_up64(int64_t v) { return (v + 1) & ~int64_t(1); }

// Best-fit free-list allocator for contribution blocks (offsets in doubles).  Lifetimes are
// "from the front's level until its parent's level has run", so memory is recycled level by
// level; the high-water mark is the workspace size.
class Arena {
public:
    int64_t alloc(int64_t size) {
        auto it = by_size_.lower_bound(size);
        if (it == by_size_.end()) {
            // extend the top (merge with a free block that touches the top, if any)
            int64_t off = top_;
            if (!by_off_.empty()) {
                auto last = std::prev(by_off_.end());
                if (last->first + last->second == top_) {
                    off = last->first;
                    erase_size(last->second, last->first);
                    by_off_.erase(last);
                }
            }
            top_ = off + size;
            high_ = std::max(high_, top_);
            return off;
        }
        const int64_t bsz = it->first, off = it->second;
        by_size_.erase(it);
        by_off_.erase(off);
        if (bsz > size) insert(off + size, bsz - size);
        return off;
    }
    void release(int64_t off, int64_t size) {
        auto nxt = by_off_.lower_bound(off);
        if (nxt != by_off_.end() && off + size == nxt->first) {
            size += nxt->second;
            erase_size(nxt->second, nxt->first);
            nxt = by_off_.erase(nxt);
        }
        if (nxt != by_off_.begin()) {
            auto prv = std::prev(nxt);
            if (prv->first + prv->second == off) {
                off = prv->first; size += prv->second;
                erase_size(prv->second, prv->first);
                by_off_.erase(prv);
            }
        }
        insert(off, size);
    }
    int64_t high_water() const { return high_; }
private:
    void insert(int64_t off, int64_t size) { by_off_[off] = size; by_size_.emplace(size, off); }
    void erase_size(int64_t size, int64_t off) {
        auto r = by_size_.equal_range(size);
        for (auto it = r.first; it != r.second; ++it) if (it->second == off) { by_size_.erase(it); return; }
    }
    std::map<int64_t, int64_t> by_off_;
    std::multimap<int64_t, int64_t> by_size_;
    int64_t top_ = 0, high_ = 0;
};

}  // namespace

// ---------------------------------------------------------------------------------------------
// Proportional mapping of the assembly tree onto `world` ranks.  Top-down: a front that owns a
// rank group [lo, lo+n) with n > 1 deals the group out to its children -- the largest children
// are seeded with one rank each, spare ranks go one by one to the child with the highest load per
// rank, children that got no rank of their own are packed onto the least loaded rank (LPT).  A
// front whose group has one rank pins its whole subtree there.  Fronts with a multi-rank group are
// factored by the group when they are at least dist_min_w columns wide, otherwise by rank `lo`.
// ---------------------------------------------------------------------------------------------
void map_fronts(int ns, const int* parent, const int* w, const int* m, int world, int dist_min_w,
                std::vector<int>& own_lo, std::vector<int>& own_n, std::vector<double>& rank_load) {
    own_lo.assign(ns, 0); own_n.assign(ns, 1); rank_load.assign(std::max(world, 1), 0.0);
    std::vector<double> work(ns), sub(ns);
    for (int s = 0; s < ns; ++s) {
        const double mm = m[s], ww = w[s];
        // sum_{k=0}^{w-1} (m-k)^2
        work[s] = ww * mm * mm - mm * ww * (ww - 1) + (ww - 1) * ww * (2 * ww - 1) / 6.0;
        sub[s] = work[s];
    }
    for (int s = 0; s < ns; ++s) if (parent[s] >= 0) sub[parent[s]] += sub[s];
    if (world <= 1) { for (int s = 0; s < ns; ++s) rank_load[0] += work[s]; return; }

    std::vector<int> cptr(ns + 2, 0), cidx;
    // slot ns = virtual root holding the roots of the forest
    for (int s = 0; s < ns; ++s) ++cptr[(parent[s] >= 0 ? parent[s] : ns) + 1];
    for (int s = 0; s <= ns; ++s) cptr[s + 1] += cptr[s];
    cidx.resize(cptr[ns + 1]);
    {
        std::vector<int> cur(cptr.begin(), cptr.end() - 1);
        for (int s = 0; s < ns; ++s) cidx[cur[parent[s] >= 0 ? parent[s] : ns]++] = s;
    }
    std::vector<int> glo(ns + 1, 0), gn(ns + 1, 1);     // group handed to each front by its parent
    glo[ns] = 0; gn[ns] = world;
    std::vector<int> kids, cnt;
    std::vector<double> bin;
    for (int s = ns; s >= 0; --s) {                       // parents before children
        const int lo = glo[s], n = gn[s];
        if (s < ns) {
            if (n > 1 && w[s] >= dist_min_w) {
                own_lo[s] = lo; own_n[s] = n;
                for (int r = 0; r < n; ++r) rank_load[lo + r] += work[s] / n;
            } else {
                own_lo[s] = lo; own_n[s] = 1;
                rank_load[lo] += work[s];
            }
        }
        const int kb = cptr[s], ke = cptr[s + 1];
        if (kb == ke) continue;
        if (n == 1) { for (int q = kb; q < ke; ++q) { glo[cidx[q]] = lo; gn[cidx[q]] = 1; } continue; }
        kids.assign(cidx.begin() + kb, cidx.begin() + ke);
        std::stable_sort(kids.begin(), kids.end(), [&](int a, int b) { return sub[a] > sub[b]; });
        const int k = static_cast<int>(kids.size()), seeds = std::min(n, k);
        cnt.assign(k, 0);
        for (int i = 0; i < seeds; ++i) cnt[i] = 1;
        for (int spare = n - seeds; spare > 0; --spare) {
            int best = 0;
            for (int i = 1; i < seeds; ++i) if (sub[kids[i]] / cnt[i] > sub[kids[best]] / cnt[best]) best = i;
            ++cnt[best];
        }
        bin.assign(n, 0.0);
        int cur = lo;
        for (int i = 0; i < seeds; ++i) {
            glo[kids[i]] = cur; gn[kids[i]] = cnt[i];
            for (int r = 0; r < cnt[i]; ++r) bin[cur - lo + r] += sub[kids[i]] / cnt[i];
            cur += cnt[i];
        }
        for (int i = seeds; i < k; ++i) {
            int best = 0;
            for (int r = 1; r < n; ++r) if (bin[r] < bin[best]) best = r;
            glo[kids[i]] = lo + best; gn[kids[i]] = 1;
            bin[best] += sub[kids[i]];
        }
    }
}

void build_plan(const Symbolic& S, const PlanOptions& opt, HostPlan& P) {
    P = HostPlan();
    const int ns = S.ns, n = S.n;
    const int rank = opt.rank, world = std::max(opt.world, 1);
    if (rank < 0 || rank >= world) throw std::runtime_error("spchol: rank outside [0, world)");
    const int nb = opt.nb, NB_base = std::max(opt.NB, opt.nb), NB_wide = std::max(opt.NB_wide, NB_base);
    const int DNB = std::max(opt.dist_NB, 1);
    P.n = n; P.ns = ns; P.nnzA = S.nnzA; P.nnzL = S.nnzL; P.flops = S.flops;
    P.rank = rank; P.world = world; P.dist_NB = DNB;
    P.first.resize(ns); P.w.resize(ns); P.m.resize(ns); P.ld.resize(ns);
    P.parent = S.sn_parent;
    P.panel_off.assign(ns, -1); P.cb_off.assign(ns, -1); P.rel_off.resize(ns + 1);
    for (int s = 0; s < ns; ++s) {
        P.first[s] = S.sn_start[s];
        P.w[s] = S.w(s);
        P.m[s] = S.m(s);
        P.ld[s] = even_up(P.m[s]);
        P.max_m = std::max(P.max_m, P.m[s]);
        P.max_w = std::max(P.max_w, P.w[s]);
    }
    map_fronts(ns, P.parent.data(), P.w.data(), P.m.data(), world, opt.dist_min_w, P.own_lo, P.own_n, P.rank_load);
    P.own_idx.assign(ns, -1);
    P.front_marked.assign(ns, 0);
    int64_t off = 0, roff = 0;
    for (int s = 0; s < ns; ++s) {
        if (rank >= P.own_lo[s] && rank < P.own_lo[s] + P.own_n[s]) {
            P.own_idx[s] = rank - P.own_lo[s];
            P.panel_off[s] = off;
            off += static_cast<int64_t>(P.ld[s]) * P.w[s];
            if (P.own_n[s] > 1) ++P.n_dist;
        }
    }
    P.lbuf_entries = off;
    // relative indices are needed where the parent is assembled, i.e. on the parent's holders
    for (int s = 0; s < ns; ++s) {
        P.rel_off[s] = roff;
        const int p = P.parent[s];
        if (p >= 0 && P.holds(p)) roff += P.m[s] - P.w[s];
    }
    P.rel_off[ns] = roff;

    // children (ascending) and height-based levels
    P.child_ptr.assign(ns + 1, 0);
    for (int s = 0; s < ns; ++s) if (P.parent[s] >= 0) ++P.child_ptr[P.parent[s] + 1];
    for (int s = 0; s < ns; ++s) P.child_ptr[s + 1] += P.child_ptr[s];
    P.child_idx.resize(P.child_ptr[ns]);
    {
        std::vector<int> cur(P.child_ptr.begin(), P.child_ptr.end() - 1);
        for (int s = 0; s < ns; ++s) if (P.parent[s] >= 0) P.child_idx[cur[P.parent[s]]++] = s;
    }
    P.level.assign(ns, 0);
    int nlev = 0;
    for (int s = 0; s < ns; ++s) {   // children have smaller ids than parents
        const int p = P.parent[s];
        if (p >= 0) {
            if (p <= s) throw std::runtime_error("spchol: assembly tree is not topologically ordered");
            P.level[p] = std::max(P.level[p], P.level[s] + 1);
        }
        nlev = std::max(nlev, P.level[s] + 1);
    }
    P.nlevels = nlev;
    std::vector<std::vector<int>> by_level(nlev);
    for (int s = 0; s < ns; ++s) by_level[P.level[s]].push_back(s);

    // relative indices: position of every below-row of s inside the parent's front
    P.rel.resize(roff);
    for (int s = 0; s < ns; ++s) {
        const int p = P.parent[s];
        const int nbr = P.m[s] - P.w[s];
        if (nbr == 0) continue;
        if (p < 0) throw std::runtime_error("spchol: root front with below rows");
        if (!P.holds(p)) continue;
        const int* rs = S.below(s);
        const int* rp = S.below(p);
        const int nbp = P.m[p] - P.w[p];
        const int pend = P.first[p] + P.w[p];
        int* out = P.rel.data() + P.rel_off[s];
        int q = 0;
        for (int t = 0; t < nbr; ++t) {
            const int r = rs[t];
            if (r < pend) { out[t] = r - P.first[p]; continue; }
            while (q < nbp && rp[q] < r) ++q;
            if (q >= nbp || rp[q] != r) throw std::runtime_error("spchol: child row missing in parent front");
            out[t] = P.w[p] + q;
        }
    }

    // A scatter map (A is upper CSC: entry (i,j), i <= j, lands at L(j,i)); -1 where the front
    // of column i is not held by this rank
    P.a_dest.resize(S.nnzA);
    const bool want_tree = opt.tree != 0 && world == 1;
    std::vector<int> a_front, a_fcol, a_fpos;          // per A entry: front, column inside it, row position inside it (tree programs)
    if (want_tree) { a_front.resize(S.nnzA); a_fcol.resize(S.nnzA); a_fpos.resize(S.nnzA); }
    for (int j = 0; j < n; ++j) {
        for (int p = S.Ap[j]; p < S.Ap[j + 1]; ++p) {
            const int i = S.Ai[p];
            const int s = S.sn_id[i];
            const int c = i - P.first[s];
            int pos;
            if (j < P.first[s] + P.w[s]) pos = j - P.first[s];
            else {
                const int* b = S.below(s);
                const int nbr = P.m[s] - P.w[s];
                const int* it = std::lower_bound(b, b + nbr, j);
                if (it == b + nbr || *it != j) throw std::runtime_error("spchol: A entry outside the symbolic pattern");
                pos = P.w[s] + static_cast<int>(it - b);
            }
            P.a_dest[p] = P.holds(s) ? P.panel_off[s] + static_cast<int64_t>(c) * P.ld[s] + pos : -1;
            if (want_tree) { a_front[p] = s; a_fcol[p] = c; a_fpos[p] = pos; }
        }
    }

    P.a_row.assign(S.Ai.begin(), S.Ai.end());
    P.a_col.resize(S.nnzA);
    for (int j = 0; j < n; ++j) for (int p = S.Ap[j]; p < S.Ap[j + 1]; ++p) P.a_col[p] = j;
    if (world > 1) {
        P.ridx_off.assign(ns, -1);
        int64_t o = 0;
        for (int s = 0; s < ns; ++s) if (P.holds(s) && P.own_idx[s] == 0) { P.ridx_off[s] = o; o += P.m[s] - P.w[s]; }
        P.ridx.resize(o);
        for (int s = 0; s < ns; ++s) if (P.ridx_off[s] >= 0) std::copy(S.below(s), S.below(s) + (P.m[s] - P.w[s]), P.ridx.begin() + P.ridx_off[s]);
    }

    // CSC export map; a panel replicated over a rank group is exported by the group's first rank
    if (opt.want_csc) {
        P.l_src.resize(S.nnzL);
        const std::vector<int>& fbelow = S.fundamental_below();
        for (int j = 0; j < n; ++j) {
            const int s = S.sn_id[j], f = S.fsn_id[j];
            int64_t q = S.Lp[j];
            if (!P.holds(s) || (P.own_n[s] > 1 && P.own_idx[s] != 0)) {
                for (; q < S.Lp[j + 1]; ++q) P.l_src[q] = -1;
                continue;
            }
            const int c = j - P.first[s];
            const int64_t colbase = P.panel_off[s] + static_cast<int64_t>(c) * P.ld[s];
            const int fend = S.fsn_start[f + 1];
            for (int r = j; r < fend; ++r) P.l_src[q++] = colbase + (r - P.first[s]);
            const int* fb = fbelow.data() + S.fbelow_ptr[f];
            const int nfb = static_cast<int>(S.fbelow_ptr[f + 1] - S.fbelow_ptr[f]);
            if (S.sn_top[s] == f) {
                for (int t = 0; t < nfb; ++t) P.l_src[q++] = colbase + P.w[s] + t;
            } else {
                const int* b = S.below(s);
                const int nbr = P.m[s] - P.w[s];
                const int send = P.first[s] + P.w[s];
                int k = 0;
                for (int t = 0; t < nfb; ++t) {
                    const int r = fb[t];
                    if (r < send) { P.l_src[q++] = colbase + (r - P.first[s]); continue; }
                    while (k < nbr && b[k] < r) ++k;
                    if (k >= nbr || b[k] != r) throw std::runtime_error("spchol: export map row missing");
                    P.l_src[q++] = colbase + P.w[s] + k;
                }
            }
        }
    }

    // ---- contribution-block workspace + launch program, level by level ----------------------
    Arena arena;
    std::vector<int64_t> ws_size(ns, 0);      // what this rank has allocated for front s's cb (own or staged copy)
    auto cb_entries = [&](int s) { const int mb = P.m[s] - P.w[s]; return static_cast<int64_t>(even_up(mb)) * mb; };
    auto ws_alloc = [&](int s) {
        ws_size[s] = even_up64(cb_entries(s));
        P.cb_off[s] = arena.alloc(ws_size[s]);
    };
    // panels in address order, to find the front an operand offset belongs to
    std::vector<std::pair<int64_t, int>> pan;
    for (int s = 0; s < ns; ++s) if (P.holds(s)) pan.push_back({P.panel_off[s], s});
    std::sort(pan.begin(), pan.end());
    std::vector<int> front_tmap(ns, -1);
    auto add_gemm = [&](Op& op, GemmTask ge) {
        {   // TMA view: both operands are row ranges of one front's panel
            auto it = std::upper_bound(pan.begin(), pan.end(), std::make_pair(ge.a_off, ns));
            if (it == pan.begin()) throw std::runtime_error("spchol: update operand outside every panel");
            const int s = std::prev(it)->second;
            const int64_t po = P.panel_off[s], ea = ge.a_off - po, eb = ge.b_off - po;
            const int ld = P.ld[s];
            if (eb < 0 || ea / ld != eb / ld || ea / ld + ge.K > P.w[s] || ge.lda != ld || ge.ldb != ld)
                throw std::runtime_error("spchol: update operands are not row ranges of one panel");
            if (front_tmap[s] < 0) { front_tmap[s] = static_cast<int>(P.tmap_fronts.size()); P.tmap_fronts.push_back(s); }
            ge.tmap = front_tmap[s]; ge.ra = static_cast<int>(ea % ld); ge.rb = static_cast<int>(eb % ld); ge.kc = static_cast<int>(ea / ld);
        }
        P.gemm_tasks.push_back(ge);
        op.gx = std::max(op.gx, ge.M); op.gy = std::max(op.gy, ge.N);
        op.threads = std::max(op.threads, ge.K);                    // OP_GEMM: largest K of the launch (profiling classes)
        P.gemm_flops += 2.0 * ge.M * ge.N * ge.K;
        // wanted entries: column j holds rows i >= j - diag
        double cnt = 0;
        const int Mq = ge.M, Nq = ge.N, dg = ge.diag;
        const int jfull = std::min(Nq, std::max(0, dg + 1));      // columns with all M rows
        cnt += static_cast<double>(jfull) * Mq;
        const int jlast = std::min(Nq, Mq + dg);                    // beyond this: empty
        if (jlast > jfull) {
            const double a = Mq - (jfull - dg), b = Mq - (jlast - 1 - dg);
            cnt += (a + b) * (jlast - jfull) / 2.0;
        }
        P.gemm_flops_alg += 2.0 * cnt * ge.K;
        P.gemm_flops_alg_k[gemm_k_class(ge.K)] += 2.0 * cnt * ge.K;
        P.gemm_bytes_min += 8.0 * ((static_cast<double>(ge.M) + ge.N) * ge.K + ((ge.c_space & 2) ? 1.0 : 2.0) * cnt);
    };
    auto add_comm = [&](int kind, int lo, int cn, int root, int space, int64_t coff, int64_t count, int lv) {
        Op op{}; op.kind = OP_COMM; op.begin = static_cast<int>(P.comm_tasks.size()); op.count = 1; op.level = lv;
        P.comm_tasks.push_back({kind, lo, cn, root, space, coff, count});
        P.ops.push_back(op);
        P.comm_doubles += count;
    };

    // stream ordering (see Op in plan.h): a record hands out a fresh event id, the wait names it
    auto ev_record = [&](int stream, int lv) {
        Op op{}; op.kind = OP_EV_RECORD; op.begin = P.n_events++; op.stream = stream; op.level = lv;
        P.ops.push_back(op);
        return op.begin;
    };
    auto ev_wait = [&](int stream, int ev, int lv) {
        if (P.n_events - ev > kEventRing) throw std::runtime_error("spchol: event waited on after its ring slot was reused");
        Op op{}; op.kind = OP_EV_WAIT; op.begin = ev; op.stream = stream; op.level = lv;
        P.ops.push_back(op);
    };

    static const int cls_m[] = {16, 24, 32, 48, 64, 96, 128, 1 << 30};
    static const int cls_thr[] = {32, 32, 64, 64, 128, 128, 256, 256};
    const int ncls = 8;

    // fronts of one level that live on this rank only: small ones one CTA each, large ones in
    // lockstep (zero cb, extend-add rounds, blocked panel factorization, Schur complement)
    auto emit_local = [&](const std::vector<int>& fronts, int lv) {
        std::vector<int> small[8], large;
        for (int s : fronts) {
            if (P.m[s] <= opt.small_max_m) {
                int c = 0;
                while (P.m[s] > cls_m[c]) ++c;
                small[c].push_back(s);
            } else large.push_back(s);
        }
        for (int c = 0; c < ncls; ++c) {
            if (small[c].empty()) continue;
            int mm = 0;
            for (int s : small[c]) mm = std::max(mm, P.m[s]);
            Op op{};
            op.kind = OP_SMALL; op.begin = static_cast<int>(P.small_list.size());
            op.count = static_cast<int>(small[c].size());
            op.threads = cls_thr[c];
            op.smem = mm * (mm + 1) / 2 * 8;             // packed lower triangle
            op.level = lv;
            P.small_list.insert(P.small_list.end(), small[c].begin(), small[c].end());
            P.ops.push_back(op);
            P.n_small += op.count;
        }
        if (large.empty()) return;
        P.n_large += static_cast<int64_t>(large.size());
        P.max_potrf_slots = std::max(P.max_potrf_slots, static_cast<int>(large.size()));
        const bool defer = opt.defer_cb_add != 0;
        if (!defer) {
            Op op{}; op.kind = OP_ZERO; op.begin = static_cast<int>(P.zero_tasks.size()); op.level = lv;
            int64_t zmax = 0;
            for (int s : large) if (P.cb_off[s] >= 0) { P.zero_tasks.push_back({P.cb_off[s], cb_entries(s)}); zmax = std::max(zmax, cb_entries(s)); }
            op.count = static_cast<int>(P.zero_tasks.size()) - op.begin;
            op.gx = static_cast<int>(std::min<int64_t>((zmax + 2047) / 2048, 1184));      // slices of 2048 doubles per CTA pass
            if (op.count) P.ops.push_back(op);
        }
        int max_children = 0;
        for (int s : large) max_children = std::max(max_children, P.child_ptr[s + 1] - P.child_ptr[s]);
        // one child per parent per round (no two CTAs of a launch touch the same parent).  mode 0: the whole block;
        // with defer_cb_add the columns landing in the parent's panel go in before the factorization (mode 1) and
        // those landing in its contribution block after the Schur complement has been stored there (mode 2)
        auto emit_extadd = [&](int mode) {
            for (int r = 0; r < max_children; ++r) {
                Op op{}; op.kind = OP_EXTADD; op.begin = static_cast<int>(P.extadd_list.size()); op.level = lv; op.threads = mode;
                int maxmb = 0;
                for (int s : large) {
                    if (P.child_ptr[s] + r >= P.child_ptr[s + 1]) continue;
                    if (mode == 2 && P.m[s] == P.w[s]) continue;
                    const int c = P.child_idx[P.child_ptr[s] + r];
                    if (P.m[c] == P.w[c]) continue;
                    P.extadd_list.push_back(c);
                    maxmb = std::max(maxmb, P.m[c] - P.w[c]);
                }
                op.count = static_cast<int>(P.extadd_list.size()) - op.begin;
                op.gx = maxmb;
                if (op.count) P.ops.push_back(op);
            }
        };
        emit_extadd(defer ? 1 : 0);
        // Blocked right-looking factorization of the panels, all fronts of the level in lockstep.
        // Outer step J covers the block column [KB, KE) of width NB:
        //   panel phase  : inner blocks of width nb -- diagonal block + inverse, panel solve, update
        //                  of the rest of the outer block (K = nb)
        //   update phase : (a) the NEXT outer block column [KE, KE2) gets the deferred update (K = NB),
        //                  (b) then the remaining columns [KE2, w); a front whose panel is complete
        //                  computes its Schur complement (K = w) instead.
        // With look-ahead the panel phase of step J+1 (stream 1) only waits for (a) of step J, so it
        // overlaps (b) on stream 0: the latency-bound panel kernels hide behind the big updates.
        int maxJ = 0, maxw = 0;
        for (int s : large) maxw = std::max(maxw, P.w[s]);
        const int NB = maxw >= opt.wide_w ? NB_wide : NB_base;
        for (int s : large) maxJ = std::max(maxJ, (P.w[s] + NB - 1) / NB);
        const bool pipe = opt.lookahead && maxJ > 1;
        const int sp = pipe ? 1 : 0;
        if (pipe) { const int e = ev_record(0, lv); ev_wait(1, e, lv); }     // assembly done before the first panel
        int ev_next = -1;
        const int max_inner = (NB + nb - 1) / nb;
        for (int J = 0; J < maxJ; ++J) {
            if (pipe && ev_next >= 0) ev_wait(1, ev_next, lv);
            for (int i = 0; i < max_inner; ++i) {
                // diagonal blocks and panel solves, one launch per size class of the block width
                for (int cls = 0; cls < 3; ++cls) {
                    const int lo_nb = cls == 0 ? 0 : (cls == 1 ? 32 : 64), hi_nb = cls == 0 ? 32 : (cls == 1 ? 64 : 1 << 30);
                    Op a{}, b{};
                    a.kind = OP_POTRF; a.begin = static_cast<int>(P.potrf_tasks.size()); a.level = lv; a.threads = 32 << cls; a.stream = sp;
                    b.kind = OP_TRSM;  b.begin = static_cast<int>(P.trsm_tasks.size());  b.level = lv; b.threads = 32 << cls; b.stream = sp;
                    for (size_t li = 0; li < large.size(); ++li) {
                        const int s = large[li];
                        const int w = P.w[s], m = P.m[s], ld = P.ld[s];
                        const int kb = J * NB + i * nb;
                        if (kb >= w || kb >= std::min((J + 1) * NB, w)) continue;
                        const int ke = std::min(std::min(kb + nb, (J + 1) * NB), w);
                        if (ke - kb <= lo_nb || ke - kb > hi_nb) continue;
                        const int64_t po = P.panel_off[s];
                        const int64_t d = po + kb + static_cast<int64_t>(kb) * ld;
                        P.potrf_tasks.push_back({d, ld, ke - kb, P.first[s] + kb, static_cast<int>(li)});
                        if (ke < m) {
                            P.trsm_tasks.push_back({po + ke + static_cast<int64_t>(kb) * ld, ld, ke - kb, m - ke, d, static_cast<int>(li)});
                            b.gx = std::max(b.gx, m - ke);
                        }
                    }
                    a.count = static_cast<int>(P.potrf_tasks.size()) - a.begin;
                    b.count = static_cast<int>(P.trsm_tasks.size()) - b.begin;
                    if (a.count) P.ops.push_back(a);
                    if (b.count) P.ops.push_back(b);
                }
                Op c{};
                c.kind = OP_GEMM; c.begin = static_cast<int>(P.gemm_tasks.size()); c.level = lv; c.stream = sp;
                for (int s : large) {
                    const int w = P.w[s], m = P.m[s], ld = P.ld[s];
                    const int KE = std::min((J + 1) * NB, w);
                    const int kb = J * NB + i * nb;
                    if (kb >= KE) continue;
                    const int ke = std::min(kb + nb, KE);
                    if (ke >= KE) continue;               // last inner block: nothing left inside the outer block
                    const int64_t po = P.panel_off[s];
                    const int64_t a2 = po + ke + static_cast<int64_t>(kb) * ld;
                    add_gemm(c, {po + ke + static_cast<int64_t>(ke) * ld, a2, a2, ld, ld, ld, m - ke, KE - ke, ke - kb, 0, 0});
                }
                c.count = static_cast<int>(P.gemm_tasks.size()) - c.begin;
                if (c.count) P.ops.push_back(c);
            }
            if (opt.export_mark_w > 0) {        // block column J of the wide fronts is final: its read-back may start
                Op ex{}; ex.kind = OP_EXPORT; ex.begin = static_cast<int>(P.export_tasks.size()); ex.level = lv; ex.stream = sp;
                for (int s : large) {
                    const int w = P.w[s], ld = P.ld[s];
                    const int KB = J * NB, KE = std::min(KB + NB, w);
                    if (w < opt.export_mark_w || KB >= w) continue;
                    P.export_tasks.push_back({P.panel_off[s] + static_cast<int64_t>(KB) * ld, static_cast<int64_t>(KE - KB) * ld});
                    P.front_marked[s] = 1;
                }
                ex.count = static_cast<int>(P.export_tasks.size()) - ex.begin;
                if (ex.count) P.ops.push_back(ex);
            }
            if (pipe) { const int e = ev_record(1, lv); ev_wait(0, e, lv); }  // panels of step J are final
            Op un{}, ur{};
            un.kind = OP_GEMM; un.begin = static_cast<int>(P.gemm_tasks.size()); un.level = lv;
            for (int s : large) {
                const int w = P.w[s], m = P.m[s], ld = P.ld[s];
                const int KB = J * NB, KE = std::min(KB + NB, w);
                if (KB >= w || KE >= w) continue;
                const int KE2 = std::min(KE + NB, w);
                const int64_t po = P.panel_off[s];
                const int64_t a2 = po + KE + static_cast<int64_t>(KB) * ld;
                add_gemm(un, {po + KE + static_cast<int64_t>(KE) * ld, a2, a2, ld, ld, ld, m - KE, KE2 - KE, KE - KB, 0, 0});
            }
            un.count = static_cast<int>(P.gemm_tasks.size()) - un.begin;
            if (un.count) P.ops.push_back(un);
            ev_next = (pipe && un.count) ? ev_record(0, lv) : -1;
            ur.kind = OP_GEMM; ur.begin = static_cast<int>(P.gemm_tasks.size()); ur.level = lv;
            for (int s : large) {
                const int w = P.w[s], m = P.m[s], ld = P.ld[s];
                const int KB = J * NB, KE = std::min(KB + NB, w);
                if (KB >= w) continue;
                const int64_t po = P.panel_off[s];
                if (KE >= w) {                           // panel complete: Schur complement into the cb, K = w
                    if (m > w) { const int64_t a2 = po + w; add_gemm(ur, {P.cb_off[s], a2, a2, even_up(m - w), ld, ld, m - w, m - w, w, 0, defer ? 3 : 1}); }
                    continue;
                }
                const int KE2 = std::min(KE + NB, w);
                if (KE2 >= w) continue;
                const int64_t a2 = po + KE2 + static_cast<int64_t>(KB) * ld;
                add_gemm(ur, {po + KE2 + static_cast<int64_t>(KE2) * ld, a2, a2, ld, ld, ld, m - KE2, w - KE2, KE - KB, 0, 0});
            }
            ur.count = static_cast<int>(P.gemm_tasks.size()) - ur.begin;
            if (ur.count) P.ops.push_back(ur);
        }
        if (defer) emit_extadd(2);
    };

    auto add_copy = [&](Op& op, int sspace, int64_t soff, int lds, int dspace, int64_t doff, int ldd, int rows, int cols) {
        P.copy_tasks.push_back({soff, doff, sspace, dspace, lds, ldd, rows, cols});
        op.gx = std::max(op.gx, cols);
        op.gy = std::max(op.gy, (rows + 2047) / 2048);          // row slices
    };
    std::vector<int64_t> tmp_allocs;          // (offset, size) pairs of transfer buffers, released at the end of the level
    auto tmp_alloc = [&](int64_t entries) {
        const int64_t sz = even_up64(std::max<int64_t>(entries, 2));
        const int64_t o = arena.alloc(sz);
        tmp_allocs.push_back(o); tmp_allocs.push_back(sz);
        return o;
    };

    // Move child c's contribution block to the holders of its parent p (called by every rank in the
    // same order; ranks that are neither source nor destination emit nothing).  The block travels
    // in column blocks of width DNB, each packed to its rows on and below the diagonal, and goes
    // only to the ranks that assemble one of its columns: with a parent factored by a group that is
    // the owner of the parent block column the child column maps to.  One grouped exchange per child.
    auto emit_transfer = [&](int c, int p, int lv) {
        const int mbc = P.m[c] - P.w[c];
        if (mbc == 0) return;
        const int ldc = even_up(mbc);
        const int cl = P.own_lo[c], cn = P.own_n[c], pl = P.own_lo[p], pn = P.own_n[p];
        if (cn == 1 && pn == 1 && cl == pl) return;               // same single rank: nothing moves
        const bool in_p = P.holds(p), in_c = P.holds(c);
        if (!in_p && !in_c) return;
        if (cn > 1 && pn > 1 && (cl < pl || cl + cn > pl + pn)) throw std::runtime_error("spchol: child group outside its parent's rank group");
        if (cn == 1 && pn > 1 && (cl < pl || cl >= pl + pn)) throw std::runtime_error("spchol: child mapped outside its parent's rank group");
        if (in_p && !in_c) ws_alloc(c);                            // staged copy of the child's block (natural layout)
        const int nbp_c = (P.w[c] + DNB - 1) / DNB, nbc_c = (mbc + DNB - 1) / DNB;
        const int wp = P.w[p], nbp_p = (wp + DNB - 1) / DNB;
        const int* r = in_p ? P.rel.data() + P.rel_off[c] : nullptr;
        Op pack{}, comm{}, unpack{};
        pack.kind = OP_COPY; pack.begin = static_cast<int>(P.copy_tasks.size()); pack.level = lv;
        std::vector<CommTask> ctasks;
        std::vector<CopyTask> unpacks;
        std::vector<char> need(pn);
        for (int jb = 0; jb < nbc_c; ++jb) {
            const int o = cn == 1 ? cl : cl + (nbp_c + jb) % cn;   // holder of this column block
            const int j0 = jb * DNB, j1 = std::min(j0 + DNB, mbc);
            const int rows = mbc - j0, cols = j1 - j0;
            const int64_t count = static_cast<int64_t>(rows) * cols;
            const int64_t nat = j0 + static_cast<int64_t>(j0) * ldc;   // natural offset of the block inside the cb
            if (rank != o && !in_p) continue;
            // destinations
            std::fill(need.begin(), need.end(), 0);
            if (pn == 1) need[0] = 1;
            else if (in_p) {
                for (int jj = j0; jj < j1; ++jj) {
                    const int cj = r[jj];
                    const int blk = cj < wp ? cj / DNB : nbp_p + (cj - wp) / DNB;
                    need[blk % pn] = 1;
                }
            }
            if (rank == o) {
                // (o holds p whenever pn > 1, so `need` is known here; for pn == 1 it is {pl})
                int64_t stage = -1;
                for (int q = 0; q < pn; ++q) {
                    if (!need[q] || pl + q == o) continue;
                    if (stage < 0) {
                        stage = tmp_alloc(count);
                        add_copy(pack, 1, P.cb_off[c] + nat, ldc, 1, stage, rows, rows, cols);
                    }
                    ctasks.push_back({COMM_SEND, 0, 0, pl + q, 1, stage, count});
                }
            } else if (in_p && need[rank - pl]) {
                const int64_t stage = tmp_alloc(count);
                ctasks.push_back({COMM_RECV, 0, 0, o, 1, stage, count});
                unpacks.push_back({stage, P.cb_off[c] + nat, 1, 1, rows, ldc, rows, cols});
            }
        }
        pack.count = static_cast<int>(P.copy_tasks.size()) - pack.begin;
        if (pack.count) P.ops.push_back(pack);
        if (!ctasks.empty()) {
            comm.kind = OP_COMM; comm.begin = static_cast<int>(P.comm_tasks.size()); comm.count = static_cast<int>(ctasks.size()); comm.level = lv;
            for (const CommTask& t : ctasks) { P.comm_tasks.push_back(t); P.comm_doubles += t.count; }
            P.ops.push_back(comm);
        }
        if (!unpacks.empty()) {
            unpack.kind = OP_COPY; unpack.begin = static_cast<int>(P.copy_tasks.size()); unpack.count = static_cast<int>(unpacks.size()); unpack.level = lv;
            for (const CopyTask& t : unpacks) { P.copy_tasks.push_back(t); unpack.gx = std::max(unpack.gx, t.cols); unpack.gy = std::max(unpack.gy, (t.rows + 2047) / 2048); }
            P.ops.push_back(unpack);
        }
    };

    // a front factored by its rank group: block columns of width DNB dealt round-robin; the owner
    // factors a block column, broadcasts it in place, every rank updates the block columns it owns
    auto emit_dist = [&](int s, int lv) {
        const int w = P.w[s], m = P.m[s], ld = P.ld[s], mb = m - w;
        const int64_t po = P.panel_off[s];
        const int G = P.own_n[s], idx = P.own_idx[s], lo = P.own_lo[s];
        const int nbp = (w + DNB - 1) / DNB, nbc = (mb + DNB - 1) / DNB;
        const int ldcb = even_up(mb);
        const int64_t stage = tmp_alloc(static_cast<int64_t>(m) * std::min(DNB, w));    // packed block column in flight
        const bool defer = opt.defer_cb_add != 0;
        if (mb > 0 && !defer) {
            Op op{}; op.kind = OP_ZERO; op.begin = static_cast<int>(P.zero_tasks.size()); op.count = 1; op.level = lv;
            op.gx = static_cast<int>(std::min<int64_t>((cb_entries(s) + 2047) / 2048, 1184));
            P.zero_tasks.push_back({P.cb_off[s], cb_entries(s)});
            P.ops.push_back(op);
        }
        auto emit_extadd = [&](int mode) {
            for (int q = P.child_ptr[s]; q < P.child_ptr[s + 1]; ++q) {
                const int c = P.child_idx[q];
                if (P.m[c] == P.w[c]) continue;
                Op op{}; op.kind = OP_EXTADD; op.begin = static_cast<int>(P.extadd_list.size()); op.count = 1; op.level = lv; op.threads = mode;
                op.gx = P.m[c] - P.w[c];
                P.extadd_list.push_back(c);
                P.ops.push_back(op);
            }
        };
        emit_extadd(defer ? 1 : 0);
        // Look-ahead (see Op in plan.h): the owner of block column kb+1 applies panel kb to that
        // block column first, then factors and broadcasts it on stream 1 while every rank is still
        // applying panel kb to the rest of its block columns on stream 0.
        const bool pipe = opt.lookahead && nbp > 1;
        const int sp = pipe ? 1 : 0;
        if (pipe) { const int e = ev_record(0, lv); ev_wait(1, e, lv); }      // assembly done before the first panel
        int ev_next = -1;
        for (int kb = 0; kb < nbp; ++kb) {
            const int c0 = kb * DNB, c1 = std::min(c0 + DNB, w);
            const int owner = kb % G;
            if (owner == idx) {
                if (pipe && ev_next >= 0) ev_wait(1, ev_next, lv);
                for (int k0 = c0; k0 < c1; k0 += nb) {
                    const int k1 = std::min(k0 + nb, c1);
                    const int64_t d = po + k0 + static_cast<int64_t>(k0) * ld;
                    const int cls_pb = (k1 - k0) <= 32 ? 32 : ((k1 - k0) <= 64 ? 64 : 128);
                    Op a{}; a.kind = OP_POTRF; a.begin = static_cast<int>(P.potrf_tasks.size()); a.count = 1; a.level = lv; a.threads = cls_pb; a.stream = sp;
                    P.potrf_tasks.push_back({d, ld, k1 - k0, P.first[s] + k0, 0});
                    P.ops.push_back(a);
                    if (k1 < m) {
                        Op b{}; b.kind = OP_TRSM; b.begin = static_cast<int>(P.trsm_tasks.size()); b.count = 1; b.level = lv;
                        b.gx = m - k1; b.threads = cls_pb; b.stream = sp;
                        P.trsm_tasks.push_back({po + k1 + static_cast<int64_t>(k0) * ld, ld, k1 - k0, m - k1, d, 0});
                        P.ops.push_back(b);
                    }
                    if (k1 < c1) {
                        Op g{}; g.kind = OP_GEMM; g.begin = static_cast<int>(P.gemm_tasks.size()); g.count = 1; g.level = lv; g.stream = sp;
                        const int64_t a2 = po + k1 + static_cast<int64_t>(k0) * ld;
                        add_gemm(g, {po + k1 + static_cast<int64_t>(k1) * ld, a2, a2, ld, ld, ld, m - k1, c1 - k1, k1 - k0, 0, 0});
                        P.ops.push_back(g);
                    }
                }
            }
            {   // the factored block column travels packed to its rows c0..m-1
                const int rows = m - c0, cols = c1 - c0;
                const int64_t src = po + c0 + static_cast<int64_t>(c0) * ld;
                if (owner == idx) {
                    Op cp{}; cp.kind = OP_COPY; cp.begin = static_cast<int>(P.copy_tasks.size()); cp.count = 1; cp.level = lv; cp.stream = sp;
                    add_copy(cp, 0, src, ld, 1, stage, rows, rows, cols);
                    P.ops.push_back(cp);
                }
                add_comm(COMM_BCAST, lo, G, lo + owner, 1, stage, static_cast<int64_t>(rows) * cols, lv);
                P.ops.back().stream = sp;
                if (owner != idx) {
                    Op cp{}; cp.kind = OP_COPY; cp.begin = static_cast<int>(P.copy_tasks.size()); cp.count = 1; cp.level = lv; cp.stream = sp;
                    add_copy(cp, 1, stage, rows, 0, src, ld, rows, cols);
                    P.ops.push_back(cp);
                }
            }
            if (pipe) { const int e = ev_record(1, lv); ev_wait(0, e, lv); }  // panel kb is final on this rank
            ev_next = -1;
            if (kb + 1 < nbp && (kb + 1) % G == idx) {    // the next block column first
                const int jb = kb + 1;
                Op g{}; g.kind = OP_GEMM; g.begin = static_cast<int>(P.gemm_tasks.size()); g.count = 1; g.level = lv;
                const int j0 = jb * DNB, j1 = std::min(j0 + DNB, w);
                const int64_t a2 = po + j0 + static_cast<int64_t>(c0) * ld;
                add_gemm(g, {po + j0 + static_cast<int64_t>(j0) * ld, a2, a2, ld, ld, ld, m - j0, j1 - j0, c1 - c0, 0, 0});
                P.ops.push_back(g);
                if (pipe) ev_next = ev_record(0, lv);
            }
            Op g{}; g.kind = OP_GEMM; g.begin = static_cast<int>(P.gemm_tasks.size()); g.level = lv;
            for (int jb = kb + 2; jb < nbp; ++jb) {
                if (jb % G != idx) continue;
                const int j0 = jb * DNB, j1 = std::min(j0 + DNB, w);
                const int64_t a2 = po + j0 + static_cast<int64_t>(c0) * ld;
                add_gemm(g, {po + j0 + static_cast<int64_t>(j0) * ld, a2, a2, ld, ld, ld, m - j0, j1 - j0, c1 - c0, 0, 0});
            }
            g.count = static_cast<int>(P.gemm_tasks.size()) - g.begin;
            if (g.count) P.ops.push_back(g);
        }
        if (mb > 0) {
            Op g{}; g.kind = OP_GEMM; g.begin = static_cast<int>(P.gemm_tasks.size()); g.level = lv;
            for (int jb = 0; jb < nbc; ++jb) {
                if ((nbp + jb) % G != idx) continue;
                const int j0 = jb * DNB, j1 = std::min(j0 + DNB, mb);
                const int64_t a2 = po + w + j0;
                add_gemm(g, {P.cb_off[s] + j0 + static_cast<int64_t>(j0) * ldcb, a2, a2, ldcb, ld, ld, mb - j0, j1 - j0, w, 0, defer ? 3 : 1});
            }
            g.count = static_cast<int>(P.gemm_tasks.size()) - g.begin;
            if (g.count) P.ops.push_back(g);
            if (defer) emit_extadd(2);
        }
    };

    // ---- the bottom of the tree: whole subtrees, one warp each (tree_plan.h); the level program starts above them.
    // The blocks of the subtree roots live in the arena from the start of the factorization until their parents ran.
    std::vector<char> in_tree(ns, 0);
    if (want_tree) {
        if (P.max_m <= TW_MAX_M) {
            if (opt.batch_group_m > 0 && opt.want_csc) build_group_program(P, opt.batch_group_m, a_front, a_fcol, a_fpos, P.tree_whole);
            else build_tree_program(P, {}, true, a_front, a_fcol, a_fpos, P.tree_whole);
        }
        const std::vector<int> roots = choose_subtrees(P, opt.tree_cap_m, opt.tree_work, in_tree);
        if (!roots.empty()) {
            for (int s : roots) if (P.m[s] > P.w[s] && P.parent[s] >= 0) ws_alloc(s);
            build_tree_program(P, roots, false, a_front, a_fcol, a_fpos, P.tree);
            Op op{}; op.kind = OP_TREE; op.begin = 0; op.count = static_cast<int>(P.tree.tasks.size()); op.level = 0;
            P.ops.push_back(op);
            P.n_tree = static_cast<int64_t>(P.tree.fronts.size());
        }
        // second tier: the subtrees of fronts up to tree2_cap_m rows that sit on top of the first tier, one 4-warp team each.
        // These fronts (2D 1024^2: 6 000 of them on ~8 levels) would otherwise take one launch per level of one CTA per front.
        if (opt.tree2_cap_m > opt.tree_cap_m) {
            std::vector<char> in2;
            const std::vector<int> roots2 = choose_subtrees(P, opt.tree2_cap_m, opt.tree2_work, in2, &in_tree);
            if (!roots2.empty()) {
                for (int s : roots2) if (P.m[s] > P.w[s] && P.parent[s] >= 0) ws_alloc(s);
                build_tree_program(P, roots2, false, a_front, a_fcol, a_fpos, P.tree2, &in_tree);
                Op op{}; op.kind = OP_TREE; op.begin = 1; op.count = static_cast<int>(P.tree2.tasks.size()); op.level = 0;
                P.ops.push_back(op);
                P.n_tree += static_cast<int64_t>(P.tree2.fronts.size());
                for (int s = 0; s < ns; ++s) if (in2[s]) in_tree[s] = 1;
            }
        }
    }
    for (int lv = 0; lv < nlev; ++lv) {
        std::vector<int> local_a, local_b, dists;
        for (int s : by_level[lv]) {
            if (!P.holds(s) || in_tree[s]) continue;
            if (P.m[s] > P.w[s]) ws_alloc(s);
            if (P.own_n[s] > 1) { dists.push_back(s); continue; }
            bool remote = false;
            for (int q = P.child_ptr[s]; q < P.child_ptr[s + 1] && !remote; ++q) {
                const int c = P.child_idx[q];
                remote = !(P.own_n[c] == 1 && P.own_lo[c] == rank);
            }
            (remote ? local_b : local_a).push_back(s);
        }
        emit_local(local_a, lv);
        if (world > 1)
            for (int p : by_level[lv])
                for (int q = P.child_ptr[p]; q < P.child_ptr[p + 1]; ++q) emit_transfer(P.child_idx[q], p, lv);
        emit_local(local_b, lv);
        for (int s : dists) emit_dist(s, lv);
        for (size_t q = 0; q + 1 < tmp_allocs.size(); q += 2) arena.release(tmp_allocs[q], tmp_allocs[q + 1]);
        tmp_allocs.clear();
        // children of this level's fronts are consumed now
        for (int s : by_level[lv]) {
            for (int q = P.child_ptr[s]; q < P.child_ptr[s + 1]; ++q) {
                const int c = P.child_idx[q];
                if (ws_size[c] > 0) { arena.release(P.cb_off[c], ws_size[c]); ws_size[c] = 0; }
            }
        }
    }
    P.ws_entries = arena.high_water();
}

}  // namespace spchol
