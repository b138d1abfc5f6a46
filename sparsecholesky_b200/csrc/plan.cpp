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

void build_plan(const Symbolic& S, const PlanOptions& opt, HostPlan& P) {
    P = HostPlan();
    const int ns = S.ns, n = S.n;
    P.n = n; P.ns = ns; P.nnzA = S.nnzA; P.nnzL = S.nnzL; P.flops = S.flops;
    P.first.resize(ns); P.w.resize(ns); P.m.resize(ns); P.ld.resize(ns);
    P.parent = S.sn_parent;
    P.panel_off.resize(ns); P.cb_off.assign(ns, -1); P.rel_off.resize(ns + 1);
    int64_t off = 0, roff = 0;
    for (int s = 0; s < ns; ++s) {
        P.first[s] = S.sn_start[s];
        P.w[s] = S.w(s);
        P.m[s] = S.m(s);
        P.ld[s] = even_up(P.m[s]);
        P.panel_off[s] = off;
        off += static_cast<int64_t>(P.ld[s]) * P.w[s];
        P.rel_off[s] = roff;
        roff += P.m[s] - P.w[s];
        P.max_m = std::max(P.max_m, P.m[s]);
        P.max_w = std::max(P.max_w, P.w[s]);
    }
    P.rel_off[ns] = roff;
    P.lbuf_entries = off;

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
        const int nb = P.m[s] - P.w[s];
        if (nb == 0) continue;
        if (p < 0) throw std::runtime_error("spchol: root front with below rows");
        const int* rs = S.below(s);
        const int* rp = S.below(p);
        const int nbp = P.m[p] - P.w[p];
        const int pend = P.first[p] + P.w[p];
        int* out = P.rel.data() + P.rel_off[s];
        int q = 0;
        for (int t = 0; t < nb; ++t) {
            const int r = rs[t];
            if (r < pend) { out[t] = r - P.first[p]; continue; }
            while (q < nbp && rp[q] < r) ++q;
            if (q >= nbp || rp[q] != r) throw std::runtime_error("spchol: child row missing in parent front");
            out[t] = P.w[p] + q;
        }
    }

    // A scatter map (A is upper CSC: entry (i,j), i <= j, lands at L(j,i))
    P.a_dest.resize(S.nnzA);
    for (int j = 0; j < n; ++j) {
        for (int p = S.Ap[j]; p < S.Ap[j + 1]; ++p) {
            const int i = S.Ai[p];
            const int s = S.sn_id[i];
            const int c = i - P.first[s];
            int pos;
            if (j < P.first[s] + P.w[s]) pos = j - P.first[s];
            else {
                const int* b = S.below(s);
                const int nb = P.m[s] - P.w[s];
                const int* it = std::lower_bound(b, b + nb, j);
                if (it == b + nb || *it != j) throw std::runtime_error("spchol: A entry outside the symbolic pattern");
                pos = P.w[s] + static_cast<int>(it - b);
            }
            P.a_dest[p] = P.panel_off[s] + static_cast<int64_t>(c) * P.ld[s] + pos;
        }
    }

    // CSC export map
    if (opt.want_csc) {
        P.l_src.resize(S.nnzL);
        for (int j = 0; j < n; ++j) {
            const int s = S.sn_id[j], f = S.fsn_id[j];
            const int c = j - P.first[s];
            const int64_t colbase = P.panel_off[s] + static_cast<int64_t>(c) * P.ld[s];
            int64_t q = S.Lp[j];
            const int fend = S.fsn_start[f + 1];
            for (int r = j; r < fend; ++r) P.l_src[q++] = colbase + (r - P.first[s]);
            const int* fb = S.fbelow.data() + S.fbelow_ptr[f];
            const int nfb = static_cast<int>(S.fbelow_ptr[f + 1] - S.fbelow_ptr[f]);
            if (S.sn_top[s] == f) {
                for (int t = 0; t < nfb; ++t) P.l_src[q++] = colbase + P.w[s] + t;
            } else {
                const int* b = S.below(s);
                const int nb = P.m[s] - P.w[s];
                const int send = P.first[s] + P.w[s];
                int k = 0;
                for (int t = 0; t < nfb; ++t) {
                    const int r = fb[t];
                    if (r < send) { P.l_src[q++] = colbase + (r - P.first[s]); continue; }
                    while (k < nb && b[k] < r) ++k;
                    if (k >= nb || b[k] != r) throw std::runtime_error("spchol: export map row missing");
                    P.l_src[q++] = colbase + P.w[s] + k;
                }
            }
        }
    }

    // contribution-block workspace + launch program, level by level
    Arena arena;
    std::vector<int64_t> cb_size(ns, 0);
    const int nb = opt.nb, NB = std::max(opt.NB, opt.nb);
    static const int cls_m[] = {16, 24, 32, 48, 64, 96, 128, 1 << 30};
    static const int cls_thr[] = {32, 32, 64, 64, 128, 128, 256, 256};
    const int ncls = 8;
    for (int lv = 0; lv < nlev; ++lv) {
        std::vector<int> small[8], large;
        for (int s : by_level[lv]) {
            if (opt.owner && opt.owner[s] != opt.rank) continue;
            const int mb = P.m[s] - P.w[s];
            if (mb > 0) {
                cb_size[s] = static_cast<int64_t>(even_up(mb)) * mb;
                P.cb_off[s] = arena.alloc(even_up64(cb_size[s]));
            }
            if (P.m[s] <= opt.small_max_m) {
                int c = 0;
                while (P.m[s] > cls_m[c]) ++c;
                small[c].push_back(s);
            } else large.push_back(s);
        }
        // --- small fronts: one CTA per front, bucketed by size class -------------------------
        for (int c = 0; c < ncls; ++c) {
            if (small[c].empty()) continue;
            int mm = 0;
            for (int s : small[c]) mm = std::max(mm, P.m[s]);
            Op op{};
            op.kind = OP_SMALL; op.begin = static_cast<int>(P.small_list.size());
            op.count = static_cast<int>(small[c].size());
            op.threads = cls_thr[c];
            op.smem = mm * (mm | 1) * 8;
            op.level = lv;
            P.small_list.insert(P.small_list.end(), small[c].begin(), small[c].end());
            P.ops.push_back(op);
            P.n_small += op.count;
        }
        // --- large fronts: zero cb, extend-add rounds, blocked panel factorization, SYRK ----
        if (!large.empty()) {
            P.n_large += static_cast<int64_t>(large.size());
            P.max_potrf_slots = std::max(P.max_potrf_slots, static_cast<int>(large.size()));
            {
                Op op{}; op.kind = OP_ZERO; op.begin = static_cast<int>(P.zero_tasks.size()); op.level = lv;
                for (int s : large) if (P.cb_off[s] >= 0) P.zero_tasks.push_back({P.cb_off[s], cb_size[s]});
                op.count = static_cast<int>(P.zero_tasks.size()) - op.begin;
                if (op.count) P.ops.push_back(op);
            }
            int max_children = 0;
            for (int s : large) max_children = std::max(max_children, P.child_ptr[s + 1] - P.child_ptr[s]);
            for (int r = 0; r < max_children; ++r) {
                Op op{}; op.kind = OP_EXTADD; op.begin = static_cast<int>(P.extadd_list.size()); op.level = lv;
                int maxmb = 0;
                for (int s : large) {
                    if (P.child_ptr[s] + r >= P.child_ptr[s + 1]) continue;
                    const int c = P.child_idx[P.child_ptr[s] + r];
                    if (P.m[c] == P.w[c]) continue;
                    P.extadd_list.push_back(c);
                    maxmb = std::max(maxmb, P.m[c] - P.w[c]);
                }
                op.count = static_cast<int>(P.extadd_list.size()) - op.begin;
                op.gx = maxmb;
                if (op.count) P.ops.push_back(op);
            }
            // per-front step sequences, merged slot by slot
            struct Step { bool has_trsm = false, has_gemm = false; PotrfTask po; TrsmTask tr; GemmTask ge; };
            std::vector<std::vector<Step>> seq(large.size());
            size_t max_steps = 0;
            for (size_t li = 0; li < large.size(); ++li) {
                const int s = large[li];
                const int w = P.w[s], m = P.m[s], ld = P.ld[s];
                const int64_t po = P.panel_off[s];
                for (int KB = 0; KB < w; KB += NB) {
                    const int KE = std::min(KB + NB, w);
                    for (int kb = KB; kb < KE; kb += nb) {
                        const int ke = std::min(kb + nb, KE);
                        Step st;
                        st.po = {po + kb + static_cast<int64_t>(kb) * ld, ld, ke - kb, P.first[s] + kb, static_cast<int>(li)};
                        if (ke < m) {
                            st.has_trsm = true;
                            st.tr = {po + ke + static_cast<int64_t>(kb) * ld, ld, ke - kb, m - ke,
                                     po + kb + static_cast<int64_t>(kb) * ld, static_cast<int>(li)};
                        }
                        if (ke < KE) {            // update the rest of the outer block, K = nb
                            st.has_gemm = true;
                            const int64_t a = po + ke + static_cast<int64_t>(kb) * ld;
                            st.ge = {po + ke + static_cast<int64_t>(ke) * ld, a, a, ld, ld, ld,
                                     m - ke, KE - ke, ke - kb, 0, 0};
                        } else if (KE < w) {      // deferred update of the remaining panel, K = NB
                            st.has_gemm = true;
                            const int64_t a = po + KE + static_cast<int64_t>(KB) * ld;
                            st.ge = {po + KE + static_cast<int64_t>(KE) * ld, a, a, ld, ld, ld,
                                     m - KE, w - KE, KE - KB, 0, 0};
                        } else if (m > w) {       // panel done: Schur complement into the cb, K = w
                            st.has_gemm = true;
                            const int64_t a = po + w;
                            st.ge = {P.cb_off[s], a, a, even_up(m - w), ld, ld, m - w, m - w, w, 0, 1};
                        }
                        seq[li].push_back(st);
                    }
                }
                max_steps = std::max(max_steps, seq[li].size());
            }
            for (size_t i = 0; i < max_steps; ++i) {
                Op a{}, b{}, c{};
                a.kind = OP_POTRF; a.begin = static_cast<int>(P.potrf_tasks.size()); a.level = lv;
                b.kind = OP_TRSM;  b.begin = static_cast<int>(P.trsm_tasks.size());  b.level = lv;
                c.kind = OP_GEMM;  c.begin = static_cast<int>(P.gemm_tasks.size());  c.level = lv;
                for (size_t li = 0; li < large.size(); ++li) {
                    if (i >= seq[li].size()) continue;
                    const Step& st = seq[li][i];
                    P.potrf_tasks.push_back(st.po);
                    if (st.has_trsm) { P.trsm_tasks.push_back(st.tr); b.gx = std::max(b.gx, st.tr.rows); }
                    if (st.has_gemm) {
                        P.gemm_tasks.push_back(st.ge);
                        c.gx = std::max(c.gx, st.ge.M); c.gy = std::max(c.gy, st.ge.N);
                        P.gemm_flops += 2.0 * st.ge.M * st.ge.N * st.ge.K;
                        {   // wanted entries: column j holds rows i >= j - diag
                            double cnt = 0;
                            const int Mq = st.ge.M, Nq = st.ge.N, dg = st.ge.diag;
                            const int jfull = std::min(Nq, std::max(0, dg + 1));      // columns with all M rows
                            cnt += static_cast<double>(jfull) * Mq;
                            const int jlast = std::min(Nq, Mq + dg);                    // beyond this: empty
                            if (jlast > jfull) {
                                const double a = Mq - (jfull - dg), b = Mq - (jlast - 1 - dg);
                                cnt += (a + b) * (jlast - jfull) / 2.0;
                            }
                            P.gemm_flops_alg += 2.0 * cnt * st.ge.K;
                        }
                    }
                }
                a.count = static_cast<int>(P.potrf_tasks.size()) - a.begin;
                b.count = static_cast<int>(P.trsm_tasks.size()) - b.begin;
                c.count = static_cast<int>(P.gemm_tasks.size()) - c.begin;
                if (a.count) P.ops.push_back(a);
                if (b.count) P.ops.push_back(b);
                if (c.count) P.ops.push_back(c);
            }
        }
        // children of this level's fronts are consumed now
        for (int s : by_level[lv]) {
            if (opt.owner && opt.owner[s] != opt.rank) continue;
            for (int q = P.child_ptr[s]; q < P.child_ptr[s + 1]; ++q) {
                const int c = P.child_idx[q];
                if (P.cb_off[c] >= 0 && cb_size[c] > 0) { arena.release(P.cb_off[c], even_up64(cb_size[c])); cb_size[c] = 0; }
            }
        }
    }
    P.ws_entries = arena.high_water();
}

}  // namespace spchol
