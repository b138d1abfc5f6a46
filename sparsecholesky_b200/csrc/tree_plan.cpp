// Host planner of the warp-per-subtree program (see tree_plan.h).
#include "plan.h"

#include <algorithm>
#include <stdexcept>

namespace spchol {

namespace {
inline int64_t tri64(int64_t x) { return x * (x + 1) / 2; }
inline int even_up(int v) { return (v + 1) & ~1; }
inline double front_flops(double w, double m) { return w * m * m - m * w * (w - 1) + (w - 1) * w * (2 * w - 1) / 6.0; }
}  // namespace

std::vector<int> choose_subtrees(const HostPlan& P, int cap_m, double work_limit, std::vector<char>& in_tree,
                                 const std::vector<char>* below) {
    const int ns = P.ns;
    cap_m = std::min(cap_m, TW_MAX_M);
    std::vector<double> sub(ns);
    std::vector<int> submax(ns);
    auto gone = [&](int s) { return below && (*below)[s]; };      // factored by an earlier tier: not part of this forest
    for (int s = 0; s < ns; ++s) { sub[s] = front_flops(P.w[s], P.m[s]); submax[s] = P.m[s]; }
    for (int s = 0; s < ns; ++s) {                    // children have smaller ids than parents
        const int p = P.parent[s];
        if (p >= 0 && !gone(s)) { sub[p] += sub[s]; submax[p] = std::max(submax[p], submax[s]); }
    }
    in_tree.assign(ns, 0);
    std::vector<int> roots;
    for (int s = 0; s < ns; ++s)
        in_tree[s] = (!gone(s) && submax[s] <= cap_m && sub[s] <= work_limit && P.holds(s) && P.own_n[s] == 1) ? 1 : 0;
    for (int s = 0; s < ns; ++s) {
        if (!in_tree[s]) continue;
        const int p = P.parent[s];
        if (p < 0 || !in_tree[p]) roots.push_back(s);
    }
    // heaviest tasks first: the CTA scheduler hands them out in order, so the long ones do not end up in the tail
    std::stable_sort(roots.begin(), roots.end(), [&](int a, int b) { return sub[a] > sub[b]; });
    return roots;
}

void build_tree_program(const HostPlan& P, const std::vector<int>& roots_in, bool whole, const std::vector<int>& a_front,
                        const std::vector<int>& a_col, const std::vector<int>& a_pos, TreeProgram& out,
                        const std::vector<char>* below) {
    out = TreeProgram();
    const int ns = P.ns;
    auto gone = [&](int s) { return below && (*below)[s]; };
    std::vector<int> roots = roots_in;
    if (whole) {
        roots.clear();
        for (int s = 0; s < ns; ++s) if (P.parent[s] < 0) roots.push_back(s);
    }
    out.in_tree.assign(ns, 0);
    // A entries by front (counting sort keeps the CSC order inside a front)
    std::vector<int> acnt(ns + 1, 0);
    for (size_t e = 0; e < a_front.size(); ++e) if (a_front[e] >= 0) ++acnt[a_front[e] + 1];
    for (int s = 0; s < ns; ++s) acnt[s + 1] += acnt[s];
    std::vector<int> aord(acnt[ns]);
    { std::vector<int> cur(acnt.begin(), acnt.end() - 1);
      for (size_t e = 0; e < a_front.size(); ++e) if (a_front[e] >= 0) aord[cur[a_front[e]]++] = static_cast<int>(e); }

    std::vector<int> order;                 // fronts of one task in postorder
    std::vector<int> kids;
    struct Frame { int s; int next; };
    std::vector<Frame> stack;
    std::vector<std::vector<int>> kid_order(0);
    int64_t stash_total = 0;
    if (whole) out.tasks.push_back({0, 0, 0});
    for (size_t ri = 0; ri < roots.size(); ++ri) {
        // postorder of the subtree below roots[ri]; children sorted so that the one with the largest block comes LAST
        // (it is the one carried in shared memory), ties by id
        order.clear();
        stack.clear();
        stack.push_back({roots[ri], 0});
        // children lists in visiting order are rebuilt on the fly (the tree is small per task)
        std::vector<std::pair<int, std::vector<int>>> local;   // (front, ordered children) in discovery order
        std::vector<int> lidx;                                  // stack frame -> index into local
        auto ordered_children = [&](int s) {
            kids.clear();
            for (int q = P.child_ptr[s]; q < P.child_ptr[s + 1]; ++q) if (!gone(P.child_idx[q])) kids.push_back(P.child_idx[q]);
            std::stable_sort(kids.begin(), kids.end(), [&](int a, int b) { return (P.m[a] - P.w[a]) < (P.m[b] - P.w[b]); });
            return kids;
        };
        local.push_back({roots[ri], ordered_children(roots[ri])});
        lidx.push_back(0);
        while (!stack.empty()) {
            Frame& f = stack.back();
            const int li = lidx.back();
            if (f.next < static_cast<int>(local[li].second.size())) {
                const int c = local[li].second[f.next++];
                local.push_back({c, ordered_children(c)});
                stack.push_back({c, 0});
                lidx.push_back(static_cast<int>(local.size()) - 1);
            } else {
                order.push_back(f.s);
                stack.pop_back(); lidx.pop_back();
            }
        }
        // emit
        TwTask task{static_cast<int>(out.fronts.size()), 0, 0};
        int64_t sp = 0, high = 0;                         // stash stack (doubles), relative to the task's base
        std::vector<std::pair<int, int64_t>> live;        // (front, offset) of stashed blocks, LIFO
        for (size_t oi = 0; oi < order.size(); ++oi) {
            const int s = order[oi];
            if (P.m[s] > TW_MAX_M) throw std::runtime_error("spchol: front too large for the warp-per-subtree program");
            if (!P.holds(s) || P.own_n[s] != 1) throw std::runtime_error("spchol: warp-per-subtree program over a distributed front");
            out.in_tree[s] = 1;
            out.cap_m = std::max(out.cap_m, P.m[s]);
            out.flops += front_flops(P.w[s], P.m[s]);
            TwFront d{};
            d.w = P.w[s]; d.m = P.m[s]; d.ld = P.ld[s]; d.first = P.first[s];
            d.panel_off = P.panel_off[s];
            d.a0 = static_cast<int>(out.a_src.size());
            for (int q = acnt[s]; q < acnt[s + 1]; ++q) {
                const int e = aord[q];
                const int jr = d.m - 1 - a_col[e], ir = d.m - 1 - a_pos[e];        // reversed column / row, ir <= jr
                out.a_src.push_back(e);
                out.a_dst.push_back(static_cast<unsigned short>(tri64(jr) + ir));
            }
            d.an = static_cast<int>(out.a_src.size()) - d.a0;
            // children: the last one in visiting order was factored just before s and is carried (if it has a block)
            kids.clear();
            for (int q = P.child_ptr[s]; q < P.child_ptr[s + 1]; ++q) if (!gone(P.child_idx[q])) kids.push_back(P.child_idx[q]);
            std::stable_sort(kids.begin(), kids.end(), [&](int a, int b) { return (P.m[a] - P.w[a]) < (P.m[b] - P.w[b]); });
            auto push_map = [&](int c) {
                const int mc = P.m[c] - P.w[c];
                const int off = static_cast<int>(out.rmap.size());
                const int* r = P.rel.data() + P.rel_off[c];
                for (int ir = 0; ir < mc; ++ir) out.rmap.push_back(d.m - 1 - r[mc - 1 - ir]);
                return off;
            };
            int carried = -1;
            if (!kids.empty() && oi > 0 && order[oi - 1] == kids.back() && P.m[kids.back()] > P.w[kids.back()]) carried = kids.back();
            d.map_lo = static_cast<int>(out.rmap.size());
            if (carried >= 0) {
                d.flags |= TW_CARRY_IN;
                d.mc = P.m[carried] - P.w[carried];
                const int fwd = push_map(carried);                      // forward map, replaced by its inverse below
                std::vector<int> inv(d.m, -1);
                for (int ir = 0; ir < d.mc; ++ir) inv[out.rmap[fwd + ir]] = ir;
                d.ident = 0;
                while (d.ident < d.mc && out.rmap[fwd + d.ident] == d.ident) ++d.ident;
                out.rmap.resize(fwd);
                d.map_off = fwd;
                out.rmap.insert(out.rmap.end(), inv.begin(), inv.end());
                out.fronts.back().flags |= TW_CARRY_OUT;
            }
            d.stash0 = static_cast<int>(out.stashes.size());
            // the other children with a block were stashed, most recent on top of the stack: pop them in LIFO order
            for (int qi = static_cast<int>(kids.size()) - 1; qi >= 0; --qi) {
                const int c = kids[qi];
                if (c == carried || P.m[c] == P.w[c]) continue;
                if (live.empty() || live.back().first != c) throw std::runtime_error("spchol: stash stack out of order");
                TwStash st{P.m[c] - P.w[c], push_map(c), live.back().second};
                out.stashes.push_back(st);
                sp = live.back().second;              // pop
                live.pop_back();
            }
            // children an earlier tier factored: their blocks wait in the level program's workspace (column-major, ld = even(mc));
            // a negative order tells the kernel
            for (int q = P.child_ptr[s]; q < P.child_ptr[s + 1]; ++q) {
                const int c = P.child_idx[q];
                if (!gone(c) || P.m[c] == P.w[c]) continue;
                if (P.cb_off[c] < 0) throw std::runtime_error("spchol: block of a lower-tier subtree root has no workspace slot");
                out.stashes.push_back(TwStash{-(P.m[c] - P.w[c]), push_map(c), P.cb_off[c]});
            }
            d.nstash = static_cast<int>(out.stashes.size()) - d.stash0;
            d.map_n = static_cast<int>(out.rmap.size()) - d.map_lo;
            // this front's own block: carried (decided when the parent is emitted), the level program's slot (task
            // roots of a subtree program), or the stash stack
            const int mb = d.m - d.w;
            d.cb_out = -1;
            const bool is_root = oi + 1 == order.size();
            if (mb > 0) {
                if (is_root) {
                    if (P.parent[s] >= 0) {
                        if (whole) throw std::runtime_error("spchol: forest root with a contribution block");
                        d.flags |= TW_CB_STD; d.cb_out = P.cb_off[s];
                    }
                } else {
                    // provisionally stashed; if the parent turns out to come next and carries it, the flag is set then
                    // and the stash slot is simply not used (the slot is popped again right away)
                    d.cb_out = sp;
                    live.push_back({s, sp});
                    sp += tri64(mb);
                    high = std::max(high, sp);
                }
            }
            out.fronts.push_back(d);
            // a carried block does not occupy its provisional slot: undo when the next front is its parent and takes it
            if (mb > 0 && !is_root && oi + 1 < order.size() && order[oi + 1] == P.parent[s]) {
                // the parent carries its LAST child in visiting order, which is this front
                live.pop_back();
                sp -= tri64(mb);
                out.fronts.back().cb_out = -1;
            }
        }
        if (!live.empty()) throw std::runtime_error("spchol: stashed blocks left over");
        if (whole) {
            out.tasks[0].nfront += static_cast<int>(order.size());
            out.tasks[0].stash_base = 0;
            stash_total = std::max(stash_total, high);
        } else {
            task.nfront = static_cast<int>(order.size());
            task.stash_base = stash_total;
            stash_total += (high + 1) & ~int64_t(1);
            out.tasks.push_back(task);
        }
    }
    out.stash_entries = stash_total;
    (void)even_up;
}

void build_group_program(const HostPlan& P, int cap_m, const std::vector<int>& a_front, const std::vector<int>& a_col,
                         const std::vector<int>& a_pos, TreeProgram& out) {
    const int ns = P.ns;
    cap_m = std::min(cap_m, TW_MAX_M);      // a front absorbs children only while the group stays within cap_m rows; fronts of
                                            // the plan that are larger than that stand alone (the caller checked them against TW_MAX_M)
    if (P.l_src.empty() && P.nnzL > 0) throw std::runtime_error("spchol: the group program needs the CSC export map");
    // ---- 1. grouping, bottom-up: a front absorbs the groups of its children (largest block first) while |U| <= cap_m
    std::vector<int> gw(ns), top(ns);                 // columns of the group whose top is s; group top of every front
    std::vector<std::vector<int>> members(ns);
    for (int s = 0; s < ns; ++s) { gw[s] = P.w[s]; top[s] = s; members[s].push_back(s); }
    std::vector<int> kids;
    for (int p = 0; p < ns; ++p) {
        if (P.m[p] > TW_MAX_M) throw std::runtime_error("spchol: front too large for the group program");
        kids.assign(P.child_idx.begin() + P.child_ptr[p], P.child_idx.begin() + P.child_ptr[p + 1]);
        std::stable_sort(kids.begin(), kids.end(), [&](int a, int b) { return (P.m[a] - P.w[a]) > (P.m[b] - P.w[b]); });
        for (int c : kids) {
            if (gw[p] + gw[c] + (P.m[p] - P.w[p]) > cap_m) continue;
            gw[p] += gw[c];
            members[p].insert(members[p].end(), members[c].begin(), members[c].end());
            std::vector<int>().swap(members[c]);
        }
    }
    std::vector<int> gid(ns, -1), tops;
    for (int s = 0; s < ns; ++s) if (!members[s].empty()) {
        std::sort(members[s].begin(), members[s].end(), [&](int a, int b) { return P.first[a] < P.first[b]; });      // elimination order
        for (int q : members[s]) { top[q] = s; gid[q] = static_cast<int>(tops.size()); }
        tops.push_back(s);
    }
    const int ng = static_cast<int>(tops.size());
    // ---- 2. position in U of every row of every member front (columns: members in elimination order; then the top's below rows)
    std::vector<int> colbase(ns, 0);
    std::vector<int64_t> pos_off(ns + 1, 0);
    for (int s = 0; s < ns; ++s) pos_off[s + 1] = pos_off[s] + P.m[s];
    std::vector<int> posU(static_cast<size_t>(pos_off[ns]));
    for (int g = 0; g < ng; ++g) {
        const int t = tops[g];
        int cb = 0;
        for (int q : members[t]) { colbase[q] = cb; cb += P.w[q]; }
        // members top-down (parents before children): descending id works, children have smaller ids
        std::vector<int> ord(members[t]);
        std::sort(ord.begin(), ord.end(), std::greater<int>());
        for (int q : ord) {
            int* pq = posU.data() + pos_off[q];
            for (int c = 0; c < P.w[q]; ++c) pq[c] = colbase[q] + c;
            const int nb = P.m[q] - P.w[q];
            if (q == t) { for (int i = 0; i < nb; ++i) pq[P.w[q] + i] = gw[t] + i; continue; }
            const int par = P.parent[q];
            const int* r = P.rel.data() + P.rel_off[q];
            const int* pp = posU.data() + pos_off[par];
            for (int i = 0; i < nb; ++i) pq[P.w[q] + i] = pp[r[i]];
        }
    }
    // ---- 3. the groups as the fronts of a virtual plan (only the fields the program builder reads)
    HostPlan V;
    V.ns = ng; V.n = P.n; V.rank = 0; V.world = 1;
    V.w.resize(ng); V.m.resize(ng); V.ld.assign(ng, 0); V.first.resize(ng); V.parent.assign(ng, -1);
    V.panel_off.assign(ng, 0); V.cb_off.assign(ng, -1); V.own_lo.assign(ng, 0); V.own_n.assign(ng, 1); V.own_idx.assign(ng, 0);
    V.rel_off.assign(ng + 1, 0);
    for (int g = 0; g < ng; ++g) {
        const int t = tops[g];
        V.w[g] = gw[t]; V.m[g] = gw[t] + (P.m[t] - P.w[t]); V.first[g] = g;      // `first` carries the group id through the builder
        V.parent[g] = P.parent[t] >= 0 ? gid[P.parent[t]] : -1;
        V.rel_off[g + 1] = V.rel_off[g] + (P.m[t] - P.w[t]);
        out.padded_flops += front_flops(V.w[g], V.m[g]);
    }
    V.rel.resize(static_cast<size_t>(V.rel_off[ng]));
    for (int g = 0; g < ng; ++g) {
        const int t = tops[g], par = P.parent[t];
        if (par < 0) continue;
        const int nb = P.m[t] - P.w[t];
        const int* r = P.rel.data() + P.rel_off[t];
        const int* pp = posU.data() + pos_off[par];
        for (int i = 0; i < nb; ++i) V.rel[V.rel_off[g] + i] = pp[r[i]];
        // the rows below a group's top must come out ascending in the parent group for the in-place spread (monotone map)
        for (int i = 1; i < nb; ++i) if (V.rel[V.rel_off[g] + i] <= V.rel[V.rel_off[g] + i - 1]) throw std::runtime_error("spchol: group map is not monotone");
    }
    V.child_ptr.assign(ng + 1, 0);
    for (int g = 0; g < ng; ++g) if (V.parent[g] >= 0) ++V.child_ptr[V.parent[g] + 1];
    for (int g = 0; g < ng; ++g) V.child_ptr[g + 1] += V.child_ptr[g];
    V.child_idx.resize(V.child_ptr[ng]);
    { std::vector<int> cur(V.child_ptr.begin(), V.child_ptr.end() - 1);
      for (int g = 0; g < ng; ++g) if (V.parent[g] >= 0) V.child_idx[cur[V.parent[g]]++] = g; }
    std::vector<int> af(a_front.size()), ac(a_front.size()), ap(a_front.size());
    for (size_t e = 0; e < a_front.size(); ++e) {
        const int s = a_front[e];
        af[e] = gid[s]; ac[e] = colbase[s] + a_col[e]; ap[e] = posU[pos_off[s] + a_pos[e]];
    }
    build_tree_program(V, {}, true, af, ac, ap, out);
    // ---- 4. output lists (true entries of L only) and the columns' global indices, attached to the emitted fronts
    std::vector<std::pair<int64_t, int>> pan;
    for (int s = 0; s < ns; ++s) pan.push_back({P.panel_off[s], s});
    std::sort(pan.begin(), pan.end());
    std::vector<std::vector<std::pair<unsigned short, int>>> lists(ng);
    for (int64_t e = 0; e < P.nnzL; ++e) {
        const int64_t off = P.l_src[e];
        if (off < 0) throw std::runtime_error("spchol: export map hole in a single-rank plan");
        const int s = std::prev(std::upper_bound(pan.begin(), pan.end(), std::make_pair(off, ns)))->second;
        const int64_t rel = off - P.panel_off[s];
        const int c = static_cast<int>(rel / P.ld[s]), t = static_cast<int>(rel % P.ld[s]);
        const int g = gid[s], mg = V.m[g];
        const int jr = mg - 1 - (colbase[s] + c), ir = mg - 1 - posU[pos_off[s] + t];
        lists[g].push_back({static_cast<unsigned short>(tri64(jr) + ir), static_cast<int>(e)});
    }
    std::vector<int> gcol0(ng), out0(ng);
    for (int g = 0; g < ng; ++g) {
        gcol0[g] = static_cast<int>(out.gcol.size());
        for (int q : members[tops[g]]) for (int c = 0; c < P.w[q]; ++c) out.gcol.push_back(P.first[q] + c);
        out0[g] = static_cast<int>(out.out_src.size());
        for (const auto& pr : lists[g]) { out.out_src.push_back(pr.first); out.out_dst.push_back(pr.second); }
    }
    for (TwFront& d : out.fronts) {
        const int g = d.first;
        d.flags |= TW_GROUP;
        d.ld = gcol0[g]; d.pad0 = out0[g]; d.pad1 = static_cast<int>(lists[g].size());
    }
    out.n_groups = ng;
    out.in_tree.assign(ns, 1);
    out.flops = 0;
    for (int s = 0; s < ns; ++s) out.flops += front_flops(P.w[s], P.m[s]);
}

}  // namespace spchol
