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

std::vector<int> choose_subtrees(const HostPlan& P, int cap_m, double work_limit, std::vector<char>& in_tree) {
    const int ns = P.ns;
    cap_m = std::min(cap_m, TW_MAX_M);
    std::vector<double> sub(ns);
    std::vector<int> submax(ns);
    for (int s = 0; s < ns; ++s) { sub[s] = front_flops(P.w[s], P.m[s]); submax[s] = P.m[s]; }
    for (int s = 0; s < ns; ++s) {                    // children have smaller ids than parents
        const int p = P.parent[s];
        if (p >= 0) { sub[p] += sub[s]; submax[p] = std::max(submax[p], submax[s]); }
    }
    in_tree.assign(ns, 0);
    std::vector<int> roots;
    for (int s = 0; s < ns; ++s) in_tree[s] = (submax[s] <= cap_m && sub[s] <= work_limit && P.holds(s) && P.own_n[s] == 1) ? 1 : 0;
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
                        const std::vector<int>& a_col, const std::vector<int>& a_pos, TreeProgram& out) {
    out = TreeProgram();
    const int ns = P.ns;
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
            for (int q = P.child_ptr[s]; q < P.child_ptr[s + 1]; ++q) kids.push_back(P.child_idx[q]);
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
            for (int q = P.child_ptr[s]; q < P.child_ptr[s + 1]; ++q) kids.push_back(P.child_idx[q]);
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

}  // namespace spchol
