#include "path_plan.h"

#include <algorithm>
#include <stdexcept>

namespace spchol {

namespace {

inline int even_up_i(int v) { return (v + 1) & ~1; }

// A group is a connected piece of the assembly tree that contains its top: its members in ascending id (= a
// topological order), the top last.  Two ways to cut the region into groups:
struct Cut {
    std::vector<std::vector<int>> groups;     // members, ascending
    std::vector<int> order;                   // emission order of the groups (children groups before parents)
};

// (0) heavy paths cut greedily from the bottom: side paths first, then the groups of a path back to back, so that
//     every link of the path hands its block over in registers
void cut_paths(const HostPlan& P, int level0, int max_nu, Cut& C) {
    const int ns = P.ns;
    auto in_region = [&](int s) { return P.level[s] >= level0; };
    std::vector<int> heavy(ns, -1);           // region child with the most rows (ties: larger id)
    for (int s = 0; s < ns; ++s) {
        if (!in_region(s)) continue;
        const int p = P.parent[s];
        if (p >= 0 && (heavy[p] < 0 || P.m[s] >= P.m[heavy[p]])) heavy[p] = s;
    }
    std::vector<std::vector<int>> paths;      // bottom to top; a bottom is a region front without a heavy child
    std::vector<int> path_of(ns, -1);
    for (int s = 0; s < ns; ++s) {
        if (!in_region(s) || heavy[s] >= 0) continue;
        std::vector<int> path;
        for (int t = s;;) {
            path.push_back(t);
            path_of[t] = static_cast<int>(paths.size());
            const int p = P.parent[t];
            if (p < 0 || heavy[p] != t) break;
            t = p;
        }
        paths.push_back(std::move(path));
    }
    std::vector<std::vector<int>> side(paths.size());          // side[q] = paths whose top's parent lies on path q
    std::vector<int> roots, path_order;
    for (size_t q = 0; q < paths.size(); ++q) {
        const int p = P.parent[paths[q].back()];
        if (p < 0) roots.push_back(static_cast<int>(q)); else side[path_of[p]].push_back(static_cast<int>(q));
    }
    auto by_top = [&](int a, int b) { return paths[a].back() < paths[b].back(); };
    for (auto& v : side) std::sort(v.begin(), v.end(), by_top);
    std::sort(roots.begin(), roots.end(), by_top);
    std::vector<std::pair<int, size_t>> stk;
    for (int r : roots) {
        stk.push_back({r, 0});
        while (!stk.empty()) {
            auto& [q, k] = stk.back();
            if (k < side[q].size()) { const int nxt = side[q][k++]; stk.push_back({nxt, 0}); }
            else { path_order.push_back(q); stk.pop_back(); }
        }
    }
    for (int q : path_order) {
        const std::vector<int>& path = paths[q];
        size_t i0 = 0;
        while (i0 < path.size()) {             // members path[i0 .. i1), |U| = widths before the top + rows of the top
            size_t i1 = i0 + 1;
            int sumw = 0;
            while (i1 < path.size() && sumw + P.w[path[i1 - 1]] + P.m[path[i1]] <= max_nu) { sumw += P.w[path[i1 - 1]]; ++i1; }
            C.order.push_back(static_cast<int>(C.groups.size()));
            C.groups.emplace_back(path.begin() + i0, path.begin() + i1);
            i0 = i1;
        }
    }
}

// (1) subtrees grown bottom-up: a front absorbs the groups of its region children, largest first, while the merged
//     row set still fits; what does not fit stays a group of its own.  Fewer groups on bushy trees (every group
//     start pays the latency of its incoming blocks), and independent branches are eliminated in place as well.
void cut_subtrees(const HostPlan& P, int level0, int max_nu, Cut& C) {
    const int ns = P.ns;
    auto in_region = [&](int s) { return P.level[s] >= level0; };
    std::vector<int> gid(ns, -1), sumw;        // group of a front; total width of a group's members
    std::vector<std::vector<int>> members;
    std::vector<char> alive;
    for (int s = 0; s < ns; ++s) {             // ascending ids: children first
        if (!in_region(s)) continue;
        std::vector<int> kids;                 // groups topped by a region child of s
        for (int cq = P.child_ptr[s]; cq < P.child_ptr[s + 1]; ++cq) { const int c = P.child_idx[cq]; if (in_region(c)) kids.push_back(gid[c]); }
        std::sort(kids.begin(), kids.end(), [&](int a, int b) { return sumw[a] != sumw[b] ? sumw[a] > sumw[b] : a < b; });
        const int g = static_cast<int>(members.size());
        members.push_back({s}); sumw.push_back(P.w[s]); alive.push_back(1);
        gid[s] = g;
        int nu = P.m[s];
        for (int k : kids) {
            if (nu + sumw[k] > max_nu) continue;
            nu += sumw[k]; sumw[g] += sumw[k];
            for (int t : members[k]) { members[g].push_back(t); gid[t] = g; }
            members[k].clear(); alive[k] = 0;
        }
    }
    std::vector<int> remap(members.size(), -1);
    for (size_t g = 0; g < members.size(); ++g) {
        if (!alive[g]) continue;
        std::sort(members[g].begin(), members[g].end());
        remap[g] = static_cast<int>(C.groups.size());
        C.groups.push_back(members[g]);
    }
    // emission: depth first over the groups, the child group with the largest block last (it is the one whose block
    // can stay in registers)
    const int ng = static_cast<int>(C.groups.size());
    std::vector<std::vector<int>> kids(ng);
    std::vector<int> roots;
    for (int g = 0; g < ng; ++g) {
        const int top = C.groups[g].back(), p = P.parent[top];
        if (p < 0) roots.push_back(g); else kids[remap[gid[p]]].push_back(g);
    }
    auto cb = [&](int g) { const int t = C.groups[g].back(); return P.m[t] - P.w[t]; };
    for (auto& v : kids) std::sort(v.begin(), v.end(), [&](int a, int b) { return cb(a) != cb(b) ? cb(a) < cb(b) : a < b; });
    std::vector<std::pair<int, size_t>> stk;
    for (int r : roots) {
        stk.push_back({r, 0});
        while (!stk.empty()) {
            auto& [g, k] = stk.back();
            if (k < kids[g].size()) { const int nxt = kids[g][k++]; stk.push_back({nxt, 0}); }
            else { C.order.push_back(g); stk.pop_back(); }
        }
    }
}

}  // namespace

bool build_path_program(const HostPlan& P, int level0, int max_nu, int64_t carry_cap, PathProgram& out, int mode) {
    out = PathProgram();
    out.level0 = level0;
    out.ws_entries = P.ws_entries;
    const int ns = P.ns;
    if (P.world != 1) throw std::runtime_error("build_path_program: single-rank plans only");
    if (static_cast<int64_t>(P.rel.size()) != P.rel_off[ns]) throw std::runtime_error("build_path_program: the plan no longer holds its relative indices");
    auto in_region = [&](int s) { return P.level[s] >= level0; };
    for (int s = 0; s < ns; ++s) {
        if (!in_region(s)) continue;
        if (P.m[s] > max_nu) return false;
        if (P.parent[s] >= 0 && !in_region(P.parent[s])) throw std::runtime_error("build_path_program: region is not closed under parent");
    }
    Cut C;
    if (mode == 1) cut_subtrees(P, level0, max_nu, C); else cut_paths(P, level0, max_nu, C);

    std::vector<int> group_of(ns, -1);          // emitted group that eliminates front s
    std::vector<int> pos_of(ns, -1);            // pos_off of the member entry of front s
    for (int cg : C.order) {
        const std::vector<int>& mem_ids = C.groups[cg];
        const int nmem = static_cast<int>(mem_ids.size());
        const int top = mem_ids.back();
        const int gi = static_cast<int>(out.groups.size());
        PathGroup g{};
        int sumw = 0;
        for (int s : mem_ids) sumw += P.w[s];
        g.nu = sumw - P.w[top] + P.m[top];
        if (g.nu > max_nu) throw std::runtime_error("build_path_program: group exceeds max_nu");
        g.member0 = static_cast<int>(out.members.size());
        g.nmember = nmem;
        g.mb = P.m[top] - P.w[top];
        g.carry_out = 0;
        g.carried_child = -1;
        g.cb_dst = -1;
        for (int s : mem_ids) group_of[s] = gi;
        // members and their row positions, top first (a member's below rows are located through its parent's list);
        // U = the members' columns in ascending order, then below(top)
        std::vector<PathMember> mem(nmem);
        int off = sumw;
        for (int i = nmem - 1; i >= 0; --i) {
            const int s = mem_ids[i];
            off -= P.w[s];
            PathMember& M = mem[i];
            M.s = s; M.off = off; M.nd = 0; M.pad2 = 0;
            M.w = P.w[s]; M.m = P.m[s]; M.ld = P.ld[s]; M.first = P.first[s]; M.panel_off = P.panel_off[s];
            M.pos_off = static_cast<int>(out.pos.size());
            pos_of[s] = M.pos_off;
            out.pos.resize(out.pos.size() + P.m[s]);
            int* ps = out.pos.data() + M.pos_off;
            for (int t = 0; t < P.w[s]; ++t) ps[t] = off + t;
            if (s == top) {
                for (int t = P.w[s]; t < P.m[s]; ++t) ps[t] = off + t;          // U ends with below(top), in order
            } else {
                const int p = P.parent[s];
                if (p < 0 || group_of[p] != gi || p <= s) throw std::runtime_error("build_path_program: group is not connected");
                const int* pp = out.pos.data() + pos_of[p];
                const int* r = P.rel.data() + P.rel_off[s];
                for (int t = P.w[s]; t < P.m[s]; ++t) ps[t] = pp[r[t - P.w[s]]];
            }
            for (int t = 1; t < P.m[s]; ++t) if (ps[t] <= ps[t - 1]) throw std::runtime_error("build_path_program: positions not increasing");
            out.max_m = std::max(out.max_m, P.m[s]);
        }
        if (off != 0) throw std::runtime_error("build_path_program: internal offset mismatch");
        for (int q = 0; q + 1 < nmem; ++q) {            // the corner of member q's update that is the next member's diagonal block
            const int* ps = out.pos.data() + mem[q].pos_off;
            const int lim = mem[q + 1].off + mem[q + 1].w;
            int nd = 0;
            while (mem[q].w + nd < mem[q].m && ps[mem[q].w + nd] < lim) ++nd;
            mem[q].nd = nd;
        }
        out.members.insert(out.members.end(), mem.begin(), mem.end());
        // contribution blocks of the fronts outside the group whose parent is a member
        g.child0 = static_cast<int>(out.children.size());
        for (int s : mem_ids) {
            for (int cq = P.child_ptr[s]; cq < P.child_ptr[s + 1]; ++cq) {
                const int c = P.child_idx[cq];
                if (group_of[c] == gi) continue;                               // eliminated in place
                const int mc = P.m[c] - P.w[c];
                if (mc == 0) continue;
                PathChild ch{};
                ch.c = c; ch.mc = mc; ch.carried = 0;
                const int* r = P.rel.data() + P.rel_off[c];
                std::vector<int> cp(mc);
                { const int* ps = out.pos.data() + pos_of[s]; for (int t = 0; t < mc; ++t) cp[t] = ps[r[t]]; }
                ch.pos_off = static_cast<int>(out.pos.size());
                out.pos.insert(out.pos.end(), cp.begin(), cp.end());
                // the block's columns grouped by the owner of the column of the frontal matrix they are added to:
                // owners never share a target column, so the blocks of a group need no barrier between them
                ch.cols_off = static_cast<int>(out.pos.size());
                ch.optr[0] = 0;
                for (int o = 0; o < PK_OWNERS; ++o) {
                    for (int t = 0; t < mc; ++t) if (cp[t] % PK_OWNERS == o) out.pos.push_back(t);
                    ch.optr[o + 1] = static_cast<int>(out.pos.size()) - ch.cols_off;
                }
                if (in_region(c)) {
                    const int gc = group_of[c];
                    if (gc < 0) throw std::runtime_error("build_path_program: child group not emitted yet");
                    PathGroup& G = out.groups[gc];
                    if (gc == gi - 1 && g.carried_child < 0 && static_cast<int64_t>(mc) * (mc + 1) / 2 <= carry_cap) {
                        G.carry_out = 1; ch.carried = 1; ch.src = -1; ++out.n_carried;
                        g.carried_child = static_cast<int>(out.children.size()) - g.child0;
                    } else {
                        G.cb_dst = out.ws_entries;
                        ch.src = G.cb_dst;
                        const int64_t cnt = static_cast<int64_t>(even_up_i(mc)) * mc;
                        out.ws_entries += (cnt + 1) & ~static_cast<int64_t>(1);
                    }
                } else {
                    ch.src = P.cb_off[c];
                    if (ch.src < 0) throw std::runtime_error("build_path_program: child without a workspace block");
                }
                out.children.push_back(ch);
            }
        }
        g.nchild = static_cast<int>(out.children.size()) - g.child0;
        out.max_nu = std::max(out.max_nu, g.nu);
        out.groups.push_back(g);
    }
    // every region front is eliminated exactly once and parents come after their children
    for (int s = 0; s < ns; ++s) {
        if (in_region(s) != (group_of[s] >= 0)) throw std::runtime_error("build_path_program: coverage mismatch");
        if (in_region(s) && P.parent[s] >= 0 && group_of[P.parent[s]] < group_of[s])
            throw std::runtime_error("build_path_program: parent scheduled before its child");
    }
    return true;
}

}  // namespace spchol
