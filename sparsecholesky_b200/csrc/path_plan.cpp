#include "path_plan.h"

#include <algorithm>
#include <stdexcept>

namespace spchol {

namespace {
inline int even_up_i(int v) { return (v + 1) & ~1; }
}

bool build_path_program(const HostPlan& P, int level0, int max_nu, int64_t carry_cap, PathProgram& out) {
    out = PathProgram();
    out.level0 = level0;
    out.ws_entries = P.ws_entries;
    const int ns = P.ns;
    if (P.world != 1) throw std::runtime_error("build_path_program: single-rank plans only");
    if (static_cast<int64_t>(P.rel.size()) != P.rel_off[ns]) throw std::runtime_error("build_path_program: the plan no longer holds its relative indices");
    auto in_region = [&](int s) { return P.level[s] >= level0; };

    // heavy child of every region front: the region child with the most rows (ties: larger id)
    std::vector<int> heavy(ns, -1);
    for (int s = 0; s < ns; ++s) {
        if (!in_region(s)) continue;
        if (P.m[s] > max_nu) { out = PathProgram(); return false; }
        const int p = P.parent[s];
        if (p < 0) continue;
        if (!in_region(p)) throw std::runtime_error("build_path_program: region is not closed under parent");
        if (heavy[p] < 0 || P.m[s] >= P.m[heavy[p]]) heavy[p] = s;
    }
    // paths, bottom to top: a bottom is a region front without a heavy child
    std::vector<std::vector<int>> paths;
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
    // emission order of the paths: every side path before the path it hangs off (explicit stack, children ascending)
    std::vector<int> order;
    {
        std::vector<std::vector<int>> side(paths.size());          // side[q] = paths whose top's parent lies on path q
        std::vector<int> roots;
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
                else { order.push_back(q); stk.pop_back(); }
            }
        }
    }

    std::vector<int> group_of(ns, -1);          // group that eliminates front s
    std::vector<int> pos_of(ns, -1);            // pos_off of the member entry of front s
    for (int q : order) {
        const std::vector<int>& path = paths[q];
        size_t i0 = 0;
        while (i0 < path.size()) {
            // greedy cut: members path[i0 .. i1), |U| = sum of widths before the top + rows of the top
            size_t i1 = i0 + 1;
            int sumw = 0;
            while (i1 < path.size() && sumw + P.w[path[i1 - 1]] + P.m[path[i1]] <= max_nu) { sumw += P.w[path[i1 - 1]]; ++i1; }
            const int top = path[i1 - 1];
            PathGroup g{};
            g.nu = sumw + P.m[top];
            g.member0 = static_cast<int>(out.members.size());
            g.nmember = static_cast<int>(i1 - i0);
            g.mb = P.m[top] - P.w[top];
            g.carry_out = 0;
            g.carried_child = -1;
            g.cb_dst = -1;
            const int gi = static_cast<int>(out.groups.size());
            // members and their row positions, top first (a member's below rows are located through its parent's list)
            int off = sumw;
            std::vector<PathMember> mem(g.nmember);
            for (size_t i = i1; i-- > i0;) {
                const int s = path[i];
                PathMember& M = mem[i - i0];
                M.s = s; M.off = off; M.pad = 0; M.pad2 = 0;
                M.w = P.w[s]; M.m = P.m[s]; M.ld = P.ld[s]; M.first = P.first[s]; M.panel_off = P.panel_off[s];
                M.pos_off = static_cast<int>(out.pos.size());
                pos_of[s] = M.pos_off;
                group_of[s] = gi;
                out.pos.resize(out.pos.size() + P.m[s]);
                int* ps = out.pos.data() + M.pos_off;
                for (int t = 0; t < P.w[s]; ++t) ps[t] = off + t;
                if (i + 1 == i1) {
                    for (int t = P.w[s]; t < P.m[s]; ++t) ps[t] = off + t;      // U ends with below(top), in order
                } else {
                    const int* pp = out.pos.data() + pos_of[path[i + 1]];
                    const int* r = P.rel.data() + P.rel_off[s];
                    for (int t = P.w[s]; t < P.m[s]; ++t) ps[t] = pp[r[t - P.w[s]]];
                }
                if (i > i0) off -= P.w[path[i - 1]];
                out.max_m = std::max(out.max_m, P.m[s]);
            }
            if (off != 0) throw std::runtime_error("build_path_program: internal offset mismatch");
            out.members.insert(out.members.end(), mem.begin(), mem.end());
            // contribution blocks of the fronts outside the group whose parent is a member
            g.child0 = static_cast<int>(out.children.size());
            for (size_t i = i0; i < i1; ++i) {
                const int s = path[i];
                const int* ps = out.pos.data() + pos_of[s];
                for (int cq = P.child_ptr[s]; cq < P.child_ptr[s + 1]; ++cq) {
                    const int c = P.child_idx[cq];
                    if (i > i0 && c == path[i - 1]) continue;                  // eliminated in place
                    const int mc = P.m[c] - P.w[c];
                    if (mc == 0) continue;
                    PathChild ch{};
                    ch.c = c; ch.mc = mc; ch.carried = 0;
                    ch.pos_off = static_cast<int>(out.pos.size());
                    const int* r = P.rel.data() + P.rel_off[c];
                    // (ps is re-read after the resize below)
                    std::vector<int> cp(mc);
                    for (int t = 0; t < mc; ++t) cp[t] = ps[r[t]];
                    out.pos.insert(out.pos.end(), cp.begin(), cp.end());
                    // the block's columns grouped by the owner of the column of the frontal matrix they are added to:
                    // owners never share a target column, so the blocks of a group need no barrier between them
                    ch.cols_off = static_cast<int>(out.pos.size());
                    ch.optr[0] = 0;
                    for (int o = 0; o < PK_OWNERS; ++o) {
                        for (int t = 0; t < mc; ++t) if (cp[t] % PK_OWNERS == o) out.pos.push_back(t);
                        ch.optr[o + 1] = static_cast<int>(out.pos.size()) - ch.cols_off;
                    }
                    ps = out.pos.data() + pos_of[s];
                    if (in_region(c)) {
                        const int gc = group_of[c];
                        if (gc < 0) throw std::runtime_error("build_path_program: child group not emitted yet");
                        PathGroup& G = out.groups[gc];
                        if (gc == gi - 1 && static_cast<int64_t>(mc) * (mc + 1) / 2 <= carry_cap) {
                            G.carry_out = 1; ch.carried = 1; ch.src = -1; ++out.n_carried;
                            g.carried_child = static_cast<int>(out.children.size()) - g.child0;
                        } else {
                            G.cb_dst = out.ws_entries;
                            ch.src = G.cb_dst;
                            int64_t cnt = static_cast<int64_t>(even_up_i(mc)) * mc;
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
            i0 = i1;
        }
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
