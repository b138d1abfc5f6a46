// TEST INFRASTRUCTURE ONLY -- never linked into libspchol.so.
//
// Runs the SOURCE of the warp-per-subtree kernel (csrc/tree_kernel.cuh, k_tree_warp) on host threads:
// 32 std::threads play the lanes of one warp, __syncwarp() is a std::barrier, shared memory is a heap
// array.  The program comes from the product's planner (csrc/tree_plan.cpp via build_plan); fronts the
// program does not cover are factored by plain loops (the semantics of k_front_small), and L is
// gathered through the plan's CSC export map (subtree mode) or written by the kernel source itself
// through the inverse map (whole-tree mode = the batched factorization).  A missing __syncwarp shows
// up as a data race between real threads (ThreadSanitizer build), a wrong map as a wrong factor.
#include <algorithm>
#include <atomic>
#include <barrier>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <limits>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

namespace spchol { namespace emu {
thread_local int t_tid = 0;
std::barrier<>* g_team = nullptr;
std::vector<std::barrier<>*> g_warp;
std::mutex g_mu;
int g_team_id = 0;                     // stands in for blockIdx.x (which warp of a team runs the serial part)
double g_fa[8][32], g_fb[8][32];      // operand fragments of the emulated tensor-core instruction, per warp
inline int tid() { return t_tid; }
inline void sync() { g_team->arrive_and_wait(); }
inline void syncwarp() { g_warp[t_tid >> 5]->arrive_and_wait(); }
inline void atomic_min(int* p, int v) { std::lock_guard<std::mutex> lk(g_mu); if (v < *p) *p = v; }
// mma.sync.m8n8k4.row.col.f64: lane l holds A[l/4][l%4], B[l%4][l/4] and C[l/4][2 (l%4)], C[l/4][2 (l%4) + 1]
inline void dmma884(double& c0, double& c1, double a, double b) {
    const int l = t_tid & 31, wq = t_tid >> 5, gq = l >> 2, q = l & 3;
    g_fa[wq][l] = a; g_fb[wq][l] = b;
    syncwarp();
    for (int k = 0; k < 4; ++k) {
        c0 = std::fma(g_fa[wq][gq * 4 + k], g_fb[wq][(2 * q) * 4 + k], c0);
        c1 = std::fma(g_fa[wq][gq * 4 + k], g_fb[wq][(2 * q + 1) * 4 + k], c1);
    }
    syncwarp();
}
} }

#define SPCHOL_HOST_EMU
using std::max; using std::min; using std::fma;
#include "../../sparsecholesky_b200/csrc/plan.h"
#include "../../sparsecholesky_b200/csrc/symbolic.h"
#include "../../sparsecholesky_b200/csrc/tree_kernel.cuh"

using namespace spchol;

namespace {
inline int even_up(int v) { return (v + 1) & ~1; }

// one front factored whole (what k_front_small does)
void run_small(const HostPlan& H, std::vector<double>& Lb, std::vector<double>& Wb, int s, int& info) {
    const int w = H.w[s], m = H.m[s], ld = H.ld[s];
    std::vector<double> F(static_cast<size_t>(m) * m, 0.0);
    double* P = Lb.data() + H.panel_off[s];
    for (int j = 0; j < w; ++j) for (int i = j; i < m; ++i) F[i + static_cast<size_t>(j) * m] = P[i + static_cast<size_t>(j) * ld];
    for (int q = H.child_ptr[s]; q < H.child_ptr[s + 1]; ++q) {
        const int c = H.child_idx[q];
        const int mc = H.m[c] - H.w[c];
        if (!mc) continue;
        const int ldc = even_up(mc);
        const double* C = Wb.data() + H.cb_off[c];
        const int* r = H.rel.data() + H.rel_off[c];
        for (int jj = 0; jj < mc; ++jj) for (int ii = jj; ii < mc; ++ii) F[r[ii] + static_cast<size_t>(r[jj]) * m] += C[ii + static_cast<size_t>(jj) * ldc];
    }
    for (int k = 0; k < w; ++k) {
        const double d = F[k + static_cast<size_t>(k) * m];
        if (d <= 0.0) info = std::min(info, H.first[s] + k + 1);
        const double piv = std::sqrt(d);
        for (int i = k + 1; i < m; ++i) F[i + static_cast<size_t>(k) * m] /= piv;
        F[k + static_cast<size_t>(k) * m] = piv;
        for (int j = k + 1; j < m; ++j) {
            const double uj = F[j + static_cast<size_t>(k) * m];
            for (int i = j; i < m; ++i) F[i + static_cast<size_t>(j) * m] -= F[i + static_cast<size_t>(k) * m] * uj;
        }
    }
    for (int j = 0; j < w; ++j) for (int i = j; i < m; ++i) P[i + static_cast<size_t>(j) * ld] = F[i + static_cast<size_t>(j) * m];
    const int mb = m - w;
    if (mb > 0) {
        double* C = Wb.data() + H.cb_off[s];
        const int ldc = even_up(mb);
        for (int j = 0; j < mb; ++j) for (int i = j; i < mb; ++i) C[i + static_cast<size_t>(j) * ldc] = F[(w + i) + static_cast<size_t>(w + j) * m];
    }
}
}  // namespace

// whole = 1: the whole tree as one task, factor written in CSC order by the kernel source (batched mode);
// whole = 0: subtree tasks (cap_m rows, `work` dense flops per task) + plain loops for the fronts above them.
// stats[6]: #tasks, #fronts in tasks, #stashes, stash doubles, cap_m, #fronts of the plan
template <int NW>
static void run_tasks(const TreeArrays& ta, const double* Ax, double* Lb, double* Wb, double* Sb, double* Lx, int* info, int cap_m, int mode) {
    const double nan = std::numeric_limits<double>::quiet_NaN();
    std::vector<double> F(static_cast<size_t>(cap_m) * (cap_m + 1) / 2 + TK_F_SLACK, nan);
    double spiv[2] = {nan, nan};
    std::vector<double> sT(64, nan);
    std::barrier<> team(32 * NW);
    emu::g_team = &team;
    std::vector<std::unique_ptr<std::barrier<>>> wb;
    emu::g_warp.clear();
    for (int q = 0; q < NW; ++q) { wb.push_back(std::make_unique<std::barrier<>>(32)); emu::g_warp.push_back(wb.back().get()); }
    std::vector<std::thread> th;
    for (int t = 0; t < 32 * NW; ++t)
        th.emplace_back([&, t] {
            emu::t_tid = t;
            for (int q = 0; q < ta.ntasks; ++q) {
                tree_team_body<NW>(ta, ta.tasks[q], Ax, Lb, Wb, Sb, Lx, info, F.data(), spiv, sT.data(), mode);
                emu::sync();
            }
        });
    for (auto& x : th) x.join();
    emu::g_team = nullptr;
}

// whole = 2: the whole-tree program over amalgamated groups of up to cap_m rows (build_group_program), else as whole = 1
extern "C" int treeemu_factor(int n, const int* Ap, const int* Ai, const double* Ax, int relax, int whole, int cap_m, double work, int nw,
                              double* Lx, int* info_out, int64_t* stats, char* err, int errlen) {
    auto fail = [&](const std::string& msg) { std::snprintf(err, errlen, "%s", msg.c_str()); return -1; };
    emu::g_team_id = std::getenv("TREEEMU_TEAM") ? std::atoi(std::getenv("TREEEMU_TEAM")) : 0;
    try {
        Symbolic S;
        RelaxParams rp; rp.enable = relax != 0;
        analyze(n, Ap, Ai, rp, S);
        HostPlan H;
        PlanOptions opt;
        opt.small_max_m = 1 << 20;          // everything above the subtrees on the whole-front path
        opt.tree = 1; opt.tree_cap_m = cap_m; opt.tree_work = work;
        opt.batch_group_m = whole == 2 ? cap_m : 0;
        // second tier of subtree tasks (whole = 0 only): TREEEMU_TIER2 = "cap_m,work,nw"
        int cap2 = 0, nw2 = 4; double work2 = 3e7;
        if (const char* v = std::getenv("TREEEMU_TIER2")) std::sscanf(v, "%d,%lf,%d", &cap2, &work2, &nw2);
        opt.tree2_cap_m = whole ? 0 : cap2; opt.tree2_work = work2;
        build_plan(S, opt, H);
        if (H.n_large > 0) return fail("pattern has fronts beyond the shared-memory front kernel");
        const TreeProgram& T = whole ? H.tree_whole : H.tree;
        if (whole && T.tasks.empty()) return fail("no whole-tree program (a front has more than TW_MAX_M rows)");
        if (stats) {
            stats[0] = static_cast<int64_t>(T.tasks.size()); stats[1] = static_cast<int64_t>(T.fronts.size());
            stats[2] = static_cast<int64_t>(T.stashes.size()); stats[3] = T.stash_entries; stats[4] = T.cap_m; stats[5] = H.ns;
            stats[6] = T.n_groups; stats[7] = static_cast<int64_t>(T.padded_flops);
            if (!whole) { stats[6] = static_cast<int64_t>(H.tree2.tasks.size()); stats[7] = static_cast<int64_t>(H.tree2.fronts.size()); }
        }
        const double nan = std::numeric_limits<double>::quiet_NaN();
        std::vector<double> Lb(static_cast<size_t>(H.lbuf_entries), 0.0);
        std::vector<double> Wb(static_cast<size_t>(std::max<int64_t>(H.ws_entries, 1)), nan);      // poison: read-before-write shows
        std::vector<int> csc_map;
        if (whole) {
            csc_map.assign(static_cast<size_t>(H.lbuf_entries), -1);
            for (int64_t e = 0; e < S.nnzL; ++e) { if (H.l_src[e] < 0) return fail("export map hole"); csc_map[H.l_src[e]] = static_cast<int>(e); }
            for (int64_t e = 0; e < S.nnzL; ++e) Lx[e] = nan;
        } else {
            for (int64_t i = 0; i < H.nnzA; ++i) if (H.a_dest[i] >= 0) Lb[H.a_dest[i]] = Ax[i];
        }
        int info = 0x7f7f7f7f;
        auto run_prog = [&](const TreeProgram& Q, int teams) -> bool {
            if (Q.tasks.empty()) return true;
            std::vector<double> Sq(static_cast<size_t>(std::max<int64_t>(Q.stash_entries, 1)), nan);
            TreeArrays ta{};
            ta.tasks = Q.tasks.data(); ta.ntasks = static_cast<int>(Q.tasks.size()); ta.fronts = Q.fronts.data(); ta.stashes = Q.stashes.data();
            ta.rmap = Q.rmap.data(); ta.a_src = Q.a_src.data(); ta.a_dst = Q.a_dst.data(); ta.csc_map = csc_map.data();
            ta.out_src = Q.out_src.data(); ta.out_dst = Q.out_dst.data(); ta.gcol = Q.gcol.data();
            switch (teams) {
            case 1: run_tasks<1>(ta, Ax, Lb.data(), Wb.data(), Sq.data(), Lx, &info, Q.cap_m, whole ? 1 : 0); break;
            case 2: run_tasks<2>(ta, Ax, Lb.data(), Wb.data(), Sq.data(), Lx, &info, Q.cap_m, whole ? 1 : 0); break;
            case 4: run_tasks<4>(ta, Ax, Lb.data(), Wb.data(), Sq.data(), Lx, &info, Q.cap_m, whole ? 1 : 0); break;
            default: return false;
            }
            return true;
        };
        if (!run_prog(T, nw)) return fail("nw must be 1, 2 or 4");
        if (!whole && !run_prog(H.tree2, nw2)) return fail("nw of the second tier must be 1, 2 or 4");
        if (!whole) {
            for (const Op& op : H.ops) {
                if (op.kind == OP_TREE) continue;
                if (op.kind != OP_SMALL) return fail("unexpected op kind in an all-small plan");
                for (int i = 0; i < op.count; ++i) run_small(H, Lb, Wb, H.small_list[op.begin + i], info);
            }
            for (int64_t i = 0; i < S.nnzL; ++i) { const int64_t q = H.l_src[i]; if (q < 0) return fail("export map hole"); Lx[i] = Lb[q]; }
        }
        *info_out = info == 0x7f7f7f7f ? 0 : info;
    } catch (const std::exception& e) { return fail(e.what()); }
    return 0;
}
