// TEST INFRASTRUCTURE ONLY -- never linked into libspchol.so.
//
// Runs the SOURCE of the batched path-group kernel (csrc/path_kernel.cuh, k_front_path) on host
// threads: 256 std::threads play the threads of one CTA, __syncthreads() is a std::barrier, shared
// memory is a heap array.  The fronts of the lower levels are factored by plain loops (the
// semantics of k_front_small, as in plan_sim.cpp), then the group program built by the product's
// planner (csrc/path_plan.cpp) runs through the kernel body, and L is gathered through the plan's
// CSC export map.  A missing barrier shows up as a data race between real threads, a wrong position
// list as a wrong factor.
#include <algorithm>
#include <atomic>
#include <barrier>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <limits>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

namespace spchol { namespace emu {
thread_local int t_tid = 0;
std::barrier<>* g_barrier = nullptr;
std::vector<std::barrier<>*> g_warp_barrier;
std::mutex g_mu;
inline int tid() { return t_tid; }
inline void sync() { g_barrier->arrive_and_wait(); }
inline void syncwarp() { g_warp_barrier[t_tid >> 5]->arrive_and_wait(); }
inline void atomic_min(int* p, int v) { std::lock_guard<std::mutex> lk(g_mu); if (v < *p) *p = v; }
} }

#define SPCHOL_HOST_EMU
using std::max; using std::min; using std::fma;
#include "../../sparsecholesky_b200/csrc/path_kernel.cuh"
#include "../../sparsecholesky_b200/csrc/symbolic.h"

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

// relax: 1 = default amalgamation, 2 = the batched setting of spchol_symbolic_create_ex.
// level0 < 0: the rule of the engine (levels from which on no level has more than 16 fronts).
// mode: 0 = path groups, 1 = subtree groups (build_path_program).
// stats[6]: level0, #groups, #members, #children, #carried, max |U|
extern "C" int pathemu_factor(int n, const int* Ap, const int* Ai, const double* Ax, int relax, int level0, int max_nu,
                              int carry_cap, int mode, double* Lx, int* info_out, int64_t* stats, char* err, int errlen) {
    auto fail = [&](const std::string& msg) { std::snprintf(err, errlen, "%s", msg.c_str()); return -1; };
    try {
        Symbolic S;
        RelaxParams rp; rp.enable = relax != 0;
        if (relax == 2) {
            rp.nrelax[0] = 32; rp.nrelax[1] = 96; rp.nrelax[2] = 128;
            rp.zrelax[0] = 0.9; rp.zrelax[1] = 0.7; rp.zrelax[2] = 0.5;
            rp.max_m = 128;
        }
        analyze(n, Ap, Ai, rp, S);
        HostPlan H;
        PlanOptions opt;
        opt.small_max_m = 1 << 20;          // every front on the whole-front path: the test may go beyond the product's 128 rows
        build_plan(S, opt, H);
        if (H.n_large > 0) return fail("pattern has fronts beyond the shared-memory front kernel");
        if (level0 < 0) {
            std::vector<int> cnt(H.nlevels + 1, 0);
            for (int s = 0; s < H.ns; ++s) ++cnt[H.level[s]];
            level0 = H.nlevels;
            while (level0 > 0 && cnt[level0 - 1] <= 16) --level0;
        }
        PathProgram prog;
        if (!build_path_program(H, level0, max_nu, carry_cap, prog, mode)) return fail("a front of the region exceeds max_nu");
        if (stats) {
            stats[0] = level0; stats[1] = static_cast<int64_t>(prog.groups.size()); stats[2] = static_cast<int64_t>(prog.members.size());
            stats[3] = static_cast<int64_t>(prog.children.size()); stats[4] = prog.n_carried; stats[5] = prog.max_nu;
        }
        std::vector<double> Lb(static_cast<size_t>(H.lbuf_entries), 0.0);
        // poison: nothing may be read from the workspace before it was written
        std::vector<double> Wb(static_cast<size_t>(prog.ws_entries), std::numeric_limits<double>::quiet_NaN());
        for (int64_t i = 0; i < H.nnzA; ++i) if (H.a_dest[i] >= 0) Lb[H.a_dest[i]] = Ax[i];
        int info = 0x7f7f7f7f;
        // lower levels, in program order
        for (const Op& op : H.ops) {
            if (op.kind != OP_SMALL) return fail("unexpected op kind in an all-small plan");
            if (op.level >= level0) continue;
            for (int i = 0; i < op.count; ++i) run_small(H, Lb, Wb, H.small_list[op.begin + i], info);
        }
        // the group program through the kernel source
        std::vector<int> tab;
        for (int ii = 0; static_cast<int>(tab.size()) < PK_MAXE * PK_T; ++ii)
            for (int jj = 0; jj <= ii && static_cast<int>(tab.size()) < PK_MAXE * PK_T; ++jj) tab.push_back((ii << 16) | jj);
        if (carry_cap > PK_MAXE * PK_T) return fail("carry_cap beyond the kernel's register budget");
        PathArrays pa{};
        pa.groups = prog.groups.data(); pa.ngroups = static_cast<int>(prog.groups.size());
        pa.members = prog.members.data(); pa.children = prog.children.data(); pa.pos = prog.pos.data();
        pa.tri_tab = tab.data();
        std::vector<double> F(static_cast<size_t>(prog.max_nu) * (prog.max_nu + 1) / 2 + 1, std::numeric_limits<double>::quiet_NaN());
        if (prog.max_m > PK_SPOS_MAX) return fail("a front of the region exceeds the staged position list");
        std::vector<int> spos(2 * PK_SPOS_MAX, -1);
        std::vector<double> srinv(PK_SPOS_MAX, std::numeric_limits<double>::quiet_NaN());
        std::barrier<> bar(PK_T);
        emu::g_barrier = &bar;
        std::vector<std::unique_ptr<std::barrier<>>> wbar;
        emu::g_warp_barrier.clear();
        for (int wq = 0; wq < PK_T / 32; ++wq) { wbar.push_back(std::make_unique<std::barrier<>>(32)); emu::g_warp_barrier.push_back(wbar.back().get()); }
        std::vector<std::thread> th;
        th.reserve(PK_T);
        for (int t = 0; t < PK_T; ++t)
            th.emplace_back([&, t] { emu::t_tid = t; front_path_body(pa, Lb.data(), Wb.data(), &info, F.data(), spos.data(), srinv.data()); });
        for (auto& x : th) x.join();
        emu::g_barrier = nullptr;
        *info_out = info == 0x7f7f7f7f ? 0 : info;
        for (int64_t i = 0; i < S.nnzL; ++i) { const int64_t q = H.l_src[i]; if (q < 0) return fail("export map hole"); Lx[i] = Lb[q]; }
    } catch (const std::exception& e) { return fail(e.what()); }
    return 0;
}
