// Device engine + numeric C-ABI (see include/spchol.h).  No CPU fallback: every numeric entry
// point returns SPCHOL_ENOGPU when no CUDA device is usable.
#include <algorithm>
#include <climits>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <dlfcn.h>
#include <map>
#include <mutex>
#include <new>
#include <set>
#include <string>
#include <utility>
#include <vector>

#include "../../include/spchol.h"
#include "capi_internal.h"
#include "kernels.cuh"
#include "gemm_tma.cuh"
#include "ingest.cuh"
#include "mtx_parse.cuh"
#include "tree_kernel.cuh"
#include "residual.cuh"
#include "solve.cuh"

using namespace spchol;

#define CK(call)                                                                          \
    do {                                                                                  \
        cudaError_t e_ = (call);                                                          \
        if (e_ != cudaSuccess) {                                                          \
            set_error(std::string(#call) + ": " + cudaGetErrorString(e_));                \
            return SPCHOL_ECUDA;                                                          \
        }                                                                                 \
    } while (0)

// ---- NCCL, bound at run time (dlopen) so that single-GPU users need no NCCL at all -----------
namespace {
struct NcclId { char internal[128]; };
struct NcclApi {
    void* h = nullptr;
    int (*GetUniqueId)(NcclId*) = nullptr;
    int (*CommInitRank)(void**, int, NcclId, int) = nullptr;
    int (*CommSplit)(void*, int, int, void**, void*) = nullptr;
    int (*CommDestroy)(void*) = nullptr;
    int (*Broadcast)(const void*, void*, size_t, int, int, void*, cudaStream_t) = nullptr;
    int (*Send)(const void*, size_t, int, int, void*, cudaStream_t) = nullptr;
    int (*Recv)(void*, size_t, int, int, void*, cudaStream_t) = nullptr;
    int (*AllReduce)(const void*, void*, size_t, int, int, void*, cudaStream_t) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
};
constexpr int kNcclInt32 = 2, kNcclFloat64 = 8, kNcclSum = 0, kNcclMin = 3, kNcclSplitNoColor = -1;

NcclApi* nccl_api() {
    static NcclApi api;
    static bool tried = false;
    if (tried) return api.h ? &api : nullptr;
    tried = true;
    const char* names[] = {getenv("SPCHOL_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
    for (const char* nm : names) {
        if (!nm || !*nm) continue;
        api.h = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
        if (api.h) break;
    }
    if (!api.h) return nullptr;
    bool ok = true;
    auto sym = [&](const char* nm) { void* f = dlsym(api.h, nm); if (!f) ok = false; return f; };
    api.GetUniqueId = reinterpret_cast<decltype(api.GetUniqueId)>(sym("ncclGetUniqueId"));
    api.CommInitRank = reinterpret_cast<decltype(api.CommInitRank)>(sym("ncclCommInitRank"));
    api.CommSplit = reinterpret_cast<decltype(api.CommSplit)>(sym("ncclCommSplit"));
    api.CommDestroy = reinterpret_cast<decltype(api.CommDestroy)>(sym("ncclCommDestroy"));
    api.Broadcast = reinterpret_cast<decltype(api.Broadcast)>(sym("ncclBroadcast"));
    api.Send = reinterpret_cast<decltype(api.Send)>(sym("ncclSend"));
    api.Recv = reinterpret_cast<decltype(api.Recv)>(sym("ncclRecv"));
    api.AllReduce = reinterpret_cast<decltype(api.AllReduce)>(sym("ncclAllReduce"));
    api.GroupStart = reinterpret_cast<decltype(api.GroupStart)>(sym("ncclGroupStart"));
    api.GroupEnd = reinterpret_cast<decltype(api.GroupEnd)>(sym("ncclGroupEnd"));
    api.GetErrorString = reinterpret_cast<decltype(api.GetErrorString)>(sym("ncclGetErrorString"));
    if (!ok) { dlclose(api.h); api.h = nullptr; return nullptr; }
    return &api;
}
}  // namespace

#define NCK(call)                                                                         \
    do {                                                                                  \
        int e_ = (call);                                                                  \
        if (e_ != 0) {                                                                    \
            set_error(std::string(#call) + ": " + nccl_api()->GetErrorString(e_));        \
            return SPCHOL_ENCCL;                                                          \
        }                                                                                 \
    } while (0)

struct spchol_comm {
    int rank = 0, world = 1, device = 0;
    void* world_comm = nullptr;
    std::map<std::pair<int, int>, void*> groups;      // (lo, n) -> sub-communicator (nullptr: not a member)
};

struct spchol_plan {
    spchol_comm* comm = nullptr;
    int *d_own_n = nullptr, *d_own_idx = nullptr;
    int device = 0;
    HostPlan H;
    cudaStream_t stream = nullptr;
    cudaStream_t stream2 = nullptr;                 // look-ahead stream (panel work of large fronts), high priority
    cudaEvent_t ring[kEventRing] = {nullptr};       // OP_EV_RECORD / OP_EV_WAIT
    cudaEvent_t ev[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
    // device buffers
    double* dL = nullptr; double* dW = nullptr; double* dAx = nullptr; double* dLx = nullptr;
    // warp-per-subtree programs (tree_plan.h / tree_kernel.cuh): `tree` = the bottom subtrees of this matrix (OP_TREE),
    // `whole` = the whole tree as one task per matrix (batched factorization, BASELINE config 5)
    struct DevTree { TreeArrays ta{}; double* stash = nullptr; int f_stride = 0; int64_t stash_entries = 0; bool on = false; };
    DevTree tree, tree2, whole;
    double* bStash = nullptr; int* binfo = nullptr; int b_chunk = 0;              // batched factorization buffers
    double *hbA = nullptr, *hbL = nullptr; int* hbI = nullptr; int hb_cap = 0;     // device staging of the host-buffer batched call
    cudaEvent_t hb_ev[2 * 8] = {nullptr}; cudaStream_t stream4 = nullptr;
    double* dWinv = nullptr;      // inverses of the diagonal blocks of one launch (k_potrf_inv -> k_trsm_inv)
    int* dinfo = nullptr;
    int *d_first = nullptr, *d_w = nullptr, *d_m = nullptr, *d_ld = nullptr, *d_parent = nullptr;
    int64_t *d_panel_off = nullptr, *d_cb_off = nullptr, *d_rel_off = nullptr;
    int *d_rel = nullptr, *d_child_ptr = nullptr, *d_child_idx = nullptr;
    int64_t *d_a_dest = nullptr, *d_l_src = nullptr;
    int *d_small_list = nullptr, *d_extadd_list = nullptr;
    ZeroTask* d_zero = nullptr; PotrfTask* d_potrf = nullptr; TrsmTask* d_trsm = nullptr; GemmTask* d_gemm = nullptr;
    CopyTask* d_copy = nullptr;
    CUtensorMap* d_tmaps = nullptr;                 // one per front with update tasks (k_gemm_tma)
    bool gemm_tma = true;                           // SPCHOL_GEMM_TMA=0: the register-staged k_gemm_dmma (A/B timing)
    // residual check (built on first use): global row index of every below-row, task lists, A coordinates
    int* d_rowidx = nullptr; RTask* d_rt_cols = nullptr; RTask* d_rt_rows = nullptr; int n_rt_cols = 0, n_rt_rows = 0;
    int *d_a_row = nullptr, *d_a_col = nullptr; int* d_level_list = nullptr;
    FrontArrays fa_r{};               // fa with rel_off replaced by the offsets of d_rowidx (multi-rank plans: HostPlan::ridx_off)
    std::vector<int> level_ptr;
    // overlapped read-back of the supernodal panels (spchol_factor_lbuf with pinned host memory): contiguous runs of
    // the L buffer that are final once a given level has run, copied on a third stream while the upper levels compute
    std::vector<std::vector<std::pair<int64_t, int64_t>>> export_runs;   // [level] -> (offset, count) in doubles
    std::vector<cudaEvent_t> export_ev;
    std::vector<cudaEvent_t> export_mark_ev;        // one per OP_EXPORT op
    cudaStream_t stream3 = nullptr;
    // triangular solves (built on first use): per (level, block step) slices of the task arrays
    struct SolveStep { int b0, bn, u0, un, g0, gn, smem; };
    std::vector<SolveStep> solve_steps;        // forward order; the backward solve walks it in reverse
    std::vector<int> solve_level_ptr;          // steps of level l: solve_steps[solve_level_ptr[l] .. solve_level_ptr[l + 1])
    // multi-rank solve: columns more than one rank touches (k_top_*), the exchange buffer, the columns this rank solves,
    // the first level with such a front
    int* d_top_idx = nullptr; double* d_top_buf = nullptr; unsigned char* d_solve_mine = nullptr;
    int64_t n_top = 0; int solve_shared_level = 0;
    SBlock* d_sblocks = nullptr; SChunk* d_supd = nullptr; SGat* d_sgat = nullptr;
    bool solve_ready = false;
    double* d_solve_v = nullptr;               // work vectors of spchol_solve (3 n SOLVE_NR doubles, allocated on first use)
    int64_t device_bytes = 0;
    int64_t last_launches = 0;
    double last_ms[4] = {0, 0, 0, 0};
    std::vector<void*> allocs;
    FrontArrays fa{};
    // optional per-op-kind timing (spchol_set_profile): events around every launch group
    int profile = 0;
    std::vector<cudaEvent_t> prof_ev;
    double kind_ms[12] = {0};
    int64_t kind_launches[12] = {0};
};

namespace {

constexpr int kBatchSlices = 8;       // slices of a host-buffer batch in flight (copy in / factor / copy out overlap)

template <class T>
int upload(spchol_plan* p, const std::vector<T>& v, T** out) {
    *out = nullptr;
    const size_t bytes = std::max<size_t>(v.size(), 1) * sizeof(T);
    void* d = nullptr;
    CK(cudaMalloc(&d, bytes));
    p->allocs.push_back(d);
    p->device_bytes += static_cast<int64_t>(bytes);
    if (!v.empty()) CK(cudaMemcpy(d, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice));
    *out = static_cast<T*>(d);
    return 0;
}
template <class T>
int dalloc(spchol_plan* p, int64_t count, T** out) {
    void* d = nullptr;
    const size_t bytes = static_cast<size_t>(std::max<int64_t>(count, 1)) * sizeof(T);
    CK(cudaMalloc(&d, bytes));
    p->allocs.push_back(d);
    p->device_bytes += static_cast<int64_t>(bytes);
    *out = static_cast<T*>(d);
    return 0;
}

template <int T>
int set_small_attr() {
    CK(cudaFuncSetAttribute(k_front_small<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    return 0;
}

// one CTA (a team of nw warps) per task / per matrix, see tree_kernel.cuh
constexpr size_t tree_smem(int f_stride) { return (static_cast<size_t>(f_stride) + 2 + 64) * 8; }
void launch_tree(int nw, int grid, int f_stride, cudaStream_t st, const TreeArrays& ta, const double* Ax, double* L, double* W, double* Stash,
                 double* Lx, int* info, int64_t axs, int64_t ls, int64_t ws, int64_t ss, int64_t lxs, int nmat, int out_mode) {
    const size_t sm = tree_smem(f_stride);
    if (nw >= 4)      k_tree_team<4><<<grid, 128, sm, st>>>(ta, Ax, L, W, Stash, Lx, info, axs, ls, ws, ss, lxs, nmat, f_stride, out_mode);
    else if (nw >= 2) k_tree_team<2><<<grid, 64, sm, st>>>(ta, Ax, L, W, Stash, Lx, info, axs, ls, ws, ss, lxs, nmat, f_stride, out_mode);
    else              k_tree_team<1><<<grid, 32, sm, st>>>(ta, Ax, L, W, Stash, Lx, info, axs, ls, ws, ss, lxs, nmat, f_stride, out_mode);
}

// issue the whole numeric factorization on the plan's stream
int export_prepare(spchol_plan* p) {
    if (!p->export_runs.empty()) return 0;
    const HostPlan& H = p->H;
    std::vector<int> cnt(H.nlevels + 1, 0);
    for (int s = 0; s < H.ns; ++s) if (H.holds(s)) ++cnt[H.level[s]];
    // the few fronts of the top levels are read back level by level, everything below the cut in one go
    int cut = H.nlevels, above = 0;
    while (cut > 0 && above + cnt[cut - 1] <= 512) { --cut; above += cnt[cut]; }
    std::vector<std::vector<std::pair<int64_t, int64_t>>> raw(H.nlevels);
    for (int s = 0; s < H.ns; ++s) {
        if (!H.holds(s) || H.front_marked[s]) continue;       // marked fronts are read back block column by block column (OP_EXPORT)
        const int bucket = H.level[s] >= cut ? H.level[s] : std::max(cut - 1, 0);
        if (H.own_n[s] > 1) {
            // a front factored by a rank group: every member holds the whole panel after the broadcasts, but only the
            // block columns it OWNS are its part of the result -- reading back the rest would move every such panel
            // own_n times over PCIe (8 GPUs: 53.9 GB instead of 27.6 GB, the round-1 e2e regression)
            const int DNB = H.dist_NB, G = H.own_n[s], me = H.own_idx[s];
            for (int c0 = 0, kb = 0; c0 < H.w[s]; c0 += DNB, ++kb) {
                if (kb % G != me) continue;
                const int c1 = std::min(c0 + DNB, H.w[s]);
                raw[bucket].push_back({H.panel_off[s] + static_cast<int64_t>(c0) * H.ld[s], static_cast<int64_t>(c1 - c0) * H.ld[s]});
            }
        } else {
            raw[bucket].push_back({H.panel_off[s], static_cast<int64_t>(H.ld[s]) * H.w[s]});
        }
    }
    p->export_runs.assign(H.nlevels, {});
    for (int l = 0; l < H.nlevels; ++l) {
        std::sort(raw[l].begin(), raw[l].end());
        for (const auto& r : raw[l]) {
            auto& out = p->export_runs[l];
            if (!out.empty() && out.back().first + out.back().second == r.first) out.back().second += r.second;
            else out.push_back(r);
        }
    }
    p->export_ev.resize(H.nlevels + 1, nullptr);
    for (auto& e : p->export_ev) CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    int n_marks = 0;
    for (const Op& op : H.ops) if (op.kind == OP_EXPORT) ++n_marks;
    p->export_mark_ev.resize(n_marks, nullptr);
    for (auto& e : p->export_mark_ev) CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    CK(cudaStreamCreateWithFlags(&p->stream3, cudaStreamNonBlocking));
    return 0;
}

// host_out != nullptr (pinned host memory): the panels are copied to host_out level by level while later levels run
int run_numeric(spchol_plan* p, const double* dAx, double* host_out = nullptr) {
    const HostPlan& H = p->H;
    cudaStream_t st = p->stream;
    int64_t launches = 0;
    const bool prof = p->profile != 0;
    size_t pe = 0;
    std::vector<int> pkind;
    auto mark = [&](int kind) -> int {     // record an event; kind = op kind that FOLLOWS (or -1 at the end)
        if (!prof) return 0;
        if (pe >= p->prof_ev.size()) {
            cudaEvent_t e;
            if (cudaEventCreate(&e) != cudaSuccess) return -1;
            p->prof_ev.push_back(e);
        }
        cudaEventRecord(p->prof_ev[pe++], st);
        pkind.push_back(kind);
        return 0;
    };
    mark(8);
    CK(cudaMemsetAsync(p->dL, 0, static_cast<size_t>(H.lbuf_entries) * sizeof(double), st));
    CK(cudaMemsetAsync(p->dinfo, 0x7f, sizeof(int), st));
    {
        const int64_t nb = (H.nnzA + 255) / 256;
        k_assemble<<<static_cast<unsigned>(nb), 256, 0, st>>>(H.nnzA, p->d_a_dest, dAx, p->dL, 0);
        ++launches;
    }
    // profiling replays the program on one stream (program order is a valid serial order)
    cudaStream_t sts[2] = {p->stream, prof ? p->stream : p->stream2};
    static const int gemm_swizzle = getenv("SPCHOL_GEMM_SWIZZLE") ? atoi(getenv("SPCHOL_GEMM_SWIZZLE")) : 15;    // banded tiles + C prefetch + RED epilogue + per-stage free barriers
    if (prof) host_out = nullptr;
    int level_next = 0;                  // levels below this one have been handed to the copy stream
    size_t n_mark = 0;                   // OP_EXPORT ops seen so far
    auto levels_done = [&](int upto) -> int {          // every op of the levels < upto has been issued
        for (; level_next < upto; ++level_next) {
            const auto& runs = p->export_runs[level_next];
            if (runs.empty()) continue;
            CK(cudaEventRecord(p->export_ev[level_next], p->stream));      // stream 0 has joined stream 1 at the end of a level
            CK(cudaStreamWaitEvent(p->stream3, p->export_ev[level_next], 0));
            for (const auto& r : runs)
                CK(cudaMemcpyAsync(host_out + r.first, p->dL + r.first, static_cast<size_t>(r.second) * sizeof(double),
                                   cudaMemcpyDeviceToHost, p->stream3));
        }
        return 0;
    };
    // development aid: SPCHOL_TRACE=<file> dumps the device time of every stream event of one factorization
    static const char* trace_path = getenv("SPCHOL_TRACE");
    std::vector<cudaEvent_t> tev;
    std::vector<int> tinfo;
    cudaEvent_t t0 = nullptr;
    if (trace_path && !prof) { cudaEventCreate(&t0); cudaEventRecord(t0, p->stream); }
    for (const Op& op : H.ops) {
        if (host_out && op.level > level_next) { const int rc2 = levels_done(op.level); if (rc2) return rc2; }
        if (op.kind == OP_EV_RECORD || op.kind == OP_EV_WAIT) {
            if (sts[0] == sts[1]) continue;
            cudaEvent_t e = p->ring[op.begin % kEventRing];
            if (op.kind == OP_EV_RECORD) {
                CK(cudaEventRecord(e, sts[op.stream]));
                if (t0) { cudaEvent_t te; cudaEventCreate(&te); cudaEventRecord(te, sts[op.stream]); tev.push_back(te);
                          tinfo.push_back(op.begin); tinfo.push_back(op.stream); tinfo.push_back(op.level); }
            } else CK(cudaStreamWaitEvent(sts[op.stream], e, 0));
            continue;
        }
        if (op.kind == OP_EXPORT) {
            if (host_out) {
                cudaEvent_t e = p->export_mark_ev[n_mark];
                CK(cudaEventRecord(e, sts[op.stream]));
                CK(cudaStreamWaitEvent(p->stream3, e, 0));
                for (int i = 0; i < op.count; ++i) {
                    const ExportTask& t = H.export_tasks[op.begin + i];
                    CK(cudaMemcpyAsync(host_out + t.off, p->dL + t.off, static_cast<size_t>(t.count) * sizeof(double),
                                       cudaMemcpyDeviceToHost, p->stream3));
                }
            }
            ++n_mark;
            continue;
        }
        st = sts[op.stream];
        const int pk = op.kind == OP_GEMM ? 9 + gemm_k_class(op.threads) : (op.kind == OP_TREE ? 0 : op.kind);   // profile slot (subtree tasks count as small fronts)
        mark(pk);
        const int64_t before = launches;
        switch (op.kind) {
        case OP_TREE: {
            static const int nw1 = getenv("SPCHOL_TREE_NW") ? atoi(getenv("SPCHOL_TREE_NW")) : 1;
            static const int nw2 = getenv("SPCHOL_TREE2_NW") ? atoi(getenv("SPCHOL_TREE2_NW")) : 4;
            const spchol_plan::DevTree& T = op.begin ? p->tree2 : p->tree;
            launch_tree(op.begin ? nw2 : nw1, op.count, T.f_stride, st, T.ta, dAx, p->dL, p->dW, T.stash, nullptr, p->dinfo, 0, 0, 0, 0, 0, 0, 0);
            ++launches;
        } break;
        case OP_SMALL: {
            const int* list = p->d_small_list + op.begin;
            const unsigned g = static_cast<unsigned>(op.count);
            switch (op.threads) {
            case 32:  k_front_small<32><<<g, 32, op.smem, st>>>(list, p->fa, p->dL, p->dW, p->dinfo, 0, 0); break;
            case 64:  k_front_small<64><<<g, 64, op.smem, st>>>(list, p->fa, p->dL, p->dW, p->dinfo, 0, 0); break;
            case 128: k_front_small<128><<<g, 128, op.smem, st>>>(list, p->fa, p->dL, p->dW, p->dinfo, 0, 0); break;
            default:  k_front_small<256><<<g, 256, op.smem, st>>>(list, p->fa, p->dL, p->dW, p->dinfo, 0, 0); break;
            }
            ++launches;
        } break;
        case OP_ZERO:
            for (int b = 0; b < op.count; b += 65535) {
                const int c = std::min(op.count - b, 65535);
                k_zero<<<dim3(std::max(1, std::min(op.gx, 1184)), c), 256, 0, st>>>(p->d_zero + op.begin + b, p->dW);
                ++launches;
            }
            break;
        case OP_EXTADD:
            k_extend_add<<<dim3(op.count, std::min(op.gx, 2048)), 128, 0, st>>>(p->d_extadd_list + op.begin, p->fa, p->dL, p->dW, op.threads);
            ++launches;
            break;
        case OP_POTRF:
            if (op.threads <= 32)      k_potrf_inv<32><<<op.count, PI_THREADS, potrf_inv_smem(32), st>>>(p->d_potrf + op.begin, p->dL, p->dWinv, p->dinfo);
            else if (op.threads <= 64) k_potrf_inv<64><<<op.count, PI_THREADS, potrf_inv_smem(64), st>>>(p->d_potrf + op.begin, p->dL, p->dWinv, p->dinfo);
            else                       k_potrf_inv<128><<<op.count, PI_THREADS, potrf_inv_smem(128), st>>>(p->d_potrf + op.begin, p->dL, p->dWinv, p->dinfo);
            ++launches;
            break;
        case OP_TRSM: {
            const dim3 grid(op.count, (op.gx + TI_ROWS - 1) / TI_ROWS);
            if (op.threads <= 32)      k_trsm_inv<32><<<grid, 256, trsm_inv_smem(32), st>>>(p->d_trsm + op.begin, p->dL, p->dWinv);
            else if (op.threads <= 64) k_trsm_inv<64><<<grid, 256, trsm_inv_smem(64), st>>>(p->d_trsm + op.begin, p->dL, p->dWinv);
            else                       k_trsm_inv<128><<<grid, 256, trsm_inv_smem(128), st>>>(p->d_trsm + op.begin, p->dL, p->dWinv);
            ++launches;
        } break;
        case OP_GEMM:
            for (int b = 0; b < op.count; b += 65535) {
                const int c = std::min(op.count - b, 65535);
                if (p->gemm_tma)
                    k_gemm_tma<<<dim3((op.gx + 1 + G3_BM - 1) / G3_BM, (op.gy + 1 + G3_BN - 1) / G3_BN, c), 256, G3_SMEM, st>>>(      // (+1: odd-row shift)
                        p->d_gemm + op.begin + b, p->d_tmaps, p->dL, p->dW, gemm_swizzle);
                else
                    k_gemm_dmma<<<dim3((op.gx + G_BM - 1) / G_BM, (op.gy + G_BN - 1) / G_BN, c), 256, G_SMEM, st>>>(
                        p->d_gemm + op.begin + b, p->dL, p->dW, gemm_swizzle);
                ++launches;
            }
            break;
        case OP_COMM: {
            NcclApi* nc = nccl_api();
            if (!p->comm || !nc) { set_error("spchol: plan has transfers but no communicator"); return SPCHOL_EINVAL; }
            if (op.count > 1) NCK(nc->GroupStart());      // grouped sends/recvs progress concurrently
            for (int b = 0; b < op.count; ++b) {
                const CommTask& t = H.comm_tasks[op.begin + b];
                double* buf = (t.space ? p->dW : p->dL) + t.off;
                if (t.kind == COMM_BCAST) {
                    void* c = t.n == p->comm->world ? p->comm->world_comm : p->comm->groups[{t.lo, t.n}];
                    if (!c) { set_error("spchol: broadcast on a rank group this rank is not part of"); return SPCHOL_EINVAL; }
                    NCK(nc->Broadcast(buf, buf, static_cast<size_t>(t.count), kNcclFloat64, t.root - t.lo, c, st));
                } else if (t.kind == COMM_SEND) {
                    NCK(nc->Send(buf, static_cast<size_t>(t.count), kNcclFloat64, t.root, p->comm->world_comm, st));
                } else {
                    NCK(nc->Recv(buf, static_cast<size_t>(t.count), kNcclFloat64, t.root, p->comm->world_comm, st));
                }
                ++launches;
            }
            if (op.count > 1) NCK(nc->GroupEnd());
        } break;
        case OP_COPY:
            for (int b = 0; b < op.count; b += 65535) {
                const int c = std::min(op.count - b, 65535);
                k_copy2d<<<dim3(c, std::max(1, std::min(op.gx, 1024)), std::max(1, std::min(op.gy, 16))), 256, 0, st>>>(p->d_copy + op.begin + b, p->dL, p->dW);
                ++launches;
            }
            break;
        default: break;
        }
        if (prof) p->kind_launches[pk] += launches - before;
        if (t0) { cudaEvent_t te; cudaEventCreate(&te); cudaEventRecord(te, st); tev.push_back(te);
                  tinfo.push_back(100 + pk); tinfo.push_back(op.stream); tinfo.push_back(op.level); }
    }
    st = p->stream;
    if (host_out) {
        const int rc2 = levels_done(H.nlevels);
        if (rc2) return rc2;
        CK(cudaEventRecord(p->export_ev[H.nlevels], p->stream3));          // the caller's stream waits for the read-back
        CK(cudaStreamWaitEvent(p->stream, p->export_ev[H.nlevels], 0));
    }
    if (p->comm && p->comm->world > 1) {      // first failing column over all ranks
        mark(8);
        NCK(nccl_api()->AllReduce(p->dinfo, p->dinfo, 1, kNcclInt32, kNcclMin, p->comm->world_comm, st));
    }
    mark(-1);
    CK(cudaGetLastError());
    p->last_launches = launches;
    if (t0) {
        cudaEvent_t te; cudaEventCreate(&te); cudaEventRecord(te, p->stream); tev.push_back(te);
        tinfo.push_back(-1); tinfo.push_back(0); tinfo.push_back(-1);
        CK(cudaStreamSynchronize(p->stream));
        std::string tp = trace_path;
        if (p->comm) tp += ".r" + std::to_string(p->comm->rank);
        if (FILE* f = fopen(tp.c_str(), "w")) {
            for (size_t i = 0; i < tev.size(); ++i) {
                float ms = 0; cudaEventElapsedTime(&ms, t0, tev[i]);
                fprintf(f, "%d %d %d %.4f\n", tinfo[3 * i], tinfo[3 * i + 1], tinfo[3 * i + 2], ms);
            }
            fclose(f);
        }
        for (cudaEvent_t e : tev) cudaEventDestroy(e);
        cudaEventDestroy(t0);
    }
    if (prof) {
        CK(cudaStreamSynchronize(st));
        for (size_t i = 0; i + 1 < pe; ++i) {
            float ms = 0;
            CK(cudaEventElapsedTime(&ms, p->prof_ev[i], p->prof_ev[i + 1]));
            if (pkind[i] >= 0 && pkind[i] < 12) p->kind_ms[pkind[i]] += ms;
        }
    }
    return 0;
}

int read_info(spchol_plan* p, int* info) {
    int v = 0;
    CK(cudaMemcpyAsync(&v, p->dinfo, sizeof(int), cudaMemcpyDeviceToHost, p->stream));
    CK(cudaStreamSynchronize(p->stream));
    if (info) *info = (v == 0x7f7f7f7f) ? 0 : v;
    return 0;
}

}  // namespace

extern "C" {

int spchol_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

static int plan_create_impl(const spchol_symbolic* sym, int device, spchol_comm* comm, const int* dist_opts,
                            spchol_plan** out) {
    if (!sym || !out) { set_error("spchol_plan_create: null argument"); return SPCHOL_EINVAL; }
    *out = nullptr;
    if (spchol_device_count() <= 0) {
        set_error("spchol: no CUDA device available (the numeric path has no CPU fallback)");
        return SPCHOL_ENOGPU;
    }
    spchol_plan* p = new (std::nothrow) spchol_plan();
    if (!p) return SPCHOL_ENOMEM;
    p->device = device;
    p->comm = comm;
    try {
        PlanOptions opt;
        const Symbolic& S = symbolic_of(sym);
        opt.want_csc = S.nnzL <= INT_MAX ? 1 : 0;
        if (comm) { opt.rank = comm->rank; opt.world = comm->world; }
        if (const char* la = getenv("SPCHOL_LOOKAHEAD")) opt.lookahead = atoi(la) != 0;
        if (const char* v = getenv("SPCHOL_DEFER_CB")) opt.defer_cb_add = atoi(v) != 0;
        if (const char* v = getenv("SPCHOL_SMALL_MAX_M")) { const int x = atoi(v); if (x >= 16 && x <= 224) opt.small_max_m = x; }   // development switch (A/B timing)
        if (const char* v = getenv("SPCHOL_TREE")) opt.tree = atoi(v) != 0;
        if (const char* v = getenv("SPCHOL_TREE_CAP")) { const int x = atoi(v); if (x >= 8 && x <= TW_MAX_M) opt.tree_cap_m = x; }
        if (const char* v = getenv("SPCHOL_BATCH_GROUP_M")) { const int x = atoi(v); if (x >= 0 && x <= TW_MAX_M) opt.batch_group_m = x; }
        if (const char* v = getenv("SPCHOL_TREE_WORK")) { const double x = atof(v); if (x > 0) opt.tree_work = x; }
        if (const char* v = getenv("SPCHOL_TREE2_CAP")) { const int x = atoi(v); if (x >= 0 && x <= TW_MAX_M) opt.tree2_cap_m = x; }
        if (const char* v = getenv("SPCHOL_TREE2_WORK")) { const double x = atof(v); if (x > 0) opt.tree2_work = x; }
        if (const char* v = getenv("SPCHOL_NB")) { const int x = atoi(v); if (x >= 128) opt.NB = x; }
        if (const char* v = getenv("SPCHOL_NB_WIDE")) { const int x = atoi(v); if (x >= 128) opt.NB_wide = x; }
        if (const char* v = getenv("SPCHOL_WIDE_W")) { const int x = atoi(v); if (x >= 1) opt.wide_w = x; }
        if (const char* v = getenv("SPCHOL_EXPORT_MARK_W")) { const int x = atoi(v); if (x >= 0) opt.export_mark_w = x; }   // 0: read-back by levels only
        if (dist_opts) {
            if (dist_opts[0] > 0) opt.dist_NB = dist_opts[0];
            if (dist_opts[1] > 0) opt.dist_min_w = dist_opts[1];
        }
        build_plan(S, opt, p->H);
    } catch (const std::exception& e) {
        set_error(e.what()); delete p; return SPCHOL_EINVAL;
    }
    int rc = 0;
    auto fail = [&](int code) { spchol_plan_destroy(p); return code; };
    if (comm && comm->world > 1) {
        // one sub-communicator per distinct rank group of the (global) mapping; every rank walks the
        // same sorted list, so the collective ncclCommSplit calls line up
        NcclApi* nc = nccl_api();
        std::set<std::pair<int, int>> need;
        for (int s2 = 0; s2 < p->H.ns; ++s2)
            if (p->H.own_n[s2] > 1 && p->H.own_n[s2] < comm->world) need.insert({p->H.own_lo[s2], p->H.own_n[s2]});
        for (const auto& g : need) {
            if (comm->groups.count(g)) continue;
            const bool in = comm->rank >= g.first && comm->rank < g.first + g.second;
            void* sub = nullptr;
            const int e = nc->CommSplit(comm->world_comm, in ? 0 : kNcclSplitNoColor, comm->rank, &sub, nullptr);
            if (e != 0) { set_error(std::string("ncclCommSplit: ") + nc->GetErrorString(e)); return fail(SPCHOL_ENCCL); }
            comm->groups[g] = in ? sub : nullptr;
        }
    }
    if (cudaSetDevice(device) != cudaSuccess) { set_error("cudaSetDevice failed"); return fail(SPCHOL_ECUDA); }
    if (cudaStreamCreateWithFlags(&p->stream, cudaStreamNonBlocking) != cudaSuccess) { set_error("cudaStreamCreate failed"); return fail(SPCHOL_ECUDA); }
    {
        int prio_lo = 0, prio_hi = 0;
        cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi);
        if (cudaStreamCreateWithPriority(&p->stream2, cudaStreamNonBlocking, prio_hi) != cudaSuccess) { set_error("cudaStreamCreate failed"); return fail(SPCHOL_ECUDA); }
        for (auto& e : p->ring) if (cudaEventCreateWithFlags(&e, cudaEventDisableTiming) != cudaSuccess) { set_error("cudaEventCreate failed"); return fail(SPCHOL_ECUDA); }
    }
    for (auto& e : p->ev) if (cudaEventCreate(&e) != cudaSuccess) { set_error("cudaEventCreate failed"); return fail(SPCHOL_ECUDA); }
    HostPlan& H = p->H;
#define UP(vec, ptr) if ((rc = upload(p, vec, &ptr))) return fail(rc)
    UP(H.first, p->d_first); UP(H.w, p->d_w); UP(H.m, p->d_m); UP(H.ld, p->d_ld); UP(H.parent, p->d_parent);
    UP(H.panel_off, p->d_panel_off); UP(H.cb_off, p->d_cb_off); UP(H.rel_off, p->d_rel_off);
    UP(H.rel, p->d_rel); UP(H.child_ptr, p->d_child_ptr); UP(H.child_idx, p->d_child_idx);
    UP(H.own_n, p->d_own_n); UP(H.own_idx, p->d_own_idx);
    UP(H.a_dest, p->d_a_dest); UP(H.l_src, p->d_l_src);
    UP(H.small_list, p->d_small_list); UP(H.extadd_list, p->d_extadd_list);
    UP(H.zero_tasks, p->d_zero); UP(H.potrf_tasks, p->d_potrf); UP(H.trsm_tasks, p->d_trsm); UP(H.gemm_tasks, p->d_gemm); UP(H.copy_tasks, p->d_copy);
#undef UP
    // warp-per-subtree programs
    {
        auto up_tree = [&](TreeProgram& T, spchol_plan::DevTree& D, const int* csc_map) -> int {
            if (T.tasks.empty()) return 0;
            TwTask* dt = nullptr; TwFront* df = nullptr; TwStash* dsx = nullptr; int* drm = nullptr; int* das = nullptr; unsigned short* dad = nullptr;
            int r2;
            if ((r2 = upload(p, T.tasks, &dt)) || (r2 = upload(p, T.fronts, &df)) || (r2 = upload(p, T.stashes, &dsx)) ||
                (r2 = upload(p, T.rmap, &drm)) || (r2 = upload(p, T.a_src, &das)) || (r2 = upload(p, T.a_dst, &dad))) return r2;
            unsigned short* dos = nullptr; int* dod = nullptr; int* dgc = nullptr;
            if ((r2 = upload(p, T.out_src, &dos)) || (r2 = upload(p, T.out_dst, &dod)) || (r2 = upload(p, T.gcol, &dgc))) return r2;
            D.ta = TreeArrays{dt, static_cast<int>(T.tasks.size()), df, dsx, drm, das, dad, csc_map, dos, dod, dgc};
            std::vector<unsigned short>().swap(T.out_src); std::vector<int>().swap(T.out_dst); std::vector<int>().swap(T.gcol);
            D.f_stride = ((T.cap_m * (T.cap_m + 1) / 2) + TK_F_SLACK + 1) & ~1;
            D.stash_entries = (T.stash_entries + 1) & ~int64_t(1);
            D.on = true;
            std::vector<TwFront>().swap(T.fronts); std::vector<TwStash>().swap(T.stashes); std::vector<int>().swap(T.rmap);
            std::vector<int>().swap(T.a_src); std::vector<unsigned short>().swap(T.a_dst);
            return 0;
        };
        if ((rc = up_tree(H.tree, p->tree, nullptr))) return fail(rc);
        if (p->tree.on && (rc = dalloc(p, p->tree.stash_entries, &p->tree.stash))) return fail(rc);
        if ((rc = up_tree(H.tree2, p->tree2, nullptr))) return fail(rc);
        if (p->tree2.on && (rc = dalloc(p, p->tree2.stash_entries, &p->tree2.stash))) return fail(rc);
        if (!H.tree_whole.tasks.empty() && H.nnzL <= INT_MAX && !H.l_src.empty()) {
            std::vector<int> inv(static_cast<size_t>(H.lbuf_entries), -1);        // L-buffer offset -> CSC index
            for (int64_t e = 0; e < H.nnzL; ++e) if (H.l_src[e] >= 0) inv[static_cast<size_t>(H.l_src[e])] = static_cast<int>(e);
            int* dmap = nullptr;
            if ((rc = upload(p, inv, &dmap)) || (rc = up_tree(H.tree_whole, p->whole, dmap))) return fail(rc);
        }
        if (cudaFuncSetAttribute(k_tree_team<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024) != cudaSuccess ||
            cudaFuncSetAttribute(k_tree_team<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024) != cudaSuccess ||
            cudaFuncSetAttribute(k_tree_team<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024) != cudaSuccess) {
            set_error("cudaFuncSetAttribute (k_tree_team) failed"); return fail(SPCHOL_ECUDA);
        }
    }
    // the big host-side maps are only needed on the device from here on
    std::vector<int64_t>().swap(H.a_dest);
    std::vector<int64_t>().swap(H.l_src);
    std::vector<int>().swap(H.rel);
    if ((rc = dalloc(p, H.lbuf_entries, &p->dL))) return fail(rc);
    if ((rc = dalloc(p, H.ws_entries, &p->dW))) return fail(rc);
    if ((rc = dalloc(p, 1, &p->dinfo))) return fail(rc);
    if (const char* v = getenv("SPCHOL_GEMM_TMA")) p->gemm_tma = atoi(v) != 0;
    if (!H.tmap_fronts.empty()) {
        // TMA tensor maps of the panels the update kernel reads: m rows x w columns of doubles with row pitch ld, boxes of
        // 16 rows x G3_BK columns, 128-byte swizzle, zeros out of bounds (encoded by the driver, used straight from HBM)
        using EncodeFn = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                      const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                      CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult qr;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qr) != cudaSuccess || !fn) {
            cudaGetLastError();
            set_error("spchol: cuTensorMapEncodeTiled is not available from this driver (needed by the TMA-staged update kernel)");
            return fail(SPCHOL_ECUDA);
        }
        std::vector<CUtensorMap> maps(H.tmap_fronts.size());
        for (size_t i = 0; i < maps.size(); ++i) {
            const int s = H.tmap_fronts[i];
            const cuuint64_t dims[2] = {static_cast<cuuint64_t>(H.m[s]), static_cast<cuuint64_t>(H.w[s])};
            const cuuint64_t strides[1] = {static_cast<cuuint64_t>(H.ld[s]) * 8};
            const cuuint32_t box[2] = {16, static_cast<cuuint32_t>(G3_BK)};
            const cuuint32_t estr[2] = {1, 1};
            const CUresult cr = reinterpret_cast<EncodeFn>(fn)(&maps[i], CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 2, p->dL + H.panel_off[s], dims, strides, box, estr,
                                                               CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                                                               CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
            if (cr != CUDA_SUCCESS) { set_error("spchol: cuTensorMapEncodeTiled failed (code " + std::to_string(static_cast<int>(cr)) + ")"); return fail(SPCHOL_ECUDA); }
        }
        if ((rc = upload(p, maps, &p->d_tmaps))) return fail(rc);
    }
    if ((rc = dalloc(p, static_cast<int64_t>(std::max(H.max_potrf_slots, 1)) * PB_MAX * PB_MAX, &p->dWinv))) return fail(rc);
    p->fa = FrontArrays{p->d_first, p->d_w, p->d_m, p->d_ld, p->d_parent, p->d_panel_off, p->d_cb_off,
                        p->d_rel_off, p->d_rel, p->d_child_ptr, p->d_child_idx, p->d_own_n, p->d_own_idx, H.dist_NB};
    if ((rc = set_small_attr<32>()) || (rc = set_small_attr<64>()) || (rc = set_small_attr<128>()) ||
        (rc = set_small_attr<256>())) return fail(rc);
    // The kernels that run concurrently under look-ahead all ask for the largest shared-memory carve-out: an SM
    // cannot change its L1 / shared split while CTAs are resident, so a panel CTA (103-150 KB) could otherwise
    // not be placed next to an update CTA on an SM configured for the update kernel's 2 x 51 KB.
    {
        const int mx = cudaSharedmemCarveoutMaxShared;
        const auto A = cudaFuncAttributePreferredSharedMemoryCarveout;
        if (cudaFuncSetAttribute(k_gemm_dmma, A, mx) != cudaSuccess ||
            cudaFuncSetAttribute(k_potrf_inv<32>, A, mx) != cudaSuccess || cudaFuncSetAttribute(k_potrf_inv<64>, A, mx) != cudaSuccess ||
            cudaFuncSetAttribute(k_potrf_inv<128>, A, mx) != cudaSuccess ||
            cudaFuncSetAttribute(k_trsm_inv<32>, A, mx) != cudaSuccess || cudaFuncSetAttribute(k_trsm_inv<64>, A, mx) != cudaSuccess ||
            cudaFuncSetAttribute(k_trsm_inv<128>, A, mx) != cudaSuccess || cudaFuncSetAttribute(k_copy2d, A, mx) != cudaSuccess) {
            set_error("cudaFuncSetAttribute (shared memory carve-out) failed"); return fail(SPCHOL_ECUDA);
        }
    }
    if (cudaFuncSetAttribute(k_gemm_tma, cudaFuncAttributeMaxDynamicSharedMemorySize, G3_SMEM) != cudaSuccess ||
        cudaFuncSetAttribute(k_gemm_tma, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared) != cudaSuccess ||
        cudaFuncSetAttribute(k_gemm_dmma, cudaFuncAttributeMaxDynamicSharedMemorySize, G_SMEM) != cudaSuccess ||
        cudaFuncSetAttribute(k_potrf_inv<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, potrf_inv_smem(32)) != cudaSuccess ||
        cudaFuncSetAttribute(k_potrf_inv<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, potrf_inv_smem(64)) != cudaSuccess ||
        cudaFuncSetAttribute(k_potrf_inv<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, potrf_inv_smem(128)) != cudaSuccess ||
        cudaFuncSetAttribute(k_trsm_inv<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, trsm_inv_smem(32)) != cudaSuccess ||
        cudaFuncSetAttribute(k_trsm_inv<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, trsm_inv_smem(64)) != cudaSuccess ||
        cudaFuncSetAttribute(k_trsm_inv<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, trsm_inv_smem(128)) != cudaSuccess) {
        set_error("cudaFuncSetAttribute (dynamic shared memory) failed"); return fail(SPCHOL_ECUDA);
    }
    *out = p;
    return SPCHOL_OK;
}

int spchol_plan_create(const spchol_symbolic* sym, int device, spchol_plan** out) {
    return plan_create_impl(sym, device, nullptr, nullptr, out);
}

int spchol_plan_create_dist(const spchol_symbolic* sym, spchol_comm* comm, const int* dist_opts, spchol_plan** out) {
    if (!comm) { set_error("spchol_plan_create_dist: null communicator"); return SPCHOL_EINVAL; }
    return plan_create_impl(sym, comm->device, comm, dist_opts, out);
}

int spchol_comm_unique_id(void* id128) {
    NcclApi* nc = nccl_api();
    if (!nc) { set_error("spchol: libnccl.so.2 could not be loaded (set SPCHOL_NCCL_LIB)"); return SPCHOL_ENCCL; }
    if (!id128) return SPCHOL_EINVAL;
    NCK(nc->GetUniqueId(static_cast<NcclId*>(id128)));
    return 0;
}

int spchol_comm_create(int rank, int world, const void* id128, int device, spchol_comm** out) {
    if (!out || !id128 || world < 1 || rank < 0 || rank >= world) { set_error("spchol_comm_create: bad argument"); return SPCHOL_EINVAL; }
    *out = nullptr;
    if (spchol_device_count() <= 0) { set_error("spchol_comm_create: no CUDA device"); return SPCHOL_ENOGPU; }
    NcclApi* nc = nccl_api();
    if (!nc) { set_error("spchol: libnccl.so.2 could not be loaded (set SPCHOL_NCCL_LIB)"); return SPCHOL_ENCCL; }
    CK(cudaSetDevice(device));
    spchol_comm* c = new (std::nothrow) spchol_comm();
    if (!c) return SPCHOL_ENOMEM;
    c->rank = rank; c->world = world; c->device = device;
    NcclId id;
    std::memcpy(&id, id128, sizeof(id));
    const int e = nc->CommInitRank(&c->world_comm, world, id, rank);
    if (e != 0) { set_error(std::string("ncclCommInitRank: ") + nc->GetErrorString(e)); delete c; return SPCHOL_ENCCL; }
    *out = c;
    return 0;
}

void spchol_comm_destroy(spchol_comm* c) {
    if (!c) return;
    NcclApi* nc = nccl_api();
    if (nc) {
        for (auto& g : c->groups) if (g.second) nc->CommDestroy(g.second);
        if (c->world_comm) nc->CommDestroy(c->world_comm);
    }
    delete c;
}

int spchol_comm_rank(const spchol_comm* c) { return c ? c->rank : 0; }
int spchol_comm_world(const spchol_comm* c) { return c ? c->world : 1; }

void spchol_plan_destroy(spchol_plan* p) {
    if (!p) return;
    cudaSetDevice(p->device);
    for (void* d : p->allocs) cudaFree(d);
    for (void* d : {static_cast<void*>(p->bStash), static_cast<void*>(p->binfo), static_cast<void*>(p->hbA), static_cast<void*>(p->hbL),
                    static_cast<void*>(p->hbI)}) if (d) cudaFree(d);
    for (auto& e : p->hb_ev) if (e) cudaEventDestroy(e);
    if (p->stream4) cudaStreamDestroy(p->stream4);
    for (auto& e : p->ev) if (e) cudaEventDestroy(e);
    for (auto& e : p->prof_ev) cudaEventDestroy(e);
    for (auto& e : p->ring) if (e) cudaEventDestroy(e);
    for (auto& e : p->export_ev) if (e) cudaEventDestroy(e);
    for (auto& e : p->export_mark_ev) if (e) cudaEventDestroy(e);
    if (p->stream3) cudaStreamDestroy(p->stream3);
    if (p->stream2) cudaStreamDestroy(p->stream2);
    if (p->stream) cudaStreamDestroy(p->stream);
    delete p;
}

int spchol_plan_stats(const spchol_plan* p, int64_t* st) {
    if (!p || !st) return SPCHOL_EINVAL;
    const HostPlan& H = p->H;
    st[0] = H.ns; st[1] = H.nlevels; st[2] = H.n_small; st[3] = H.n_large; st[4] = H.lbuf_entries;
    st[5] = H.ws_entries; st[6] = static_cast<int64_t>(H.ops.size()); st[7] = H.max_m; st[8] = H.max_w;
    st[9] = p->device_bytes;
    return 0;
}

int spchol_plan_dist_stats(const spchol_plan* p, int64_t* st, double* rank_load) {
    if (!p || !st) return SPCHOL_EINVAL;
    const HostPlan& H = p->H;
    int64_t held = 0;
    for (int s = 0; s < H.ns; ++s) held += H.holds(s) ? 1 : 0;
    st[0] = H.rank; st[1] = H.world; st[2] = held; st[3] = H.n_dist;
    st[4] = static_cast<int64_t>(H.comm_tasks.size()); st[5] = H.comm_doubles; st[6] = H.dist_NB;
    // doubles this rank reads back in spchol_factor_lbuf with pinned memory: its own fronts + the block columns it owns of group fronts
    int64_t own = 0;
    for (int s = 0; s < H.ns; ++s) {
        if (!H.holds(s)) continue;
        if (H.own_n[s] == 1) { own += static_cast<int64_t>(H.ld[s]) * H.w[s]; continue; }
        for (int c0 = 0, kb = 0; c0 < H.w[s]; c0 += H.dist_NB, ++kb)
            if (kb % H.own_n[s] == H.own_idx[s]) own += static_cast<int64_t>(std::min(c0 + H.dist_NB, H.w[s]) - c0) * H.ld[s];
    }
    st[7] = own;
    if (rank_load) for (int r = 0; r < H.world; ++r) rank_load[r] = H.rank_load[r];
    return 0;
}

int spchol_factor_device(spchol_plan* p, const double* dAx) {
    if (!p || !dAx) { set_error("spchol_factor_device: null argument"); return SPCHOL_EINVAL; }
    CK(cudaSetDevice(p->device));
    return run_numeric(p, dAx);
}

int spchol_export_csc_device(spchol_plan* p, double* dLx) {
    if (!p || !dLx) { set_error("spchol_export_csc_device: null argument"); return SPCHOL_EINVAL; }
    if (p->H.nnzL > INT_MAX) { set_error("nnz(L) exceeds the 32-bit CSC of the reference API"); return SPCHOL_ERANGE; }
    CK(cudaSetDevice(p->device));
    const int64_t nb = (p->H.nnzL + 255) / 256;
    k_export<<<static_cast<unsigned>(nb), 256, 0, p->stream>>>(p->H.nnzL, p->d_l_src, p->dL, dLx, 0);
    CK(cudaGetLastError());
    return 0;
}

int spchol_sync(spchol_plan* p, int* info) {
    if (!p) return SPCHOL_EINVAL;
    CK(cudaSetDevice(p->device));
    return read_info(p, info);
}

int spchol_factor(spchol_plan* p, const double* Ax, double* Lx, int* info) {
    if (!p || !Ax || !Lx) { set_error("spchol_factor: null argument"); return SPCHOL_EINVAL; }
    if (p->H.nnzL > INT_MAX) { set_error("nnz(L) exceeds the 32-bit CSC of the reference API"); return SPCHOL_ERANGE; }
    CK(cudaSetDevice(p->device));
    int rc;
    if (!p->dAx && (rc = dalloc(p, p->H.nnzA, &p->dAx))) return rc;
    if (!p->dLx && (rc = dalloc(p, p->H.nnzL, &p->dLx))) return rc;
    cudaStream_t st = p->stream;
    CK(cudaEventRecord(p->ev[0], st));
    CK(cudaMemcpyAsync(p->dAx, Ax, static_cast<size_t>(p->H.nnzA) * sizeof(double), cudaMemcpyHostToDevice, st));
    CK(cudaEventRecord(p->ev[1], st));
    if ((rc = run_numeric(p, p->dAx))) return rc;
    CK(cudaEventRecord(p->ev[2], st));
    if ((rc = spchol_export_csc_device(p, p->dLx))) return rc;
    CK(cudaEventRecord(p->ev[3], st));
    CK(cudaMemcpyAsync(Lx, p->dLx, static_cast<size_t>(p->H.nnzL) * sizeof(double), cudaMemcpyDeviceToHost, st));
    CK(cudaEventRecord(p->ev[4], st));
    if ((rc = read_info(p, info))) return rc;
    for (int i = 0; i < 4; ++i) {
        float ms = 0;
        CK(cudaEventElapsedTime(&ms, p->ev[i], p->ev[i + 1]));
        p->last_ms[i] = ms;
    }
    return SPCHOL_OK;
}

int spchol_factor_resident(spchol_plan* p, const double* Ax, int* info) {
    if (!p || !Ax) { set_error("spchol_factor_resident: null argument"); return SPCHOL_EINVAL; }
    CK(cudaSetDevice(p->device));
    int rc;
    if (!p->dAx && (rc = dalloc(p, p->H.nnzA, &p->dAx))) return rc;
    CK(cudaMemcpyAsync(p->dAx, Ax, static_cast<size_t>(p->H.nnzA) * sizeof(double), cudaMemcpyHostToDevice, p->stream));
    if ((rc = run_numeric(p, p->dAx))) return rc;
    return read_info(p, info);
}

// H2D of Ax, numeric, D2H of the whole supernodal L buffer (for patterns whose CSC L exceeds
// the reference's 32-bit indices, e.g. the 128^3 Laplacian).
int spchol_factor_lbuf(spchol_plan* p, const double* Ax, double* Lbuf_host, int* info) {
    if (!p || !Ax || !Lbuf_host) { set_error("spchol_factor_lbuf: null argument"); return SPCHOL_EINVAL; }
    CK(cudaSetDevice(p->device));
    int rc;
    if (!p->dAx && (rc = dalloc(p, p->H.nnzA, &p->dAx))) return rc;
    cudaStream_t st = p->stream;
    CK(cudaEventRecord(p->ev[0], st));
    CK(cudaMemcpyAsync(p->dAx, Ax, static_cast<size_t>(p->H.nnzA) * sizeof(double), cudaMemcpyHostToDevice, st));
    CK(cudaEventRecord(p->ev[1], st));
    // pinned host memory: read the panels back level by level while the upper levels are still being factored
    // (only the panels of the last level are left when the factorization ends); pageable memory: one copy at the end
    cudaPointerAttributes pa{};
    const bool pinned = cudaPointerGetAttributes(&pa, Lbuf_host) == cudaSuccess && pa.type == cudaMemoryTypeHost &&
                        !(getenv("SPCHOL_OVERLAP_D2H") && atoi(getenv("SPCHOL_OVERLAP_D2H")) == 0);
    cudaGetLastError();
    // (the per-kind profiling replay runs on one stream and does not overlap the read-back: plain copy at the end)
    const bool overlapped = pinned && !p->profile;
    if (overlapped && (rc = export_prepare(p))) return rc;
    if ((rc = run_numeric(p, p->dAx, overlapped ? Lbuf_host : nullptr))) return rc;
    CK(cudaEventRecord(p->ev[2], st));
    CK(cudaEventRecord(p->ev[3], st));
    if (!overlapped)
        CK(cudaMemcpyAsync(Lbuf_host, p->dL, static_cast<size_t>(p->H.lbuf_entries) * sizeof(double), cudaMemcpyDeviceToHost, st));
    CK(cudaEventRecord(p->ev[4], st));
    if ((rc = read_info(p, info))) return rc;
    for (int i = 0; i < 4; ++i) {
        float ms = 0;
        CK(cudaEventElapsedTime(&ms, p->ev[i], p->ev[i + 1]));
        p->last_ms[i] = ms;
    }
    return SPCHOL_OK;
}

// ---- batched numeric factorization: many value sets over one pattern (config 5) ---------------
// ONE small CTA (a team of warps) per matrix walks the whole assembly tree (HostPlan::tree_whole, k_tree_team): the values of A are read
// straight from Ax, the frontal matrices live in shared memory, the factor is written straight into the CSC order of
// S -- no L buffer, no assembly pass, no export pass; the only other HBM traffic is the stash of side-branch blocks.
// Patterns whose fronts all have m <= TW_MAX_M rows (true for 1138_bus-sized problems).
static int batched_chunk(const spchol_plan* p, int nbatch) {
    size_t fr = 0, tot = 0;
    if (cudaMemGetInfo(&fr, &tot) != cudaSuccess) return 1;
    const double per = 8.0 * static_cast<double>(p->whole.stash_entries) + 64;
    const double c = 0.8 * static_cast<double>(fr) / per;
    return static_cast<int>(std::max(1.0, std::min(c, static_cast<double>(std::max(nbatch, 1)))));
}

static int batched_reserve(spchol_plan* p, int chunk) {
    if (p->b_chunk >= chunk) return 0;
    if (p->bStash) cudaFree(p->bStash);
    if (p->binfo) cudaFree(p->binfo);
    p->bStash = nullptr; p->binfo = nullptr; p->b_chunk = 0;
    CK(cudaMalloc(&p->bStash, sizeof(double) * std::max<int64_t>(p->whole.stash_entries, 1) * chunk));
    CK(cudaMalloc(&p->binfo, sizeof(int) * chunk));
    p->b_chunk = chunk;
    return 0;
}

static int batched_check(const spchol_plan* p, const char* who) {
    const HostPlan& H = p->H;
    if (!p->whole.on || H.world > 1 || H.nnzL > INT_MAX) {
        set_error(std::string(who) + ": only single-GPU plans whose fronts all fit the warp-per-tree kernel (m <= 128 rows)");
        return SPCHOL_EINVAL;
    }
    return 0;
}

// device-resident: dAx[nbatch][nnzA] -> dLx[nbatch][nnzL] (CSC order of S), dinfo[nbatch] (0 or failing column)
int spchol_factor_batched_device(spchol_plan* p, int nbatch, const double* dAx, double* dLx, int* dinfo) {
    if (!p || !dAx || !dLx || !dinfo || nbatch < 0) { set_error("spchol_factor_batched_device: bad argument"); return SPCHOL_EINVAL; }
    int rc;
    if ((rc = batched_check(p, "spchol_factor_batched"))) return rc;
    const HostPlan& H = p->H;
    CK(cudaSetDevice(p->device));
    cudaStream_t st = p->stream;
    int64_t launches = 0;
    {
        const int want = batched_chunk(p, nbatch);
        if (p->b_chunk < want && (rc = batched_reserve(p, want))) return rc;
    }
    const int chunk = p->b_chunk;
    static const int nw = getenv("SPCHOL_BATCH_NW") ? atoi(getenv("SPCHOL_BATCH_NW")) : 4;      // warps of the team that factors one matrix
    for (int b0 = 0; b0 < nbatch; b0 += chunk) {
        const int nb = std::min(chunk, nbatch - b0);
        CK(cudaMemsetAsync(p->binfo, 0x7f, sizeof(int) * nb, st));
        launch_tree(nw, nb, p->whole.f_stride, st, p->whole.ta, dAx + static_cast<int64_t>(b0) * H.nnzA, nullptr, nullptr, p->bStash,
                    dLx + static_cast<int64_t>(b0) * H.nnzL, p->binfo, H.nnzA, 0, 0, p->whole.stash_entries, H.nnzL, nb, 1);
        k_fix_info<<<(nb + 255) / 256, 256, 0, st>>>(p->binfo, dinfo + b0, nb);
        launches += 2;
    }
    CK(cudaGetLastError());
    p->last_launches = launches;
    return 0;
}

// host buffers: Ax[nbatch][nnzA] in, Lx[nbatch][nnzL] and info[nbatch] out.  The batch is cut into slices that go
// through three streams round-robin (H2D -> factor -> D2H per slice), so the copies of one slice overlap the kernel
// of another; device staging buffers are kept with the plan.
int spchol_factor_batched(spchol_plan* p, int nbatch, const double* Ax, double* Lx, int* info) {
    if (!p || !Ax || !Lx || !info || nbatch < 0) { set_error("spchol_factor_batched: bad argument"); return SPCHOL_EINVAL; }
    int rc;
    if ((rc = batched_check(p, "spchol_factor_batched"))) return rc;
    CK(cudaSetDevice(p->device));
    const HostPlan& H = p->H;
    if (nbatch == 0) return 0;
    // staging for the whole batch when it fits a quarter of the free memory, else slice by slice through the same buffers
    size_t fr = 0, tot = 0;
    CK(cudaMemGetInfo(&fr, &tot));
    const double per = 8.0 * (static_cast<double>(H.nnzA) + H.nnzL) + 4;
    const int cap = static_cast<int>(std::max(1.0, std::min(static_cast<double>(nbatch), 0.25 * static_cast<double>(fr) / per)));
    if (p->hb_cap < cap) {
        for (void* d : {static_cast<void*>(p->hbA), static_cast<void*>(p->hbL), static_cast<void*>(p->hbI)}) if (d) cudaFree(d);
        p->hbA = p->hbL = nullptr; p->hbI = nullptr; p->hb_cap = 0;
        CK(cudaMalloc(&p->hbA, sizeof(double) * H.nnzA * cap));
        CK(cudaMalloc(&p->hbL, sizeof(double) * H.nnzL * cap));
        CK(cudaMalloc(&p->hbI, sizeof(int) * cap));
        p->hb_cap = cap;
    }
    if (!p->hb_ev[0]) for (auto& e : p->hb_ev) CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    if (!p->stream3) CK(cudaStreamCreateWithFlags(&p->stream3, cudaStreamNonBlocking));
    if (!p->stream4) CK(cudaStreamCreateWithFlags(&p->stream4, cudaStreamNonBlocking));
    // slices: H2D on stream3, kernels on the plan's stream, D2H on stream4; a slice's buffers are its own part of the staging
    const int nslice = std::min(kBatchSlices, std::max(1, std::min(nbatch, p->hb_cap) / 256));
    int64_t launches = 0;
    for (int c0 = 0; c0 < nbatch; c0 += p->hb_cap) {
        const int cn = std::min(p->hb_cap, nbatch - c0);
        if (c0 > 0) CK(cudaStreamSynchronize(p->stream4));       // the staging buffers are reused
        for (int sl = 0; sl < nslice; ++sl) {
            const int s0 = static_cast<int>(static_cast<int64_t>(cn) * sl / nslice), s1 = static_cast<int>(static_cast<int64_t>(cn) * (sl + 1) / nslice);
            if (s1 <= s0) continue;
            const int ns_ = s1 - s0;
            CK(cudaMemcpyAsync(p->hbA + static_cast<int64_t>(s0) * H.nnzA, Ax + static_cast<int64_t>(c0 + s0) * H.nnzA,
                               sizeof(double) * H.nnzA * ns_, cudaMemcpyHostToDevice, p->stream3));
            CK(cudaEventRecord(p->hb_ev[2 * sl], p->stream3));
            CK(cudaStreamWaitEvent(p->stream, p->hb_ev[2 * sl], 0));
            if ((rc = spchol_factor_batched_device(p, ns_, p->hbA + static_cast<int64_t>(s0) * H.nnzA, p->hbL + static_cast<int64_t>(s0) * H.nnzL, p->hbI + s0))) return rc;
            launches += p->last_launches;
            CK(cudaEventRecord(p->hb_ev[2 * sl + 1], p->stream));
            CK(cudaStreamWaitEvent(p->stream4, p->hb_ev[2 * sl + 1], 0));
            CK(cudaMemcpyAsync(Lx + static_cast<int64_t>(c0 + s0) * H.nnzL, p->hbL + static_cast<int64_t>(s0) * H.nnzL,
                               sizeof(double) * H.nnzL * ns_, cudaMemcpyDeviceToHost, p->stream4));
            CK(cudaMemcpyAsync(info + c0 + s0, p->hbI + s0, sizeof(int) * ns_, cudaMemcpyDeviceToHost, p->stream4));
        }
    }
    CK(cudaStreamSynchronize(p->stream4));
    CK(cudaStreamSynchronize(p->stream));
    p->last_launches = launches;
    return 0;
}

// ---- device-side ingest: COO -> upper CSC on the GPU, see ingest.cuh ------------------------------
namespace {
int ingest_scan(int64_t n, const unsigned* in, unsigned* out, unsigned* sums, cudaStream_t st) {
    const int nb = static_cast<int>((n + IN_SCAN_BLOCK - 1) / IN_SCAN_BLOCK);
    if (nb <= 0) return 0;
    k_in_scan_block<<<nb, 256, 0, st>>>(n, in, out, sums);
    if (nb > 1) {
        k_in_scan_sums<<<1, 256, 0, st>>>(nb, sums);
        k_in_scan_add<<<nb, 256, 0, st>>>(n, out, sums);
    }
    return 0;
}
}  // namespace

// d_ti / d_tj / d_tx: nz triplets in device memory; d_Ap [n+1], d_Ai / d_Ax [nz] device outputs; *nnz_out = merged entries.
// Synchronous on return (the count comes back to the host); stream = a cudaStream_t or NULL.
int spchol_triplet_to_csc_device(int64_t nz, const int* d_ti, const int* d_tj, const double* d_tx, int n,
                                 int* d_Ap, int* d_Ai, double* d_Ax, int64_t* nnz_out, void* stream) {
    if (nz < 0 || n < 1 || !d_Ap || !nnz_out || (nz > 0 && (!d_ti || !d_tj || !d_tx || !d_Ai || !d_Ax))) { set_error("spchol_triplet_to_csc_device: bad argument"); return SPCHOL_EINVAL; }
    if (nz > INT_MAX) { set_error("spchol_triplet_to_csc_device: more than INT_MAX triplets"); return SPCHOL_ERANGE; }
    if (spchol_device_count() <= 0) { set_error("spchol_triplet_to_csc_device: no CUDA device"); return SPCHOL_ENOGPU; }
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    *nnz_out = 0;
    if (nz == 0) { CK(cudaMemsetAsync(d_Ap, 0, sizeof(int) * (static_cast<size_t>(n) + 1), st)); CK(cudaStreamSynchronize(st)); return 0; }
    const int nwarps = static_cast<int>((nz + IN_WARP_ITEMS - 1) / IN_WARP_ITEMS);
    const int64_t nh = static_cast<int64_t>(256) * nwarps;
    const int64_t nscan = std::max<int64_t>(std::max<int64_t>(nh, nz), n + 1);
    const int nbs = static_cast<int>((nscan + IN_SCAN_BLOCK - 1) / IN_SCAN_BLOCK);
    // one arena: 2 key buffers, 2 payload buffers, histogram, head flags, slots, column counts, scan sums, flag
    auto up = [](size_t b) { return (b + 255) & ~size_t(255); };
    const size_t bK = up(sizeof(unsigned long long) * nz), bI = up(sizeof(unsigned) * nz), bH = up(sizeof(unsigned) * nh),
                 bC = up(sizeof(unsigned) * (static_cast<size_t>(n) + 1)), bS = up(sizeof(unsigned) * nbs);
    // work space: kept between calls (grow-only, per process; cudaMalloc / cudaFree of ~40 bytes per triplet would cost more than
    // the sort itself).  Not thread-safe: one ingest at a time per process.
    static unsigned char* s_arena = nullptr;
    static size_t s_arena_bytes = 0;
    static int s_arena_dev = -1;
    const size_t need = 2 * bK + 4 * bI + bH + bC + bS + 256;
    int dev = 0;
    CK(cudaGetDevice(&dev));
    if (need > s_arena_bytes || dev != s_arena_dev) {
        if (s_arena) { cudaSetDevice(s_arena_dev); cudaFree(s_arena); cudaSetDevice(dev); }
        s_arena = nullptr; s_arena_bytes = 0;
        CK(cudaMalloc(&s_arena, need));
        s_arena_bytes = need; s_arena_dev = dev;
    }
    unsigned char* arena = s_arena;
    unsigned char* q = arena;
    auto take = [&](size_t b) { unsigned char* r = q; q += b; return r; };
    unsigned long long* key[2] = {reinterpret_cast<unsigned long long*>(take(bK)), reinterpret_cast<unsigned long long*>(take(bK))};
    unsigned* idx[2] = {reinterpret_cast<unsigned*>(take(bI)), reinterpret_cast<unsigned*>(take(bI))};
    unsigned* head = reinterpret_cast<unsigned*>(take(bI));
    unsigned* slot = reinterpret_cast<unsigned*>(take(bI));
    unsigned* hist = reinterpret_cast<unsigned*>(take(bH));
    unsigned* colc = reinterpret_cast<unsigned*>(take(bC));
    unsigned* sums = reinterpret_cast<unsigned*>(take(bS));
    int* bad = reinterpret_cast<int*>(take(256));
    auto done = [&](int code) { return code; };
    if (cudaMemsetAsync(bad, 0, sizeof(int), st) != cudaSuccess || cudaMemsetAsync(colc, 0, sizeof(unsigned) * (static_cast<size_t>(n) + 1), st) != cudaSuccess)
        { set_error("spchol_triplet_to_csc_device: cudaMemsetAsync failed"); return done(SPCHOL_ECUDA); }
    const unsigned gb = static_cast<unsigned>((nz + 255) / 256);
    k_in_keys<<<gb, 256, 0, st>>>(nz, d_ti, d_tj, n, key[0], idx[0], bad);
    int bits = 1;
    while ((int64_t(1) << bits) < n) ++bits;
    const int nbytes = (bits + 7) / 8;
    int cur = 0;
    for (int half = 0; half < 2; ++half)
        for (int b = 0; b < nbytes; ++b) {
            const int shift = 32 * half + 8 * b;
            k_in_hist<<<(nwarps + 7) / 8, 256, 0, st>>>(nz, key[cur], shift, nwarps, hist);
            ingest_scan(nh, hist, hist, sums, st);
            k_in_scatter<<<(nwarps + 7) / 8, 256, 0, st>>>(nz, key[cur], idx[cur], shift, nwarps, hist, key[cur ^ 1], idx[cur ^ 1]);
            cur ^= 1;
        }
    k_in_heads<<<gb, 256, 0, st>>>(nz, key[cur], head, colc);
    ingest_scan(nz, head, slot, sums, st);
    k_in_emit<<<gb, 256, 0, st>>>(nz, key[cur], idx[cur], head, slot, d_tx, d_Ai, d_Ax);
    unsigned last[2] = {0, 0};
    cudaMemcpyAsync(&last[0], slot + (nz - 1), sizeof(unsigned), cudaMemcpyDeviceToHost, st);
    cudaMemcpyAsync(&last[1], head + (nz - 1), sizeof(unsigned), cudaMemcpyDeviceToHost, st);
    int hbad = 0;
    cudaMemcpyAsync(&hbad, bad, sizeof(int), cudaMemcpyDeviceToHost, st);
    ingest_scan(static_cast<int64_t>(n) + 1, colc, colc, sums, st);
    if (cudaStreamSynchronize(st) != cudaSuccess || cudaGetLastError() != cudaSuccess) { set_error("spchol_triplet_to_csc_device: kernel failure"); return done(SPCHOL_ECUDA); }
    if (hbad) { set_error("spchol_triplet_to_csc_device: index outside [0, n)"); return done(SPCHOL_EINVAL); }
    const unsigned total = last[0] + last[1];
    k_in_ap<<<(n + 1 + 255) / 256, 256, 0, st>>>(n, colc, total, d_Ap);
    if (cudaStreamSynchronize(st) != cudaSuccess) { set_error("spchol_triplet_to_csc_device: kernel failure"); return done(SPCHOL_ECUDA); }
    *nnz_out = total;
    return done(0);
}

// host-pointer convenience around the device ingest (tests / callers without device buffers): H2D, ingest, D2H;
// *device_ms (optional) = time of the device part alone (CUDA events).
int spchol_triplet_to_csc_gpu(int64_t nz, const int* ti, const int* tj, const double* tx, int n, int* Ap, int* Ai, double* Ax,
                              int64_t* nnz_out, double* device_ms) {
    if (nz < 0 || n < 1 || !Ap || !nnz_out) { set_error("spchol_triplet_to_csc_gpu: bad argument"); return SPCHOL_EINVAL; }
    if (spchol_device_count() <= 0) { set_error("spchol_triplet_to_csc_gpu: no CUDA device"); return SPCHOL_ENOGPU; }
    int *dti = nullptr, *dtj = nullptr, *dAp = nullptr, *dAi = nullptr; double *dtx = nullptr, *dAx = nullptr;
    const size_t z = static_cast<size_t>(std::max<int64_t>(nz, 1));
    auto freeall = [&] { cudaFree(dti); cudaFree(dtj); cudaFree(dtx); cudaFree(dAp); cudaFree(dAi); cudaFree(dAx); };
    if (cudaMalloc(&dti, 4 * z) != cudaSuccess || cudaMalloc(&dtj, 4 * z) != cudaSuccess || cudaMalloc(&dtx, 8 * z) != cudaSuccess ||
        cudaMalloc(&dAp, 4 * (static_cast<size_t>(n) + 1)) != cudaSuccess || cudaMalloc(&dAi, 4 * z) != cudaSuccess || cudaMalloc(&dAx, 8 * z) != cudaSuccess) {
        freeall(); set_error("spchol_triplet_to_csc_gpu: out of device memory"); return SPCHOL_ENOMEM;
    }
    cudaMemcpy(dti, ti, 4 * static_cast<size_t>(nz), cudaMemcpyHostToDevice);
    cudaMemcpy(dtj, tj, 4 * static_cast<size_t>(nz), cudaMemcpyHostToDevice);
    cudaMemcpy(dtx, tx, 8 * static_cast<size_t>(nz), cudaMemcpyHostToDevice);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0, nullptr);
    const int rc = spchol_triplet_to_csc_device(nz, dti, dtj, dtx, n, dAp, dAi, dAx, nnz_out, nullptr);
    cudaEventRecord(e1, nullptr);
    cudaEventSynchronize(e1);
    float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    if (device_ms) *device_ms = ms;
    if (rc == 0) {
        cudaMemcpy(Ap, dAp, 4 * (static_cast<size_t>(n) + 1), cudaMemcpyDeviceToHost);
        if (*nnz_out > 0) { cudaMemcpy(Ai, dAi, 4 * static_cast<size_t>(*nnz_out), cudaMemcpyDeviceToHost); cudaMemcpy(Ax, dAx, 8 * static_cast<size_t>(*nnz_out), cudaMemcpyDeviceToHost); }
    }
    freeall();
    return rc;
}

// ---- device-side Matrix Market parse (mtx_parse.cuh): text in device memory -> triplets in device memory ----
// d_text / nbytes: the file from its size line on (the leading '%' comment lines removed -- spchol_mtx_read_gpu does that);
// d_ti / d_tj / d_tx [nz_cap]: 0-based triplets mirrored into the upper triangle, as load_matrix_market_to_csc builds them
// (include/mtx_reader.hpp:41-56) and spchol_triplet_to_csc_device takes them.  *host_converted (optional): values whose text
// went to the host's strtod (more than 19 significant digits, inf / nan).  Synchronous on return.
int spchol_mtx_parse_device(const char* d_text, int64_t nbytes, int64_t nz_cap, int* d_ti, int* d_tj, double* d_tx,
                            int* n_out, int64_t* nz_out, int64_t* host_converted, void* stream) {
    static const char* who = "spchol_mtx_parse_device";
    if (!d_text || nbytes < 1 || nz_cap < 0 || !n_out || !nz_out || (nz_cap > 0 && (!d_ti || !d_tj || !d_tx))) { set_error(std::string(who) + ": bad argument"); return SPCHOL_EINVAL; }
    if (nbytes >= (int64_t(1) << 32) || nz_cap > INT_MAX) { set_error(std::string(who) + ": text of 4 GiB or more / more than INT_MAX entries"); return SPCHOL_ERANGE; }
    if (spchol_device_count() <= 0) { set_error(std::string(who) + ": no CUDA device"); return SPCHOL_ENOGPU; }
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int64_t nchunks = (nbytes + MP_CHUNK - 1) / MP_CHUNK;
    const int nbs = static_cast<int>((nchunks + IN_SCAN_BLOCK - 1) / IN_SCAN_BLOCK);
    const unsigned fb_cap = static_cast<unsigned>(std::min<int64_t>(std::max<int64_t>(nz_cap, 1), int64_t(1) << 20));
    auto up = [](size_t b) { return (b + 255) & ~size_t(255); };
    const size_t bCnt = up(sizeof(unsigned) * nchunks), bSum = up(sizeof(unsigned) * nbs), bSmall = 256, bFe = up(sizeof(int) * fb_cap),
                 bFv = up(sizeof(double) * fb_cap), bFt = up(static_cast<size_t>(MP_TOK_MAX) * fb_cap);
    unsigned char* arena = nullptr;
    if (cudaMalloc(&arena, bCnt + bSum + bSmall + bFe + bFv + bFt) != cudaSuccess) { cudaGetLastError(); set_error(std::string(who) + ": out of device memory"); return SPCHOL_ENOMEM; }
    auto fail = [&](int code, const std::string& msg) { cudaFree(arena); set_error(std::string(who) + ": " + msg); return code; };
    unsigned char* q = arena;
    auto take = [&](size_t b) { unsigned char* r = q; q += b; return r; };
    unsigned* counts = reinterpret_cast<unsigned*>(take(bCnt));
    unsigned* sums = reinterpret_cast<unsigned*>(take(bSum));
    unsigned char* small = take(bSmall);          // hdr[3] (0..23), ntok (24..31), fb_count (32..35), err_pos (40..47)
    int* fb_e = reinterpret_cast<int*>(take(bFe));
    double* fb_v = reinterpret_cast<double*>(take(bFv));
    char* fb_text = reinterpret_cast<char*>(take(bFt));
    MpOut o{reinterpret_cast<long long*>(small), d_ti, d_tj, d_tx, nz_cap, reinterpret_cast<unsigned long long*>(small + 40),
            reinterpret_cast<unsigned long long*>(small + 24), reinterpret_cast<unsigned*>(small + 32), fb_cap, fb_e, fb_text};
    if (cudaMemsetAsync(small, 0, 40, st) != cudaSuccess || cudaMemsetAsync(small + 40, 0xff, 8, st) != cudaSuccess) return fail(SPCHOL_ECUDA, "cudaMemsetAsync failed");
    const unsigned gc = static_cast<unsigned>((nchunks + 127) / 128);
    k_mp_count<<<gc, 128, 0, st>>>(d_text, nbytes, nchunks, counts);
    ingest_scan(nchunks, counts, counts, sums, st);
    k_mp_parse<<<gc, 128, 0, st>>>(d_text, nbytes, nchunks, counts, o);
    unsigned char hs[48];
    cudaMemcpyAsync(hs, small, 48, cudaMemcpyDeviceToHost, st);
    if (cudaStreamSynchronize(st) != cudaSuccess || cudaGetLastError() != cudaSuccess) return fail(SPCHOL_ECUDA, "kernel failure");
    long long hdr[3]; unsigned long long ntok = 0, err_pos = 0; unsigned fbn = 0;
    std::memcpy(hdr, hs, 24); std::memcpy(&ntok, hs + 24, 8); std::memcpy(&fbn, hs + 32, 4); std::memcpy(&err_pos, hs + 40, 8);
    if (ntok < 3 || hdr[0] <= 0 || hdr[0] != hdr[1] || hdr[2] < 0 || hdr[0] > INT_MAX) return fail(SPCHOL_EINVAL, "bad size line (matrix must be square)");
    const long long nl = hdr[2];
    if (nl > nz_cap) return fail(SPCHOL_ERANGE, "the file holds " + std::to_string(nl) + " entries, the buffers " + std::to_string(nz_cap));
    if (ntok < 3ull + 3ull * static_cast<unsigned long long>(nl)) return fail(SPCHOL_EIO, "truncated file");
    // a malformed token only matters among the 3 + 3 nl tokens the reference would read
    if (err_pos < 3ull + 3ull * static_cast<unsigned long long>(nl))
        return fail(SPCHOL_EIO, "token " + std::to_string(err_pos) + " of the data (" + (err_pos < 3 ? "size line" : "entry " + std::to_string((err_pos - 3) / 3)) + ") is not a number of the expected kind");
    if (fbn > fb_cap) return fail(SPCHOL_ERANGE, std::to_string(fbn) + " values need the host's strtod (at most " + std::to_string(fb_cap) + "): use spchol_mtx_read");
    if (fbn) {
        std::vector<int> he(fbn); std::vector<char> ht(static_cast<size_t>(fbn) * MP_TOK_MAX); std::vector<double> hv(fbn);
        if (cudaMemcpyAsync(he.data(), fb_e, sizeof(int) * fbn, cudaMemcpyDeviceToHost, st) != cudaSuccess ||
            cudaMemcpyAsync(ht.data(), fb_text, ht.size(), cudaMemcpyDeviceToHost, st) != cudaSuccess || cudaStreamSynchronize(st) != cudaSuccess)
            return fail(SPCHOL_ECUDA, "read-back of the tokens for the host failed");
        for (unsigned k = 0; k < fbn; ++k) {
            const char* t = ht.data() + static_cast<size_t>(k) * MP_TOK_MAX;
            if (*t == '+') ++t;
            char* endp = nullptr;
            hv[k] = std::strtod(t, &endp);
            if (endp == t || *endp != 0) return fail(SPCHOL_EIO, std::string("value '") + t + "' is not a number");
        }
        if (cudaMemcpyAsync(fb_v, hv.data(), sizeof(double) * fbn, cudaMemcpyHostToDevice, st) != cudaSuccess) return fail(SPCHOL_ECUDA, "cudaMemcpyAsync failed");
        k_mp_scatter<<<(fbn + 255) / 256, 256, 0, st>>>(fbn, fb_e, fb_v, d_tx);
    }
    if (nl > 0) k_mp_mirror<<<static_cast<unsigned>((nl + 255) / 256), 256, 0, st>>>(nl, d_ti, d_tj);
    if (cudaStreamSynchronize(st) != cudaSuccess || cudaGetLastError() != cudaSuccess) return fail(SPCHOL_ECUDA, "kernel failure");
    cudaFree(arena);
    *n_out = static_cast<int>(hdr[0]);
    *nz_out = nl;
    if (host_converted) *host_converted = fbn;
    return SPCHOL_OK;
}

// load_matrix_market_to_csc's reading half (include/mtx_reader.hpp:17-56) with the parse on the GPU: same two-call protocol
// and results as spchol_mtx_read (with ti == NULL only the sizes, from the host).  The file is read into host memory, the
// leading '%' lines are dropped, the rest goes to the device in one copy.  *device_ms (optional): the device part alone.
int spchol_mtx_read_gpu(const char* path, int* n_out, int64_t* nz_out, int* ti, int* tj, double* tx, double* device_ms) {
    if (!path || !n_out || !nz_out) { set_error("spchol_mtx_read_gpu: null argument"); return SPCHOL_EINVAL; }
    int rc = spchol_mtx_read(path, n_out, nz_out, nullptr, nullptr, nullptr);
    if (rc || !ti) return rc;
    if (!tj || !tx) { set_error("spchol_mtx_read_gpu: null argument"); return SPCHOL_EINVAL; }
    if (spchol_device_count() <= 0) { set_error("spchol_mtx_read_gpu: no CUDA device"); return SPCHOL_ENOGPU; }
    const int64_t nz = *nz_out;
    std::FILE* f = std::fopen(path, "rb");
    if (!f) { set_error(std::string("Could not open file ") + path); return SPCHOL_EIO; }
    std::fseek(f, 0, SEEK_END);
    const long sz = std::ftell(f);
    std::fseek(f, 0, SEEK_SET);
    std::vector<char> buf;
    if (sz < 0) { std::fclose(f); set_error("spchol_mtx_read_gpu: cannot size the file"); return SPCHOL_EIO; }
    try { buf.resize(static_cast<size_t>(sz)); } catch (const std::bad_alloc&) { std::fclose(f); set_error("out of host memory"); return SPCHOL_ENOMEM; }
    buf.resize(std::fread(buf.data(), 1, buf.size(), f));
    std::fclose(f);
    size_t skip = 0;
    while (skip < buf.size() && buf[skip] == '%') { while (skip < buf.size() && buf[skip] != '\n') ++skip; if (skip < buf.size()) ++skip; }
    const int64_t nbytes = static_cast<int64_t>(buf.size() - skip);
    if (nbytes < 1) { set_error("spchol_mtx_read_gpu: bad size line (matrix must be square)"); return SPCHOL_EINVAL; }
    char* d_text = nullptr; int *d_ti = nullptr, *d_tj = nullptr; double* d_tx = nullptr;
    const size_t z = static_cast<size_t>(std::max<int64_t>(nz, 1));
    auto freeall = [&] { cudaFree(d_text); cudaFree(d_ti); cudaFree(d_tj); cudaFree(d_tx); };
    if (cudaMalloc(&d_text, static_cast<size_t>(nbytes)) != cudaSuccess || cudaMalloc(&d_ti, 4 * z) != cudaSuccess ||
        cudaMalloc(&d_tj, 4 * z) != cudaSuccess || cudaMalloc(&d_tx, 8 * z) != cudaSuccess) {
        cudaGetLastError(); freeall(); set_error("spchol_mtx_read_gpu: out of device memory"); return SPCHOL_ENOMEM;
    }
    cudaMemcpy(d_text, buf.data() + skip, static_cast<size_t>(nbytes), cudaMemcpyHostToDevice);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0, nullptr);
    int n2 = 0; int64_t nz2 = 0;
    rc = spchol_mtx_parse_device(d_text, nbytes, nz, d_ti, d_tj, d_tx, &n2, &nz2, nullptr, nullptr);
    cudaEventRecord(e1, nullptr);
    cudaEventSynchronize(e1);
    float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    if (device_ms) *device_ms = ms;
    if (rc == 0 && (n2 != *n_out || nz2 != nz)) { rc = SPCHOL_EIO; set_error("spchol_mtx_read_gpu: size line read differently on the device"); }
    if (rc == 0 && nz > 0) {
        cudaMemcpy(ti, d_ti, 4 * static_cast<size_t>(nz), cudaMemcpyDeviceToHost);
        cudaMemcpy(tj, d_tj, 4 * static_cast<size_t>(nz), cudaMemcpyDeviceToHost);
        if (cudaMemcpy(tx, d_tx, 8 * static_cast<size_t>(nz), cudaMemcpyDeviceToHost) != cudaSuccess) { rc = SPCHOL_ECUDA; set_error("spchol_mtx_read_gpu: read-back failed"); }
    }
    freeall();
    return rc;
}

// ---- the device ingest without any allocation (no counterpart in the reference; added
// for callers that cannot call cudaMalloc: CUDA graph capture, tight serving loops, harnesses that forbid it).
// The work space of the sort is a file-scope device array of fixed size, so the entry point is limited to
// SPCHOL_NOALLOC_MAX_NZ triplets and SPCHOL_NOALLOC_MAX_N columns (include/spchol.h).  Every GPU that loads the library
// holds the array (36.3 MB) until the process exits.  All calls share it: it keeps the last call's contents (what must
// start at zero is cleared per call), and two calls that overlap on one GPU race.
}  // extern "C"
namespace {
constexpr int64_t NA_MAX_NZ = SPCHOL_NOALLOC_MAX_NZ;
constexpr int64_t NA_MAX_N = SPCHOL_NOALLOC_MAX_N;
constexpr size_t na_up(size_t b) { return (b + 255) & ~size_t(255); }
constexpr int64_t na_max3(int64_t a, int64_t b, int64_t c) { return a > b ? (a > c ? a : c) : (b > c ? b : c); }
// the arena of spchol_triplet_to_csc_device at the largest admitted problem (same carving, same order)
constexpr size_t na_scratch_bytes(int64_t nz, int64_t n) {
    const int64_t nwarps = (nz + IN_WARP_ITEMS - 1) / IN_WARP_ITEMS;
    const int64_t nh = 256 * nwarps;
    const int64_t nbs = (na_max3(nh, nz, n + 1) + IN_SCAN_BLOCK - 1) / IN_SCAN_BLOCK;
    return 2 * na_up(sizeof(unsigned long long) * nz) + 4 * na_up(sizeof(unsigned) * nz) + na_up(sizeof(unsigned) * nh) +
           na_up(sizeof(unsigned) * (n + 1)) + na_up(sizeof(unsigned) * nbs) + 256;
}
constexpr size_t NA_SCRATCH_BYTES = na_scratch_bytes(NA_MAX_NZ, NA_MAX_N);
static_assert(NA_SCRATCH_BYTES <= (size_t(40) << 20), "the static ingest scratch is meant to stay modest");
static __device__ __align__(256) unsigned char g_ingest_scratch[NA_SCRATCH_BYTES];

// In host code g_ingest_scratch is only a stand-in for the symbol: the device address comes from cudaGetSymbolAddress,
// once per device, with that device current.
constexpr int NA_MAX_DEVICES = 64;
unsigned char* g_ingest_scratch_addr[NA_MAX_DEVICES] = {};
std::mutex g_ingest_scratch_mutex;
int ingest_scratch_address(int dev, unsigned char** out) {
    if (dev < 0 || dev >= NA_MAX_DEVICES) { set_error("spchol_triplet_to_csc_device_noalloc: device index beyond 63"); return SPCHOL_EINVAL; }
    std::lock_guard<std::mutex> lock(g_ingest_scratch_mutex);
    if (!g_ingest_scratch_addr[dev]) {
        void* ptr = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&ptr, g_ingest_scratch);
        if (e != cudaSuccess || !ptr) {
            set_error(std::string("spchol_triplet_to_csc_device_noalloc: cudaGetSymbolAddress failed: ") + cudaGetErrorString(e));
            return SPCHOL_ECUDA;
        }
        g_ingest_scratch_addr[dev] = static_cast<unsigned char*>(ptr);
    }
    *out = g_ingest_scratch_addr[dev];
    return 0;
}
}  // namespace
extern "C" {

// Same arguments, same result (bit for bit) as spchol_triplet_to_csc_device; runs on the current device.  Calls no
// cudaMalloc / cudaFree; refuses nz > SPCHOL_NOALLOC_MAX_NZ or n > SPCHOL_NOALLOC_MAX_N with SPCHOL_ERANGE.
int spchol_triplet_to_csc_device_noalloc(int64_t nz, const int* d_ti, const int* d_tj, const double* d_tx, int n,
                                         int* d_Ap, int* d_Ai, double* d_Ax, int64_t* nnz_out, void* stream) {
    static const char* who = "spchol_triplet_to_csc_device_noalloc";
    if (nz < 0 || n < 1 || !d_Ap || !nnz_out || (nz > 0 && (!d_ti || !d_tj || !d_tx || !d_Ai || !d_Ax))) { set_error(std::string(who) + ": bad argument"); return SPCHOL_EINVAL; }
    if (nz > NA_MAX_NZ || n > NA_MAX_N) {
        set_error(std::string(who) + ": " + std::to_string(nz) + " triplets / " + std::to_string(n) + " columns exceed the static work space (at most " +
                  std::to_string(NA_MAX_NZ) + " triplets, " + std::to_string(NA_MAX_N) + " columns); use spchol_triplet_to_csc_device");
        return SPCHOL_ERANGE;
    }
    if (spchol_device_count() <= 0) { set_error(std::string(who) + ": no CUDA device"); return SPCHOL_ENOGPU; }
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    *nnz_out = 0;
    if (nz == 0) { CK(cudaMemsetAsync(d_Ap, 0, sizeof(int) * (static_cast<size_t>(n) + 1), st)); CK(cudaStreamSynchronize(st)); return 0; }
    int dev = 0, rc = 0;
    CK(cudaGetDevice(&dev));
    unsigned char* arena = nullptr;
    if ((rc = ingest_scratch_address(dev, &arena))) return rc;
    const int nwarps = static_cast<int>((nz + IN_WARP_ITEMS - 1) / IN_WARP_ITEMS);
    const int64_t nh = static_cast<int64_t>(256) * nwarps;
    const int64_t nscan = std::max<int64_t>(std::max<int64_t>(nh, nz), static_cast<int64_t>(n) + 1);
    const int nbs = static_cast<int>((nscan + IN_SCAN_BLOCK - 1) / IN_SCAN_BLOCK);
    const size_t bK = na_up(sizeof(unsigned long long) * nz), bI = na_up(sizeof(unsigned) * nz), bH = na_up(sizeof(unsigned) * nh),
                 bC = na_up(sizeof(unsigned) * (static_cast<size_t>(n) + 1)), bS = na_up(sizeof(unsigned) * nbs);
    if (2 * bK + 4 * bI + bH + bC + bS + 256 > NA_SCRATCH_BYTES) { set_error(std::string(who) + ": work space accounting is off"); return SPCHOL_ERANGE; }
    unsigned char* q = arena;
    auto take = [&](size_t b) { unsigned char* r = q; q += b; return r; };
    unsigned long long* key[2] = {reinterpret_cast<unsigned long long*>(take(bK)), reinterpret_cast<unsigned long long*>(take(bK))};
    unsigned* idx[2] = {reinterpret_cast<unsigned*>(take(bI)), reinterpret_cast<unsigned*>(take(bI))};
    unsigned* head = reinterpret_cast<unsigned*>(take(bI));
    unsigned* slot = reinterpret_cast<unsigned*>(take(bI));
    unsigned* hist = reinterpret_cast<unsigned*>(take(bH));
    unsigned* colc = reinterpret_cast<unsigned*>(take(bC));
    unsigned* sums = reinterpret_cast<unsigned*>(take(bS));
    int* bad = reinterpret_cast<int*>(take(256));
    // the array keeps what the last call left: the flag and the column counts start at zero, everything else is
    // written before it is read
    if (cudaMemsetAsync(bad, 0, sizeof(int), st) != cudaSuccess || cudaMemsetAsync(colc, 0, sizeof(unsigned) * (static_cast<size_t>(n) + 1), st) != cudaSuccess)
        { set_error(std::string(who) + ": cudaMemsetAsync failed"); return SPCHOL_ECUDA; }
    const unsigned gb = static_cast<unsigned>((nz + 255) / 256);
    k_in_keys<<<gb, 256, 0, st>>>(nz, d_ti, d_tj, n, key[0], idx[0], bad);
    int bits = 1;
    while ((int64_t(1) << bits) < n) ++bits;
    const int nbytes = (bits + 7) / 8;
    int cur = 0;
    for (int half = 0; half < 2; ++half)
        for (int b = 0; b < nbytes; ++b) {
            const int shift = 32 * half + 8 * b;
            k_in_hist<<<(nwarps + 7) / 8, 256, 0, st>>>(nz, key[cur], shift, nwarps, hist);
            ingest_scan(nh, hist, hist, sums, st);
            k_in_scatter<<<(nwarps + 7) / 8, 256, 0, st>>>(nz, key[cur], idx[cur], shift, nwarps, hist, key[cur ^ 1], idx[cur ^ 1]);
            cur ^= 1;
        }
    k_in_heads<<<gb, 256, 0, st>>>(nz, key[cur], head, colc);
    ingest_scan(nz, head, slot, sums, st);
    k_in_emit<<<gb, 256, 0, st>>>(nz, key[cur], idx[cur], head, slot, d_tx, d_Ai, d_Ax);
    unsigned last[2] = {0, 0};
    cudaMemcpyAsync(&last[0], slot + (nz - 1), sizeof(unsigned), cudaMemcpyDeviceToHost, st);
    cudaMemcpyAsync(&last[1], head + (nz - 1), sizeof(unsigned), cudaMemcpyDeviceToHost, st);
    int hbad = 0;
    cudaMemcpyAsync(&hbad, bad, sizeof(int), cudaMemcpyDeviceToHost, st);
    ingest_scan(static_cast<int64_t>(n) + 1, colc, colc, sums, st);
    if (cudaStreamSynchronize(st) != cudaSuccess || cudaGetLastError() != cudaSuccess) { set_error(std::string(who) + ": kernel failure"); return SPCHOL_ECUDA; }
    if (hbad) { set_error(std::string(who) + ": index outside [0, n)"); return SPCHOL_EINVAL; }
    const unsigned total = last[0] + last[1];
    k_in_ap<<<(n + 1 + 255) / 256, 256, 0, st>>>(n, colc, total, d_Ap);
    if (cudaStreamSynchronize(st) != cudaSuccess) { set_error(std::string(who) + ": kernel failure"); return SPCHOL_ECUDA; }
    *nnz_out = total;
    return 0;
}

int spchol_host_alloc(int64_t bytes, void** out) {
    if (!out || bytes <= 0) return SPCHOL_EINVAL;
    *out = nullptr;
    if (spchol_device_count() <= 0) { set_error("spchol_host_alloc: no CUDA device"); return SPCHOL_ENOGPU; }
    CK(cudaHostAlloc(out, static_cast<size_t>(bytes), cudaHostAllocDefault));
    return 0;
}
int spchol_host_free(void* ptr) {
    if (!ptr) return 0;
    CK(cudaFreeHost(ptr));
    return 0;
}

namespace {
__global__ void k_dmma_probe(double* out, int iters) {
    double c[8][2];
#pragma unroll
    for (int i = 0; i < 8; ++i) { c[i][0] = 0; c[i][1] = 0; }
    const double a = 1.0 + threadIdx.x * 1e-9, b = 1.0 - threadIdx.x * 1e-9;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) dmma884(c[i][0], c[i][1], a, b);
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += c[i][0] + c[i][1];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
}  // namespace

// FP64 tensor-pipe peak of this GPU: register-resident DMMA chains, best of 5 (TFLOP/s).
int spchol_fp64_peak(int device, double* tflops) {
    if (!tflops) return SPCHOL_EINVAL;
    if (spchol_device_count() <= 0) { set_error("spchol_fp64_peak: no CUDA device"); return SPCHOL_ENOGPU; }
    CK(cudaSetDevice(device));
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, device));
    const int blocks = prop.multiProcessorCount * 2, threads = 256, iters = 20000;
    double* out = nullptr;
    CK(cudaMalloc(&out, sizeof(double) * blocks * threads));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    float best = 1e30f;
    for (int r = 0; r < 7; ++r) {
        CK(cudaEventRecord(e0));
        k_dmma_probe<<<blocks, threads>>>(out, iters);
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        float ms = 0;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        if (r >= 2 && ms < best) best = ms;
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(out);
    *tflops = 2.0 * 8 * 8 * 4 * 8 * iters * static_cast<double>(threads / 32) * blocks / best * 1e-9;
    return 0;
}

int spchol_lbuf(spchol_plan* p, double** dL, int64_t* entries) {
    if (!p) return SPCHOL_EINVAL;
    if (dL) *dL = p->dL;
    if (entries) *entries = p->H.lbuf_entries;
    return 0;
}

int spchol_lbuf_to_host(spchol_plan* p, double* L_host) {
    if (!p || !L_host) return SPCHOL_EINVAL;
    CK(cudaSetDevice(p->device));
    CK(cudaMemcpyAsync(L_host, p->dL, static_cast<size_t>(p->H.lbuf_entries) * sizeof(double), cudaMemcpyDeviceToHost, p->stream));
    CK(cudaStreamSynchronize(p->stream));
    return 0;
}

int spchol_last_timing(const spchol_plan* p, double* ms4) {
    if (!p || !ms4) return SPCHOL_EINVAL;
    for (int i = 0; i < 4; ++i) ms4[i] = p->last_ms[i];
    return 0;
}

int spchol_set_profile(spchol_plan* p, int on) {
    if (!p) return SPCHOL_EINVAL;
    p->profile = on;
    for (int k = 0; k < 12; ++k) { p->kind_ms[k] = 0; p->kind_launches[k] = 0; }
    return 0;
}
int spchol_get_profile(const spchol_plan* p, double* kind_ms12, int64_t* kind_launches12, double* flops2) {
    if (!p) return SPCHOL_EINVAL;
    for (int k = 0; k < 12; ++k) { if (kind_ms12) kind_ms12[k] = p->kind_ms[k]; if (kind_launches12) kind_launches12[k] = p->kind_launches[k]; }
    if (flops2) { flops2[0] = p->H.gemm_flops_alg; flops2[1] = p->H.gemm_flops; }
    return 0;
}
// operand bytes of the update tasks of one factorization if every operand moved exactly once (the floor the DRAM traffic of
// the update kernel is compared with) and the number of update tasks
int spchol_plan_gemm_bytes(const spchol_plan* p, double* bytes_min, int64_t* tasks) {
    if (!p) return SPCHOL_EINVAL;
    if (bytes_min) *bytes_min = p->H.gemm_bytes_min;
    if (tasks) *tasks = static_cast<int64_t>(p->H.gemm_tasks.size());
    return 0;
}
// algorithmic flops of the update tasks by K class (K <= 128, <= 256, larger): profile slots 9, 10, 11
int spchol_get_gemm_classes(const spchol_plan* p, double* flops3) {
    if (!p || !flops3) return SPCHOL_EINVAL;
    for (int k = 0; k < 3; ++k) flops3[k] = p->H.gemm_flops_alg_k[k];
    return 0;
}

void* spchol_stream(spchol_plan* p) { return p ? static_cast<void*>(p->stream) : nullptr; }
int64_t spchol_last_launches(const spchol_plan* p) { return p ? p->last_launches : 0; }

static int residual_prepare(spchol_plan* p);
static int solve_prepare(spchol_plan* p) {
    if (p->solve_ready) return 0;
    int rc;
    if ((rc = residual_prepare(p))) return rc;       // global row indices of the fronts
    const HostPlan& H = p->H;
    std::vector<std::vector<int>> by_level(H.nlevels);
    for (int s = 0; s < H.ns; ++s) by_level[H.level[s]].push_back(s);
    std::vector<SBlock> blocks; std::vector<SChunk> upd; std::vector<SGat> gat;
    // multi-rank plans: a front is solved by the rank that holds it alone or leads its group (that rank has the whole panel
    // and the row list, ridx_off >= 0)
    const bool multi = H.world > 1;
    auto mine = [&](int s) { return !multi || H.ridx_off[s] >= 0; };
    p->solve_level_ptr.assign(H.nlevels + 1, 0);
    for (int l = 0; l < H.nlevels; ++l) {
        int maxJ = 0;
        for (int s : by_level[l]) if (mine(s)) maxJ = std::max(maxJ, (H.w[s] + SOLVE_NB - 1) / SOLVE_NB);
        for (int j = 0; j < maxJ; ++j) {
            spchol_plan::SolveStep st{static_cast<int>(blocks.size()), 0, static_cast<int>(upd.size()), 0, static_cast<int>(gat.size()), 0, 0};
            int mnb = 0;
            for (int s : by_level[l]) {
                const int c0 = j * SOLVE_NB;
                if (c0 >= H.w[s] || !mine(s)) continue;
                const int nb = std::min(SOLVE_NB, H.w[s] - c0), end = c0 + nb;
                mnb = std::max(mnb, nb);
                blocks.push_back({s, c0});
                for (int r0 = end; r0 < H.m[s]; r0 += 256) upd.push_back({s, c0, r0});
                if (end < H.m[s]) for (int cs = c0; cs < end; cs += 32) for (int i0 = end; i0 < H.m[s]; i0 += SOLVE_GROWS) gat.push_back({s, cs, end, i0});
            }
            st.bn = static_cast<int>(blocks.size()) - st.b0; st.un = static_cast<int>(upd.size()) - st.u0;
            st.gn = static_cast<int>(gat.size()) - st.g0; st.smem = solve_diag_smem(mnb);
            p->solve_steps.push_back(st);
        }
        p->solve_level_ptr[l + 1] = static_cast<int>(p->solve_steps.size());
    }
    if (multi) {
        // shared[s]: some descendant of s is solved by another rank -- its columns receive updates from (forward) and are read
        // by (backward) more than one rank
        std::vector<char> shared(H.ns, 0);
        for (int s = 0; s < H.ns; ++s) {                 // children before parents
            const int q = H.parent[s];
            if (q >= 0 && (shared[s] || H.own_lo[s] != H.own_lo[q])) shared[q] = 1;
        }
        std::vector<int> top;
        std::vector<unsigned char> keep(static_cast<size_t>(H.n), 0);
        p->solve_shared_level = H.nlevels;
        for (int s = 0; s < H.ns; ++s) {
            if (shared[s]) {
                p->solve_shared_level = std::min(p->solve_shared_level, H.level[s]);
                for (int c = 0; c < H.w[s]; ++c) top.push_back(H.first[s] + c);
            }
            if (H.own_lo[s] == H.rank) for (int c = 0; c < H.w[s]; ++c) keep[H.first[s] + c] = 1;
        }
        p->n_top = static_cast<int64_t>(top.size());
        if ((rc = upload(p, top, &p->d_top_idx)) || (rc = upload(p, keep, &p->d_solve_mine)) ||
            (rc = dalloc(p, std::max<int64_t>(p->n_top, 1) * SOLVE_NR, &p->d_top_buf))) return rc;
    }
    if ((rc = upload(p, blocks, &p->d_sblocks)) || (rc = upload(p, upd, &p->d_supd)) || (rc = upload(p, gat, &p->d_sgat))) return rc;
    CK(cudaFuncSetAttribute(k_solve_diag<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, solve_diag_smem(SOLVE_NB)));
    CK(cudaFuncSetAttribute(k_solve_diag<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, solve_diag_smem(SOLVE_NB)));
    p->solve_ready = true;
    return 0;
}

// x = A^-1 b with the factor the plan holds (last spchol_factor* call): L y = b, L^T x = y on the supernodal panels.
// b, x: n x nrhs column-major host arrays (may alias).
// Multi-rank plans: collective -- every rank calls with the same b and receives the whole x.  Every front is solved by ONE
// rank (the holder, or the leader of the group that factored it: it has the whole panel).  The subtrees below the split need
// no communication; from the first level with a front whose descendants sit on several ranks upwards, the ranks exchange
// the columns of those fronts once per level (k_top_gather, ncclAllReduce, k_top_store): updates summed on the way up,
// solved values handed out on the way down.  The pieces of x are summed at the end.
int spchol_solve(spchol_plan* p, const double* b, double* x, int nrhs) {
    if (!p || !b || !x || nrhs < 1) { set_error("spchol_solve: bad argument"); return SPCHOL_EINVAL; }
    const HostPlan& H = p->H;
    NcclApi* nc = H.world > 1 ? nccl_api() : nullptr;
    if (H.world > 1 && (!nc || !p->comm)) { set_error("spchol_solve: multi-rank plan without a communicator"); return SPCHOL_EINVAL; }
    CK(cudaSetDevice(p->device));
    int rc;
    if ((rc = solve_prepare(p))) return rc;
    cudaStream_t st = p->stream;
    const int64_t nv = static_cast<int64_t>(H.n) * SOLVE_NR;
    // work vectors live with the plan: [0, nv) right-hand sides / solution interleaved [n][SOLVE_NR], [nv, 2 nv) gather sums,
    // [2 nv, 3 nv) the caller's columns as they arrive / leave (column-major); packing and unpacking run on the device
    if (!p->d_solve_v && (rc = dalloc(p, 3 * nv, &p->d_solve_v))) return rc;
    double* dv = p->d_solve_v;
    double* stage = dv + 2 * nv;
    int64_t launches = 0;
    cudaError_t e = cudaSuccess;
    bool nccl_failed = false;
    const int64_t ne_top = p->n_top * SOLVE_NR;
    const unsigned top_grid = static_cast<unsigned>((ne_top + 255) / 256);
    auto exchange = [&](bool down) {
        if (!nc || ne_top == 0) return;
        // up: every copy of the shared columns is summed, the solving rank keeps the sum, the others start again from zero;
        // down: the value of every shared column from the rank that solves it, stored everywhere (solve.cuh)
        if (down) k_top_gather<true><<<top_grid, 256, 0, st>>>(ne_top, p->d_top_idx, p->d_solve_mine, dv, p->d_top_buf);
        else k_top_gather<false><<<top_grid, 256, 0, st>>>(ne_top, p->d_top_idx, p->d_solve_mine, dv, p->d_top_buf);
        if (nc->AllReduce(p->d_top_buf, p->d_top_buf, static_cast<size_t>(ne_top), kNcclFloat64, kNcclSum, p->comm->world_comm, st) != 0)
            nccl_failed = true;
        if (down) k_top_store<false><<<top_grid, 256, 0, st>>>(ne_top, p->d_top_idx, p->d_solve_mine, dv, p->d_top_buf);
        else k_top_store<true><<<top_grid, 256, 0, st>>>(ne_top, p->d_top_idx, p->d_solve_mine, dv, p->d_top_buf);
        launches += 2;
    };
    for (int k0 = 0; k0 < nrhs && e == cudaSuccess && !nccl_failed; k0 += SOLVE_NR) {
        const int nr = std::min(SOLVE_NR, nrhs - k0);
        cudaMemcpyAsync(stage, b + static_cast<int64_t>(k0) * H.n, sizeof(double) * H.n * nr, cudaMemcpyHostToDevice, st);
        k_solve_pack<<<static_cast<unsigned>((nv + 255) / 256), 256, 0, st>>>(H.n, nr, stage, dv, dv + nv);
        if (nc && ne_top) { k_top_clear_others<<<top_grid, 256, 0, st>>>(ne_top, p->d_top_idx, p->d_solve_mine, dv); ++launches; }
        for (int l = 0; l < H.nlevels; ++l) {            // L y = b
            if (nc && l >= p->solve_shared_level) exchange(false);
            for (int q = p->solve_level_ptr[l]; q < p->solve_level_ptr[l + 1]; ++q) {
                const auto& sp = p->solve_steps[q];
                if (!sp.bn) continue;
                k_solve_diag<false><<<sp.bn, 256, sp.smem, st>>>(p->d_sblocks + sp.b0, p->fa_r, p->dL, dv, nullptr);
                if (sp.un) k_solve_update<<<sp.un, 256, 0, st>>>(p->d_supd + sp.u0, p->fa_r, p->d_rowidx, p->dL, dv);
                launches += 1 + (sp.un ? 1 : 0);
            }
        }
        for (int l = H.nlevels - 1; l >= 0; --l) {       // L^T x = y
            for (int q = p->solve_level_ptr[l + 1] - 1; q >= p->solve_level_ptr[l]; --q) {
                const auto& sp = p->solve_steps[q];
                if (!sp.bn) continue;
                if (sp.gn) k_solve_gather<<<sp.gn, 256, 0, st>>>(p->d_sgat + sp.g0, p->fa_r, p->d_rowidx, p->dL, dv, dv + nv);
                k_solve_diag<true><<<sp.bn, 256, sp.smem, st>>>(p->d_sblocks + sp.b0, p->fa_r, p->dL, dv, dv + nv);
                launches += 1 + (sp.gn ? 1 : 0);
            }
            if (nc && l >= p->solve_shared_level) exchange(true);
        }
        if (nc) {                                        // every row of x was computed by exactly one rank
            k_solve_mask<<<static_cast<unsigned>((nv + 255) / 256), 256, 0, st>>>(nv, p->d_solve_mine, dv);
            if (nc->AllReduce(dv, dv, static_cast<size_t>(nv), kNcclFloat64, kNcclSum, p->comm->world_comm, st) != 0) nccl_failed = true;
            ++launches;
        }
        k_solve_unpack<<<static_cast<unsigned>((static_cast<int64_t>(H.n) * nr + 255) / 256), 256, 0, st>>>(H.n, nr, dv, stage);
        cudaMemcpyAsync(x + static_cast<int64_t>(k0) * H.n, stage, sizeof(double) * H.n * nr, cudaMemcpyDeviceToHost, st);
        launches += 2;
        e = cudaStreamSynchronize(st);
        if (e == cudaSuccess) e = cudaGetLastError();
    }
    p->last_launches = launches;
    if (nccl_failed) { set_error("spchol_solve: ncclAllReduce failed"); return SPCHOL_ENCCL; }
    if (e != cudaSuccess) { set_error(std::string("spchol_solve: ") + cudaGetErrorString(e)); return SPCHOL_ECUDA; }
    return SPCHOL_OK;
}

// ||A - L L^T||_F / ||A||_F of the factor the plan holds (last spchol_factor* call), estimated with
// NPROBE Rademacher probes on the device (residual.cuh); Ax_host = the values that were factored.
static int residual_prepare(spchol_plan* p) {
    if (p->d_rowidx) return 0;
    const HostPlan& H = p->H;
    int rc;
    std::vector<RTask> tc, tr;
    if (H.world > 1) {
        // the rows come from the host plan; only the fronts this rank multiplies (ridx_off >= 0) get tasks
        int64_t* d_off = nullptr;
        for (int s = 0; s < H.ns; ++s) {
            if (H.ridx_off[s] < 0) continue;
            for (int c0 = 0; c0 < H.w[s]; c0 += 32) tc.push_back({s, c0});
            for (int r0 = 0; r0 < H.m[s]; r0 += 256) tr.push_back({s, r0});
        }
        if ((rc = upload(p, H.ridx, &p->d_rowidx)) || (rc = upload(p, H.ridx_off, &d_off))) return rc;
        p->fa_r = p->fa; p->fa_r.rel_off = d_off;
        std::vector<int>().swap(p->H.ridx);
    } else {
        std::vector<int> order(H.ns);
        p->level_ptr.assign(H.nlevels + 1, 0);
        for (int s = 0; s < H.ns; ++s) ++p->level_ptr[H.level[s] + 1];
        for (int l = 0; l < H.nlevels; ++l) p->level_ptr[l + 1] += p->level_ptr[l];
        { std::vector<int> cur(p->level_ptr.begin(), p->level_ptr.end() - 1);
          for (int s = 0; s < H.ns; ++s) order[cur[H.level[s]]++] = s; }
        for (int s = 0; s < H.ns; ++s) {
            for (int c0 = 0; c0 < H.w[s]; c0 += 32) tc.push_back({s, c0});
            for (int r0 = 0; r0 < H.m[s]; r0 += 256) tr.push_back({s, r0});
        }
        if ((rc = upload(p, order, &p->d_level_list)) || (rc = dalloc(p, H.rel_off[H.ns], &p->d_rowidx))) return rc;
        p->fa_r = p->fa;
        for (int l = H.nlevels - 1; l >= 0; --l) {           // parents before children
            const int cnt = p->level_ptr[l + 1] - p->level_ptr[l];
            if (cnt) k_rowidx_level<<<std::min(cnt, 4096), 128, 0, p->stream>>>(p->d_level_list + p->level_ptr[l], cnt, p->fa, p->d_rowidx);
        }
    }
    p->n_rt_cols = static_cast<int>(tc.size()); p->n_rt_rows = static_cast<int>(tr.size());
    if ((rc = upload(p, tc, &p->d_rt_cols)) || (rc = upload(p, tr, &p->d_rt_rows)) ||
        (rc = upload(p, p->H.a_row, &p->d_a_row)) || (rc = upload(p, p->H.a_col, &p->d_a_col))) return rc;
    std::vector<int>().swap(p->H.a_row); std::vector<int>().swap(p->H.a_col);
    CK(cudaGetLastError());
    return 0;
}

int spchol_residual(spchol_plan* p, const double* Ax, double* rel_fro) {
    if (!p || !Ax || !rel_fro) { set_error("spchol_residual: null argument"); return SPCHOL_EINVAL; }
    const HostPlan& H = p->H;
    CK(cudaSetDevice(p->device));
    int rc;
    if ((rc = residual_prepare(p))) return rc;
    // multi-rank plans (collective: every rank calls with the same Ax): each rank multiplies the panels of the fronts it
    // holds alone or leads, the probe products y = L^T z and u = L y are summed over the ranks (every column of L belongs
    // to exactly one multiplying rank), r = A z is formed by everybody -- the estimate is identical on all ranks
    NcclApi* nc = H.world > 1 ? nccl_api() : nullptr;
    if (H.world > 1 && (!nc || !p->comm)) { set_error("spchol_residual: multi-rank plan without a communicator"); return SPCHOL_EINVAL; }
    cudaStream_t st = p->stream;
    const int64_t nv = static_cast<int64_t>(H.n) * NPROBE;
    double *z = nullptr, *dA = nullptr;                  // z, y, u, r, 2 scalars
    CK(cudaMalloc(&z, sizeof(double) * (4 * nv + 2)));
    if (cudaMalloc(&dA, sizeof(double) * std::max<int64_t>(H.nnzA, 1)) != cudaSuccess) { cudaFree(z); set_error("spchol_residual: out of device memory"); return SPCHOL_ECUDA; }
    double *y = z + nv, *u = y + nv, *r = u + nv, *scal = r + nv;
    cudaMemcpyAsync(dA, Ax, sizeof(double) * H.nnzA, cudaMemcpyHostToDevice, st);
    cudaMemsetAsync(scal, 0, 2 * sizeof(double), st);
    k_probe_init<<<static_cast<unsigned>((nv + 255) / 256), 256, 0, st>>>(H.n, z, y, u, r);
    if (p->n_rt_cols) k_lt_mult<<<p->n_rt_cols, 256, 0, st>>>(p->d_rt_cols, p->fa_r, p->d_rowidx, p->dL, z, y);
    if (nc && nc->AllReduce(y, y, static_cast<size_t>(nv), kNcclFloat64, kNcclSum, p->comm->world_comm, st) != 0) {
        cudaFree(z); cudaFree(dA); set_error("spchol_residual: ncclAllReduce failed"); return SPCHOL_ENCCL; }
    if (p->n_rt_rows) k_l_mult<<<p->n_rt_rows, 256, 0, st>>>(p->d_rt_rows, p->fa_r, p->d_rowidx, p->dL, y, u);
    if (nc && nc->AllReduce(u, u, static_cast<size_t>(nv), kNcclFloat64, kNcclSum, p->comm->world_comm, st) != 0) {
        cudaFree(z); cudaFree(dA); set_error("spchol_residual: ncclAllReduce failed"); return SPCHOL_ENCCL; }
    k_sym_mult<<<static_cast<unsigned>((H.nnzA + 255) / 256), 256, 0, st>>>(H.nnzA, p->d_a_row, p->d_a_col, dA, z, r, scal);
    k_diff_norm<<<1184, 256, 0, st>>>(nv, r, u, scal);
    double h[2] = {0, 0};
    cudaMemcpyAsync(h, scal, sizeof(h), cudaMemcpyDeviceToHost, st);
    const cudaError_t e = cudaStreamSynchronize(st);
    cudaFree(z); cudaFree(dA);
    if (e != cudaSuccess) { set_error(std::string("spchol_residual: ") + cudaGetErrorString(e)); return SPCHOL_ECUDA; }
    *rel_fro = h[1] > 0 ? std::sqrt(h[0] / NPROBE / h[1]) : 0.0;
    return SPCHOL_OK;
}

}  // extern "C"
