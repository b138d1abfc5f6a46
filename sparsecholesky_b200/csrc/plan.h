// Numeric plan: the symbolic result flattened into supernode ("front") descriptors, relative
// index maps, a level schedule and a static launch program.  Built once on the host
// (plan.cpp), uploaded once (engine.cu), reused by every numeric factorization with the same
// pattern -- this is the state `chol(A, S)` (reference README.md:26-27) shares through S.
#pragma once
#include <cstdint>
#include <vector>

#include "symbolic.h"

namespace spchol {

// One dense C -= A * B^T task (all operands column-major, offsets in doubles).
struct GemmTask {
    int64_t c_off, a_off, b_off;
    int ldc, lda, ldb;
    int M, N, K;
    int diag;        // element (i,j) of C is wanted iff i + diag >= j   (lower-trapezoid mask)
    int c_space;     // 0: C lives in the L buffer, 1: in the contribution workspace
};
struct PotrfTask { int64_t d_off; int ld; int nb; int col0; int slot; };         // diag block in L buffer
struct TrsmTask  { int64_t b_off; int ld; int nb; int rows; int64_t d_off; int slot; };  // rows below the block
struct ZeroTask  { int64_t off; int64_t count; };

enum OpKind : int { OP_SMALL = 0, OP_ZERO = 1, OP_EXTADD = 2, OP_POTRF = 3, OP_TRSM = 4, OP_GEMM = 5 };
struct Op {
    int kind;
    int begin, count;   // slice of the task array of that kind
    int gx, gy;         // grid extents besides the task dimension
    int threads;        // OP_SMALL: CTA size class
    int smem;           // OP_SMALL: dynamic shared memory bytes
    int level;
};

struct PlanOptions {
    int small_max_m = 160;     // fronts with m <= this are factored whole in shared memory
    int nb = 64;               // diagonal block of the blocked panel factorization
    int NB = 256;              // outer block (K of the deferred panel update)
    int want_csc = 1;          // build the CSC export map (needs nnz(L) <= INT_MAX)
    // owner filter for multi-GPU: only fronts with owner[s] == rank are scheduled (nullptr: all)
    const int* owner = nullptr;
    int rank = 0;
};

struct HostPlan {
    int n = 0, ns = 0, nlevels = 0;
    int64_t nnzA = 0, nnzL = 0;
    double flops = 0;
    // per front
    std::vector<int> first, w, m, parent, level;
    std::vector<int64_t> panel_off, cb_off, rel_off;   // cb_off = -1 when m == w
    std::vector<int> ld;                               // leading dimension of the panel (>= m, even)
    std::vector<int> rel;                              // per below-row position in the parent front
    std::vector<int> child_ptr, child_idx;
    int64_t lbuf_entries = 0, ws_entries = 0;
    // maps
    std::vector<int64_t> a_dest;     // nnzA  : L-buffer offset of each stored A entry
    std::vector<int64_t> l_src;      // nnzL  : L-buffer offset of each CSC entry of L (want_csc)
    // schedule
    std::vector<int> small_list;
    std::vector<int> extadd_list;    // child front ids
    std::vector<ZeroTask> zero_tasks;
    std::vector<PotrfTask> potrf_tasks;
    std::vector<TrsmTask> trsm_tasks;
    std::vector<GemmTask> gemm_tasks;
    std::vector<Op> ops;
    int max_potrf_slots = 0;
    // statistics
    int64_t n_small = 0, n_large = 0;
    int max_m = 0, max_w = 0;
    double gemm_flops = 0;           // flops of the full M x N x K rectangles of OP_GEMM tasks
    double gemm_flops_alg = 0;       // flops of the wanted (lower-trapezoid) part only
};

void build_plan(const Symbolic& S, const PlanOptions& opt, HostPlan& P);

}  // namespace spchol
