// Numeric plan: the symbolic result flattened into supernode ("front") descriptors, relative
// index maps, a level schedule and a static launch program.  Built once on the host
// (plan.cpp), uploaded once (engine.cu), reused by every numeric factorization with the same
// pattern -- this is the state `chol(A, S)` (reference README.md:26-27) shares through S.
#pragma once
#include <cstdint>
#include <vector>

#include "symbolic.h"
#include "tree_plan.h"

namespace spchol {

// One dense C -= A * B^T task (all operands column-major, offsets in doubles).
struct GemmTask {
    int64_t c_off, a_off, b_off;
    int ldc, lda, ldb;
    int M, N, K;
    int diag;        // element (i,j) of C is wanted iff i + diag >= j   (lower-trapezoid mask)
    int c_space;     // bit 0: C lives in the L buffer (0) / the contribution workspace (1); bit 1: C = -A B^T (the target
                     // holds no data yet: Schur complement into a fresh contribution block) instead of C -= A B^T
    // TMA view of the operands (k_gemm_tma): A and B are row ranges of ONE front's panel, described by that front's tensor
    // map (HostPlan::tmap_fronts[tmap]: m rows x w columns, leading dimension ld); A starts at row ra, B at row rb, both at column kc
    int tmap, ra, rb, kc;
};
struct PotrfTask { int64_t d_off; int ld; int nb; int col0; int slot; };         // diag block in L buffer; slot: scratch for its inverse
struct TrsmTask  { int64_t b_off; int ld; int nb; int rows; int64_t d_off; int slot; };  // rows below the block
struct ZeroTask  { int64_t off; int64_t count; };
// One stream-ordered transfer between ranks (multi-GPU plans only).  Broadcasts run inside the
// contiguous rank group [lo, lo + n); send/recv are point to point.  Ranks that do not take part
// do not get the task at all, so the sequence of tasks is consistent on every participant.
enum CommKind : int { COMM_BCAST = 0, COMM_SEND = 1, COMM_RECV = 2 };
struct CommTask {
    int kind;
    int lo, n;          // COMM_BCAST: rank group
    int root;           // COMM_BCAST: absolute root rank; COMM_SEND / COMM_RECV: absolute peer rank
    int space;          // 0: L buffer, 1: contribution workspace
    int64_t off, count; // doubles
};

// Strided rectangle copy (rows x cols, column-major) between / inside the two device spaces:
// packs a block column into a contiguous transfer buffer and unpacks it on the receiver, so that
// only the rows below the diagonal travel.
struct CopyTask {
    int64_t src_off, dst_off;
    int src_space, dst_space;   // 0: L buffer, 1: contribution workspace
    int lds, ldd;
    int rows, cols;
};

// The program runs on two streams: stream 0 carries everything by default; with look-ahead the
// panel work of a large front (diagonal blocks, panel solves, inner updates, and for a front
// factored by a rank group the packing / broadcast / unpacking of the factored block column) runs
// on stream 1 while stream 0 is still applying the previous panel to the rest of the front.
// OP_EV_RECORD / OP_EV_WAIT order the two (begin = event id, unique per record; a wait refers to
// the record with the same id, which precedes it in program order).  Program order is itself a
// valid serial order, so a single-stream replay (profiling, the host simulator) is correct.
enum OpKind : int { OP_SMALL = 0, OP_ZERO = 1, OP_EXTADD = 2, OP_POTRF = 3, OP_TRSM = 4, OP_GEMM = 5, OP_COMM = 6, OP_COPY = 7,
                    OP_EV_RECORD = 8, OP_EV_WAIT = 9,
                    OP_EXPORT = 10,     // ranges of the L buffer that are final from here on (read-back of wide fronts, see ExportTask)
                    OP_TREE = 11 };     // the warp-per-subtree program (HostPlan::tree, tree_plan.h): begin = tier (0: HostPlan::tree, 1: tree2), count = tasks
constexpr int kEventRing = 64;      // the engine keeps this many CUDA events; event id -> slot id % kEventRing
struct Op {
    int kind;
    int begin, count;   // slice of the task array of that kind (OP_EV_*: begin = event id)
    int gx, gy;         // grid extents besides the task dimension
    int threads;        // OP_EXTADD: 0 = every column of the child block, 1 = only those landing in the parent's panel,
                        // 2 = only those landing in the parent's contribution block; OP_SMALL: CTA size class; OP_POTRF / OP_TRSM: block-width class (32 / 64 / 128); OP_GEMM: largest K
    int smem;           // OP_SMALL: dynamic shared memory bytes
    int level;
    int stream;         // 0 or 1
};

// A block column of a wide front is final as soon as its panel phase has run, long before the front (let alone its
// level) is complete: spchol_factor_lbuf starts its read-back there, so that the copy of the top fronts (2.2 GB for
// the root of 128^3) overlaps their own trailing updates instead of trailing the factorization.
struct ExportTask { int64_t off, count; };

struct PlanOptions {
    int small_max_m = 128;     // fronts with m <= this are factored whole in shared memory (160 -> 128: 2D 1024^2 9.4 -> 9.0 ms)
    int nb = 128;              // diagonal block of the blocked panel factorization (<= 128, the potrf kernel's limit)
    int NB = 256;              // outer block (K of the deferred panel update)
    int NB_wide = 512;         // outer block of the levels whose widest front has >= wide_w columns: the deferred
    int wide_w = 6144;         //   updates run at a longer K (measured 29 vs 23.6 TFLOP/s); narrower fronts keep NB
                               //   because their panel phase would no longer hide behind the trailing update
    int want_csc = 1;          // build the CSC export map (needs nnz(L) <= INT_MAX)
    int defer_cb_add = 1;      // 1: large fronts never zero their contribution block: the Schur complement is stored (C = -A B^T)
                               // and the children's contributions to it are extend-added afterwards
    int lookahead = 1;         // 1: panel work of large fronts on stream 1, overlapped with the trailing update (see Op)
    int tree = 1;              // 1: the bottom of the tree goes to the warp-per-subtree kernel (tree_plan.h) instead of one CTA per front and level
    int tree_cap_m = 48;       // largest front inside a subtree task (48 rows: 9.4 KB packed; the teams are register-limited to 16 per SM).
                               // Measured on 2D 1024^2 (small-front time, final kernel): 2.71 ms at 48 rows / 1 warp, 2.97 at 64 / 1, 3.05 at 64 / 2,
                               // 3.03 ms without subtree tasks
    double tree_work = 2.5e6;  // dense flops of a subtree task (the unit of load balance)
    int tree2_cap_m = 128;     // second tier of subtree tasks on top of the first (4-warp teams; 0: none): fronts up to this many rows
    double tree2_work = 3e7;   // dense flops of a second-tier task
    int batch_group_m = 117;   // batched whole-tree program: fronts are amalgamated into groups of up to this many rows (tree_plan.h,
                               // build_group_program; 117 rows = 55 KB packed = four teams per SM); 0: one step per front of the plan
    int export_mark_w = 1024;  // > 0: fronts at least this wide announce every finished block column (OP_EXPORT): spchol_factor_lbuf reads
                               // them back while the front's trailing update still runs (128^3 e2e 1146 -> 1073 ms); 0: read-back by levels only
    // multi-GPU (one plan per rank, every rank builds the same global mapping):
    int rank = 0, world = 1;
    int dist_NB = 256;         // block-column width of a front that is factored by a rank group
    int dist_min_w = 1024;     // fronts narrower than this stay on one rank even if their subtree was split
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
    // multi-GPU mapping (global, identical on every rank): front s is held by ranks
    // [own_lo[s], own_lo[s] + own_n[s]); own_n > 1 means block columns of width dist_NB are dealt
    // round-robin to the group (block b of the front -> rank own_lo + b % own_n; the blocks of
    // the contribution block continue the count after the panel's blocks).
    int rank = 0, world = 1, dist_NB = 256;
    std::vector<int> own_lo, own_n;
    std::vector<int> own_idx;                          // this rank's index in the group, -1 if not a holder
    std::vector<double> rank_load;                     // estimated flops per rank (mapping statistic)
    int64_t lbuf_entries = 0, ws_entries = 0;
    // maps
    std::vector<int64_t> a_dest;     // nnzA  : L-buffer offset of each stored A entry
    std::vector<int64_t> l_src;      // nnzL  : L-buffer offset of each CSC entry of L (want_csc)
    std::vector<int> a_row, a_col;   // nnzA  : coordinates of the stored entries of triu(A) (residual check)
    // multi-rank plans: global row index of every below-row of the fronts whose panel THIS rank multiplies in the residual
    // check (fronts it holds alone, and group fronts where it is the group's first rank), ridx_off[s] = -1 elsewhere.
    // (Single-rank plans rebuild the same thing on the device from the relative indices.)
    std::vector<int64_t> ridx_off;
    std::vector<int> ridx;
    // schedule
    std::vector<int> small_list;
    std::vector<int> extadd_list;    // child front ids
    std::vector<ZeroTask> zero_tasks;
    std::vector<PotrfTask> potrf_tasks;
    std::vector<TrsmTask> trsm_tasks;
    std::vector<GemmTask> gemm_tasks;
    std::vector<int> tmap_fronts;    // fronts whose panel needs a TMA tensor map (GemmTask::tmap indexes this list)
    std::vector<CommTask> comm_tasks;
    std::vector<CopyTask> copy_tasks;
    std::vector<ExportTask> export_tasks;
    std::vector<char> front_marked;  // per front: 1 if all of its panel is announced by OP_EXPORT ops (not by its level)
    std::vector<Op> ops;
    TreeProgram tree;                // subtrees at the bottom of the tree, one warp each (single-rank plans; empty: none)
    TreeProgram tree2;               // second tier: subtrees of mid-size fronts on top of `tree` (their leaves read the first tier's blocks)
    TreeProgram tree_whole;          // the whole tree as ONE task (batched factorization; only when every front has m <= TW_MAX_M)
    int n_events = 0;                // event ids handed out (OP_EV_RECORD count)
    int max_potrf_slots = 0;
    // statistics
    int64_t n_small = 0, n_large = 0, n_tree = 0;     // fronts on the one-CTA-per-front path, the blocked path, inside subtree tasks
    int max_m = 0, max_w = 0;
    double gemm_flops = 0;           // flops of the full M x N x K rectangles of OP_GEMM tasks
    double gemm_flops_alg = 0;       // flops of the wanted (lower-trapezoid) part only
    double gemm_flops_alg_k[3] = {0, 0, 0};   // the same split by K class of the task (gemm_k_class)
    double gemm_bytes_min = 0;       // operand bytes of the update tasks if every operand moved once: (M + N) K read, wanted C entries read + written (stored only: written)
    int64_t n_dist = 0;              // fronts factored by a rank group
    int64_t comm_doubles = 0;        // doubles this rank sends or receives per factorization
    bool holds(int s) const { return own_idx[s] >= 0; }
};

// K classes of the update tasks (profiling only): inner block updates, deferred panel updates, Schur complements
inline int gemm_k_class(int K) { return K <= 128 ? 0 : (K <= 256 ? 1 : 2); }

// proportional mapping of the assembly tree onto `world` ranks (deterministic; used by build_plan)
void map_fronts(int ns, const int* parent, const int* w, const int* m, int world, int dist_min_w,
                std::vector<int>& own_lo, std::vector<int>& own_n, std::vector<double>& rank_load);

void build_plan(const Symbolic& S, const PlanOptions& opt, HostPlan& P);

}  // namespace spchol
