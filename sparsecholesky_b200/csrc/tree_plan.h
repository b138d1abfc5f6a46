// Team-per-subtree program (host planner) for k_tree_team (tree_kernel.cuh).
//
// ONE small CTA (a team of 1..4 warps) factors a whole subtree of the assembly tree, front after front
// in postorder, with the frontal matrix in shared memory; parallelism comes from the number of
// subtrees (single matrix: the bottom of the tree, where 2D 1024^2 has 145 000 fronts with m <= 128)
// or from the number of matrices (batched factorization over one pattern, BASELINE config 5: one
// team per matrix walks the whole tree).  This is the fused dense chain of the reference's
// factorize_sn + apply_update (include/chol.hpp:1232-1335, :1196-1216) without ever leaving the SM.
//
// Layout of a front inside the warp's shared memory: packed lower triangle over REVERSED indices
// (i' = m-1-i): entry (i, j), i >= j, sits at tri(j') + i' with tri(x) = x (x + 1) / 2.  The position
// depends on (i', j') only, not on m, so after the w pivot columns (the highest reversed indices)
// have been eliminated, the contribution block is exactly the standard packed triangle of order
// m - w at the start of the buffer: handing it to the parent costs nothing when the parent comes
// next in the program ("carried"); the parent's rows are then spread in place by a monotone map
// (new position >= old position, processed from the top down).  Contribution blocks of the other
// children wait in a per-task stack in HBM/L2 ("stash"), in the same packed form.
#pragma once
#include <cstdint>
#include <vector>

namespace spchol {

struct HostPlan;

constexpr int TW_MAX_M = 128;                 // rows of the largest front a program may hold (static position map)
enum : int { TW_CARRY_IN = 1,                 // the front's first child left its block in shared memory
             TW_CARRY_OUT = 2,                // this front's block stays in shared memory for the next front (its parent)
             TW_CB_STD = 4,                   // block goes to the level program's workspace slot (column-major, ld = even(mb))
             TW_GROUP = 8 };                  // an amalgamated group of fronts (batched whole-tree program): columns are not one
                                              // contiguous range -- `ld` indexes gcol[], (pad0, pad1) = (first, count) of its output list

struct alignas(16) TwFront {                  // 80 bytes; the kernel keeps two descriptors ahead in registers
    int w, m, ld, first;                      // columns, rows, leading dimension of the panel in the L buffer, first global column
    int flags, mc, map_off, nstash;           // carried child: order of its block; rmap[map_off .. +m) is the INVERSE map of the spread
                                              // (source reversed row of every reversed row of this front, -1: a new row); stashed children to add
    int stash0, a0, an, ident;                // stashes[stash0 .. +nstash); A entries a_src/a_dst[a0 .. +an); carried child: its first
                                              // `ident` reversed rows keep their place (the map is the identity there)
    int map_lo, map_n, pad0, pad1;            // rmap[map_lo .. +map_n): every map this front reads (prefetched one front ahead)
    int64_t panel_off;                        // panel in the L buffer
    int64_t cb_out;                           // TW_CB_STD: workspace offset; else (no carry, mb > 0): offset in the stash area
};
struct alignas(16) TwStash { int mc, map_off; int64_t src; };     // a stashed child block and its (forward) map into the parent;
                                                                  // mc < 0: a block of order -mc in the level program's workspace
                                                                  // (column-major lower, ld = even(-mc)) left by an earlier tier
struct TwTask { int front0, nfront; int64_t stash_base; };         // stash_base: start of the task's private stack in the stash area

struct TreeProgram {
    int cap_m = 0;                            // largest front (shared memory per warp = tri(cap_m) doubles)
    int64_t stash_entries = 0;                // doubles of the stash area (all tasks)
    std::vector<TwTask> tasks;
    std::vector<TwFront> fronts;
    std::vector<TwStash> stashes;
    std::vector<int> rmap;                    // maps in reversed indices: forward (child row -> parent row) for stashed blocks, inverse for carried ones
    std::vector<int> a_src;                   // index into Ax
    std::vector<unsigned short> a_dst;        // packed reversed index inside the front
    std::vector<unsigned short> out_src;      // groups: packed position in F of every TRUE entry of L of the group's columns ...
    std::vector<int> out_dst;                 // ... and its index in the CSC values of L (ascending per group)
    std::vector<int> gcol;                    // groups: global column of every column of a group
    int n_groups = 0;                         // 0: the fronts of the program are the fronts of the plan
    double padded_flops = 0;                  // groups: dense flops including the explicit zeros of the amalgamation
    std::vector<char> in_tree;                // per front of the plan: 1 if a task of this program factors it
    double flops = 0;                         // dense flops of the fronts covered
    int64_t n_fronts() const { return static_cast<int64_t>(fronts.size()); }
};

// The subtrees a single-matrix plan hands to k_tree_warp: maximal subtrees whose fronts all have m <= cap_m rows and
// whose dense flops stay below work_limit (tasks are the unit of load balance).  in_tree[s] marks the fronts covered;
// returns the roots (ascending).  Subtrees with fewer than min_fronts fronts are not worth a task of their own only if
// they are single fronts whose parent is not covered either -- those stay with the level program (min_fronts = 1 keeps all).
// below (optional): fronts an earlier tier of subtree tasks factors -- they are left out of the forest, so the subtrees chosen
// here sit on top of them (their blocks come in through the level program's workspace).
std::vector<int> choose_subtrees(const HostPlan& P, int cap_m, double work_limit, std::vector<char>& in_tree,
                                 const std::vector<char>* below = nullptr);

// Builds the program: one task per root of `roots` (whole = false), or ONE task that walks every tree of the forest
// (whole = true; roots ignored) -- the batched factorization.  a_front/a_col/a_pos: front, column inside the front and
// row position inside the front of every stored entry of A (from build_plan).  std_cb: blocks of task roots go to
// P.cb_off (the level program consumes them).
void build_tree_program(const HostPlan& P, const std::vector<int>& roots, bool whole, const std::vector<int>& a_front,
                        const std::vector<int>& a_col, const std::vector<int>& a_pos, TreeProgram& out,
                        const std::vector<char>* below = nullptr);

// The batched whole-tree program over AMALGAMATED GROUPS: a group is a connected piece of the assembly tree (a top front
// plus descendants, grown bottom-up while its row set U = columns of the members ++ rows below the top stays within cap_m
// rows) that is factored as ONE dense front over U.  Entries of U x U outside the members' true fronts are explicit zeros
// and stay exact zeros through the elimination (no true entry of L depends on them), so the factor is the same; what is
// bought with the extra flops -- cheap on the tensor cores -- is a program of ~50 big regular steps instead of ~500 small
// ones (1138_bus: 529 fronts -> 52 groups at 117 rows, 2.6x the flops).  Needs P.l_src (the CSC export map): the TRUE
// entries of every group leave through an explicit (position in F, CSC index) list.
void build_group_program(const HostPlan& P, int cap_m, const std::vector<int>& a_front, const std::vector<int>& a_col,
                         const std::vector<int>& a_pos, TreeProgram& out);

}  // namespace spchol
