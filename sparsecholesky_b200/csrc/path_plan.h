// Batched factorization, upper part of the assembly tree as "path groups" (host planner).
//
// Near the root the assembly tree of a 1138_bus-like pattern is one long path with short side
// branches: 174 of 529 fronts, w ~ 4, m ~ 45..117, 99 % of the flops.  Along a path the parent front
// is the child's contribution block plus a few new rows, so a run of consecutive path fronts
// s_1 .. s_k shares ONE frontal matrix over the row set
//     U = cols(s_1) ++ cols(s_2) ++ ... ++ cols(s_k) ++ below(s_k)          (ascending global rows)
// that stays in shared memory while s_1 .. s_k are eliminated in place: no per-front zeroing,
// extend-add of the carried block or write-out.  Each front only touches its TRUE rows (positions
// pos[] inside U), so no explicit zeros are factored -- the group is not an amalgamated supernode.
// Contribution blocks of fronts outside the group (side branches, the lower levels) are added at
// the start of the group; the block of s_k goes to the next group in registers when that group comes
// next in the program, else through a dedicated workspace area (the level-order arena of plan.cpp
// cannot be used: the groups run in depth-first, not level order).
//
// Consumed by k_front_path (path_kernel.cuh); tests/cpp/path_emu.cpp runs the same kernel source
// on host threads against the oracle.
#pragma once
#include <cstdint>
#include <vector>

#include "plan.h"

namespace spchol {

constexpr int PK_OWNERS = 8;    // columns of a group's frontal matrix are dealt round-robin to 8 owners (one warp each at 256 threads)

struct alignas(16) PathMember {   // a front eliminated inside a group (self-contained: one descriptor load per front)
    int s;                 // front id
    int off;               // position in U of its first column
    int pos_off;           // pos[pos_off .. pos_off + m): position in U of every row of the front
    int w, m, ld;          // the front's columns, rows, leading dimension of its panel in the L buffer
    int first;             // global index of its first column
    int nd;                // rows of this front that are columns of the NEXT member of the group (0 for the last member)
    int64_t panel_off;     // its panel in the L buffer
    int64_t pad2;
};
struct alignas(16) PathChild {    // a contribution block added into the group's frontal matrix
    int c;                 // child front
    int mc;                // order of its block
    int pos_off;           // pos[pos_off .. pos_off + mc): position in U of the block's rows
    int carried;           // 1: the block sits in the registers of the CTA (top of the previous group)
    int64_t src;           // else: workspace offset (column-major lower, ld = even(mc))
    int cols_off;          // pos[cols_off .. cols_off + mc): the block's columns jj grouped by owner of their target column
    int pad;
    int optr[PK_OWNERS + 1];   // owner o handles cols[optr[o] .. optr[o + 1])
    int pad3[3];
};
struct alignas(16) PathGroup {
    int nu;                // |U|
    int member0, nmember;
    int child0, nchild;
    int mb;                // order of the top front's contribution block = nu - sum of the members' widths
    int carry_out;         // 1: that block is handed to the next group in registers
    int carried_child;     // index (relative to child0) of the child that arrives in registers, -1: none
    int64_t cb_dst;        // else: workspace offset it is written to (-1: mb == 0)
    int64_t pad2;
};

struct PathProgram {
    int level0 = 0;                       // fronts of the levels >= level0 are covered
    int max_nu = 0;                       // largest |U| (sizes the shared memory)
    int max_m = 0;                        // largest front inside (sizes the staged position list)
    int64_t ws_entries = 0;               // workspace doubles per matrix INCLUDING the dedicated area (>= HostPlan::ws_entries)
    std::vector<PathGroup> groups;
    std::vector<PathMember> members;
    std::vector<PathChild> children;
    std::vector<int> pos;
    int n_carried = 0;
};

// Builds the program for the fronts of the levels >= level0 of a single-rank plan.  max_nu bounds |U|
// (shared memory: max_nu (max_nu + 1) / 2 doubles), carry_cap the entries of a block kept in registers.
// Returns false (program left empty) when a front of the region has more than max_nu rows.
// mode 0: groups are runs of consecutive fronts of the heavy paths (what the measurements of DESIGN.md 4c use);
// mode 1: groups are subtrees grown bottom-up (a group is any connected piece of the tree that contains its top --
//         the kernel only needs the members in a topological order), fewer groups on bushy trees.
bool build_path_program(const HostPlan& P, int level0, int max_nu, int64_t carry_cap, PathProgram& out, int mode = 0);

}  // namespace spchol
