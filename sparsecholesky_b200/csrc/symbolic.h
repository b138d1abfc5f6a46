// Host-side symbolic analysis and planner for the B200 multifrontal numeric Cholesky.
//
// Replaces (does not port) the symbolic pieces the reference's chol()/schol() rebuild on every
// call: etree (reference include/chol.hpp:378-410), post_order (:466-499), col_count (:567-622),
// schol's row pattern (:874-946), compute_levels / compute_supernodes / atree (src/chol.cpp).
// Everything here is near-linear in nnz(A) + sum_s m_s and uses 64-bit offsets, so the
// 128^3 Laplacian (nnz(L) = 2.83e9 > INT_MAX) can be planned although the reference cannot run it.
#pragma once
#include <cstdint>
#include <memory>
#include <mutex>
#include <string>
#include <vector>

namespace spchol {

// -- standalone helpers with the reference's semantics (public, also used by the KAT tests) ----
void etree(int n, const int* Ap, const int* Ai, int* parent);
void post_order(int n, const int* parent, int* post);
void col_count(int n, const int* Ap, const int* Ai, const int* parent, const int* post, int* cc);
int  ereach(int n, const int* Ap, const int* Ai, const double* Ax, int k, const int* parent,
            int* s, int* w, double* x, int top);
int  compute_levels(int n, const int* parent, int* level_ptr, int* level_nodes);
int  compute_supernodes(int n, const int* parent, const int64_t* Lp, int* sn_id, int* sn_start);
void atree_from_etree(int n, const int* parent, const int* sn_id, int ns, const int* sn_start,
                      int* super_parent);
int64_t triplet_to_csc(int64_t nz, const int* ti, const int* tj, const double* tx, int n,
                       int* Ap, int* Ai, double* Ax);

struct RelaxParams {
    // chain amalgamation (child = immediately preceding supernode whose assembly-tree parent is
    // the next one).  Merge while the merged width <= nrelax[k] and the fraction of explicit
    // zeros in the merged panel <= zrelax[k]; past nrelax[2] the bound is zrelax[2].
    int    nrelax[3] = {4, 16, 48};
    double zrelax[3] = {0.8, 0.1, 0.05};
    bool   enable = true;
    int    max_m = 0;          // > 0: never grow a merged front beyond this many rows (keeps fronts on the shared-memory kernel)
    int    par_min_w = 0;      // > 0 (multi-GPU plans): a chain of >= this many columns is not merged into a parent that
                               // has other children -- the merge is free in storage but turns two balanced subtrees
                               // into one wide front plus unequal children, which proportional mapping cannot split evenly
};

// Result of the symbolic phase (what the reference's SChol carries, plus supernodes).
struct Symbolic {
    int n = 0;
    int64_t nnzA = 0;
    int64_t nnzL = 0;                 // sum of column counts (true pattern)
    double  flops = 0;                // sum_j cc_j^2 (the metric's flop convention)
    std::vector<int> Ap, Ai;          // pattern of triu(A) (kept: the plan needs it)
    std::vector<int> parent, post, colcount;
    std::vector<int64_t> Lp;          // n+1
    // fundamental supernodes (reference rule, src/chol.cpp:75-85)
    int nsf = 0;
    std::vector<int> fsn_start;       // nsf+1
    std::vector<int> fsn_id;          // n
    std::vector<int> fsn_parent;      // nsf  (assembly tree = atree())
    std::vector<int64_t> fbelow_ptr;  // nsf+1 : offsets into fundamental_below() (counts come from colcount)
    // relaxed supernodes = the fronts the numeric phase factors
    int ns = 0;
    std::vector<int> sn_start;        // ns+1
    std::vector<int> sn_id;           // n
    std::vector<int> sn_parent;       // ns
    std::vector<int> sn_top;          // ns : last fundamental supernode inside (owns the below-rows)
    std::vector<int64_t> below_ptr;   // ns+1
    std::vector<int> below_rows;      // sum_s (m_s - w_s) : rows of front s strictly below its own columns, ascending
    std::vector<char> sn_padded;      // ns : 1 if amalgamation introduced explicit zeros
    int64_t panel_entries = 0;        // sum m_s * w_s
    int64_t sum_m = 0;                // sum_s m_s
    int64_t padded_zeros = 0;

    int w(int s) const { return sn_start[s + 1] - sn_start[s]; }
    int below_count(int s) const { return (int)(below_ptr[s + 1] - below_ptr[s]); }
    const int* below(int s) const { return below_rows.data() + below_ptr[s]; }
    int m(int s) const { return w(s) + below_count(s); }

    // rows strictly below every FUNDAMENTAL supernode (the unpadded pattern of L), sum_f (m_f - w_f)
    // entries indexed by fbelow_ptr; built on first use (thread-safe), see symbolic.cpp
    const std::vector<int>& fundamental_below() const;
    struct Lazy { std::once_flag once; std::vector<int> rows; };
    std::shared_ptr<Lazy> lazy;

    // fills Li[nnzL] with the reference's layout (diagonal first, rows ascending; :931-941)
    void fill_Li(int* Li) const;
};

// throws std::runtime_error on malformed input
void analyze(int n, const int* Ap, const int* Ai, const RelaxParams& rp, Symbolic& S);

}  // namespace spchol
