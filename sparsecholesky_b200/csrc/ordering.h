// Fill-reducing orderings (host).  The reference ships no ordering code at all (SURVEY.md 8f-2:
// thesis 5.4 discusses minimum degree on paper only), so everything here is additive: the numeric
// path factors whatever permuted matrix it is handed.
//   * nd_perm_grid : the geometric nested dissection of SURVEY.md 8d the benchmark matrices use
//   * order_amd    : approximate minimum degree on the quotient graph (Amestoy/Davis/Duff 1996)
//   * permute_upper: triu(P A P^T) of a symmetric matrix stored as its upper triangle
#pragma once
#include <cstdint>
#include <vector>

namespace spchol {

// perm[new] = old grid id (id = x + nx*(y + ny*z)); leaf boxes of volume <= 8 or longest side < 3
void nd_perm_grid(int nx, int ny, int nz, int64_t* perm);

// perm[new] = old.  A: upper triangle CSC pattern (diagonal optional), n >= 0.
void order_amd(int n, const int* Ap, const int* Ai, int* perm);

// B = triu(P A P^T), perm[new] = old; rows ascending per column.  Bp[n+1], Bi/Bx[nnz(A)];
// map (optional, nnz(A)): position in B of every entry of A (for re-permuting new values on the same pattern)
void permute_upper(int n, const int* Ap, const int* Ai, const double* Ax, const int* perm,
                   int* Bp, int* Bi, double* Bx, int64_t* map);

}  // namespace spchol
