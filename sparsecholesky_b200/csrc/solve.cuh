// Triangular solves on the device-resident supernodal factor (included only by engine.cu):
//   L y = b (fronts bottom-up, level by level), L^T x = y (top-down).
// The reference has no solve at all; a factor that cannot be applied is of little
// use, and the solve gives an end-to-end check ||A x - b|| / ||b|| where nnz(L) > INT_MAX rules out
// the reference's CSC container.  Right-hand sides are processed SOLVE_NR at a time, stored
// [n][SOLVE_NR] row-major on the device.
//
// Every front is cut into column blocks of SOLVE_NB: per block one CTA solves with the dense
// diagonal block (staged in shared memory), then many CTAs stream the rows below the block
// (HBM bound: L is read exactly once per solve direction).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

#include "kernels.cuh"
#include "residual.cuh"

namespace spchol {

constexpr int SOLVE_NR = NPROBE;      // 8 right-hand sides per pass
constexpr int SOLVE_NB = 128;         // column block
struct SBlock { int s, c0; };         // front, first column of the block
struct SChunk { int s, c0, r0; };     // + first row of a chunk of 256 rows (forward update)
struct SGat { int s, c0, r0, i0; };   // backward gather: 32 columns from c0 of the block ending at r0, rows [i0, i0 + SOLVE_GROWS)
constexpr int SOLVE_GROWS = 256;      // rows per gather CTA: a tall front (16 447 rows at the root of 128^3) is cut into chunks whose
                                      // partial sums meet in tsum by atomicAdd -- one CTA per 32 columns streamed the whole panel (1.3 ms a launch)
constexpr int solve_diag_smem(int nb) { return (nb * (nb | 1) + nb * SOLVE_NR + nb) * 8; }

// column-major right-hand sides (n x nr) -> interleaved [n][SOLVE_NR] (missing columns zero), gather sums cleared; and back
__global__ void k_solve_pack(int n, int nr, const double* __restrict__ cols, double* __restrict__ v, double* __restrict__ tsum) {
    const int64_t e = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
    if (e >= static_cast<int64_t>(n) * SOLVE_NR) return;
    const int64_t i = e / SOLVE_NR; const int k = static_cast<int>(e % SOLVE_NR);
    v[e] = k < nr ? cols[static_cast<int64_t>(k) * n + i] : 0.0;
    tsum[e] = 0.0;
}
// Multi-rank solve (engine.cu, spchol_solve): the entries of the work vector that more than one rank touches -- the columns of
// the fronts at the top of the tree that have descendants on another rank -- are listed in idx[]; mine[i] != 0 where this
// rank solves column i.  On the way up only the solving rank holds the VALUE of such a column (b minus the updates that
// have arrived), the other ranks hold the updates they have produced since the last exchange, starting from zero: an
// exchange sums the copies (ncclAllReduce between k_top_gather and k_top_store), the solver keeps the sum, the others
// start again from zero -- the same terms as the sequential substitution, added in another order, no differences taken.
// On the way down a column is REPLACED everywhere by the value its solver holds (k_top_gather with mine: the others
// contribute zero).
__global__ void k_top_clear_others(int64_t ne, const int* __restrict__ idx, const unsigned char* __restrict__ mine, double* __restrict__ v) {
    const int64_t e = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
    if (e >= ne) return;
    const int i = idx[e / SOLVE_NR];
    if (!mine[i]) v[static_cast<int64_t>(i) * SOLVE_NR + e % SOLVE_NR] = 0.0;
}
template <bool ONLY_MINE>
__global__ void k_top_gather(int64_t ne, const int* __restrict__ idx, const unsigned char* __restrict__ mine,
                             const double* __restrict__ v, double* __restrict__ out) {
    const int64_t e = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
    if (e >= ne) return;
    const int i = idx[e / SOLVE_NR];
    out[e] = (!ONLY_MINE || mine[i]) ? v[static_cast<int64_t>(i) * SOLVE_NR + e % SOLVE_NR] : 0.0;
}
template <bool ONLY_MINE>
__global__ void k_top_store(int64_t ne, const int* __restrict__ idx, const unsigned char* __restrict__ mine, double* __restrict__ v,
                            const double* __restrict__ in) {
    const int64_t e = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
    if (e >= ne) return;
    const int i = idx[e / SOLVE_NR];
    v[static_cast<int64_t>(i) * SOLVE_NR + e % SOLVE_NR] = (!ONLY_MINE || mine[i]) ? in[e] : 0.0;
}
// keeps the rows of the solution this rank computed (mine[i] != 0), clears the others: the sum over the ranks is the solution
__global__ void k_solve_mask(int64_t ne, const unsigned char* __restrict__ mine, double* __restrict__ v) {
    const int64_t e = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
    if (e < ne && !mine[e / SOLVE_NR]) v[e] = 0.0;
}
__global__ void k_solve_unpack(int n, int nr, const double* __restrict__ v, double* __restrict__ cols) {
    const int64_t e = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
    if (e >= static_cast<int64_t>(n) * nr) return;
    const int64_t i = e % n; const int k = static_cast<int>(e / n);
    cols[e] = v[i * SOLVE_NR + k];
}

// diagonal block solve, forward (TRANS = false): y = L_jj^-1 x   or backward (TRANS = true): x = L_jj^-T (y - t)
template <bool TRANS>
__global__ void __launch_bounds__(256) k_solve_diag(const SBlock* __restrict__ tasks, FrontArrays fa, const double* __restrict__ L,
                                                    double* __restrict__ b, const double* __restrict__ tsum) {
    extern __shared__ double sm[];
    const SBlock t = tasks[blockIdx.x];
    const int s = t.s, w = fa.w[s], ld = fa.ld[s];
    const int nb = min(SOLVE_NB, w - t.c0), lds = nb | 1;
    double* S = sm;                       // S[i + j*lds] = L[c0 + i][c0 + j], i >= j
    double* X = sm + nb * lds;            // X[i*SOLVE_NR + k]
    double* RD = X + nb * SOLVE_NR;       // reciprocals of the diagonal (one IEEE division per column, all in parallel, instead of
                                          // four divisions on the serial path of every column)
    const double* __restrict__ D = L + fa.panel_off[s] + t.c0 + static_cast<int64_t>(t.c0) * ld;
    const int tid = threadIdx.x;
    for (int idx = tid; idx < nb * nb; idx += 256) {
        const int i = idx % nb, j = idx / nb;
        if (i >= j) S[i + j * lds] = D[i + static_cast<int64_t>(j) * ld];
    }
    const int64_t g0 = static_cast<int64_t>(fa.first[s] + t.c0) * SOLVE_NR;
    for (int idx = tid; idx < nb * SOLVE_NR; idx += 256) X[idx] = b[g0 + idx] - (TRANS ? tsum[g0 + idx] : 0.0);
    for (int c = tid; c < nb; c += 256) RD[c] = 1.0 / D[c + static_cast<int64_t>(c) * ld];
    __syncthreads();
    // thread (i, h): row i of the block, right-hand sides 4h .. 4h+3
    const int i = tid >> 1, h = (tid & 1) * 4;
    double x[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) x[k] = i < nb ? X[i * SOLVE_NR + h + k] : 0.0;
    // Columns in micro blocks of 8: the 8 rows of a micro block sit in 16 consecutive threads of ONE warp (tid = 2 i + h), so
    // the substitution inside a micro block needs warp barriers only; the CTA meets once per micro block, before the rows
    // outside it take their update (16 CTA barriers per 128 columns instead of 128).
    if (!TRANS) {
        for (int c0 = 0; c0 < nb; c0 += 8) {
            const int c1 = min(c0 + 8, nb);
            if (i >= c0 && i < c0 + 8) {                       // warp-uniform per half warp; both halves of the warp take the barriers
                for (int c = c0; c < c1; ++c) {
                    if (i == c) {
                        const double rd = RD[c];
#pragma unroll
                        for (int k = 0; k < 4; ++k) { x[k] = x[k] * rd; X[c * SOLVE_NR + h + k] = x[k]; }
                    }
                    __syncwarp(0xffffu << (threadIdx.x & 16));
                    if (i > c && i < c1) {
                        const double l = S[i + c * lds];
#pragma unroll
                        for (int k = 0; k < 4; ++k) x[k] = fma(-l, X[c * SOLVE_NR + h + k], x[k]);
                    }
                }
            }
            __syncthreads();
            if (i >= c1 && i < nb) {
                for (int c = c0; c < c1; ++c) {
                    const double l = S[i + c * lds];
#pragma unroll
                    for (int k = 0; k < 4; ++k) x[k] = fma(-l, X[c * SOLVE_NR + h + k], x[k]);
                }
            }
        }
    } else {
        for (int lo = (nb - 1) & ~7; lo >= 0; lo -= 8) {       // micro blocks aligned to 8 from the left, last one first
            const int hi = min(lo + 8, nb);
            if (i >= lo && i < lo + 8) {
                for (int c = hi - 1; c >= lo; --c) {
                    if (i == c) {
                        const double rd = RD[c];
#pragma unroll
                        for (int k = 0; k < 4; ++k) { x[k] = x[k] * rd; X[c * SOLVE_NR + h + k] = x[k]; }
                    }
                    __syncwarp(0xffffu << (threadIdx.x & 16));
                    if (i < c && i >= lo) {
                        const double l = S[c + i * lds];          // L[c][i]
#pragma unroll
                        for (int k = 0; k < 4; ++k) x[k] = fma(-l, X[c * SOLVE_NR + h + k], x[k]);
                    }
                }
            }
            __syncthreads();
            if (i < lo) {
                for (int c = hi - 1; c >= lo; --c) {
                    const double l = S[c + i * lds];
#pragma unroll
                    for (int k = 0; k < 4; ++k) x[k] = fma(-l, X[c * SOLVE_NR + h + k], x[k]);
                }
            }
        }
    }
    __syncthreads();
    for (int idx = tid; idx < nb * SOLVE_NR; idx += 256) b[g0 + idx] = X[idx];
}

// forward update: b[row(i)] -= sum_{c in block} L[i][c] y[c]   for the rows i >= block end of the front
__global__ void __launch_bounds__(256) k_solve_update(const SChunk* __restrict__ tasks, FrontArrays fa, const int* __restrict__ rowidx,
                                                      const double* __restrict__ L, double* __restrict__ b) {
    __shared__ double Y[SOLVE_NB * SOLVE_NR];
    const SChunk t = tasks[blockIdx.x];
    const int s = t.s, w = fa.w[s], m = fa.m[s], ld = fa.ld[s];
    const int nb = min(SOLVE_NB, w - t.c0);
    const int64_t g0 = static_cast<int64_t>(fa.first[s] + t.c0) * SOLVE_NR;
    for (int idx = threadIdx.x; idx < nb * SOLVE_NR; idx += 256) Y[idx] = b[g0 + idx];
    __syncthreads();
    const int i = t.r0 + threadIdx.x;
    if (i >= m) return;
    const double* __restrict__ P = L + fa.panel_off[s] + i + static_cast<int64_t>(t.c0) * ld;
    double acc[SOLVE_NR];
#pragma unroll
    for (int k = 0; k < SOLVE_NR; ++k) acc[k] = 0.0;
    for (int c = 0; c < nb; c += 4) {
        double v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) v[u] = c + u < nb ? P[static_cast<int64_t>(c + u) * ld] : 0.0;
#pragma unroll
        for (int u = 0; u < 4; ++u)
            if (c + u < nb) {
#pragma unroll
                for (int k = 0; k < SOLVE_NR; ++k) acc[k] = fma(v[u], Y[(c + u) * SOLVE_NR + k], acc[k]);
            }
    }
    if (i < w) {          // a later column of this front: only this thread touches it
        double* dst = b + static_cast<int64_t>(fa.first[s] + i) * SOLVE_NR;
#pragma unroll
        for (int k = 0; k < SOLVE_NR; ++k) dst[k] -= acc[k];
    } else {              // a row of an ancestor: other fronts of this level may update it too
        double* dst = b + static_cast<int64_t>(rowidx[fa.rel_off[s] + i - w]) * SOLVE_NR;
#pragma unroll
        for (int k = 0; k < SOLVE_NR; ++k) atomicAdd(dst + k, -acc[k]);
    }
}

// backward gather: tsum[first + c] = sum_{i >= block end} L[i][c] x[row(i)]   for 32 columns c0 .. c0+31 of a block
// (task.r0 = end of the SOLVE_NB block the columns belong to)
__global__ void __launch_bounds__(256) k_solve_gather(const SGat* __restrict__ tasks, FrontArrays fa, const int* __restrict__ rowidx,
                                                      const double* __restrict__ L, const double* __restrict__ x, double* __restrict__ tsum) {
    __shared__ double Z[128][SOLVE_NR + 1];
    const SGat t = tasks[blockIdx.x];
    const int s = t.s, w = fa.w[s], m = fa.m[s], ld = fa.ld[s];
    const double* __restrict__ P = L + fa.panel_off[s];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int cb = t.c0 + warp * 4;
    double acc[4][SOLVE_NR];
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int k = 0; k < SOLVE_NR; ++k) acc[a][k] = 0.0;
    const int iend = min(m, t.i0 + SOLVE_GROWS);
    for (int i0 = t.i0; i0 < iend; i0 += 128) {
        __syncthreads();
        for (int e = tid; e < 128 * SOLVE_NR; e += 256) {
            const int i = i0 + e / SOLVE_NR;
            Z[e / SOLVE_NR][e % SOLVE_NR] = i < iend ? x[static_cast<int64_t>(front_row(fa, rowidx, s, w, i)) * SOLVE_NR + e % SOLVE_NR] : 0.0;
        }
        __syncthreads();
        for (int ii = lane; ii < 128; ii += 32) {
            const int i = i0 + ii;
            if (i >= iend) break;
#pragma unroll
            for (int a = 0; a < 4; ++a) {
                const int c = cb + a;
                if (c < w && c < t.r0) {
                    const double v = P[i + static_cast<int64_t>(c) * ld];
#pragma unroll
                    for (int k = 0; k < SOLVE_NR; ++k) acc[a][k] = fma(v, Z[ii][k], acc[a][k]);
                }
            }
        }
    }
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int k = 0; k < SOLVE_NR; ++k) {
            double v = acc[a][k];
            for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
            if (lane == 0 && cb + a < w && cb + a < t.r0) atomicAdd(tsum + static_cast<int64_t>(fa.first[s] + cb + a) * SOLVE_NR + k, v);
        }
}

}  // namespace spchol
