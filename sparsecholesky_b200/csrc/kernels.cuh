// sm_100a kernels of the multifrontal numeric Cholesky (included only by engine.cu).
//
// Dense sub-operations the reference's supernodal attempt hands to CPU BLAS/LAPACK
// (dpotrf_ include/chol.hpp:1263, cblas_dtrsm :1292, cblas_dsyrk :1322, apply_update :1196)
// are implemented here from scratch for the GPU; numerically they produce the same L as the
// reference's working simplicial chol() (:750-863) up to fp64 summation order.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

#include "plan.h"

namespace spchol {

struct FrontArrays {          // device pointers, one entry per front
    const int* first; const int* w; const int* m; const int* ld; const int* parent;
    const int64_t* panel_off; const int64_t* cb_off; const int64_t* rel_off;
    const int* rel; const int* child_ptr; const int* child_idx;
    // multi-GPU: size of the rank group that factors the front (1 = this rank alone), this rank's
    // index in it, and the block-column width the group deals round-robin (see plan.h)
    const int* own_n; const int* own_idx; int dist_NB;
};

__device__ __forceinline__ int even_up_d(int v) { return (v + 1) & ~1; }

// ---------------------------------------------------------------------------------------------
// assemble: scatter the stored entries of triu(A) to their slots in the (zeroed) L buffer.
// Replaces the csc_matrix(A, S) constructor (reference include/chol.hpp:190-217) + extract_panel
// (:1118-1159).  HBM bound: 8 B value + 8 B map read, 8 B scattered write per entry.
// ---------------------------------------------------------------------------------------------
// Batched use (many value sets over one pattern): blockIdx.y = matrix, Ax / L advance by one matrix.
__global__ void k_assemble(int64_t nnzA, const int64_t* __restrict__ a_dest,
                           const double* __restrict__ Ax, double* __restrict__ L, int64_t lstride) {
    Ax += blockIdx.y * nnzA; L += blockIdx.y * lstride;
    const int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
    if (i < nnzA) { const int64_t d = a_dest[i]; if (d >= 0) L[d] = Ax[i]; }   // d < 0: front held by another rank
}

// gather the CSC values of L out of the supernodal panels (replaces
// scatter_panel_column_into_L, reference include/chol.hpp:1178-1189)
__global__ void k_export(int64_t nnzL, const int64_t* __restrict__ l_src,
                         const double* __restrict__ L, double* __restrict__ Lx, int64_t lstride) {
    L += blockIdx.y * lstride; Lx += blockIdx.y * nnzL;
    const int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
    if (i < nnzL) { const int64_t q = l_src[i]; Lx[i] = q >= 0 ? L[q] : 0.0; }   // q < 0: exported by another rank
}

// batched factorization: per-matrix info words (atomicMin targets preset to 0x7f7f7f7f) -> 0 or failing column
__global__ void k_fix_info(const int* __restrict__ raw, int* __restrict__ out, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = raw[i] == 0x7f7f7f7f ? 0 : raw[i];
}

// grid = (slices, tasks): 16-byte stores, four in flight per thread (offsets and counts are even: the arena
// hands out even-sized blocks and a contribution block has an even leading dimension)
__global__ void __launch_bounds__(256) k_zero(const ZeroTask* __restrict__ tasks, double* __restrict__ W) {
    const ZeroTask t = tasks[blockIdx.y];
    double2* p = reinterpret_cast<double2*>(W + t.off);
    const int64_t n2 = t.count >> 1;
    const int64_t stride = static_cast<int64_t>(gridDim.x) * 256;
    const double2 z = make_double2(0.0, 0.0);
    int64_t i = blockIdx.x * static_cast<int64_t>(256) + threadIdx.x;
    for (; i + 3 * stride < n2; i += 4 * stride) { p[i] = z; p[i + stride] = z; p[i + 2 * stride] = z; p[i + 3 * stride] = z; }
    for (; i < n2; i += stride) p[i] = z;
    if ((t.count & 1) && blockIdx.x == 0 && threadIdx.x == 0) W[t.off + t.count - 1] = 0.0;
}

// pack / unpack of block columns for the inter-GPU transfers: rows x cols rectangle, column-major,
// between the L buffer and the workspace.  grid = (tasks, column slices).  HBM bound (16 B/entry).
__global__ void __launch_bounds__(256) k_copy2d(const CopyTask* __restrict__ tasks, double* __restrict__ L,
                                                double* __restrict__ W) {
    const CopyTask t = tasks[blockIdx.x];
    const double* __restrict__ src = (t.src_space ? W : L) + t.src_off;
    double* __restrict__ dst = (t.dst_space ? W : L) + t.dst_off;
    // grid.y = column slices, grid.z = row slices of 2048 (the packed block columns on the critical path of a
    // distributed front are tall: rows x 256); four independent 8-byte copies in flight per thread
    for (int j = blockIdx.y; j < t.cols; j += gridDim.y) {
        const double* s = src + static_cast<int64_t>(j) * t.lds;
        double* d = dst + static_cast<int64_t>(j) * t.ldd;
        for (int i0 = blockIdx.z * 2048; i0 < t.rows; i0 += gridDim.z * 2048) {
            const int i = i0 + threadIdx.x;
            double v[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) v[u] = (i + u * 256 < t.rows) ? s[i + u * 256] : 0.0;
#pragma unroll
            for (int u = 0; u < 8; ++u) if (i + u * 256 < t.rows) d[i + u * 256] = v[u];
        }
    }
}

// ---------------------------------------------------------------------------------------------
// Small fronts: one CTA factors one whole front in shared memory.
//   F (m x m, lower) = [A-panel | 0]  (+)= extend-add of every child's contribution block
//   partial Cholesky of the first w columns (right-looking, rank-1 updates in smem)
//   write the m x w panel back to the L buffer and the trailing (m-w)^2 lower block to the
//   contribution workspace.
// Fuses dpotrf_ + dtrsm + dsyrk + apply_update of the reference's factorize_sn for a front
// (include/chol.hpp:1232-1335, :1196-1216).  Bound: HBM / latency (flops per byte ~ w).
// ---------------------------------------------------------------------------------------------
// The front is held packed (lower triangle only: column j at F + j m - j (j - 1) / 2 - j, so that col[i] is valid
// for i >= j), m (m + 1) / 2 doubles: twice as many fronts of the larger classes fit an SM.
template <int T>
__global__ void __launch_bounds__(T) k_front_small(const int* __restrict__ list, FrontArrays fa,
                                                   double* __restrict__ L, double* __restrict__ W,
                                                   int* __restrict__ info, int64_t lstride, int64_t wstride) {
    extern __shared__ double F[];
    // batched use: blockIdx.y = matrix; every matrix has its own L buffer, workspace and info word
    L += blockIdx.y * lstride; W += blockIdx.y * wstride; info += blockIdx.y;
    const int s = list[blockIdx.x];
    const int w = fa.w[s], m = fa.m[s], ld = fa.ld[s];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr int NW = T / 32;
    double* __restrict__ P = L + fa.panel_off[s];
    auto colp = [&](int j) { return F + (j * m - j * (j - 1) / 2 - j); };

    for (int j = warp; j < m; j += NW) {
        double* col = colp(j);
        if (j < w) {
            const double* src = P + static_cast<int64_t>(j) * ld;
            for (int i = j + lane; i < m; i += 32) col[i] = src[i];
        } else {
            for (int i = j + lane; i < m; i += 32) col[i] = 0.0;
        }
    }
    __syncthreads();

    // extend-add the children, one child at a time (deterministic summation order)
    for (int q = fa.child_ptr[s]; q < fa.child_ptr[s + 1]; ++q) {
        const int c = fa.child_idx[q];
        const int mc = fa.m[c] - fa.w[c];
        if (mc == 0) continue;
        const int ldc = even_up_d(mc);
        const double* __restrict__ C = W + fa.cb_off[c];
        const int* __restrict__ r = fa.rel + fa.rel_off[c];
        for (int jj = warp; jj < mc; jj += NW) {
            double* col = colp(r[jj]);
            const double* src = C + static_cast<int64_t>(jj) * ldc;
            for (int ii = jj + lane; ii < mc; ii += 32) col[r[ii]] += src[ii];
        }
        __syncthreads();
    }

    // partial factorization of columns 0..w-1
    for (int k = 0; k < w; ++k) {
        double* colk = colp(k);
        const double d = colk[k];
        // reference test: d <= T(0) (chol.hpp:849); info = smallest failing 1-based column
        if (d <= 0.0 && tid == 0) atomicMin(info, fa.first[s] + k + 1);
        const double piv = sqrt(d);
        __syncthreads();                               // everyone has read the pivot before it is overwritten
        for (int i = k + 1 + tid; i < m; i += T) colk[i] = colk[i] / piv;
        if (tid == 0) colk[k] = piv;
        __syncthreads();
        for (int j = k + 1 + warp; j < m; j += NW) {
            const double uj = colk[j];
            double* col = colp(j);
            for (int i = j + lane; i < m; i += 32) col[i] -= colk[i] * uj;
        }
        __syncthreads();
    }

    for (int j = warp; j < m; j += NW) {
        const double* col = colp(j);
        if (j < w) {
            double* dst = P + static_cast<int64_t>(j) * ld;
            for (int i = j + lane; i < m; i += 32) dst[i] = col[i];
        } else {
            const int mb = m - w;
            double* dst = W + fa.cb_off[s] + static_cast<int64_t>(j - w) * even_up_d(mb) - w;
            for (int i = j + lane; i < m; i += 32) dst[i] = col[i];
        }
    }
}

// ---------------------------------------------------------------------------------------------
// Large fronts, step 1: extend-add one child per parent per launch ("rounds": no two CTAs of a
// launch touch the same parent, so no atomics and a fixed summation order).
// grid = (tasks, column slices); each CTA walks child columns jj = blockIdx.y, += gridDim.y.
// Replaces apply_update (reference include/chol.hpp:1196-1216): relative-index scatter instead
// of a binary search per entry.
// ---------------------------------------------------------------------------------------------
// mode 0: every column of the child's block; 1: only the columns that land in the parent's panel; 2: only those
// that land in the parent's contribution block (relative indices are increasing, so the two sets are a prefix and
// the remaining suffix of the child's columns)
__global__ void __launch_bounds__(128) k_extend_add(const int* __restrict__ list, FrontArrays fa,
                                                    double* __restrict__ L, double* __restrict__ W, int mode) {
    const int c = list[blockIdx.x];
    const int p = fa.parent[c];
    const int mc = fa.m[c] - fa.w[c];
    const int ldc = even_up_d(mc);
    const int wp = fa.w[p], ldp = fa.ld[p];
    const int ldcb = even_up_d(fa.m[p] - wp);
    const double* __restrict__ C = W + fa.cb_off[c];
    const int* __restrict__ r = fa.rel + fa.rel_off[c];
    double* __restrict__ Pp = L + fa.panel_off[p];
    double* __restrict__ Cp = (fa.cb_off[p] >= 0) ? W + fa.cb_off[p] : nullptr;
    // parent factored by a rank group: only the block columns this rank owns are assembled here
    const int G = fa.own_n[p], me = fa.own_idx[p], DNB = fa.dist_NB;
    const int nbp = (wp + DNB - 1) / DNB;
    for (int jj = blockIdx.y; jj < mc; jj += gridDim.y) {
        const int cj = r[jj];
        if (mode == 1 && cj >= wp) break;
        if (mode == 2 && cj < wp) continue;
        if (G > 1) {
            const int blk = cj < wp ? cj / DNB : nbp + (cj - wp) / DNB;
            if (blk % G != me) continue;
        }
        const double* src = C + static_cast<int64_t>(jj) * ldc;
        double* dst = cj < wp ? Pp + static_cast<int64_t>(cj) * ldp : Cp + static_cast<int64_t>(cj - wp) * ldcb - wp;
        // the additions go out as RED.ADD.F64 (atomicAdd with the result unused): performed at the L2, nothing to wait for on
        // the SM side; every target has exactly one writer per launch (one child per parent and round), so the sums and
        // their order are the same as with a read-modify-write
        const int T = blockDim.x;
        int ii = jj + threadIdx.x;
        for (; ii + 3 * T < mc; ii += 4 * T) {
            const int r0 = r[ii], r1 = r[ii + T], r2 = r[ii + 2 * T], r3 = r[ii + 3 * T];
            const double s0 = src[ii], s1 = src[ii + T], s2 = src[ii + 2 * T], s3 = src[ii + 3 * T];
            atomicAdd(dst + r0, s0); atomicAdd(dst + r1, s1); atomicAdd(dst + r2, s2); atomicAdd(dst + r3, s3);
        }
        for (; ii < mc; ii += T) atomicAdd(dst + r[ii], src[ii]);
    }
}

__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                 : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

// ---------------------------------------------------------------------------------------------
// Large fronts, step 2a: Cholesky of one nb x nb (nb <= 128) diagonal block in shared memory AND
// the explicit inverse of its triangular factor (written to a scratch slot), so that the panel
// solve below the block becomes a DMMA matrix product (k_trsm_inv) instead of a substitution.
// (dpotrf_ of the reference, include/chol.hpp:1263, one block of the blocked algorithm.)
// Right-looking over 8-wide micro panels: warp 0 factors the 8x8 micro block in registers (row
// per lane, shuffles) and inverts it; all threads apply the micro inverse to the rows below (no
// divisions) and do the rank-8 update of the trailing block.  W = L^-1 is then assembled from
// the micro inverses, one warp per block column.
// ---------------------------------------------------------------------------------------------
constexpr int PB_MAX = 128;           // largest diagonal block; one scratch slot holds PB_MAX^2 doubles
constexpr int PI_THREADS = 256;
constexpr int PI_XLD = 9;             // row stride of the micro-panel copy (odd: conflict-free row walks)
constexpr int potrf_inv_smem(int pb) { return (pb * (pb + 1) + 16 * 64 + pb * PI_XLD) * 8; }

// PB = size class (32 / 64 / 128): bounds the shared memory so that the many narrow fronts of the
// lower levels still get several CTAs per SM.  The scratch slot stores W with leading dimension PB.
template <int PB>
__global__ void __launch_bounds__(PI_THREADS, 2) k_potrf_inv(const PotrfTask* __restrict__ tasks,
                                                          double* __restrict__ L, double* __restrict__ Winv,
                                                          int* __restrict__ info) {
    extern __shared__ double sm[];
    constexpr int PB_LD = PB + 1;            // odd leading dimension: conflict-free row and column walks
    double* S = sm;                          // S[i + j*PB_LD]: lower = L; strict upper receives W^T at the end
    double* T = sm + PB * PB_LD;             // 16 micro inverses (8x8 lower), T[mp*64 + i + 8*j]
    double* X = T + 16 * 64;                 // current micro panel by rows: X[rr*PI_XLD + k], rr = row - (p+8)
    const PotrfTask t = tasks[blockIdx.x];
    const int nb = t.nb, ld = t.ld;
    double* __restrict__ D = L + t.d_off;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr int NW = PI_THREADS / 32;
    const int nmp = (nb + 7) >> 3;           // micro panels; the padded tail of the last one is an identity
    const int np8 = nmp * 8;

    // stage the block: 8 independent global loads in flight per thread (the block was just written by
    // the previous update kernel, so every load is an L2 / HBM round trip)
    for (int base = 0; base < np8 * PB; base += PI_THREADS * 8) {
        double v[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            const int idx = base + u * PI_THREADS + tid;
            const int i = idx % PB, j = idx / PB;
            v[u] = (j < nb && i >= j && i < nb) ? D[i + static_cast<int64_t>(j) * ld] : ((i == j) ? 1.0 : 0.0);
        }
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            const int idx = base + u * PI_THREADS + tid;
            const int i = idx % PB, j = idx / PB;
            if (j < np8 && i < np8) S[i + j * PB_LD] = v[u];
        }
    }
    __syncthreads();

    // 8x8 micro Cholesky + inverse of the diagonal micro block at p, entirely in registers of ONE
    // warp (operands are shared-memory broadcasts, no shuffles); results go to S (factor) and T
    // (inverse).  sqrt and reciprocal of a pivot come from one rsqrt seed refined by Goldschmidt steps (see below).
    auto micro_factor = [&](int mp) {
        const int p = mp * 8;
        double l[8][8], tm[8][8], rinv[8];
#pragma unroll
        for (int i = 0; i < 8; ++i)
#pragma unroll
            for (int j = 0; j <= i; ++j) l[i][j] = S[(p + i) + (p + j) * PB_LD];
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const double d = l[k][k];
            if (d <= 0.0 && lane == 0 && p + k < nb) atomicMin(info, t.col0 + p + k + 1);   // reference test (chol.hpp:849)
            // sqrt(d) and 1/sqrt(d) together: Goldschmidt iteration from the reciprocal-square-root seed, then one residual
            // correction each (square root correctly rounded up to rare last-bit cases, reciprocal within one ulp) -- no
            // software sqrt / division on the serial path of the block
            double y = rsqrt(d);
            double gq = d * y, hq = 0.5 * y;
            double r = fma(-gq, hq, 0.5);
            gq = fma(gq, r, gq); hq = fma(hq, r, hq);
            r = fma(-gq, hq, 0.5);
            gq = fma(gq, r, gq); hq = fma(hq, r, hq);
            const double piv = fma(fma(-gq, gq, d), hq, gq);
            y = hq + hq;
            y = fma(fma(-piv, y, 1.0), y, y);
            rinv[k] = y;
            l[k][k] = piv;
#pragma unroll
            for (int i = k + 1; i < 8; ++i) l[i][k] *= y;
#pragma unroll
            for (int j = k + 1; j < 8; ++j)
#pragma unroll
                for (int i = j; i < 8; ++i) l[i][j] = fma(-l[i][k], l[j][k], l[i][j]);
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) {                  // column j of the inverse: L t = e_j
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                if (i < j) { tm[i][j] = 0.0; continue; }
                double acc = (i == j) ? 1.0 : 0.0;
#pragma unroll
                for (int k = j; k < i; ++k) acc = fma(-l[i][k], tm[k][j], acc);
                tm[i][j] = acc * rinv[i];
            }
        }
        if (lane < 8) {                                  // lane i stores row i
#pragma unroll
            for (int i = 0; i < 8; ++i)
                if (i == lane) {
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        if (j <= i) S[(p + i) + (p + j) * PB_LD] = l[i][j];
                        T[mp * 64 + i + 8 * j] = tm[i][j];
                    }
                }
        }
    };
    // rank-8 update of one strip of 4 columns (c0 relative to p+8) for rows rr = r_begin + lane, +32, ...
    auto update_strip = [&](int p, int c0, int r_begin, int r_end) {
        double xj[4][8];
#pragma unroll
        for (int b2 = 0; b2 < 4; ++b2)
#pragma unroll
            for (int k = 0; k < 8; ++k) xj[b2][k] = X[(c0 + b2) * PI_XLD + k];
        for (int rr = r_begin + lane; rr < r_end; rr += 32) {
            double xi[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) xi[k] = X[rr * PI_XLD + k];
#pragma unroll
            for (int b2 = 0; b2 < 4; ++b2) {
                double acc = 0.0;
#pragma unroll
                for (int k = 0; k < 8; ++k) acc = fma(xi[k], xj[b2][k], acc);
                S[(p + 8 + rr) + (p + 8 + c0 + b2) * PB_LD] -= acc;   // above-diagonal entries of the strip are scratch
            }
        }
    };

    if (warp == 0) micro_factor(0);
    __syncthreads();
    for (int mp = 0; mp < nmp; ++mp) {
        const int p = mp * 8;
        const int below = np8 - (p + 8);
        if (below == 0) break;
        // ---- rows below the micro block: X = S[rows, p:p+8] * T^T (PB <= 128 < threads: a row per thread)
        if (tid < below) {
            const int row = p + 8 + tid;
            double b[8], x[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) b[k] = S[row + (p + k) * PB_LD];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                double acc = 0.0;
#pragma unroll
                for (int k = 0; k <= j; ++k) acc = fma(b[k], T[mp * 64 + j + 8 * k], acc);
                x[j] = acc;
            }
#pragma unroll
            for (int j = 0; j < 8; ++j) { S[row + (p + j) * PB_LD] = x[j]; X[tid * PI_XLD + j] = x[j]; }
        }
        __syncthreads();
        // ---- the next micro panel (8 columns) first, by all warps: 2 strips x up to 4 row chunks of 32
        {
            const int c0 = (warp & 1) * 4, chunk = warp >> 1;
            const int rb = c0 + chunk * 32;
            update_strip(p, c0, rb, min(rb + 32, below));
        }
        __syncthreads();
        // ---- look-ahead: warp 0 factors the next diagonal micro block while the other warps finish the
        // rank-8 update of the columns to its right (a warp owns strips of 4 columns, lanes walk down rows)
        if (warp == 0) micro_factor(mp + 1);
        else {
            const int ncs = below >> 2;
            for (int cs = 2 + (warp - 1); cs < ncs; cs += NW - 1) update_strip(p, cs * 4, cs * 4, below);
        }
        __syncthreads();
    }

    // write the factor back (lower triangle only)
    for (int j = warp; j < nb; j += NW)
        for (int i = j + lane; i < nb; i += 32) D[i + static_cast<int64_t>(j) * ld] = S[i + j * PB_LD];
    __syncthreads();      // the strict upper triangle of S is reused for W^T below

    // ---- W = L^-1, one warp per block column q: W_qq = T_q, W_pq = -T_p * sum_{k=q}^{p-1} L_pk W_kq.
    // W[i][j] (i > j, different micro blocks) lives at S[j + i*PB_LD]; lanes hold entries (i0, j), (i0+4, j).
    double* __restrict__ Wg = Winv + static_cast<int64_t>(t.slot) * PB_MAX * PB_MAX;
    for (int q = warp; q < nmp; q += NW) {
        const int j = lane & 7, i0 = lane >> 3;
        for (int pb = q + 1; pb < nmp; ++pb) {
            double acc0 = 0.0, acc1 = 0.0, acc2 = 0.0, acc3 = 0.0;      // two partial sums per entry (ILP)
#pragma unroll
            for (int k = 0; k < 8; ++k) {                 // k in block q: W_kq = T_q
                const double wkj = (k >= j) ? T[q * 64 + k + 8 * j] : 0.0;
                acc0 += S[(pb * 8 + i0) + (q * 8 + k) * PB_LD] * wkj;
                acc1 += S[(pb * 8 + i0 + 4) + (q * 8 + k) * PB_LD] * wkj;
            }
            for (int kb = q + 1; kb < pb; ++kb)
#pragma unroll
                for (int k = 0; k < 8; k += 2) {
                    const double wkj = S[(q * 8 + j) + (kb * 8 + k) * PB_LD];
                    const double wkj2 = S[(q * 8 + j) + (kb * 8 + k + 1) * PB_LD];
                    acc0 += S[(pb * 8 + i0) + (kb * 8 + k) * PB_LD] * wkj;
                    acc1 += S[(pb * 8 + i0 + 4) + (kb * 8 + k) * PB_LD] * wkj;
                    acc2 += S[(pb * 8 + i0) + (kb * 8 + k + 1) * PB_LD] * wkj2;
                    acc3 += S[(pb * 8 + i0 + 4) + (kb * 8 + k + 1) * PB_LD] * wkj2;
                }
            acc0 += acc2; acc1 += acc3;
            double w0 = 0.0, w1 = 0.0;
#pragma unroll
            for (int k = 0; k < 8; ++k) {                 // Acc[k][j] sits in lane ((k&3)<<3)|j as acc0 (k<4) / acc1
                const double ak = __shfl_sync(0xffffffffu, (k < 4) ? acc0 : acc1, ((k & 3) << 3) | j);
                if (k <= i0) w0 -= T[pb * 64 + i0 + 8 * k] * ak;
                if (k <= i0 + 4) w1 -= T[pb * 64 + (i0 + 4) + 8 * k] * ak;
            }
            S[(q * 8 + j) + (pb * 8 + i0) * PB_LD] = w0;
            S[(q * 8 + j) + (pb * 8 + i0 + 4) * PB_LD] = w1;
            __syncwarp();
        }
    }
    __syncthreads();
    // scratch slot: full PB x PB column-major W (zeros above the diagonal, identity in the padded tail)
    for (int c = warp; c < PB; c += NW)
        for (int r = lane; r < PB; r += 32) {
            double v;
            if (r >= np8 || c >= np8) v = (r == c) ? 1.0 : 0.0;
            else if (r < c) v = 0.0;
            else if ((r >> 3) == (c >> 3)) v = T[(c >> 3) * 64 + (r & 7) + 8 * (c & 7)];
            else v = S[c + r * PB_LD];
            Wg[r + c * PB] = v;
        }
}

// ---------------------------------------------------------------------------------------------
// Large fronts, step 2b: B <- B * L11^-T for the rows below a diagonal block (cblas_dtrsm Right/
// Lower/Trans, reference include/chol.hpp:1292) as a DMMA product with the explicit inverse
// W = L11^-1 from k_potrf_inv: X[i][j] = sum_{k<=j} B[i][k] W[j][k].  One CTA owns a slab of 64
// rows: the slab and W are staged in shared memory, the product runs on the FP64 tensor cores,
// the result replaces the slab in place (each row depends on that row only).
// ---------------------------------------------------------------------------------------------
constexpr int TI_ROWS = 64;
constexpr int TI_LDB = TI_ROWS + 4;   // == 4 (mod 16): conflict-free fragment loads (see G_LDA)
constexpr int TI_KC = 32;             // W is streamed through shared memory in chunks of 32 columns (k)
// the slab stays resident, W does not: 103 KB at PB = 128, so that a CTA of this kernel fits next to a CTA of
// the DMMA update kernel -- with look-ahead the two run concurrently (stream 1 / stream 0)
constexpr int trsm_inv_smem(int pb) { return (pb * TI_LDB + (pb < TI_KC ? pb : TI_KC) * (pb + 4)) * 8; }

template <int PB>
__global__ void __launch_bounds__(256) k_trsm_inv(const TrsmTask* __restrict__ tasks, double* __restrict__ L,
                                                  const double* __restrict__ Winv) {
    extern __shared__ double sm[];
    const TrsmTask t = tasks[blockIdx.x];
    const int row0 = blockIdx.y * TI_ROWS;
    if (row0 >= t.rows) return;
    constexpr int TI_LDW = PB + 4;
    constexpr int KC = PB < TI_KC ? PB : TI_KC;
    double* Bs = sm;                        // Bs[k*TI_LDB + r]  : B[row0 + r][k]
    double* Ws = sm + PB * TI_LDB;          // Ws[(k - k0)*TI_LDW + n]  : W[n][k], k in the current chunk
    const int nb = t.nb, ld = t.ld;
    const int nk = (nb + 7) & ~7;           // padded inner/outer extent (W is identity there, B is loaded as 0)
    double* __restrict__ B = L + t.b_off + row0;
    const double* __restrict__ Wg = Winv + static_cast<int64_t>(t.slot) * PB_MAX * PB_MAX;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int nrows = min(TI_ROWS, t.rows - row0);
    // stage the slab with 8 independent global loads in flight per thread
    for (int base = 0; base < nk * TI_ROWS; base += 256 * 8) {
        double v[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            const int idx = base + u * 256 + tid;
            const int r = idx & (TI_ROWS - 1), k = idx >> 6;
            v[u] = (r < nrows && k < nb) ? B[r + static_cast<int64_t>(k) * ld] : 0.0;
        }
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            const int idx = base + u * 256 + tid;
            const int r = idx & (TI_ROWS - 1), k = idx >> 6;
            if (k < nk) Bs[k * TI_LDB + r] = v[u];
        }
    }
    // 8 warps = 2 (rows) x 4 (cols): warp tile 32 x 32
    const int g = lane >> 2, tq = lane & 3;
    const int wm = warp & 1, wn = warp >> 1;
    const bool active = wn * 32 < nk && wn * 32 < PB;
    double acc[4][4][2];
#pragma unroll
    for (int mi = 0; mi < 4; ++mi)
#pragma unroll
        for (int ni = 0; ni < 4; ++ni) { acc[mi][ni][0] = 0.0; acc[mi][ni][1] = 0.0; }
    for (int k0 = 0; k0 < nk; k0 += KC) {
        __syncthreads();                                  // slab staged (first pass) / previous chunk consumed
        for (int base = 0; base < KC * PB; base += 256 * 8) {
            double v[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                const int idx = base + u * 256 + tid;
                const int n = idx % PB, k = k0 + idx / PB;
                v[u] = (idx < KC * PB && k < nk && n < nk && n >= k) ? Wg[n + k * PB] : 0.0;   // W is lower triangular
            }
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                const int idx = base + u * 256 + tid;
                const int n = idx % PB, kc = idx / PB;
                if (idx < KC * PB) Ws[kc * TI_LDW + n] = v[u];
            }
        }
        __syncthreads();
        if (active && k0 < wn * 32 + 32) {                // W[n][k] = 0 for k > n
            const int kend = min(min(nk, wn * 32 + 32), k0 + KC);
            const double* a = Bs + wm * 32 + g;
            const double* b = Ws + wn * 32 + g;
            for (int kk = k0; kk < kend; kk += 4) {
                double af[4], bf[4];
#pragma unroll
                for (int mi = 0; mi < 4; ++mi) af[mi] = a[(kk + tq) * TI_LDB + mi * 8];
#pragma unroll
                for (int ni = 0; ni < 4; ++ni) bf[ni] = b[(kk - k0 + tq) * TI_LDW + ni * 8];
#pragma unroll
                for (int mi = 0; mi < 4; ++mi)
#pragma unroll
                    for (int ni = 0; ni < 4; ++ni) dmma884(acc[mi][ni][0], acc[mi][ni][1], af[mi], bf[ni]);
            }
        }
    }
    __syncthreads();                                      // everyone is done reading the slab
    if (active) {
#pragma unroll
        for (int mi = 0; mi < 4; ++mi)
#pragma unroll
            for (int ni = 0; ni < 4; ++ni)
#pragma unroll
                for (int e = 0; e < 2; ++e)
                    Bs[(wn * 32 + ni * 8 + 2 * tq + e) * TI_LDB + wm * 32 + mi * 8 + g] = acc[mi][ni][e];
    }
    __syncthreads();
    for (int idx = tid; idx < nb * TI_ROWS; idx += 256) {
        const int r = idx & (TI_ROWS - 1), k = idx >> 6;
        if (r < nrows) B[r + static_cast<int64_t>(k) * ld] = Bs[k * TI_LDB + r];
    }
}

// ---------------------------------------------------------------------------------------------
// Large fronts, step 2c / 3: C -= A * B^T on the FP64 tensor cores (DMMA, mma.sync m8n8k4.f64).
// A: M x K, B: N x K, C: M x N, all column-major; only elements with i + diag >= j are written
// (lower trapezoid) and tiles entirely above that line are skipped.  Used for the panel updates
// of the blocked factorization and for the Schur complement L21 * L21^T (cblas_dsyrk, reference
// include/chol.hpp:1322) whose output is the front's contribution block.
// CTA tile 128 x 64, K step 16, 8 warps (4 x 2), warp tile 32 x 32 = 4 x 4 DMMA tiles.
// grid = (M tiles, N tiles, tasks).
// ---------------------------------------------------------------------------------------------
constexpr int G_BM = 128, G_BN = 64, G_BK = 16;
// leading dimensions == 4 (mod 16) doubles: the 64-bit fragment loads of a half-warp (4 k-rows x 4 consecutive
// elements) then cover 16 distinct 8-byte bank pairs (with == 8 the k-rows tq and tq+2 collide: measured 2x the
// shared-memory wavefronts, -4.5 % on the largest Schur complements)
#ifndef SPCHOL_G_PAD
#define SPCHOL_G_PAD 4
#endif
constexpr int G_LDA = G_BM + SPCHOL_G_PAD, G_LDB = G_BN + SPCHOL_G_PAD;
constexpr int G_SMEM = 2 * G_BK * (G_LDA + G_LDB) * 8;

constexpr int G_SWZ = 8;      // CTAs are dealt to tiles in bands of 8 tile rows (see below)
__global__ void __launch_bounds__(256) k_gemm_dmma(const GemmTask* __restrict__ tasks,
                                                   double* __restrict__ L, double* __restrict__ W, int swizzle /* bit 0: banded tile order, bit 1: C prefetch */) {
    extern __shared__ double sm[];
    const GemmTask t = tasks[blockIdx.z];
    // Tile order.  CTAs start in linear blockIdx order (x fastest); taken literally one wave of 296 CTAs covers
    // ~2 tile columns over ALL tile rows, so every wave streams the whole A panel again (ncu, 128^3 Schur
    // complement with K = 8192: 81 GB of DRAM reads for 2.1 GB of operands).  Dealing the linear index out in
    // bands of G_SWZ tile rows makes a wave cover 8 x ~37 tiles: A and B rows touched per wave drop ~5x.
    int bx = blockIdx.x, by = blockIdx.y;
    if (swizzle & 1) {
        const int lin = blockIdx.x + blockIdx.y * gridDim.x;
        const int band = lin / (G_SWZ * gridDim.y);
        const int rows = min(G_SWZ, static_cast<int>(gridDim.x) - band * G_SWZ);
        const int within = lin - band * G_SWZ * gridDim.y;
        bx = band * G_SWZ + within % rows;
        by = within / rows;
    }
    const int i0 = bx * G_BM, j0 = by * G_BN;
    if (i0 >= t.M || j0 >= t.N) return;
    if (i0 + G_BM - 1 + t.diag < j0) return;          // tile entirely above the wanted trapezoid
    double* As = sm;                                   // [2][G_BK][G_LDA]
    double* Bs = sm + 2 * G_BK * G_LDA;                // [2][G_BK][G_LDB]
    const double* __restrict__ A = L + t.a_off;
    const double* __restrict__ B = L + t.b_off;
    double* __restrict__ C = ((t.c_space & 1) ? W : L) + t.c_off;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int g = lane >> 2, tq = lane & 3;
    const int wm = warp & 3, wn = warp >> 2;
    const int K = t.K, lda = t.lda, ldb = t.ldb;

    if ((swizzle & 2) && K <= 1024 && !(t.c_space & 2)) {
        // short K: the read-modify-write of the C tile at the end is a visible part of the CTA's life and its loads
        // miss L2 (C is the trailing matrix, streamed from HBM) -- pull the tile towards L2 now, 2 lines per thread
        const int pc = tid >> 2, pr = (tid & 3) * 32;                  // column of the tile, 32-row segment (2 x 128 B)
        if (j0 + pc < t.N && i0 + pr < t.M) {
            const double* cp = C + (i0 + pr) + static_cast<int64_t>(j0 + pc) * t.ldc;
            asm volatile("prefetch.global.L2 [%0];" ::"l"(cp));
            if (i0 + pr + 16 < t.M) asm volatile("prefetch.global.L2 [%0];" ::"l"(cp + 16));
        }
    }
    const int arow = tid & 127, ak = tid >> 7;         // A: 8 loads per thread, k = ak + 2 r
    const int brow = tid & 63, bk = tid >> 6;          // B: 4 loads per thread, k = bk + 4 r
    const bool a_ok = (i0 + arow) < t.M, b_ok = (j0 + brow) < t.N;
    const double* Ag = A + (i0 + arow);
    const double* Bg = B + (j0 + brow);
    double ra[8], rb[4];
    auto gload = [&](int kt) {
#pragma unroll
        for (int r = 0; r < 8; ++r) {
            const int k = kt + ak + 2 * r;
            ra[r] = (a_ok && k < K) ? Ag[static_cast<int64_t>(k) * lda] : 0.0;
        }
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            const int k = kt + bk + 4 * r;
            rb[r] = (b_ok && k < K) ? Bg[static_cast<int64_t>(k) * ldb] : 0.0;
        }
    };
    auto sstore = [&](int buf) {
        double* a = As + buf * G_BK * G_LDA;
        double* b = Bs + buf * G_BK * G_LDB;
#pragma unroll
        for (int r = 0; r < 8; ++r) a[(ak + 2 * r) * G_LDA + arow] = ra[r];
#pragma unroll
        for (int r = 0; r < 4; ++r) b[(bk + 4 * r) * G_LDB + brow] = rb[r];
    };

    double acc[4][4][2];
#pragma unroll
    for (int mi = 0; mi < 4; ++mi)
#pragma unroll
        for (int ni = 0; ni < 4; ++ni) { acc[mi][ni][0] = 0.0; acc[mi][ni][1] = 0.0; }

    gload(0);
    sstore(0);
    __syncthreads();
    int buf = 0;
    for (int kt = 0; kt < K; kt += G_BK) {
        const bool more = kt + G_BK < K;
        if (more) gload(kt + G_BK);
        const double* a = As + buf * G_BK * G_LDA + wm * 32 + g;
        const double* b = Bs + buf * G_BK * G_LDB + wn * 32 + g;
#pragma unroll
        for (int kk = 0; kk < G_BK; kk += 4) {
            double af[4], bf[4];
#pragma unroll
            for (int mi = 0; mi < 4; ++mi) af[mi] = a[(kk + tq) * G_LDA + mi * 8];
#pragma unroll
            for (int ni = 0; ni < 4; ++ni) bf[ni] = b[(kk + tq) * G_LDB + ni * 8];
#pragma unroll
            for (int mi = 0; mi < 4; ++mi)
#pragma unroll
                for (int ni = 0; ni < 4; ++ni) dmma884(acc[mi][ni][0], acc[mi][ni][1], af[mi], bf[ni]);
        }
        if (more) { sstore(buf ^ 1); }
        __syncthreads();
        buf ^= 1;
    }

    const int ldc = t.ldc;
    const bool beta0 = (t.c_space & 2) != 0;
#pragma unroll
    for (int mi = 0; mi < 4; ++mi) {
        const int row = i0 + wm * 32 + mi * 8 + g;
        if (row >= t.M) continue;
#pragma unroll
        for (int ni = 0; ni < 4; ++ni) {
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                const int col = j0 + wn * 32 + ni * 8 + 2 * tq + e;
                if (col < t.N && row + t.diag >= col) {
                    double* c = C + row + static_cast<int64_t>(col) * ldc;
                    if (beta0) *c = -acc[mi][ni][e]; else *c -= acc[mi][ni][e];
                }
            }
        }
    }
}

}  // namespace spchol
