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
__global__ void k_assemble(int64_t nnzA, const int64_t* __restrict__ a_dest,
                           const double* __restrict__ Ax, double* __restrict__ L) {
    const int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
    if (i < nnzA) { const int64_t d = a_dest[i]; if (d >= 0) L[d] = Ax[i]; }   // d < 0: front held by another rank
}

// gather the CSC values of L out of the supernodal panels (replaces
// scatter_panel_column_into_L, reference include/chol.hpp:1178-1189)
__global__ void k_export(int64_t nnzL, const int64_t* __restrict__ l_src,
                         const double* __restrict__ L, double* __restrict__ Lx) {
    const int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
    if (i < nnzL) { const int64_t q = l_src[i]; Lx[i] = q >= 0 ? L[q] : 0.0; }   // q < 0: exported by another rank
}

__global__ void k_zero(const ZeroTask* __restrict__ tasks, double* __restrict__ W) {
    const ZeroTask t = tasks[blockIdx.y];
    double* p = W + t.off;
    for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < t.count;
         i += static_cast<int64_t>(gridDim.x) * blockDim.x)
        p[i] = 0.0;
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
template <int T>
__global__ void __launch_bounds__(T) k_front_small(const int* __restrict__ list, FrontArrays fa,
                                                   double* __restrict__ L, double* __restrict__ W,
                                                   int* __restrict__ info) {
    extern __shared__ double F[];
    const int s = list[blockIdx.x];
    const int w = fa.w[s], m = fa.m[s], ld = fa.ld[s];
    const int ldf = m | 1;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr int NW = T / 32;
    double* __restrict__ P = L + fa.panel_off[s];

    for (int j = warp; j < m; j += NW) {
        double* col = F + j * ldf;
        if (j < w) {
            const double* src = P + static_cast<int64_t>(j) * ld;
            for (int i = lane; i < m; i += 32) col[i] = (i >= j) ? src[i] : 0.0;
        } else {
            for (int i = lane; i < m; i += 32) col[i] = 0.0;
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
            double* col = F + r[jj] * ldf;
            const double* src = C + static_cast<int64_t>(jj) * ldc;
            for (int ii = jj + lane; ii < mc; ii += 32) col[r[ii]] += src[ii];
        }
        __syncthreads();
    }

    // partial factorization of columns 0..w-1
    for (int k = 0; k < w; ++k) {
        const double d = F[k + k * ldf];
        // reference test: d <= T(0) (chol.hpp:849); info = smallest failing 1-based column
        if (d <= 0.0 && tid == 0) atomicMin(info, fa.first[s] + k + 1);
        const double piv = sqrt(d);
        double* colk = F + k * ldf;
        for (int i = k + 1 + tid; i < m; i += T) colk[i] = colk[i] / piv;
        __syncthreads();
        if (tid == 0) colk[k] = piv;
        for (int j = k + 1 + warp; j < m; j += NW) {
            const double uj = colk[j];
            double* col = F + j * ldf;
            for (int i = j + lane; i < m; i += 32) col[i] -= colk[i] * uj;
        }
        __syncthreads();
    }

    for (int j = warp; j < m; j += NW) {
        const double* col = F + j * ldf;
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
__global__ void __launch_bounds__(128) k_extend_add(const int* __restrict__ list, FrontArrays fa,
                                                    double* __restrict__ L, double* __restrict__ W) {
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
        if (G > 1) {
            const int blk = cj < wp ? cj / DNB : nbp + (cj - wp) / DNB;
            if (blk % G != me) continue;
        }
        const double* src = C + static_cast<int64_t>(jj) * ldc;
        if (cj < wp) {
            double* dst = Pp + static_cast<int64_t>(cj) * ldp;
            for (int ii = jj + threadIdx.x; ii < mc; ii += blockDim.x) dst[r[ii]] += src[ii];
        } else {
            double* dst = Cp + static_cast<int64_t>(cj - wp) * ldcb - wp;
            for (int ii = jj + threadIdx.x; ii < mc; ii += blockDim.x) dst[r[ii]] += src[ii];
        }
    }
}

// ---------------------------------------------------------------------------------------------
// Large fronts, step 2a: Cholesky of one nb x nb (nb <= 64) diagonal block, in shared memory.
// (dpotrf_ of the reference, include/chol.hpp:1263, one block of the blocked algorithm.)
// ---------------------------------------------------------------------------------------------
constexpr int POTRF_NB = 64;
__global__ void __launch_bounds__(256) k_potrf_diag(const PotrfTask* __restrict__ tasks,
                                                    double* __restrict__ L, int* __restrict__ info) {
    __shared__ double S[POTRF_NB * (POTRF_NB + 1)];
    const PotrfTask t = tasks[blockIdx.x];
    const int nb = t.nb, ld = t.ld;
    constexpr int lds = POTRF_NB + 1;
    double* __restrict__ D = L + t.d_off;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int j = warp; j < nb; j += 8)
        for (int i = j + lane; i < nb; i += 32) S[i + j * lds] = D[i + static_cast<int64_t>(j) * ld];
    __syncthreads();
    for (int k = 0; k < nb; ++k) {
        const double d = S[k + k * lds];
        if (d <= 0.0 && tid == 0) atomicMin(info, t.col0 + k + 1);
        const double piv = sqrt(d);
        double* colk = S + k * lds;
        for (int i = k + 1 + tid; i < nb; i += 256) colk[i] = colk[i] / piv;
        __syncthreads();
        if (tid == 0) colk[k] = piv;
        for (int j = k + 1 + warp; j < nb; j += 8) {
            const double uj = colk[j];
            for (int i = j + lane; i < nb; i += 32) S[i + j * lds] -= colk[i] * uj;
        }
        __syncthreads();
    }
    for (int j = warp; j < nb; j += 8)
        for (int i = j + lane; i < nb; i += 32) D[i + static_cast<int64_t>(j) * ld] = S[i + j * lds];
}

// ---------------------------------------------------------------------------------------------
// Large fronts, step 2b: B <- B * D^{-T} for the rows below a diagonal block (cblas_dtrsm Right/
// Lower/Trans, reference include/chol.hpp:1292).  One thread owns one row (64 values in
// registers, coalesced column accesses across the CTA); D is broadcast from shared memory.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128) k_trsm(const TrsmTask* __restrict__ tasks, double* __restrict__ L) {
    __shared__ double S[POTRF_NB * (POTRF_NB + 1)];
    const TrsmTask t = tasks[blockIdx.x];
    const int nb = t.nb, ld = t.ld;
    constexpr int lds = POTRF_NB + 1;
    const int row0 = blockIdx.y * 128;
    if (row0 >= t.rows) return;
    const double* __restrict__ D = L + t.d_off;
    for (int idx = threadIdx.x; idx < POTRF_NB * POTRF_NB; idx += 128) {
        const int i = idx % POTRF_NB, j = idx / POTRF_NB;
        double v = (i == j) ? 1.0 : 0.0;             // identity padding when nb < 64
        if (i < nb && j < nb && i >= j) v = D[i + static_cast<int64_t>(j) * ld];
        S[i + j * lds] = v;
    }
    __syncthreads();
    const int row = row0 + threadIdx.x;
    if (row >= t.rows) return;
    double* __restrict__ B = L + t.b_off + row;
    double r[POTRF_NB];
#pragma unroll
    for (int j = 0; j < POTRF_NB; ++j) r[j] = (j < nb) ? B[static_cast<int64_t>(j) * ld] : 0.0;
#pragma unroll
    for (int k = 0; k < POTRF_NB; ++k) {
        const double v = r[k] / S[k + k * lds];
        r[k] = v;
#pragma unroll
        for (int j = k + 1; j < POTRF_NB; ++j) r[j] -= v * S[j + k * lds];
    }
#pragma unroll
    for (int j = 0; j < POTRF_NB; ++j) if (j < nb) B[static_cast<int64_t>(j) * ld] = r[j];
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
constexpr int G_LDA = G_BM + 8, G_LDB = G_BN + 8;   // == 8 (mod 16) doubles: conflict-free fragment loads
constexpr int G_SMEM = 2 * G_BK * (G_LDA + G_LDB) * 8;

__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                 : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

__global__ void __launch_bounds__(256) k_gemm_dmma(const GemmTask* __restrict__ tasks,
                                                   double* __restrict__ L, double* __restrict__ W) {
    extern __shared__ double sm[];
    const GemmTask t = tasks[blockIdx.z];
    const int i0 = blockIdx.x * G_BM, j0 = blockIdx.y * G_BN;
    if (i0 >= t.M || j0 >= t.N) return;
    if (i0 + G_BM - 1 + t.diag < j0) return;          // tile entirely above the wanted trapezoid
    double* As = sm;                                   // [2][G_BK][G_LDA]
    double* Bs = sm + 2 * G_BK * G_LDA;                // [2][G_BK][G_LDB]
    const double* __restrict__ A = L + t.a_off;
    const double* __restrict__ B = L + t.b_off;
    double* __restrict__ C = (t.c_space ? W : L) + t.c_off;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int g = lane >> 2, tq = lane & 3;
    const int wm = warp & 3, wn = warp >> 2;
    const int K = t.K, lda = t.lda, ldb = t.ldb;

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
                    *c -= acc[mi][ni][e];
                }
            }
        }
    }
}

}  // namespace spchol
