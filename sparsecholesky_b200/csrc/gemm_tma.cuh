// k_gemm_tma: C -= A * B^T on the FP64 tensor cores (DMMA, mma.sync m8n8k4.f64 -- tcgen05 has no f64 kind), operands
// staged by TMA: cp.async.bulk.tensor.2d loads of K-slabs into a ring of 128B-swizzled shared-memory stages, completion
// through mbarriers (complete_tx), no register staging.  Same contract as k_gemm_dmma (kernels.cuh): A: M x K, B: N x K,
// C: M x N column-major, only elements with i + diag >= j are written, tiles entirely above that line are skipped.
// Replaces cblas_dsyrk of the reference (include/chol.hpp:1322-1331) and the blocked panel updates.
//
// A and B are row ranges of ONE front's panel; each front with update tasks has a tensor map (m rows x w columns of
// doubles, row pitch ld): rows past m and columns past w are out of bounds and arrive as zeros, so ragged tiles need no
// predicates on the load side.  A K-slab of a tile is moved as boxes of 16 rows x G3_BK columns; inside a box the 128-byte
// line of one k holds 16 consecutive rows, and the hardware 128B swizzle XORs the 16-byte chunk index with (k mod 8).
// The 8 rows of a DMMA fragment are taken as rows {0,1,8,9,2,3,10,11} (+4 for the second fragment) of a box: with that
// choice the 16 lanes of a half-warp (4 rows x 4 k) hit 16 distinct 8-byte bank pairs -- conflict-free without padding.
#pragma once
#include <cstdint>
#include <type_traits>
#include <cuda.h>
#include <cuda_runtime.h>

#include "plan.h"

namespace spchol {

constexpr int G3_BM = 128, G3_BN = 64, G3_BK = 16, G3_ST = 4;      // CTA tile, K-slab, pipeline stages
constexpr int G3_BOX = 16 * G3_BK * 8;                             // bytes of one 16-row box
constexpr int G3_STAGE = (G3_BM + G3_BN) / 16 * G3_BOX;            // bytes per stage (A boxes, then B boxes)
constexpr int G3_SMEM = G3_ST * G3_STAGE + 1024 + 64;              // + alignment slack + barriers (full and empty per stage)
constexpr int G3_SWZ = 8;                                          // banded tile order, as k_gemm_dmma

__device__ __forceinline__ unsigned g3_smem_u32(const void* p) { return static_cast<unsigned>(__cvta_generic_to_shared(p)); }
__device__ __forceinline__ void g3_mbar_init(unsigned bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void g3_mbar_expect_tx(unsigned bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void g3_mbar_arrive(unsigned bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void g3_mbar_wait(unsigned bar, unsigned parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void g3_tma_load_2d(unsigned dst, const CUtensorMap* map, int row, int col, unsigned bar) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                 ::"r"(dst), "l"(map), "r"(row), "r"(col), "r"(bar) : "memory");
}
__device__ __forceinline__ void g3_dmma(double& c0, double& c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                 : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
// fragment element from a stage: an explicit ld.shared with a 32-bit address (a plain dereference compiles to generic
// 64-bit loads here: the compiler does not see through the ring arithmetic that the pointer is shared memory)
__device__ __forceinline__ double g3_lds(unsigned addr) {
    double v;
    asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(addr));
    return v;
}
// row (0..15) inside a box of fragment row g (0..7) of the h-th (0/1) 8-row fragment
__device__ __forceinline__ int g3_box_row(int h, int g) { return 4 * h + (g & 1) + 2 * (g >> 2) + 8 * ((g >> 1) & 1); }

// grid = (M tiles, N tiles, tasks), 256 threads: 8 warps = 4 (rows) x 2 (cols), warp tile 32 x 32 = 4 x 4 DMMA tiles
__global__ void __launch_bounds__(256, 2) k_gemm_tma(const GemmTask* __restrict__ tasks, const CUtensorMap* __restrict__ tmaps,
                                                     double* __restrict__ L, double* __restrict__ W, int swizzle) {
    extern __shared__ unsigned char g3_raw[];
    const GemmTask t = tasks[blockIdx.z];
    int bx = blockIdx.x, by = blockIdx.y;
    if (swizzle & 1) {                                 // bands of G3_SWZ tile rows: a wave of CTAs shares its A and B slabs in L2
        const int lin = blockIdx.x + blockIdx.y * gridDim.x;
        const int band = lin / (G3_SWZ * gridDim.y);
        const int rows = min(G3_SWZ, static_cast<int>(gridDim.x) - band * G3_SWZ);
        const int within = lin - band * G3_SWZ * gridDim.y;
        bx = band * G3_SWZ + within % rows;
        by = within / rows;
    }
    // TMA moves 16-byte granules along the contiguous dimension: the first row of a box must be even.  A range that
    // starts at an odd panel row (w odd: the Schur complement starts at row w) is tiled from the row above instead -- the
    // tile grid is shifted by one, the extra leading row (-1) is computed and never stored.
    const int i0 = bx * G3_BM - (t.ra & 1), j0 = by * G3_BN - (t.rb & 1);
    if (i0 >= t.M || j0 >= t.N) return;
    if (i0 + G3_BM - 1 + t.diag < j0) return;          // tile entirely above the wanted trapezoid
    // stages: 1024-byte aligned (the 128B swizzle pattern repeats every 8 lines); barriers behind them
    unsigned char* stage0 = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(g3_raw) + 1023) & ~uintptr_t(1023));
    const unsigned sbase = g3_smem_u32(stage0);
    const unsigned bar0 = sbase + G3_ST * G3_STAGE;
    const unsigned ebar0 = bar0 + 8 * G3_ST;           // "stage is free again" barriers (swizzle & 8), one arrival per warp
    const bool free_bars = (swizzle & 8) != 0;
    const CUtensorMap* map = tmaps + t.tmap;
    double* __restrict__ C = ((t.c_space & 1) ? W : L) + t.c_off;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int g = lane >> 2, tq = lane & 3;
    const int wm = warp & 3, wn = warp >> 2;
    const int K = t.K;
    const int nk = (K + G3_BK - 1) / G3_BK;

    auto issue = [&](int s, int kt) {                  // one thread: arm the stage's barrier, then 8 + 4 box loads
        const unsigned bar = bar0 + 8 * s, dst = sbase + s * G3_STAGE;
        g3_mbar_expect_tx(bar, G3_STAGE);
        const int kc = t.kc + kt * G3_BK;
#pragma unroll
        for (int b = 0; b < G3_BM / 16; ++b) g3_tma_load_2d(dst + b * G3_BOX, map, t.ra + i0 + 16 * b, kc, bar);
#pragma unroll
        for (int b = 0; b < G3_BN / 16; ++b) g3_tma_load_2d(dst + (G3_BM / 16 + b) * G3_BOX, map, t.rb + j0 + 16 * b, kc, bar);
    };
    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < G3_ST; ++s) { g3_mbar_init(bar0 + 8 * s, 1); g3_mbar_init(ebar0 + 8 * s, 8); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        for (int s = 0; s < G3_ST && s < nk; ++s) issue(s, s);
    }
    if ((swizzle & 2) && K <= 1024 && !(t.c_space & 2)) {
        // short K: the read-modify-write of the C tile at the end is a visible part of the CTA's life -- pull it towards L2 now
        const int pc = tid >> 2, pr = (tid & 3) * 32;
        if (j0 + pc >= 0 && j0 + pc < t.N && i0 + pr >= 0 && i0 + pr < t.M) {
            const double* cp = C + (i0 + pr) + static_cast<int64_t>(j0 + pc) * t.ldc;
            asm volatile("prefetch.global.L2 [%0];" ::"l"(cp));
            if (i0 + pr + 16 < t.M) asm volatile("prefetch.global.L2 [%0];" ::"l"(cp + 16));
        }
    }
    __syncthreads();                                   // barriers initialised before anyone waits on them

    // byte offset of this lane's fragment element inside a box, per fragment h and k-group q (k = 4 q + tq)
    int foff[2][G3_BK / 4];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const int r = g3_box_row(h, g);
#pragma unroll
        for (int q = 0; q < G3_BK / 4; ++q) {
            const int k = 4 * q + tq;
            foff[h][q] = k * 128 + ((((r >> 1) ^ (k & 7))) << 4) + (r & 1) * 8;
        }
    }
    double acc[4][4][2];
#pragma unroll
    for (int mi = 0; mi < 4; ++mi)
#pragma unroll
        for (int ni = 0; ni < 4; ++ni) { acc[mi][ni][0] = 0.0; acc[mi][ni][1] = 0.0; }

    // one K-slab of the ring: TAIL = the last slab of a K that is not a multiple of G3_BK (the panel goes on beyond K, so
    // the surplus columns are masked; full slabs carry no predicates)
    auto slab = [&](int kt, auto tail_c) {
        constexpr bool TAIL = decltype(tail_c)::value;
        const int s = kt % G3_ST;
        g3_mbar_wait(bar0 + 8 * s, (kt / G3_ST) & 1);
        const unsigned as = sbase + s * G3_STAGE + (wm * 2) * G3_BOX;
        const unsigned bs = sbase + s * G3_STAGE + (G3_BM / 16 + wn * 2) * G3_BOX;
        const int krem = K - kt * G3_BK;
#pragma unroll
        for (int q = 0; q < G3_BK / 4; ++q) {
            double af[4], bf[4];
            const bool kon = !TAIL || 4 * q + tq < krem;
#pragma unroll
            for (int mi = 0; mi < 4; ++mi) af[mi] = kon ? g3_lds(as + (mi >> 1) * G3_BOX + foff[mi & 1][q]) : 0.0;
#pragma unroll
            for (int ni = 0; ni < 4; ++ni) bf[ni] = kon ? g3_lds(bs + (ni >> 1) * G3_BOX + foff[ni & 1][q]) : 0.0;
#pragma unroll
            for (int mi = 0; mi < 4; ++mi)
#pragma unroll
                for (int ni = 0; ni < 4; ++ni) g3_dmma(acc[mi][ni][0], acc[mi][ni][1], af[mi], bf[ni]);
        }
        if (!free_bars) {
            __syncthreads();                           // everyone is done with stage s: refill it with the slab G3_ST ahead
            if (tid == 0 && kt + G3_ST < nk) issue(s, kt + G3_ST);
        } else {
            // no CTA-wide barrier in the loop: every warp reports the stage free, and the refill runs ONE slab behind (the
            // stage of slab kt - 1, which the other warps left long ago), so the issuing thread hardly ever waits and the
            // warps may drift up to two slabs apart
            __syncwarp();
            if (lane == 0) g3_mbar_arrive(ebar0 + 8 * s);
            if (tid == 0 && kt >= 1 && kt - 1 + G3_ST < nk) {
                const int p = kt - 1, sp = p % G3_ST;
                g3_mbar_wait(ebar0 + 8 * sp, (p / G3_ST) & 1);
                issue(sp, p + G3_ST);
            }
        }
    };
    const int nk_full = K / G3_BK;
    for (int kt = 0; kt < nk_full; ++kt) slab(kt, std::false_type{});
    if (nk_full < nk) slab(nk_full, std::true_type{});

    const int ldc = t.ldc;
    const bool beta0 = (t.c_space & 2) != 0;
#pragma unroll
    for (int mi = 0; mi < 4; ++mi) {
        const int row = i0 + wm * 32 + (mi >> 1) * 16 + g3_box_row(mi & 1, g);
        if (row < 0 || row >= t.M) continue;
#pragma unroll
        for (int ni = 0; ni < 4; ++ni) {
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                const int col = j0 + wn * 32 + (ni >> 1) * 16 + g3_box_row(ni & 1, 2 * tq + e);
                if (col >= 0 && col < t.N && row + t.diag >= col) {
                    double* c = C + row + static_cast<int64_t>(col) * ldc;
                    if (beta0) *c = -acc[mi][ni][e];
                    else if (swizzle & 4) atomicAdd(c, -acc[mi][ni][e]);      // result unused: RED.ADD.F64 at the L2, no load to wait for
                    else *c -= acc[mi][ni][e];                                  // (every element has exactly one writer per launch: same value either way)
                }
            }
        }
    }
}

}  // namespace spchol
