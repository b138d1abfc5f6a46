// k_tree_warp: ONE WARP factors a whole subtree of the assembly tree (or, batched, the whole tree of one
// matrix), front after front, with the frontal matrix in shared memory; see tree_plan.h for the program and
// the reversed packed layout that lets a contribution block pass to the parent without moving.
// Fuses dpotrf_ + dtrsm + dsyrk + apply_update of the reference's supernodal chain
// (include/chol.hpp:1232-1335, :1196-1216) for every front of the subtree; nothing but the factor
// (and the blocks of side branches) leaves the SM.  No CTA barrier: only __syncwarp.
//
// The body is written against a few macros so that tests/cpp/tree_emu.cpp runs the SAME source on
// 32 host threads per warp (SPCHOL_HOST_EMU) against the oracle, without a GPU.
#pragma once
#include <cmath>
#include <cstdint>

#include "tree_plan.h"

#ifdef SPCHOL_HOST_EMU
#define TK_DEVICE inline
#define TK_LANE (::spchol::emu::tid() & 31)
#define TK_SYNCWARP() ::spchol::emu::syncwarp()
#define TK_LDG(p) (*(p))
#define TK_ATOMIC_MIN(p, v) ::spchol::emu::atomic_min((p), (v))
#define TK_RESTRICT
#define TK_SQRT(x) std::sqrt(x)
#else
#define TK_DEVICE __device__ __forceinline__
#define TK_LANE static_cast<int>(threadIdx.x & 31)
#define TK_SYNCWARP() __syncwarp()
#define TK_LDG(p) __ldg(p)
#define TK_ATOMIC_MIN(p, v) atomicMin((p), (v))
#define TK_RESTRICT __restrict__
#define TK_SQRT(x) sqrt(x)
#endif

namespace spchol {

struct TreeArrays {                   // device pointers
    const TwTask* tasks; int ntasks;
    const TwFront* fronts;
    const TwStash* stashes;
    const int* rmap;
    const int* a_src;
    const unsigned short* a_dst;
    const int* csc_map;               // out_mode 1: L-buffer offset of a panel entry -> index into Lx (CSC order), -1 for padding
};

TK_DEVICE int tk_tri(int x) { return (x * (x + 1)) >> 1; }

// rank-WK update of the contribution block (reversed columns 0..mb-1) by the pivot columns k0..k0+WK-1:
// 4 target columns per step share the multipliers of a 32-row chunk; every entry is owned by lane (row % 32).
template <int WK>
TK_DEVICE void tk_update(double* F, int m, int mb, int k0, int lane) {
    const double* pc[WK];
#pragma unroll
    for (int q = 0; q < WK; ++q) pc[q] = F + tk_tri(m - 1 - k0 - q);
    for (int c0 = 0; c0 < mb; c0 += 4) {
        double cm[4][WK];
        int bc[4];
#pragma unroll
        for (int t = 0; t < 4; ++t) {
            const int cc = c0 + t < mb ? c0 + t : mb - 1;
            bc[t] = tk_tri(cc);
#pragma unroll
            for (int q = 0; q < WK; ++q) cm[t][q] = pc[q][cc];
        }
        const int cmax = c0 + 3 < mb ? c0 + 3 : mb - 1;
        for (int r0 = 0; r0 <= cmax; r0 += 32) {
            const int ir = r0 + lane;
            double lr[WK];
#pragma unroll
            for (int q = 0; q < WK; ++q) lr[q] = ir <= cmax ? pc[q][ir] : 0.0;
            if (r0 + 31 <= c0 && c0 + 3 < mb) {           // warp-uniform: the chunk lies inside all four columns
#pragma unroll
                for (int t = 0; t < 4; ++t) {
                    double acc = F[bc[t] + ir];
#pragma unroll
                    for (int q = 0; q < WK; ++q) acc = fma(-lr[q], cm[t][q], acc);
                    F[bc[t] + ir] = acc;
                }
            } else {
#pragma unroll
                for (int t = 0; t < 4; ++t) {
                    if (c0 + t < mb && ir <= c0 + t) {
                        double acc = F[bc[t] + ir];
#pragma unroll
                        for (int q = 0; q < WK; ++q) acc = fma(-lr[q], cm[t][q], acc);
                        F[bc[t] + ir] = acc;
                    }
                }
            }
        }
    }
}

// One task = fronts [front0, front0 + nfront) in postorder.  F: tri(cap_m) doubles, sinv: TW_MAX_M ints, both private
// to the warp.  out_mode 0: panels go to the L buffer; 1: entries go straight to Lx in CSC order through csc_map.
TK_DEVICE void tree_warp_body(const TreeArrays& ta, const TwTask task, const double* TK_RESTRICT Ax, double* TK_RESTRICT L,
                              double* TK_RESTRICT W, double* TK_RESTRICT Stash, double* TK_RESTRICT Lx, int* TK_RESTRICT info,
                              double* F, int* sinv, int out_mode) {
    const int lane = TK_LANE;
    Stash += task.stash_base;
    for (int f = 0; f < task.nfront; ++f) {
        const TwFront d = ta.fronts[task.front0 + f];
        const int m = d.m, w = d.w, mb = m - w;
        // ---- 1. the frontal matrix: spread the carried block in place, or start from zero
        if (d.flags & TW_CARRY_IN) {
            const int* TK_RESTRICT rm = ta.rmap + d.map_off;
            for (int t = lane; t < m; t += 32) sinv[t] = -1;
            TK_SYNCWARP();
            for (int i = lane; i < d.mc; i += 32) sinv[TK_LDG(rm + i)] = i;
            TK_SYNCWARP();
            for (int J = m - 1; J >= 0; --J) {
                const int j = sinv[J];
                if (j == J) break;                         // the map is the identity from here down: nothing moves
                double v[4];
                const int bj = j >= 0 ? tk_tri(j) : 0;
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const int t = lane + 32 * u;
                    v[u] = 0.0;
                    if (t <= J && j >= 0) { const int src = sinv[t]; if (src >= 0) v[u] = F[bj + src]; }
                }
                TK_SYNCWARP();                             // column j is read before column J (which may overlap it) is written
                const int bJ = tk_tri(J);
#pragma unroll
                for (int u = 0; u < 4; ++u) { const int t = lane + 32 * u; if (t <= J) F[bJ + t] = v[u]; }
            }
        } else {
            const int ne = tk_tri(m);
            for (int e = lane; e < ne; e += 32) F[e] = 0.0;
        }
        TK_SYNCWARP();
        // ---- 2. the front's entries of A (distinct targets)
        for (int e = lane; e < d.an; e += 32) F[TK_LDG(ta.a_dst + d.a0 + e)] += Ax[TK_LDG(ta.a_src + d.a0 + e)];
        // ---- 3. blocks of the other children, from the stash (most recent first)
        for (int q = 0; q < d.nstash; ++q) {
            TK_SYNCWARP();
            const TwStash st = ta.stashes[d.stash0 + q];
            const int* TK_RESTRICT rm = ta.rmap + st.map_off;
            for (int i = lane; i < st.mc; i += 32) sinv[i] = TK_LDG(rm + i);
            TK_SYNCWARP();
            const double* TK_RESTRICT src = Stash + st.src;
            for (int jr = 0; jr < st.mc; ++jr) {
                const int bJ = tk_tri(sinv[jr]);
                const double* col = src + tk_tri(jr);
                for (int ir = lane; ir <= jr; ir += 32) F[bJ + sinv[ir]] += col[ir];
            }
        }
        TK_SYNCWARP();
        // ---- 4. the w pivot columns (reversed columns m-1 .. m-w), right-looking inside the panel
        for (int k = 0; k < w; ++k) {
            const int Kr = m - 1 - k;
            double* colK = F + tk_tri(Kr);
            const double dd = colK[Kr];
            if (dd <= 0.0 && lane == 0) TK_ATOMIC_MIN(info, d.first + k + 1);      // reference test (chol.hpp:849)
            const double piv = TK_SQRT(dd);
            const double rinv = 1.0 / piv;
            TK_SYNCWARP();                                 // every lane has read the pivot
            for (int i = lane; i < Kr; i += 32) colK[i] *= rinv;
            if (lane == 0) colK[Kr] = piv;
            TK_SYNCWARP();
            for (int Jr = Kr - 1; Jr >= mb; --Jr) {
                const double uj = colK[Jr];
                double* colJ = F + tk_tri(Jr);
                for (int i = lane; i <= Jr; i += 32) colJ[i] = fma(-colK[i], uj, colJ[i]);
            }
            TK_SYNCWARP();
        }
        // ---- 5. contribution block -= L21 L21^T, eight pivots at a time
        if (mb > 0) {
            for (int k0 = 0; k0 < w; k0 += 8) {
                switch (w - k0 < 8 ? w - k0 : 8) {
                case 1: tk_update<1>(F, m, mb, k0, lane); break;
                case 2: tk_update<2>(F, m, mb, k0, lane); break;
                case 3: tk_update<3>(F, m, mb, k0, lane); break;
                case 4: tk_update<4>(F, m, mb, k0, lane); break;
                case 5: tk_update<5>(F, m, mb, k0, lane); break;
                case 6: tk_update<6>(F, m, mb, k0, lane); break;
                case 7: tk_update<7>(F, m, mb, k0, lane); break;
                default: tk_update<8>(F, m, mb, k0, lane); break;
                }
            }
        }
        TK_SYNCWARP();
        // ---- 6. the factored columns leave the SM
        for (int c = 0; c < w; ++c) {
            const int Jr = m - 1 - c;
            const double* col = F + tk_tri(Jr);
            const int64_t dst0 = d.panel_off + static_cast<int64_t>(c) * d.ld + (m - 1);
            if (out_mode == 0) {
                for (int ir = lane; ir <= Jr; ir += 32) L[dst0 - ir] = col[ir];
            } else {
                for (int ir = lane; ir <= Jr; ir += 32) { const int idx = TK_LDG(ta.csc_map + dst0 - ir); if (idx >= 0) Lx[idx] = col[ir]; }
            }
        }
        // ---- 7. the contribution block: stays (the parent is next), or goes to the level program's slot / the stash
        if (mb > 0 && !(d.flags & TW_CARRY_OUT) && d.cb_out >= 0) {
            if (d.flags & TW_CB_STD) {
                const int ldcb = (mb + 1) & ~1;
                double* TK_RESTRICT C = W + d.cb_out;
                for (int jr = 0; jr < mb; ++jr) {
                    const double* col = F + tk_tri(jr);
                    double* dcol = C + static_cast<int64_t>(mb - 1 - jr) * ldcb + (mb - 1);
                    for (int ir = lane; ir <= jr; ir += 32) dcol[-ir] = col[ir];
                }
            } else {
                double* TK_RESTRICT dst = Stash + d.cb_out;
                const int ne = tk_tri(mb);
                for (int e = lane; e < ne; e += 32) dst[e] = F[e];
            }
        }
        TK_SYNCWARP();
    }
}

#ifndef SPCHOL_HOST_EMU
// blockDim.x = 32 * warps per CTA; dynamic shared memory = warps * (f_stride doubles + TW_MAX_M ints).
// nmat > 0 (batched): warp g factors matrix g with task 0, all pointers advance by their per-matrix strides;
// nmat == 0: warp g runs task g of a single matrix.
__global__ void k_tree_warp(TreeArrays ta, const double* __restrict__ Ax, double* __restrict__ L, double* __restrict__ W,
                            double* __restrict__ Stash, double* __restrict__ Lx, int* __restrict__ info, int64_t ax_stride,
                            int64_t l_stride, int64_t w_stride, int64_t s_stride, int64_t lx_stride, int nmat, int f_stride,
                            int out_mode) {
    extern __shared__ double tk_smem[];
    const int wpc = blockDim.x >> 5, warp = threadIdx.x >> 5;
    double* F = tk_smem + static_cast<size_t>(warp) * f_stride;
    int* sinv = reinterpret_cast<int*>(tk_smem + static_cast<size_t>(wpc) * f_stride) + warp * TW_MAX_M;
    const int g = blockIdx.x * wpc + warp;
    if (nmat > 0) {
        if (g >= nmat) return;
        tree_warp_body(ta, ta.tasks[0], Ax + g * ax_stride, L + g * l_stride, W + g * w_stride, Stash + g * s_stride, Lx + g * lx_stride,
                       info + g, F, sinv, out_mode);
    } else {
        if (g >= ta.ntasks) return;
        tree_warp_body(ta, ta.tasks[g], Ax, L, W, Stash, Lx, info, F, sinv, out_mode);
    }
}
#endif

}  // namespace spchol
