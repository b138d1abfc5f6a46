// k_tree_team: a TEAM of NW warps (one CTA) factors a whole subtree of the assembly tree (or, batched, the whole
// tree of one matrix), front after front, with the frontal matrix in shared memory; see tree_plan.h for the
// program and the reversed packed layout that lets a contribution block pass to the parent without moving.
// Fuses dpotrf_ + dtrsm + dsyrk + apply_update of the reference's supernodal chain
// (include/chol.hpp:1232-1335, :1196-1216) for every front of the subtree; nothing but the factor
// (and the blocks of side branches) leaves the SM.  The rank-w update of the contribution block runs on the
// FP64 tensor cores (DMMA m8n8k4) in 8 x 8 tiles dealt to the warps of the team.
//
// The body is written against a few macros so that tests/cpp/tree_emu.cpp runs the SAME source on
// 32 NW host threads per team (SPCHOL_HOST_EMU) against the oracle, without a GPU.
#pragma once
#include <cmath>
#include <cstdint>

#include "tree_plan.h"

#ifdef SPCHOL_HOST_EMU
#define TK_DEVICE inline
#define TK_TID ::spchol::emu::tid()
#define TK_SYNCWARP() ::spchol::emu::syncwarp()
#define TK_SYNCTEAM() ::spchol::emu::sync()
#define TK_LDG(p) (*(p))
#define TK_ATOMIC_MIN(p, v) ::spchol::emu::atomic_min((p), (v))
#define TK_RESTRICT
#define TK_SQRT(x) std::sqrt(x)
#define TK_RSQRT(x) (1.0 / std::sqrt(x))
#define TK_DMMA(c0, c1, a, b) ::spchol::emu::dmma884((c0), (c1), (a), (b))
#define TK_TEAM_ID ::spchol::emu::g_team_id
#else
#define TK_DEVICE __device__ __forceinline__
#define TK_TID static_cast<int>(threadIdx.x)
#define TK_SYNCWARP() __syncwarp()
#define TK_SYNCTEAM() __syncthreads()
#define TK_LDG(p) __ldg(p)
#define TK_ATOMIC_MIN(p, v) atomicMin((p), (v))
#define TK_RESTRICT __restrict__
#define TK_SQRT(x) sqrt(x)
#define TK_RSQRT(x) rsqrt(x)
#define TK_TEAM_ID static_cast<int>(blockIdx.x)
#define TK_DMMA(c0, c1, a, b) asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b))
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
    const unsigned short* out_src;    // amalgamated groups (TW_GROUP): packed position in F of every true entry of L ...
    const int* out_dst;               // ... and its index in the CSC values of L
    const int* gcol;                  // ... global column of every column of a group (reporting a non-positive pivot)
};

constexpr int TK_F_SLACK = 8;         // doubles past tri(cap_m): fragment loads of edge tiles may read (never store) there

TK_DEVICE int tk_tri(int x) { return (x * (x + 1)) >> 1; }

// pull the 128-byte lines of [base + b0, base + b1) towards L1, one line per thread and pass
TK_DEVICE void tk_prefetch_range(const void* base, int64_t b0, int64_t b1, int tid, int nt) {
#ifndef SPCHOL_HOST_EMU
    const char* p = static_cast<const char*>(base);
    for (int64_t off = (b0 & ~int64_t(127)) + static_cast<int64_t>(tid) * 128; off < b1; off += static_cast<int64_t>(nt) * 128)
        asm volatile("prefetch.global.L1 [%0];" ::"l"(p + off));
#else
    (void)base; (void)b0; (void)b1; (void)tid; (void)nt;
#endif
}

// One task = fronts [front0, front0 + nfront) in postorder, factored by a team of NW warps (tid = 0 .. 32 NW - 1).
// F: tri(cap_m) + TK_F_SLACK doubles, spiv: 2 doubles, sT: 64 doubles (reciprocal pivots of a micro block), all private to the team.
// out_mode 0: panels go to the L buffer; 1: entries go straight to Lx in CSC order through csc_map.
//
// Every front starts with a chain of dependent loads (descriptor -> index lists -> values); with four teams per SM there
// is nothing to hide them behind, so the kernel runs two descriptors ahead: while front f is eliminated, the descriptor
// of f+2 is in flight and the lists of f+1 (maps, A entry lists, its slice of the CSC map, its values of A) are being
// pulled into L1.
template <int NW>
TK_DEVICE void tree_team_body(const TreeArrays& ta, const TwTask task, const double* TK_RESTRICT Ax, double* TK_RESTRICT L,
                              double* TK_RESTRICT W, double* TK_RESTRICT Stash, double* TK_RESTRICT Lx, int* TK_RESTRICT info,
                              double* F, double* spiv, double* sT, int out_mode) {
    constexpr int NT = 32 * NW;
    constexpr int MC = 4;                                   // columns a warp moves per round of the in-place spread
    const int tid = TK_TID, lane = tid & 31, warp = tid >> 5;
    const int g = lane >> 2, tq = lane & 3;                 // DMMA fragment coordinates of the lane
    // the warp that runs the serial 8 x 8 factorizations: rotated over the teams of an SM, so that the serial chains of the
    // resident teams do not all land on the same scheduler (warp w of every CTA sits on scheduler w mod 4)
    const int sw = NW > 1 ? (TK_TEAM_ID % NW) : 0;
    const int ow = NW > 1 ? (warp - sw - 1 + NW) % NW : 0;  // index of this warp among the other NW - 1
    Stash += task.stash_base;
    const TwFront* TK_RESTRICT fr = ta.fronts + task.front0;
    TwFront d = fr[0];
    TwFront dn = fr[task.nfront > 1 ? 1 : 0];
    for (int f = 0; f < task.nfront; ++f) {
        const TwFront dnn = fr[f + 2 < task.nfront ? f + 2 : task.nfront - 1];
        int ax_next = -1;                                   // index of one value of A the next front needs (prefetched further down)
        if (f + 1 < task.nfront) {
            if (warp == NW - 1) {                           // one warp issues the prefetches (a line per lane and pass)
                tk_prefetch_range(ta.rmap, 4 * static_cast<int64_t>(dn.map_lo), 4 * static_cast<int64_t>(dn.map_lo + dn.map_n), lane, 32);
                tk_prefetch_range(ta.a_dst, 2 * static_cast<int64_t>(dn.a0), 2 * static_cast<int64_t>(dn.a0 + dn.an), lane, 32);
                tk_prefetch_range(ta.stashes, 16 * static_cast<int64_t>(dn.stash0), 16 * static_cast<int64_t>(dn.stash0 + dn.nstash), lane, 32);
                if (dn.flags & TW_GROUP) {
                    tk_prefetch_range(ta.out_src, 2 * static_cast<int64_t>(dn.pad0), 2 * static_cast<int64_t>(dn.pad0 + dn.pad1), lane, 32);
                    tk_prefetch_range(ta.out_dst, 4 * static_cast<int64_t>(dn.pad0), 4 * static_cast<int64_t>(dn.pad0 + dn.pad1), lane, 32);
                } else if (out_mode == 1) tk_prefetch_range(ta.csc_map, 4 * dn.panel_off, 4 * (dn.panel_off + static_cast<int64_t>(dn.ld) * dn.w), lane, 32);
            }
            if (tid < dn.an) ax_next = TK_LDG(ta.a_src + dn.a0 + tid);
        }
        const int m = d.m, w = d.w, mb = m - w;
        // ---- 1. the frontal matrix: spread the carried block in place, or start from zero
        if (d.flags & TW_CARRY_IN) {
            // inv[t] = reversed row of the child's block that becomes reversed row t of this front, -1 for a new row.
            // Columns below d.ident are where they belong already.  The others move from the top down, MC NW columns per
            // round: every warp reads its columns into registers, the team meets, every warp writes.  A column's new
            // place never reaches below the old place of a column of a later round, so one barrier per round is enough.
            const int* TK_RESTRICT inv = ta.rmap + d.map_off;
            for (int J0 = m - 1; J0 >= d.ident; J0 -= MC * NW) {
                double v[MC][4];
#pragma unroll
                for (int h = 0; h < MC; ++h) {
                    const int J = J0 - (MC * warp + h);
                    const int j = J >= d.ident ? TK_LDG(inv + J) : -2;      // -1: a new column (all zero), -2: nothing to do
                    const int bj = j >= 0 ? tk_tri(j) : 0;
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        const int t = lane + 32 * u;
                        v[h][u] = 0.0;
                        if (t <= J && j >= 0) { const int src = TK_LDG(inv + t); if (src >= 0) v[h][u] = F[bj + src]; }
                    }
                }
                TK_SYNCTEAM();
#pragma unroll
                for (int h = 0; h < MC; ++h) {
                    const int J = J0 - (MC * warp + h);
                    if (J < d.ident) continue;
                    const int bJ = tk_tri(J);
#pragma unroll
                    for (int u = 0; u < 4; ++u) { const int t = lane + 32 * u; if (t <= J) F[bJ + t] = v[h][u]; }
                }
            }
        } else {
            const int ne = tk_tri(m);
            for (int e = tid; e < ne; e += NT) F[e] = 0.0;
        }
        TK_SYNCTEAM();
        // ---- 2. the front's entries of A (distinct targets)
        for (int e = tid; e < d.an; e += NT) F[TK_LDG(ta.a_dst + d.a0 + e)] += Ax[TK_LDG(ta.a_src + d.a0 + e)];
        // ---- 3. blocks of the other children, from the stash (most recent first); columns are dealt to the warps
        for (int q = 0; q < d.nstash; ++q) {
            TK_SYNCTEAM();                                 // two sources may hit the same entry from different threads
            const TwStash st = ta.stashes[d.stash0 + q];
            const int* TK_RESTRICT rm = ta.rmap + st.map_off;
            if (st.mc < 0) {                               // left in the workspace by an earlier tier: column-major lower, ld = even(mc)
                const int mc = -st.mc, ldc = (mc + 1) & ~1;
                const double* src = W + st.src;
                for (int jr = warp; jr < mc; jr += NW) {
                    const int bJ = tk_tri(TK_LDG(rm + jr));
                    const double* col = src + static_cast<int64_t>(mc - 1 - jr) * ldc + (mc - 1);     // reversed row ir sits at col[-ir]
                    for (int ir = lane; ir <= jr; ir += 32) F[bJ + TK_LDG(rm + ir)] += col[-ir];
                }
                continue;
            }
            const double* TK_RESTRICT src = Stash + st.src;
            for (int jr = warp; jr < st.mc; jr += NW) {
                const int bJ = tk_tri(TK_LDG(rm + jr));
                const double* col = src + tk_tri(jr);
                for (int ir = lane; ir <= jr; ir += 32) F[bJ + TK_LDG(rm + ir)] += col[ir];
            }
        }
        TK_SYNCTEAM();
        if (ax_next >= 0) tk_prefetch_range(Ax, 8 * static_cast<int64_t>(ax_next), 8 * static_cast<int64_t>(ax_next) + 8, 0, 1);
        // ---- 4. elimination of the w pivot columns, right-looking in micro-panels of 8 columns over the WHOLE front (the rest
        //         of the panel and the contribution block alike).  Per micro-panel: (a) warp 0 factors the 8 x 8 diagonal block in
        //         registers and inverts it, (b) one thread per row below multiplies the row by the inverse (no division, no
        //         dependence between rows), (c) rank-8 update of everything behind it on the FP64 tensor cores: 8 x 8 tiles (row
        //         block R <= column block C of the reversed triangle) dealt round-robin to the warps; lane (g, tq) feeds
        //         A[g][k] = -L(8R + g, pivot k) and B[k][g] = L(8C + g, pivot k) for k = tq, 4 + tq and owns C(8R + g, 8C + 2 tq)
        //         and C(8R + g, 8C + 2 tq + 1).  Look-ahead: the tiles that hold the NEXT micro-panel's columns go first (all
        //         warps), then warp 0 factors that micro-panel while the other warps finish the update -- the serial 8-pivot chain
        //         hides behind tensor-core work.  Three team barriers per 8 pivots.
        auto micro_factor = [&](int k0, int nk) {              // one warp (sw) only
            const int Kt = m - 1 - k0;
            double l[8][8], rinv[8];
#pragma unroll
            for (int i = 0; i < 8; ++i)
#pragma unroll
                for (int j = 0; j <= i; ++j) l[i][j] = (i < nk && j < nk) ? F[tk_tri(Kt - j) + (Kt - i)] : (i == j ? 1.0 : 0.0);
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                rinv[k] = 1.0;
                if (k < nk) {
                    const double dd = l[k][k];
                    if (dd <= 0.0 && lane == 0)                // reference test (chol.hpp:849)
                        TK_ATOMIC_MIN(info, ((d.flags & TW_GROUP) ? TK_LDG(ta.gcol + d.ld + k0 + k) : d.first + k0 + k) + 1);
                    // sqrt(d) and 1/sqrt(d) together (Goldschmidt from the reciprocal-square-root seed, then one residual
                    // correction each): the square root comes out correctly rounded except for rare last-bit cases, the
                    // reciprocal within one ulp -- six dependent FMAs instead of a software sqrt followed by a division
                    double y = TK_RSQRT(dd);
                    double gq = dd * y, hq = 0.5 * y;
                    double r = fma(-gq, hq, 0.5);
                    gq = fma(gq, r, gq); hq = fma(hq, r, hq);
                    r = fma(-gq, hq, 0.5);
                    gq = fma(gq, r, gq); hq = fma(hq, r, hq);
                    const double piv = fma(fma(-gq, gq, dd), hq, gq);          // g + (d - g^2) / (2 g)
                    y = hq + hq;
                    y = fma(fma(-piv, y, 1.0), y, y);                          // 1 / piv
                    if (!(dd > 0.0)) { y = 1.0 / TK_SQRT(dd); }                // non-positive pivot: the reference's sqrt / division semantics (NaN / inf)
                    rinv[k] = y;
                    l[k][k] = (dd > 0.0) ? piv : TK_SQRT(dd);
#pragma unroll
                    for (int i = k + 1; i < 8; ++i) l[i][k] *= y;
#pragma unroll
                    for (int j = k + 1; j < 8; ++j)
#pragma unroll
                        for (int i = j; i < 8; ++i) l[i][j] = fma(-l[i][k], l[j][k], l[i][j]);
                }
            }
            TK_SYNCWARP();                                     // every lane has read the block before it is overwritten
            if (lane < 8) {                                    // lane j writes row j of the factor and its reciprocal pivot
#pragma unroll
                for (int j = 0; j < 8; ++j)
                    if (lane == j) {
                        sT[j] = rinv[j];
#pragma unroll
                        for (int c = 0; c <= j; ++c) if (j < nk) F[tk_tri(Kt - c) + (Kt - j)] = l[j][c];
                    }
            }
        };
        auto rows_below = [&](int k0, int nk) {                // x L11^T = b by substitution, one row per thread (rows are independent)
            const int Kt = m - 1 - k0, mt = m - k0 - nk;
            if (tid >= mt && NT >= mt) return;
            double l[8][8], ri[8];                             // the factored diagonal block and the reciprocal pivots (shared-memory broadcasts)
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                ri[i] = sT[i];
#pragma unroll
                for (int j = 0; j < i; ++j) l[i][j] = (i < nk) ? F[tk_tri(Kt - j) + (Kt - i)] : 0.0;
            }
            for (int ir = tid; ir < mt; ir += NT) {
                double x[8];
#pragma unroll
                for (int q = 0; q < 8; ++q) x[q] = q < nk ? F[tk_tri(Kt - q) + ir] : 0.0;
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    double acc = x[j];
#pragma unroll
                    for (int q = 0; q < j; ++q) acc = fma(-x[q], l[j][q], acc);
                    x[j] = acc * ri[j];
                }
#pragma unroll
                for (int q = 0; q < 8; ++q) if (q < nk) F[tk_tri(Kt - q) + ir] = x[q];
            }
        };
        // rank-nk update by the micro-panel at k0 of the tiles with column block in [cb_lo, cb_hi).  The unit of work is a PAIR of
        // row tiles (2U, 2U + 1) of one column block -- they share the B fragments and the column bases; units `first`,
        // `first + step`, ... of the range are this warp's
        auto update_tiles = [&](int k0, int nk, int cb_lo, int cb_hi, int first, int step) {
            const int Kt = m - 1 - k0, mt = m - k0 - nk;
            const int pbA = tk_tri(Kt - tq), pbB = tk_tri(Kt - 4 - tq);      // the lane's pivot columns of the two k-steps
            const bool onA = tq < nk, onB = 4 + tq < nk;
            int Cb = cb_lo, U = first;
            while (Cb < cb_hi && U > (Cb >> 1)) { U -= (Cb >> 1) + 1; ++Cb; }
            int Cb_cur = -1, c0 = 0, b0 = 0, b1 = 0;
            double bfA = 0.0, bfB = 0.0;
            while (Cb < cb_hi) {
                if (Cb != Cb_cur) {
                    Cb_cur = Cb;
                    c0 = 8 * Cb + 2 * tq;
                    b0 = tk_tri(c0); b1 = b0 + c0 + 1;
                    bfA = onA ? F[pbA + 8 * Cb + g] : 0.0;
                    bfB = onB ? F[pbB + 8 * Cb + g] : 0.0;
                }
                const int ir0 = 16 * U + g, ir1 = ir0 + 8;
                const bool two = 2 * U + 1 <= Cb;              // the second row tile of the pair exists (warp-uniform)
                const bool on00 = ir0 <= c0 && c0 < mt, on01 = ir0 <= c0 + 1 && c0 + 1 < mt;      // entries inside the triangle
                const bool on10 = two && ir1 <= c0 && c0 < mt, on11 = two && ir1 <= c0 + 1 && c0 + 1 < mt;
                double a00 = on00 ? F[b0 + ir0] : 0.0, a01 = on01 ? F[b1 + ir0] : 0.0;
                double a10 = on10 ? F[b0 + ir1] : 0.0, a11 = on11 ? F[b1 + ir1] : 0.0;
                const double fA0 = onA ? -F[pbA + ir0] : 0.0;
                const double fA1 = (onA && two) ? -F[pbA + ir1] : 0.0;
                TK_DMMA(a00, a01, fA0, bfA);
                if (two) TK_DMMA(a10, a11, fA1, bfA);
                if (nk > 4) {
                    const double fB0 = onB ? -F[pbB + ir0] : 0.0;
                    const double fB1 = (onB && two) ? -F[pbB + ir1] : 0.0;
                    TK_DMMA(a00, a01, fB0, bfB);
                    if (two) TK_DMMA(a10, a11, fB1, bfB);
                }
                if (on00) F[b0 + ir0] = a00;
                if (on01) F[b1 + ir0] = a01;
                if (on10) F[b0 + ir1] = a10;
                if (on11) F[b1 + ir1] = a11;
                U += step;
                while (Cb < cb_hi && U > (Cb >> 1)) { U -= (Cb >> 1) + 1; ++Cb; }
            }
        };
        if (w > 0) {
            if (warp == sw) micro_factor(0, w < 8 ? w : 8);
            TK_SYNCTEAM();
            rows_below(0, w < 8 ? w : 8);
            TK_SYNCTEAM();
        }
        for (int k0 = 0; k0 < w; k0 += 8) {
            const int nk = w - k0 < 8 ? w - k0 : 8;
            const int mt = m - k0 - nk;                    // order of what remains behind the micro-panel
            const int nblk = (mt + 7) >> 3;
            const int kn = k0 + 8;                         // next micro-panel (if any)
            if (kn < w) {
                const int nkn = w - kn < 8 ? w - kn : 8;
                const int cbA = (mt - nkn) >> 3;           // column blocks from here up hold the next micro-panel's columns
                update_tiles(k0, nk, cbA, nblk, warp, NW);
                TK_SYNCTEAM();
                if (NW == 1) { update_tiles(k0, nk, 0, cbA, 0, 1); micro_factor(kn, nkn); }
                else if (warp == sw) micro_factor(kn, nkn);
                else update_tiles(k0, nk, 0, cbA, ow, NW - 1);
                TK_SYNCTEAM();
                rows_below(kn, nkn);
            } else if (mt > 0) {
                update_tiles(k0, nk, 0, nblk, warp, NW);
            }
            TK_SYNCTEAM();
        }
        // ---- 6. the factored columns leave the SM (columns dealt to the warps)
        if (d.flags & TW_GROUP) {
            // an amalgamated group of fronts (batched whole-tree program): the TRUE entries of its columns by list,
            // (packed position in F, index in the CSC values of L); the explicit zeros of the amalgamation are skipped
            for (int e = tid; e < d.pad1; e += NT) Lx[TK_LDG(ta.out_dst + d.pad0 + e)] = F[TK_LDG(ta.out_src + d.pad0 + e)];
        } else
        for (int c = warp; c < w; c += NW) {
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
                for (int jr = warp; jr < mb; jr += NW) {
                    const double* col = F + tk_tri(jr);
                    double* dcol = C + static_cast<int64_t>(mb - 1 - jr) * ldcb + (mb - 1);
                    for (int ir = lane; ir <= jr; ir += 32) dcol[-ir] = col[ir];
                }
            } else {
                double* TK_RESTRICT dst = Stash + d.cb_out;
                const int ne = tk_tri(mb);
                for (int e = tid; e < ne; e += NT) dst[e] = F[e];
            }
        }
        TK_SYNCTEAM();
        d = dn; dn = dnn;
    }
}

#ifndef SPCHOL_HOST_EMU
// One CTA = one team of NW warps.  Dynamic shared memory = (f_stride + 2 + 64) doubles.
// nmat > 0 (batched): team b factors matrix b with task 0, all pointers advance by their per-matrix strides;
// nmat == 0: team b runs task b of a single matrix.
template <int NW>
__global__ void __launch_bounds__(32 * NW, 16 / NW) k_tree_team(TreeArrays ta, const double* __restrict__ Ax, double* __restrict__ L,
                                                      double* __restrict__ W, double* __restrict__ Stash, double* __restrict__ Lx,
                                                      int* __restrict__ info, int64_t ax_stride, int64_t l_stride, int64_t w_stride,
                                                      int64_t s_stride, int64_t lx_stride, int nmat, int f_stride, int out_mode) {
    extern __shared__ double tk_smem[];
    double* F = tk_smem;
    double* spiv = tk_smem + f_stride;
    double* sT = spiv + 2;
    const int64_t b = blockIdx.x;
    if (nmat > 0)
        tree_team_body<NW>(ta, ta.tasks[0], Ax + b * ax_stride, L + b * l_stride, W + b * w_stride, Stash + b * s_stride, Lx + b * lx_stride,
                           info + b, F, spiv, sT, out_mode);
    else
        tree_team_body<NW>(ta, ta.tasks[b], Ax, L, W, Stash, Lx, info, F, spiv, sT, out_mode);
}
#endif

}  // namespace spchol
