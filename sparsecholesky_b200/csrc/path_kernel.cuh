// k_front_path: batched factorization of the upper part of the assembly tree as path groups
// (see path_plan.h).  ONE CTA per matrix runs the whole group program; the frontal matrix of a
// group lives in shared memory (packed lower triangle over the group's row set U) while its
// member fronts are eliminated in place.
//
// The body is written against five macros (thread id, CTA barrier, warp barrier, read-only load, atomic min) so
// that tests/cpp/path_emu.cpp can run the SAME source on host threads (SPCHOL_HOST_EMU) and
// compare the factor with the oracle without a GPU.
#pragma once
#include <cmath>
#include <cstdint>

#include "path_plan.h"

#ifdef SPCHOL_HOST_EMU
#define PK_DEVICE inline
#define PK_TID ::spchol::emu::tid()
#define PK_SYNC() ::spchol::emu::sync()
#define PK_SYNCWARP() ::spchol::emu::syncwarp()
#define PK_LDG(p) (*(p))
#define PK_RSQRT(x) (1.0 / std::sqrt(x))
#define PK_PREFETCH_L2(p) ((void)(p))
#define PK_ATOMIC_MIN(p, v) ::spchol::emu::atomic_min((p), (v))
#define PK_RESTRICT
#else
#define PK_DEVICE __device__ __forceinline__
#define PK_TID static_cast<int>(threadIdx.x)
#define PK_SYNC() __syncthreads()
#define PK_SYNCWARP() __syncwarp()
#define PK_LDG(p) __ldg(p)
#define PK_RSQRT(x) rsqrt(x)
#define PK_PREFETCH_L2(p) asm volatile("prefetch.global.L2 [%0];" ::"l"(p))
#define PK_ATOMIC_MIN(p, v) atomicMin((p), (v))
#define PK_RESTRICT __restrict__
#endif

namespace spchol {

#ifndef PK_THREADS                    // the host emulation also builds a 64-thread variant (fewer OS threads per barrier)
#define PK_THREADS 256
#define PK_CARRY_PER_THREAD 28
#endif
// Compile-time variants of the kernel that are validated on host threads (tests/test_path_emu.py builds PK_OPT = 0 and 15)
// but not yet timed on a GPU; the shipped build is PK_OPT = 0.  `SPCHOL_NVCC_DEFS=-DPK_OPT=<mask>` rebuilds with
//   1: size of the next member's corner from the planner (PathMember::nd) instead of a serial scan per thread
//   2: diagonal block with ONE warp barrier per pivot (the remaining columns are updated from the unscaled column)
//   4: unmasked update for 32-row chunks that lie entirely inside the block (warp-uniform choice)
//   8: L2 prefetch of everything a group will add (panels of A, blocks from outside) while the group is being zeroed
#ifndef PK_OPT
#define PK_OPT 0
#endif
constexpr int PK_T = PK_THREADS;              // threads per CTA
constexpr int PK_MAXE = PK_CARRY_PER_THREAD;  // carried contribution-block entries per thread (same budget as k_front_chain)

struct PathArrays {                   // device pointers
    const PathGroup* groups; int ngroups;
    const PathMember* members;
    const PathChild* children;
    const int* pos;
    const int* tri_tab;               // tri_tab[e] = (ii << 16) | jj, pairs ii >= jj row by row (see k_front_chain)
};

constexpr int PK_SPOS_MAX = 256;      // rows of the largest front a program may hold
constexpr int PK_NPF = PK_SPOS_MAX / PK_T;   // position-list entries a thread prefetches for the next front

// column c of the packed lower triangle of order nu starts at pk_base(c, nu): entry (r, c), r >= c, is F[base + r]
PK_DEVICE int pk_base(int c, int nu) { return c * nu - (c * (c - 1)) / 2 - c; }

// Cholesky of the w x w diagonal block of member M, in place, by ONE warp (only __syncwarp inside); the reciprocals of
// the pivots go to rinv[0..w).  sqrt and reciprocal come from one rsqrt refined by a Newton step each (as k_potrf_inv):
// no division on the serial path of a front.
PK_DEVICE void pk_diag_factor(const PathMember& M, double* F, int nu, double* rinv, const int* PK_RESTRICT tri_tab,
                              int* PK_RESTRICT info, int lane) {
    const int w = M.w, off = M.off;
    int bk = pk_base(off, nu);
    for (int k = 0; k < w; ++k) {
        double* colk = F + bk;
        const double d = colk[off + k];
        if (d <= 0.0 && lane == 0) PK_ATOMIC_MIN(info, M.first + k + 1);   // reference test (chol.hpp:849)
        double y = PK_RSQRT(d);
        double piv = d * y;
        piv = fma(fma(-piv, piv, d), 0.5 * y, piv);          // sqrt(d), Newton-corrected
        y = fma(fma(-piv, y, 1.0), y, y);                    // 1 / piv, Newton-corrected
#if PK_OPT & 2
        // update of the remaining columns from the UNSCALED column k: (a_i y)(a_j y) are exactly the products the scaled
        // column would give, and column k itself is not read again before the CTA barrier -- so its scaling below needs
        // no barrier of its own
        const int nt = w - k - 1;
        if (nt <= 32) {                                      // few columns: one pair per lane and step
            const int np = nt * (nt + 1) / 2;
            for (int e = lane; e < np; e += 32) {
                const int t = PK_LDG(tri_tab + e);
                const int i = off + k + 1 + (t >> 16), j = off + k + 1 + (t & 0xffff);
                F[pk_base(j, nu) + i] -= (colk[i] * y) * (colk[j] * y);
            }
        } else {                                             // wide front: column by column, lanes down the rows
            for (int j = off + k + 1; j < off + w; ++j) {
                const double uj = colk[j] * y;
                double* colj = F + pk_base(j, nu);
                for (int i = j + lane; i < off + w; i += 32) colj[i] -= (colk[i] * y) * uj;
            }
        }
        PK_SYNCWARP();                                       // column k + 1 is complete; every lane is done reading column k
        for (int i = k + 1 + lane; i < w; i += 32) colk[off + i] = colk[off + i] * y;
        if (lane == 0) { colk[off + k] = piv; rinv[k] = y; }
#else
        PK_SYNCWARP();                                       // every lane has read the pivot
        for (int i = k + 1 + lane; i < w; i += 32) colk[off + i] = colk[off + i] * y;
        if (lane == 0) { colk[off + k] = piv; rinv[k] = y; }
        PK_SYNCWARP();
        const int nt = w - k - 1;
        if (nt <= 32) {                                      // few columns: one pair per lane and step
            const int np = nt * (nt + 1) / 2;
            for (int e = lane; e < np; e += 32) {
                const int t = PK_LDG(tri_tab + e);
                const int i = off + k + 1 + (t >> 16), j = off + k + 1 + (t & 0xffff);
                F[pk_base(j, nu) + i] -= colk[i] * colk[j];
            }
        } else {                                             // wide front: column by column, lanes down the rows
            for (int j = off + k + 1; j < off + w; ++j) {
                const double uj = colk[j];
                double* colj = F + pk_base(j, nu);
                for (int i = j + lane; i < off + w; i += 32) colj[i] -= colk[i] * uj;
            }
        }
        PK_SYNCWARP();
#endif
        bk += nu - (off + k) - 1;
    }
}

// F: nu_max (nu_max + 1) / 2 doubles of shared memory; spos: 2 * PK_SPOS_MAX ints, srinv: PK_SPOS_MAX doubles of shared memory
//
// Barriers per group: 5 + 2 per member front.  Everything that is added into the frontal matrix before the
// eliminations (panels of A, blocks from outside) is dealt to the PK_OWNERS column owners, so no two warps touch
// the same column and the additions of one column happen in program order inside one warp.
PK_DEVICE void front_path_body(const PathArrays pa, double* PK_RESTRICT L, double* PK_RESTRICT W, int* PK_RESTRICT info,
                               double* F, int* spos, double* srinv) {
    const int tid = PK_TID, lane = tid & 31, warp = tid >> 5;
    constexpr int NW = PK_T / 32;
    double carry[PK_MAXE];
#pragma unroll
    for (int u = 0; u < PK_MAXE; ++u) carry[u] = 0.0;
    for (int gq = 0; gq < pa.ngroups; ++gq) {
        const PathGroup g = pa.groups[gq];
        const int nu = g.nu;
        const int ne_u = nu * (nu + 1) / 2;
        const PathMember* PK_RESTRICT mem = pa.members + g.member0;
        // 1. zero the frontal matrix of the group; stage the position list of the first member
#if PK_OPT & 8
        // everything step 3 will read from HBM is known now: one 128-byte line per thread and pass
        for (int q = 0; q < g.nmember; ++q) {
            const PathMember M = mem[q];
            const char* base = reinterpret_cast<const char*>(L + M.panel_off);
            const int64_t bytes = static_cast<int64_t>(M.ld) * M.w * 8;
            for (int64_t b = static_cast<int64_t>(tid) * 128; b < bytes; b += static_cast<int64_t>(PK_T) * 128) PK_PREFETCH_L2(base + b);
        }
        for (int q = 0; q < g.nchild; ++q) {
            if (q == g.carried_child) continue;
            const PathChild* PK_RESTRICT chp = pa.children + g.child0 + q;
            const char* base = reinterpret_cast<const char*>(W + chp->src);
            const int64_t bytes = static_cast<int64_t>((chp->mc + 1) & ~1) * chp->mc * 8;
            for (int64_t b = static_cast<int64_t>(tid) * 128; b < bytes; b += static_cast<int64_t>(PK_T) * 128) PK_PREFETCH_L2(base + b);
        }
#endif
        for (int e = tid; e < ne_u; e += PK_T) F[e] = 0.0;
        {
            const PathMember M0 = mem[0];
            for (int t = tid; t < M0.m; t += PK_T) spos[t] = PK_LDG(pa.pos + M0.pos_off + t);
        }
        PK_SYNC();
        // 2. the block that arrives in registers (all threads, any column)
        if (g.carried_child >= 0) {
            const PathChild ch = pa.children[g.child0 + g.carried_child];
            const int* PK_RESTRICT cp = pa.pos + ch.pos_off;
            const int ne = ch.mc * (ch.mc + 1) / 2;
#pragma unroll
            for (int u = 0; u < PK_MAXE; ++u) {
                if (u * PK_T >= ne) break;
                const int e = tid + u * PK_T;
                if (e < ne) {
                    const int t = PK_LDG(pa.tri_tab + e);
                    F[pk_base(PK_LDG(cp + (t & 0xffff)), nu) + PK_LDG(cp + (t >> 16))] += carry[u];
                }
            }
            PK_SYNC();
        }
        // 3. by column owner: the members' columns of A (scattered into their panels of the L buffer by k_assemble, zeros
        //    elsewhere), then the blocks from outside the group in program order (fixed summation order per column)
        for (int o = warp; o < PK_OWNERS; o += NW) {
            for (int q = 0; q < g.nmember; ++q) {
                const PathMember M = mem[q];
                const double* PK_RESTRICT P = L + M.panel_off;
                const int* PK_RESTRICT ps = pa.pos + M.pos_off;
                for (int j = 0; j < M.w; ++j) {
                    if ((M.off + j) % PK_OWNERS != o) continue;
                    double* col = F + pk_base(M.off + j, nu);
                    const double* src = P + static_cast<int64_t>(j) * M.ld;
                    for (int t = j + lane; t < M.m; t += 32) col[PK_LDG(ps + t)] += src[t];
                }
            }
            for (int q = 0; q < g.nchild; ++q) {
                PK_SYNCWARP();                                   // an entry is added to by different lanes for different sources
                if (q == g.carried_child) continue;
                const PathChild* PK_RESTRICT chp = pa.children + g.child0 + q;
                const int mc = chp->mc, c_lo = chp->optr[o], c_hi = chp->optr[o + 1];
                if (c_lo == c_hi) continue;
                const int* PK_RESTRICT cp = pa.pos + chp->pos_off;
                const int* PK_RESTRICT cols = pa.pos + chp->cols_off;
                const int ldc = (mc + 1) & ~1;
                const double* PK_RESTRICT C = W + chp->src;
                for (int cq = c_lo; cq < c_hi; ++cq) {
                    const int jj = PK_LDG(cols + cq);
                    double* col = F + pk_base(PK_LDG(cp + jj), nu);
                    const double* src = C + static_cast<int64_t>(jj) * ldc;
                    for (int ii = jj + lane; ii < mc; ii += 32) col[PK_LDG(cp + ii)] += src[ii];
                }
            }
        }
        PK_SYNC();
        // 4. eliminate the members in place, bottom to top.  Per member: (b) one thread per row solves against the
        //    diagonal block, barrier, (c) rank-w update of its contribution block by warps 1.. WHILE warp 0 updates
        //    and factors the diagonal block of the NEXT member (the serial part of a front, off the critical path),
        //    barrier.
        if (warp == 0) pk_diag_factor(mem[0], F, nu, srinv, pa.tri_tab, info, lane);
        PK_SYNC();
        for (int q = 0; q < g.nmember; ++q) {
            const PathMember M = mem[q];
            const int w = M.w, m = M.m, off = M.off;
            const int* sp = spos + (q & 1) * PK_SPOS_MAX;
            const bool last = q + 1 == g.nmember;
            const int b0 = pk_base(off, nu);
            // descriptor and position list of the next member: loaded now, used behind the arithmetic of this one
            PathMember Mn = M;
            int nx[PK_NPF], nxm = 0;
            if (!last) {
                Mn = mem[q + 1];
                nxm = Mn.m;
#pragma unroll
                for (int u = 0; u < PK_NPF; ++u) { const int t = tid + u * PK_T; nx[u] = t < nxm ? PK_LDG(pa.pos + Mn.pos_off + t) : 0; }
            }
            // 4b. rows below the diagonal block: x L11^T = a, one row per thread (a thread only touches its own row)
            for (int t = w + tid; t < m; t += PK_T) {
                const int r = sp[t];
                int bk = b0;
                for (int k = 0; k < w; ++k) {
                    double v = F[bk + r];
                    int bj = b0;
                    for (int j = 0; j < k; ++j) { v -= F[bj + r] * F[bj + off + k]; bj += nu - (off + j) - 1; }
                    F[bk + r] = v * srinv[k];
                    bk += nu - (off + k) - 1;
                }
            }
            PK_SYNC();
            // 4c. rank-w update.  nd = rows of this front that are columns of the next member: that corner is the next
            //     member's diagonal block and belongs to warp 0
#if PK_OPT & 1
            const int nd = M.nd;
#else
            int nd = 0;
            if (!last) { while (w + nd < m && sp[w + nd] < Mn.off + Mn.w) ++nd; }
#endif
            if (warp == 0 && !last) {
                const int np = nd * (nd + 1) / 2;
                if (nd <= 32) {
                    for (int e = lane; e < np; e += 32) {
                        const int tt = PK_LDG(pa.tri_tab + e);
                        const int r = sp[w + (tt >> 16)], cj = sp[w + (tt & 0xffff)];
                        double acc = F[pk_base(cj, nu) + r];
                        int bk = b0;
                        for (int k = 0; k < w; ++k) { acc -= F[bk + r] * F[bk + cj]; bk += nu - (off + k) - 1; }
                        F[pk_base(cj, nu) + r] = acc;
                    }
                } else {
                    for (int j = w; j < w + nd; ++j) {
                        const int cj = sp[j], bc = pk_base(cj, nu);
                        for (int t = j + lane; t < w + nd; t += 32) {
                            const int r = sp[t];
                            double acc = F[bc + r];
                            int bk = b0;
                            for (int k = 0; k < w; ++k) { acc -= F[bk + r] * F[bk + cj]; bk += nu - (off + k) - 1; }
                            F[bc + r] = acc;
                        }
                    }
                }
                PK_SYNCWARP();
                pk_diag_factor(Mn, F, nu, srinv, pa.tri_tab, info, lane);
            } else {
                // the rest: groups of 4 columns per warp step, lanes down the rows; a loaded L(r, k) serves 4 updates
                const int nwb = last ? NW : NW - 1, wb = last ? warp : warp - 1;
                for (int j0 = w + 4 * wb; j0 < m; j0 += 4 * nwb) {
                    int cj[4], bc[4];
#pragma unroll
                    for (int c = 0; c < 4; ++c) { cj[c] = sp[min(j0 + c, m - 1)]; bc[c] = pk_base(cj[c], nu); }
                    const int t_begin = (j0 < w + nd) ? max(j0, w + nd - 3) : j0;      // rows above belong to warp 0's corner
                    const bool full_cols = j0 + 3 < m;
                    const int t_full = max(j0 + 3, w + nd);      // from this row on all four entries exist and are ours
                    for (int t = t_begin + lane; t < m; t += 32) {
                        const int r = sp[t];
                        double acc[4];
                        if ((PK_OPT & 4) && full_cols && t - lane >= t_full) {   // warp-uniform: the whole 32-row chunk is interior
#pragma unroll
                            for (int c = 0; c < 4; ++c) acc[c] = F[bc[c] + r];
                            int bk = b0;
                            for (int k = 0; k < w; ++k) {
                                const double lr = F[bk + r];
#pragma unroll
                                for (int c = 0; c < 4; ++c) acc[c] -= lr * F[bk + cj[c]];
                                bk += nu - (off + k) - 1;
                            }
#pragma unroll
                            for (int c = 0; c < 4; ++c) F[bc[c] + r] = acc[c];
                        } else {
                            bool on[4];
#pragma unroll
                            for (int c = 0; c < 4; ++c) {
                                // entry (t, j0 + c) exists if t >= j0 + c, and is warp 0's if both indices are below w + nd
                                on[c] = j0 + c < m && t >= j0 + c && !(t < w + nd && j0 + c < w + nd);
                                acc[c] = on[c] ? F[bc[c] + r] : 0.0;
                            }
                            int bk = b0;
                            for (int k = 0; k < w; ++k) {
                                const double lr = F[bk + r];
#pragma unroll
                                for (int c = 0; c < 4; ++c) acc[c] -= lr * F[bk + cj[c]];
                                bk += nu - (off + k) - 1;
                            }
#pragma unroll
                            for (int c = 0; c < 4; ++c) if (on[c]) F[bc[c] + r] = acc[c];
                        }
                    }
                }
            }
            if (!last) {
                int* spn = spos + ((q + 1) & 1) * PK_SPOS_MAX;
#pragma unroll
                for (int u = 0; u < PK_NPF; ++u) { const int t = tid + u * PK_T; if (t < nxm) spn[t] = nx[u]; }
            }
            PK_SYNC();
        }
        // 5. panels -> L buffer; block of the top front -> registers (the next group consumes it) or workspace
        for (int q = 0; q < g.nmember; ++q) {
            const PathMember M = mem[q];
            double* PK_RESTRICT P = L + M.panel_off;
            const int* PK_RESTRICT ps = pa.pos + M.pos_off;
            for (int j = warp; j < M.w; j += NW) {
                const double* col = F + pk_base(M.off + j, nu);
                double* dst = P + static_cast<int64_t>(j) * M.ld;
                for (int t = j + lane; t < M.m; t += 32) dst[t] = col[PK_LDG(ps + t)];
            }
        }
        const int mb = g.mb, c0 = nu - mb;
        if (mb > 0) {
            if (g.carry_out) {
                const int ne = mb * (mb + 1) / 2;
#pragma unroll
                for (int u = 0; u < PK_MAXE; ++u) {
                    if (u * PK_T >= ne) break;
                    const int e = tid + u * PK_T;
                    if (e < ne) { const int t = PK_LDG(pa.tri_tab + e); carry[u] = F[pk_base(c0 + (t & 0xffff), nu) + c0 + (t >> 16)]; }
                }
            } else if (g.cb_dst >= 0) {
                const int ldcb = (mb + 1) & ~1;
                double* PK_RESTRICT C = W + g.cb_dst;
                for (int j = warp; j < mb; j += NW) {
                    const double* col = F + pk_base(c0 + j, nu) + c0;
                    double* dst = C + static_cast<int64_t>(j) * ldcb;
                    for (int i = j + lane; i < mb; i += 32) dst[i] = col[i];
                }
            }
        }
        PK_SYNC();                                               // F is rebuilt for the next group
    }
}

#ifndef SPCHOL_HOST_EMU
// grid = matrices; dynamic shared memory = max_nu (max_nu + 1) / 2 doubles
__global__ void __launch_bounds__(PK_T, 2) k_front_path(PathArrays pa, double* __restrict__ L, double* __restrict__ W,
                                                        int* __restrict__ info, int64_t lstride, int64_t wstride) {
    extern __shared__ double pk_F[];
    __shared__ int pk_spos[2 * PK_SPOS_MAX];
    __shared__ double pk_rinv[PK_SPOS_MAX];
    front_path_body(pa, L + blockIdx.x * lstride, W + blockIdx.x * wstride, info + blockIdx.x, pk_F, pk_spos, pk_rinv);
}
#endif

}  // namespace spchol
