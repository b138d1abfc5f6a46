// k_front_path: batched factorization of the upper part of the assembly tree as path groups
// (see path_plan.h).  ONE CTA per matrix runs the whole group program; the frontal matrix of a
// group lives in shared memory (packed lower triangle over the group's row set U) while its
// member fronts are eliminated in place.
//
// The body is written against four macros (thread id, CTA barrier, read-only load, atomic min) so
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
#define PK_LDG(p) (*(p))
#define PK_ATOMIC_MIN(p, v) ::spchol::emu::atomic_min((p), (v))
#define PK_RESTRICT
#else
#define PK_DEVICE __device__ __forceinline__
#define PK_TID static_cast<int>(threadIdx.x)
#define PK_SYNC() __syncthreads()
#define PK_LDG(p) __ldg(p)
#define PK_ATOMIC_MIN(p, v) atomicMin((p), (v))
#define PK_RESTRICT __restrict__
#endif

namespace spchol {

#ifndef PK_THREADS                    // the host emulation also builds a 64-thread variant (fewer OS threads per barrier)
#define PK_THREADS 256
#define PK_CARRY_PER_THREAD 28
#endif
constexpr int PK_T = PK_THREADS;              // threads per CTA
constexpr int PK_MAXE = PK_CARRY_PER_THREAD;  // carried contribution-block entries per thread (same budget as k_front_chain)

struct PathArrays {                   // device pointers
    const PathGroup* groups; int ngroups;
    const PathMember* members;
    const PathChild* children;
    const int* pos;
    const int* tri_tab;               // tri_tab[e] = (ii << 16) | jj, pairs ii >= jj row by row (see k_front_chain)
    const int* first; const int* w; const int* m; const int* ld;
    const int64_t* panel_off;
};

// column c of the packed lower triangle of order nu starts at pk_base(c, nu): entry (r, c), r >= c, is F[base + r]
PK_DEVICE int pk_base(int c, int nu) { return c * nu - (c * (c - 1)) / 2 - c; }

// F: nu_max (nu_max + 1) / 2 doubles of shared memory; spos: max_m ints of shared memory
PK_DEVICE void front_path_body(const PathArrays pa, double* PK_RESTRICT L, double* PK_RESTRICT W, int* PK_RESTRICT info,
                               double* F, int* spos) {
    const int tid = PK_TID, lane = tid & 31, warp = tid >> 5;
    constexpr int NW = PK_T / 32;
    double carry[PK_MAXE];
#pragma unroll
    for (int u = 0; u < PK_MAXE; ++u) carry[u] = 0.0;
    for (int gq = 0; gq < pa.ngroups; ++gq) {
        const PathGroup g = pa.groups[gq];
        const int nu = g.nu;
        const int ne_u = nu * (nu + 1) / 2;
        // 1. zero the frontal matrix of the group
        for (int e = tid; e < ne_u; e += PK_T) F[e] = 0.0;
        PK_SYNC();
        // 2. the members' columns of A (already scattered into their panels of the L buffer, zeros elsewhere)
        for (int q = 0; q < g.nmember; ++q) {
            const PathMember M = pa.members[g.member0 + q];
            const int w = pa.w[M.s], m = pa.m[M.s], ld = pa.ld[M.s];
            const double* PK_RESTRICT P = L + pa.panel_off[M.s];
            const int* PK_RESTRICT ps = pa.pos + M.pos_off;
            for (int j = warp; j < w; j += NW) {
                double* col = F + pk_base(M.off + j, nu);
                const double* src = P + static_cast<int64_t>(j) * ld;
                for (int t = j + lane; t < m; t += 32) col[PK_LDG(ps + t)] = src[t];
            }
        }
        PK_SYNC();
        // 3. contribution blocks from outside the group, in program order (fixed summation order)
        for (int q = 0; q < g.nchild; ++q) {
            const PathChild ch = pa.children[g.child0 + q];
            const int mc = ch.mc;
            const int* PK_RESTRICT cp = pa.pos + ch.pos_off;
            if (ch.carried) {
                const int ne = mc * (mc + 1) / 2;
#pragma unroll
                for (int u = 0; u < PK_MAXE; ++u) {
                    if (u * PK_T >= ne) break;
                    const int e = tid + u * PK_T;
                    if (e < ne) {
                        const int t = PK_LDG(pa.tri_tab + e);
                        F[pk_base(PK_LDG(cp + (t & 0xffff)), nu) + PK_LDG(cp + (t >> 16))] += carry[u];
                    }
                }
            } else {
                const int ldc = (mc + 1) & ~1;
                const double* PK_RESTRICT C = W + ch.src;
                for (int jj = warp; jj < mc; jj += NW) {
                    double* col = F + pk_base(PK_LDG(cp + jj), nu);
                    const double* src = C + static_cast<int64_t>(jj) * ldc;
                    for (int ii = jj + lane; ii < mc; ii += 32) col[PK_LDG(cp + ii)] += src[ii];
                }
            }
            PK_SYNC();
        }
        // 4. eliminate the members in place, bottom to top
        for (int q = 0; q < g.nmember; ++q) {
            const PathMember M = pa.members[g.member0 + q];
            const int w = pa.w[M.s], m = pa.m[M.s];
            for (int t = tid; t < m; t += PK_T) spos[t] = PK_LDG(pa.pos + M.pos_off + t);
            PK_SYNC();
            // 4a. the front's own columns, right-looking inside the panel (rows = the front's true rows only)
            for (int k = 0; k < w; ++k) {
                double* colk = F + pk_base(M.off + k, nu);
                const double d = colk[M.off + k];
                if (d <= 0.0 && tid == 0) PK_ATOMIC_MIN(info, pa.first[M.s] + k + 1);   // reference test (chol.hpp:849)
                const double piv = sqrt(d);
                PK_SYNC();                                       // everyone has read the pivot
                for (int t = k + 1 + tid; t < m; t += PK_T) { const int r = spos[t]; colk[r] = colk[r] / piv; }
                if (tid == 0) colk[M.off + k] = piv;
                PK_SYNC();
                for (int j = k + 1 + warp; j < w; j += NW) {
                    const double uj = colk[M.off + j];
                    double* col = F + pk_base(M.off + j, nu);
                    for (int t = j + lane; t < m; t += 32) { const int r = spos[t]; col[r] -= colk[r] * uj; }
                }
                PK_SYNC();
            }
            // 4b. rank-w update of the front's contribution block, which is part of the group's matrix
            for (int j = w + warp; j < m; j += NW) {
                const int cj = spos[j];
                double* col = F + pk_base(cj, nu);
                for (int t = j + lane; t < m; t += 32) {
                    const int r = spos[t];
                    double acc = col[r];
                    for (int k = 0; k < w; ++k) { const double* ck = F + pk_base(M.off + k, nu); acc -= ck[r] * ck[cj]; }
                    col[r] = acc;
                }
            }
            PK_SYNC();
        }
        // 5. panels -> L buffer; block of the top front -> registers (the next group consumes it) or workspace
        for (int q = 0; q < g.nmember; ++q) {
            const PathMember M = pa.members[g.member0 + q];
            const int w = pa.w[M.s], m = pa.m[M.s], ld = pa.ld[M.s];
            double* PK_RESTRICT P = L + pa.panel_off[M.s];
            const int* PK_RESTRICT ps = pa.pos + M.pos_off;
            for (int j = warp; j < w; j += NW) {
                const double* col = F + pk_base(M.off + j, nu);
                double* dst = P + static_cast<int64_t>(j) * ld;
                for (int t = j + lane; t < m; t += 32) dst[t] = col[PK_LDG(ps + t)];
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
constexpr int PK_SPOS_MAX = 256;      // rows of the largest front a program may hold (build_path_program's max_nu <= this)
// grid = matrices; dynamic shared memory = max_nu (max_nu + 1) / 2 doubles
__global__ void __launch_bounds__(PK_T, 2) k_front_path(PathArrays pa, double* __restrict__ L, double* __restrict__ W,
                                                        int* __restrict__ info, int64_t lstride, int64_t wstride) {
    extern __shared__ double pk_F[];
    __shared__ int pk_spos[PK_SPOS_MAX];
    front_path_body(pa, L + blockIdx.x * lstride, W + blockIdx.x * wstride, info + blockIdx.x, pk_F, pk_spos);
}
#endif

}  // namespace spchol
