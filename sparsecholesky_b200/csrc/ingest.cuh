// Device-side ingest: COO triplets -> upper CSC on the GPU.  Replaces the host sort of
// triplet_to_csc_matrix (reference include/chol.hpp:309-369: mirror to the upper triangle, sort by (column, row), sum
// duplicates, keep explicit zeros) for callers whose triplets already live in HBM; the result stays on the device and
// feeds spchol_factor_device directly.  Duplicates are summed in INPUT order (the radix sort is stable), which is also
// what the host path does; the reference's std::sort leaves that order unspecified.
//
// Pipeline (all HBM bound, 16 B per entry and pass):
//   k_in_keys      : key = (max(i,j) << 32) | min(i,j), payload = entry index, range check
//   radix passes   : stable LSD radix sort of (key, payload), 8 bits per pass, only over the bytes that can be non-zero
//                    (2 * ceil(log2 n / 8) passes); per pass: per-warp digit histograms -> exclusive scan in
//                    digit-major order -> stable scatter (a warp walks its 4096 entries in order, ranks inside a
//                    32-entry tile by __match_any_sync)
//   k_in_heads     : first entry of every (row, column) run -> flag; scan -> output slot; per-column counts
//   k_in_emit      : one thread per run: Ai, Ax = sum of the run in order, Ap from the scanned column counts
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace spchol {

constexpr int IN_WARP_ITEMS = 4096;          // entries one warp sorts per pass (128 tiles of 32)
constexpr int IN_SCAN_BLOCK = 2048;          // elements per CTA of the scan (256 threads x 8)

__global__ void k_in_keys(int64_t nz, const int* __restrict__ ti, const int* __restrict__ tj, int n,
                          unsigned long long* __restrict__ key, unsigned* __restrict__ idx, int* __restrict__ bad) {
    const int64_t k = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
    if (k >= nz) return;
    int i = ti[k], j = tj[k];
    if (i < 0 || i >= n || j < 0 || j >= n) { *bad = 1; i = 0; j = 0; }
    if (j < i) { const int t = i; i = j; j = t; }
    key[k] = (static_cast<unsigned long long>(static_cast<unsigned>(j)) << 32) | static_cast<unsigned>(i);
    idx[k] = static_cast<unsigned>(k);
}

// per-warp histogram of one 8-bit digit: hist[digit * nwarps + warp]
__global__ void __launch_bounds__(256) k_in_hist(int64_t nz, const unsigned long long* __restrict__ key, int shift, int nwarps,
                                                 unsigned* __restrict__ hist) {
    __shared__ unsigned h[8][256];
    const int lane = threadIdx.x & 31, wl = threadIdx.x >> 5;
    const int64_t gw = blockIdx.x * static_cast<int64_t>(8) + wl;
    for (int d = lane; d < 256; d += 32) h[wl][d] = 0;
    __syncwarp();
    if (gw < nwarps) {
        const int64_t b = gw * IN_WARP_ITEMS, e = min(b + IN_WARP_ITEMS, nz);
        for (int64_t q = b + lane; q < e; q += 32) atomicAdd(&h[wl][(key[q] >> shift) & 255u], 1u);
        __syncwarp();
        for (int d = lane; d < 256; d += 32) hist[static_cast<int64_t>(d) * nwarps + gw] = h[wl][d];
    }
}

// stable scatter of one pass: off[digit * nwarps + warp] = first output slot of that (digit, warp)
__global__ void __launch_bounds__(256) k_in_scatter(int64_t nz, const unsigned long long* __restrict__ key, const unsigned* __restrict__ idx,
                                                    int shift, int nwarps, const unsigned* __restrict__ off,
                                                    unsigned long long* __restrict__ key_out, unsigned* __restrict__ idx_out) {
    __shared__ unsigned o[8][256];
    const int lane = threadIdx.x & 31, wl = threadIdx.x >> 5;
    const int64_t gw = blockIdx.x * static_cast<int64_t>(8) + wl;
    if (gw >= nwarps) return;
    for (int d = lane; d < 256; d += 32) o[wl][d] = off[static_cast<int64_t>(d) * nwarps + gw];
    __syncwarp();
    const int64_t b = gw * IN_WARP_ITEMS, e = min(b + IN_WARP_ITEMS, nz);
    const unsigned lt = (1u << lane) - 1u;
    for (int64_t q0 = b; q0 < e; q0 += 32) {
        const int64_t q = q0 + lane;
        const bool on = q < e;
        const unsigned long long kq = on ? key[q] : 0ull;
        const unsigned pay = on ? idx[q] : 0u;
        const unsigned d = on ? static_cast<unsigned>((kq >> shift) & 255u) : 256u + lane;      // idle lanes: private digits, no peers
        const unsigned peers = __match_any_sync(0xffffffffu, d);
        const unsigned rank = __popc(peers & lt);
        unsigned base = 0;
        if (on) base = o[wl][d];
        __syncwarp();
        if (on) {
            key_out[base + rank] = kq; idx_out[base + rank] = pay;
            if (rank == 0) o[wl][d] = base + __popc(peers);          // the first lane of the group advances the cursor
        }
        __syncwarp();
    }
}

// ---- exclusive scan of unsigned counters (two levels: 2048-element CTAs, then their sums) -------------------------
__global__ void __launch_bounds__(256) k_in_scan_block(int64_t n, const unsigned* __restrict__ in, unsigned* __restrict__ out,
                                                       unsigned* __restrict__ sums) {
    __shared__ unsigned wsum[8];
    const int64_t base = blockIdx.x * static_cast<int64_t>(IN_SCAN_BLOCK) + threadIdx.x * 8;
    unsigned v[8], tot = 0;
#pragma unroll
    for (int u = 0; u < 8; ++u) { v[u] = base + u < n ? in[base + u] : 0u; tot += v[u]; }
    const int lane = threadIdx.x & 31, wl = threadIdx.x >> 5;
    unsigned inc = tot;
    for (int o = 1; o < 32; o <<= 1) { const unsigned t = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += t; }
    if (lane == 31) wsum[wl] = inc;
    __syncthreads();
    unsigned wbase = 0;
    for (int q = 0; q < wl; ++q) wbase += wsum[q];
    unsigned run = wbase + inc - tot;
#pragma unroll
    for (int u = 0; u < 8; ++u) { if (base + u < n) out[base + u] = run; run += v[u]; }
    if (threadIdx.x == 255 && sums) sums[blockIdx.x] = wbase + inc;
}
__global__ void k_in_scan_sums(int nb, unsigned* __restrict__ sums) {          // one CTA: exclusive scan of the block sums, in place
    __shared__ unsigned carry;
    __shared__ unsigned wsum[8];
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (int b0 = 0; b0 < nb; b0 += 256) {
        const int i = b0 + threadIdx.x;
        const unsigned v = i < nb ? sums[i] : 0u;
        const int lane = threadIdx.x & 31, wl = threadIdx.x >> 5;
        unsigned inc = v;
        for (int o = 1; o < 32; o <<= 1) { const unsigned t = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += t; }
        if (lane == 31) wsum[wl] = inc;
        __syncthreads();
        unsigned wbase = carry;
        for (int q = 0; q < wl; ++q) wbase += wsum[q];
        if (i < nb) sums[i] = wbase + inc - v;
        __syncthreads();
        if (threadIdx.x == 255) carry = wbase + inc;
        __syncthreads();
    }
}
__global__ void __launch_bounds__(256) k_in_scan_add(int64_t n, unsigned* __restrict__ out, const unsigned* __restrict__ sums) {
    const unsigned add = sums[blockIdx.x];
    const int64_t base = blockIdx.x * static_cast<int64_t>(IN_SCAN_BLOCK) + threadIdx.x * 8;
#pragma unroll
    for (int u = 0; u < 8; ++u) if (base + u < n) out[base + u] += add;
}

// head[q] = 1 where a new (row, column) starts; per-column counts of merged entries
__global__ void k_in_heads(int64_t nz, const unsigned long long* __restrict__ key, unsigned* __restrict__ head, unsigned* __restrict__ colcnt) {
    const int64_t q = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
    if (q >= nz) return;
    const bool h = q == 0 || key[q] != key[q - 1];
    head[q] = h ? 1u : 0u;
    if (h) atomicAdd(&colcnt[key[q] >> 32], 1u);
}
// one thread per run: slot = scanned head flag; the run's values are added in sorted (= input) order
__global__ void k_in_emit(int64_t nz, const unsigned long long* __restrict__ key, const unsigned* __restrict__ idx,
                          const unsigned* __restrict__ head, const unsigned* __restrict__ slot, const double* __restrict__ tx,
                          int* __restrict__ Ai, double* __restrict__ Ax) {
    const int64_t q = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x;
    if (q >= nz || !head[q]) return;
    const unsigned long long k = key[q];
    double s = tx[idx[q]];
    for (int64_t r = q + 1; r < nz && key[r] == k; ++r) s += tx[idx[r]];
    const unsigned m = slot[q];
    Ai[m] = static_cast<int>(k & 0xffffffffull);
    Ax[m] = s;
}
__global__ void k_in_ap(int n, const unsigned* __restrict__ colscan, unsigned total, int* __restrict__ Ap) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j < n) Ap[j] = static_cast<int>(colscan[j]);
    if (j == n) Ap[n] = static_cast<int>(total);
}

}  // namespace spchol
