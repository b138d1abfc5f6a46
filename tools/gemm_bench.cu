// Stand-alone timing + cross-check of the two DMMA update kernels (C -= A * B^T, lower trapezoid) on the shapes the
// multifrontal plan produces: k_gemm_dmma (register-staged, round 1) against k_gemm_tma (TMA + mbarrier ring, round 2).
// Builds against csrc/ so the measured kernels are the product's.  Usage: gemm_bench [shape-filter]
//   nvcc -O3 -std=c++17 -lineinfo -gencode arch=compute_100a,code=sm_100a -I sparsecholesky_b200/csrc tools/gemm_bench.cu -o tools/gemm_bench
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "kernels.cuh"
#include "gemm_tma.cuh"

using namespace spchol;

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1); } } while (0)

// a task: panel of m = r0 + M rows and wpan >= kc + K columns; A = rows [r0, r0 + M), B = rows [r0, r0 + N), columns [kc, kc + K)
struct Shape { const char* name; int M, N, K, diag, ntask, r0, kc, extra_cols; };

static double alg_flops(const Shape& s) {
    double cnt = 0;
    for (int j = 0; j < s.N; ++j) { int rows = s.M - std::max(0, j - s.diag); if (rows > 0) cnt += rows; }
    return 2.0 * cnt * s.K * s.ntask;
}
template <class F>
static float time_it(F f, int reps) {
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    f(); f();
    CK(cudaDeviceSynchronize());
    float best = 1e30f;
    for (int r = 0; r < reps; ++r) {
        CK(cudaEventRecord(e0)); f(); CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1)); best = std::min(best, ms);
    }
    CK(cudaGetLastError());
    return best;
}
__global__ void k_fill(double* p, int64_t n) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        p[i] = ((i * 2654435761u) & 1023) * (1.0 / 1024) - 0.5;
}
__global__ void k_diff(const double* a, const double* b, int64_t n, double* out) {
    double m = 0;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        m = fmax(m, fabs(a[i] - b[i]));
    for (int o = 16; o; o >>= 1) m = fmax(m, __shfl_xor_sync(~0u, m, o));
    if ((threadIdx.x & 31) == 0 && m > 0) atomicMax(reinterpret_cast<unsigned long long*>(out), __double_as_longlong(m));
}

int main(int argc, char** argv) {
    const char* filt = argc > 1 ? argv[1] : "";
    using EncodeFn = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    void* fn = nullptr; cudaDriverEntryPointQueryResult qr;
    CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qr));
    CK(cudaFuncSetAttribute(k_gemm_dmma, cudaFuncAttributeMaxDynamicSharedMemorySize, G_SMEM));
    CK(cudaFuncSetAttribute(k_gemm_tma, cudaFuncAttributeMaxDynamicSharedMemorySize, G3_SMEM));
    const Shape shapes[] = {
        {"schur_8k_k2k", 8192, 8192, 2048, 0, 1, 2048, 0, 0},
        {"schur_8k_k2k_oddrow", 8191, 8191, 2047, 0, 1, 2047, 0, 0},
        {"schur_12k_k4k", 12288, 12288, 4096, 0, 1, 4097, 0, 0},
        {"schur_16k_k8k_x2", 16256, 16256, 8192, 0, 2, 8192, 0, 0},
        {"defer_16k_n512_k512", 16384, 512, 512, 0, 1, 1024, 512, 1024},
        {"panel_16k_n256_k256", 16384, 256, 256, 0, 1, 512, 256, 512},
        {"panel_16k_n1792_k256", 16384, 1792, 256, 0, 1, 512, 256, 2048},
        {"inner_16k_n128_k128", 16384, 128, 128, 0, 1, 256, 128, 256},
        {"inner_x64_m3000_n128_k128", 3000, 128, 128, 0, 64, 256, 128, 200},
        {"mid_64x_m1500_k200", 1500, 1500, 200, 0, 64, 201, 0, 0},
        {"mid_512x_m400_k100", 400, 400, 100, 0, 512, 101, 0, 0},
        {"tail_k37_m777_n333", 777, 333, 37, 0, 8, 55, 19, 70},
        {"dist_24k_n256_k256_x4", 24576, 256, 256, 0, 4, 256, 0, 512},
    };
    for (const Shape& s : shapes) {
        if (*filt && !strstr(s.name, filt)) continue;
        const int m = s.r0 + s.M, wpan = s.kc + s.K + s.extra_cols;
        const int ld = (m + 1) & ~1;
        const int64_t per_task_L = (int64_t)ld * wpan;
        const int ldc = (s.M + 1) & ~1;
        const int64_t per_task_W = (int64_t)ldc * s.N;
        double *L, *W, *W0, *Wref, *dmax;
        CK(cudaMalloc(&L, per_task_L * s.ntask * 8));
        CK(cudaMalloc(&W, per_task_W * s.ntask * 8));
        CK(cudaMalloc(&W0, per_task_W * s.ntask * 8));
        CK(cudaMalloc(&Wref, per_task_W * s.ntask * 8));
        CK(cudaMalloc(&dmax, 8));
        k_fill<<<1024, 256>>>(L, per_task_L * s.ntask);
        k_fill<<<1024, 256>>>(W0, per_task_W * s.ntask);
        std::vector<GemmTask> ht(s.ntask);
        std::vector<CUtensorMap> maps(s.ntask);
        for (int t = 0; t < s.ntask; ++t) {
            const int64_t a = t * per_task_L + s.r0 + (int64_t)s.kc * ld;
            ht[t] = {t * per_task_W, a, a, ldc, ld, ld, s.M, s.N, s.K, s.diag, 1, t, s.r0, s.r0, s.kc};
            const cuuint64_t dims[2] = {(cuuint64_t)m, (cuuint64_t)wpan};
            const cuuint64_t strides[1] = {(cuuint64_t)ld * 8};
            const cuuint32_t box[2] = {16, (cuuint32_t)G3_BK};
            const cuuint32_t estr[2] = {1, 1};
            const CUresult cr = reinterpret_cast<EncodeFn>(fn)(&maps[t], CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 2, L + t * per_task_L, dims, strides, box, estr,
                                                               CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                                                               CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
            if (cr != CUDA_SUCCESS) { printf("cuTensorMapEncodeTiled failed: %d\n", (int)cr); return 1; }
        }
        GemmTask* dt; CK(cudaMalloc(&dt, sizeof(GemmTask) * s.ntask));
        CK(cudaMemcpy(dt, ht.data(), sizeof(GemmTask) * s.ntask, cudaMemcpyHostToDevice));
        CUtensorMap* dm; CK(cudaMalloc(&dm, sizeof(CUtensorMap) * s.ntask));
        CK(cudaMemcpy(dm, maps.data(), sizeof(CUtensorMap) * s.ntask, cudaMemcpyHostToDevice));
        const double fl = alg_flops(s);
        bool have_ref = false;
        auto run = [&](const char* vname, auto f) {
            CK(cudaMemcpy(W, W0, per_task_W * s.ntask * 8, cudaMemcpyDeviceToDevice));
            f();
            CK(cudaDeviceSynchronize());
            double diff = -1;
            if (!have_ref) { CK(cudaMemcpy(Wref, W, per_task_W * s.ntask * 8, cudaMemcpyDeviceToDevice)); have_ref = true; }
            else {
                CK(cudaMemset(dmax, 0, 8));
                k_diff<<<1024, 256>>>(W, Wref, per_task_W * s.ntask, dmax);
                CK(cudaMemcpy(&diff, dmax, 8, cudaMemcpyDeviceToHost));
            }
            const float ms = time_it(f, 5);
            printf("{\"shape\": \"%s\", \"M\": %d, \"N\": %d, \"K\": %d, \"tasks\": %d, \"variant\": \"%s\", \"ms\": %.4f, \"tflops\": %.2f, \"maxdiff_vs_first\": %.3g}\n",
                   s.name, s.M, s.N, s.K, s.ntask, vname, ms, fl / ms * 1e-9, diff);
            fflush(stdout);
        };
        const dim3 g1((s.M + G_BM - 1) / G_BM, (s.N + G_BN - 1) / G_BN, s.ntask), g3((s.M + 1 + G3_BM - 1) / G3_BM, (s.N + 1 + G3_BN - 1) / G3_BN, s.ntask);
        run("r1_regstage", [&] { k_gemm_dmma<<<g1, 256, G_SMEM>>>(dt, L, W, 3); });
        run("r2_tma", [&] { k_gemm_tma<<<g3, 256, G3_SMEM>>>(dt, dm, L, W, 3); });
        run("r2_tma_red", [&] { k_gemm_tma<<<g3, 256, G3_SMEM>>>(dt, dm, L, W, 7); });
        run("r2_tma_red_freebars", [&] { k_gemm_tma<<<g3, 256, G3_SMEM>>>(dt, dm, L, W, 15); });
        cudaFree(L); cudaFree(W); cudaFree(W0); cudaFree(Wref); cudaFree(dmax); cudaFree(dt); cudaFree(dm);
    }
    return 0;
}
