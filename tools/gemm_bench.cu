// Stand-alone timing of the DMMA update kernels (C -= A * B^T, lower trapezoid) on the shapes the
// multifrontal plan produces.  Builds against csrc/kernels.cuh so the measured kernels are the
// product's.  Usage: gemm_bench [variant-filter]
//   nvcc -O3 -std=c++17 -lineinfo -gencode arch=compute_100a,code=sm_100a -I sparsecholesky_b200/csrc \
//        tools/gemm_bench.cu -o tools/gemm_bench
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "kernels.cuh"

using namespace spchol;

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1); } } while (0)

struct Shape { const char* name; int M, N, K, diag, ntask; int aoff_odd; };

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

template <int BM, int BN, int WGM, int WGN, int ST>
static void launch2(const GemmTask* dt, int nt, int gx, int gy, double* L, double* W) {
    static bool init = false;
    if (!init) { CK(cudaFuncSetAttribute(k_gemm2<BM, BN, WGM, WGN, ST>, cudaFuncAttributeMaxDynamicSharedMemorySize, gemm2_smem<BM, BN, ST>())); init = true; }
    k_gemm2<BM, BN, WGM, WGN, ST><<<dim3((gx + BM - 1) / BM, (gy + BN - 1) / BN, nt), WGM * WGN * 32, gemm2_smem<BM, BN, ST>()>>>(dt, L, W);
}
static void launch1(const GemmTask* dt, int nt, int gx, int gy, double* L, double* W, int swz = 0) {
    static bool init = false;
    if (!init) { CK(cudaFuncSetAttribute(k_gemm_dmma, cudaFuncAttributeMaxDynamicSharedMemorySize, G_SMEM)); init = true; }
    k_gemm_dmma<<<dim3((gx + G_BM - 1) / G_BM, (gy + G_BN - 1) / G_BN, nt), 256, G_SMEM>>>(dt, L, W, swz);
}

int main(int argc, char** argv) {
    const char* filt = argc > 1 ? argv[1] : "";
    const Shape shapes[] = {
        {"schur_8k_k2k", 8192, 8192, 2048, 0, 1, 0},
        {"schur_8k_k2k_odd", 8192, 8192, 2048, 0, 1, 1},
        {"schur_12k_k4k", 12288, 12288, 4096, 0, 1, 1},
        {"schur_16k_k8k_x2", 16256, 16256, 8192, 0, 2, 0},
        {"panel_16k_n256_k256", 16384, 256, 256, 0, 1, 0},
        {"panel_16k_n1792_k256", 16384, 1792, 256, 0, 1, 0},
        {"inner_16k_n128_k128", 16384, 128, 128, 0, 1, 0},
        {"mid_64x_m1500_k200", 1500, 1500, 200, 0, 64, 1},
        {"mid_512x_m400_k100", 400, 400, 100, 0, 512, 1},
        {"dist_24k_n256_k256_x4", 24576, 256, 256, 0, 4, 0},
    };
    for (const Shape& s : shapes) {
        // layout per task: a panel of (K cols) with ld = M + K + pad, A = B = rows [K.., ) ; C separate in W
        const int ld = ((s.M + s.K + 2) + 1) & ~1;
        const int64_t per_task_L = (int64_t)ld * s.K + 2;
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
        for (int t = 0; t < s.ntask; ++t) {
            const int64_t a = t * per_task_L + s.K + (s.aoff_odd ? 1 : 0);
            ht[t] = {t * per_task_W, a, a, ldc, ld, ld, s.M, s.N, s.K, s.diag, 1};
        }
        GemmTask* dt; CK(cudaMalloc(&dt, sizeof(GemmTask) * s.ntask));
        CK(cudaMemcpy(dt, ht.data(), sizeof(GemmTask) * s.ntask, cudaMemcpyHostToDevice));
        const double fl = alg_flops(s);
        bool have_ref = false;
        auto run = [&](const char* vname, auto fn) {
            if (*filt && !strstr(vname, filt)) return;
            CK(cudaMemcpy(W, W0, per_task_W * s.ntask * 8, cudaMemcpyDeviceToDevice));
            fn();
            CK(cudaDeviceSynchronize());
            double diff = -1;
            if (!have_ref) { CK(cudaMemcpy(Wref, W, per_task_W * s.ntask * 8, cudaMemcpyDeviceToDevice)); have_ref = true; }
            else {
                CK(cudaMemset(dmax, 0, 8));
                k_diff<<<1024, 256>>>(W, Wref, per_task_W * s.ntask, dmax);
                CK(cudaMemcpy(&diff, dmax, 8, cudaMemcpyDeviceToHost));
            }
            const float ms = time_it(fn, 5);
            printf("{\"shape\": \"%s\", \"variant\": \"%s\", \"ms\": %.4f, \"tflops\": %.2f, \"maxdiff_vs_first\": %.3g}\n",
                   s.name, vname, ms, fl / ms * 1e-9, diff);
            fflush(stdout);
        };
        run("v1_128x64_regstage", [&] { launch1(dt, s.ntask, s.M, s.N, L, W); });
        run("v1s_128x64_regstage_swizzle8", [&] { launch1(dt, s.ntask, s.M, s.N, L, W, 1); });
        run("v2_128x64_w4x2_s3", [&] { launch2<128, 64, 4, 2, 3>(dt, s.ntask, s.M, s.N, L, W); });
        run("v2_128x64_w4x2_s4", [&] { launch2<128, 64, 4, 2, 4>(dt, s.ntask, s.M, s.N, L, W); });
        run("v2_128x128_w2x4_s3", [&] { launch2<128, 128, 2, 4, 3>(dt, s.ntask, s.M, s.N, L, W); });
        run("v2_128x128_w4x2_s3", [&] { launch2<128, 128, 4, 2, 3>(dt, s.ntask, s.M, s.N, L, W); });
        run("v2_128x128_w2x4_s4", [&] { launch2<128, 128, 2, 4, 4>(dt, s.ntask, s.M, s.N, L, W); });
        run("v2_128x128_w4x4_s3", [&] { launch2<128, 128, 4, 4, 3>(dt, s.ntask, s.M, s.N, L, W); });
        run("v2_256x64_w4x2_s3", [&] { launch2<256, 64, 4, 2, 3>(dt, s.ntask, s.M, s.N, L, W); });
        run("v2_64x64_w2x2_s4", [&] { launch2<64, 64, 2, 2, 4>(dt, s.ntask, s.M, s.N, L, W); });
        cudaFree(L); cudaFree(W); cudaFree(W0); cudaFree(Wref); cudaFree(dmax); cudaFree(dt);
    }
    return 0;
}
