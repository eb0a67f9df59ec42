// FP32 FMA-pipe probe for sm_100a: how fast do scalar FFMA and packed FFMA2 issue for the
// operand patterns the DAU kernels use?  Prints warp-instructions / clk / SMSP and TFLOP/s.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o fma_probe fma_probe.cu
#include <cstdio>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA %s at %d\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)

// A: 16 independent chains, every FFMA reads three distinct, changing registers
__global__ void __launch_bounds__(256) k_scalar_distinct(float* sink, int iters) {
    float a[16], x[16], y[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) { a[j] = threadIdx.x + j; x[j] = 1.f + 1e-7f * (threadIdx.x + j); y[j] = 1e-9f * j * threadIdx.x; }
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int rep = 0; rep < 4; ++rep)
#pragma unroll
            for (int j = 0; j < 16; ++j) a[j] = fmaf(x[j], y[(j + rep) & 15], a[j]);
    }
    float t = 0.f;
#pragma unroll
    for (int j = 0; j < 16; ++j) t += a[j];
    if (t == 123.456f) sink[0] = t;
}

// B: the gradient kernel's pattern: acc[u][k] += e[u] * xk[k]  (e reused by 4 consecutive FFMAs)
__global__ void __launch_bounds__(256) k_scalar_outer(float* sink, int iters) {
    float acc[8][4], e[8], xk[4];
#pragma unroll
    for (int u = 0; u < 8; ++u) { e[u] = 1.f + 1e-7f * (threadIdx.x + u);
#pragma unroll
        for (int k = 0; k < 4; ++k) acc[u][k] = u + k; }
#pragma unroll
    for (int k = 0; k < 4; ++k) xk[k] = 1e-9f * (k + 1) * threadIdx.x;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int rep = 0; rep < 2; ++rep)
#pragma unroll
            for (int u = 0; u < 8; ++u)
#pragma unroll
                for (int k = 0; k < 4; ++k) acc[u][k] = fmaf(e[u], xk[k], acc[u][k]);
    }
    float t = 0.f;
#pragma unroll
    for (int u = 0; u < 8; ++u)
#pragma unroll
        for (int k = 0; k < 4; ++k) t += acc[u][k];
    if (t == 123.456f) sink[0] = t;
}

// C: FFMA2, three distinct register pairs per instruction
__global__ void __launch_bounds__(256) k_packed_distinct(float* sink, int iters) {
    float2 a[16], x[16], y[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) {
        a[j] = make_float2(threadIdx.x + j, j);
        x[j] = make_float2(1.f + 1e-7f * (threadIdx.x + j), 1.f - 1e-7f * j);
        y[j] = make_float2(1e-9f * j * threadIdx.x, 1e-9f * j);
    }
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int rep = 0; rep < 4; ++rep)
#pragma unroll
            for (int j = 0; j < 16; ++j) a[j] = __ffma2_rn(x[j], y[(j + rep) & 15], a[j]);
    }
    float t = 0.f;
#pragma unroll
    for (int j = 0; j < 16; ++j) t += a[j].x + a[j].y;
    if (t == 123.456f) sink[0] = t;
}

// D: FFMA2 outer product: acc2[u][k] += e2[u] * xk2[k]
__global__ void __launch_bounds__(256) k_packed_outer(float* sink, int iters) {
    float2 acc[8][4], e[8], xk[4];
#pragma unroll
    for (int u = 0; u < 8; ++u) { e[u] = make_float2(1.f + 1e-7f * (threadIdx.x + u), 1.f - 1e-7f * u);
#pragma unroll
        for (int k = 0; k < 4; ++k) acc[u][k] = make_float2(u + k, u - k); }
#pragma unroll
    for (int k = 0; k < 4; ++k) xk[k] = make_float2(1e-9f * (k + 1) * threadIdx.x, 1e-9f * k);
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int rep = 0; rep < 2; ++rep)
#pragma unroll
            for (int u = 0; u < 8; ++u)
#pragma unroll
                for (int k = 0; k < 4; ++k) acc[u][k] = __ffma2_rn(e[u], xk[k], acc[u][k]);
    }
    float t = 0.f;
#pragma unroll
    for (int u = 0; u < 8; ++u)
#pragma unroll
        for (int k = 0; k < 4; ++k) t += acc[u][k].x + acc[u][k].y;
    if (t == 123.456f) sink[0] = t;
}

// E: scalar, multiplicand and addend fixed (the bench.py peak probe's pattern)
__global__ void __launch_bounds__(256) k_scalar_const(float* sink, int iters) {
    float a[16];
    const float m = 1.0000001f, c = 1e-9f * (float)threadIdx.x;
#pragma unroll
    for (int j = 0; j < 16; ++j) a[j] = (float)(threadIdx.x + j);
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int rep = 0; rep < 4; ++rep)
#pragma unroll
            for (int j = 0; j < 16; ++j) a[j] = fmaf(a[j], m, c);
    }
    float t = 0.f;
#pragma unroll
    for (int j = 0; j < 16; ++j) t += a[j];
    if (t == 123.456f) sink[0] = t;
}

// F/G: the gather kernel's pattern acc2[j] += w2[t] * v2[j]: weight pair warp-uniform.
// F keeps the weights in the constant bank (ptxas: FFMA2 R, R, UR, R), G in vector registers.
struct WPar { float2 w[8]; };
__global__ void __launch_bounds__(256) k_packed_uniform_w(const __grid_constant__ WPar p, float* sink, int iters) {
    float2 acc[16], v[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) { acc[j] = make_float2(threadIdx.x + j, j); v[j] = make_float2(1e-9f * j * threadIdx.x, 1e-9f * j); }
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int t = 0; t < 4; ++t)
#pragma unroll
            for (int j = 0; j < 16; ++j) acc[j] = __ffma2_rn(p.w[t], v[(j + t) & 15], acc[j]);
    }
    float t = 0.f;
#pragma unroll
    for (int j = 0; j < 16; ++j) t += acc[j].x + acc[j].y;
    if (t == 123.456f) sink[0] = t;
}
__global__ void __launch_bounds__(256) k_packed_vector_w(const float2* wsrc, float* sink, int iters) {
    float2 acc[16], v[16], w[4];
#pragma unroll
    for (int t = 0; t < 4; ++t) w[t] = wsrc[t * 256 + threadIdx.x];
#pragma unroll
    for (int j = 0; j < 16; ++j) { acc[j] = make_float2(threadIdx.x + j, j); v[j] = make_float2(1e-9f * j * threadIdx.x, 1e-9f * j); }
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int t = 0; t < 4; ++t)
#pragma unroll
            for (int j = 0; j < 16; ++j) acc[j] = __ffma2_rn(w[t], v[(j + t) & 15], acc[j]);
    }
    float t = 0.f;
#pragma unroll
    for (int j = 0; j < 16; ++j) t += acc[j].x + acc[j].y;
    if (t == 123.456f) sink[0] = t;
}
// H: scalar version of F (FFMA R, R, UR, R)
__global__ void __launch_bounds__(256) k_scalar_uniform_w(const __grid_constant__ WPar p, float* sink, int iters) {
    float acc[16], v[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) { acc[j] = threadIdx.x + j; v[j] = 1e-9f * j * threadIdx.x; }
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int t = 0; t < 4; ++t)
#pragma unroll
            for (int j = 0; j < 16; ++j) acc[j] = fmaf(p.w[t].x, v[(j + t) & 15], acc[j]);
    }
    float t = 0.f;
#pragma unroll
    for (int j = 0; j < 16; ++j) t += acc[j];
    if (t == 123.456f) sink[0] = t;
}

typedef void (*kern_t)(float*, int);

int main() {
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, 0));
    int clk_khz = 0;
    CK(cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0));
    float* sink; CK(cudaMalloc(&sink, 4));
    struct { const char* name; kern_t k; int fma_per_inst; } tests[] = {
        {"scalar FFMA, 3 distinct regs      ", k_scalar_distinct, 1},
        {"scalar FFMA, outer e[u]*xk[k]     ", k_scalar_outer, 1},
        {"FFMA2, 3 distinct reg pairs       ", k_packed_distinct, 2},
        {"FFMA2, outer e2[u]*xk2[k]         ", k_packed_outer, 2},
        {"scalar FFMA, a=a*m+c (peak probe) ", k_scalar_const, 1},
    };
    const int iters = 20000, per_iter = 64;
    for (int warps_per_smsp = 1; warps_per_smsp <= 4; warps_per_smsp *= 2) {
        const int blocks = prop.multiProcessorCount * warps_per_smsp / 2;   // 256 thr = 8 warps = 2 / SMSP
        printf("-- %d blocks of 256 threads (%d warp(s) per SMSP)\n", blocks > 0 ? blocks : 1, warps_per_smsp);
        for (auto& t : tests) {
            const int nb = warps_per_smsp == 1 ? prop.multiProcessorCount : blocks;
            const int nthreads = warps_per_smsp == 1 ? 128 : 256;
            cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
            t.k<<<nb, nthreads>>>(sink, 100);
            CK(cudaDeviceSynchronize());
            CK(cudaEventRecord(e0));
            t.k<<<nb, nthreads>>>(sink, iters);
            CK(cudaEventRecord(e1));
            CK(cudaDeviceSynchronize());
            float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
            const double inst = (double)nb * (nthreads / 32) * iters * per_iter;
            const double tflops = inst * 32 * t.fma_per_inst * 2 / (ms * 1e-3) / 1e12;
            const double per_clk = inst / (ms * 1e-3 * clk_khz * 1e3) / (prop.multiProcessorCount * 4.0);
            printf("%s %8.3f ms  %6.2f TFLOP/s  %.3f warp-inst/clk/SMSP (at %d MHz nominal)\n", t.name, ms, tflops,
                   per_clk, clk_khz / 1000);
        }
    }
    {
        WPar wp; for (int t = 0; t < 8; ++t) wp.w[t] = make_float2(1.f + 1e-7f * t, 1.f - 1e-7f * t);
        float2* wsrc; CK(cudaMalloc(&wsrc, 4 * 256 * sizeof(float2))); CK(cudaMemset(wsrc, 0, 4 * 256 * sizeof(float2)));
        const int nb = prop.multiProcessorCount * 2, nthreads = 256;
        for (int which = 0; which < 3; ++which) {
            cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
            for (int rep = 0; rep < 2; ++rep) {
                const int it = rep == 0 ? 100 : iters;
                if (rep == 1) CK(cudaEventRecord(e0));
                if (which == 0) k_packed_uniform_w<<<nb, nthreads>>>(wp, sink, it);
                else if (which == 1) k_packed_vector_w<<<nb, nthreads>>>(wsrc, sink, it);
                else k_scalar_uniform_w<<<nb, nthreads>>>(wp, sink, it);
                if (rep == 1) CK(cudaEventRecord(e1));
                CK(cudaDeviceSynchronize());
            }
            float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
            const double inst = (double)nb * (nthreads / 32) * iters * 64;
            const int fpi = which == 2 ? 1 : 2;
            printf("%s %8.3f ms  %6.2f TFLOP/s  %.3f warp-inst/clk/SMSP\n",
                   which == 0 ? "FFMA2 acc+=w*v, w in uniform regs " : which == 1 ? "FFMA2 acc+=w*v, w in vector regs  " : "FFMA  acc+=w*v, w in uniform regs ",
                   ms, inst * 32 * fpi * 2 / (ms * 1e-3) / 1e12, inst / (ms * 1e-3 * clk_khz * 1e3) / (prop.multiProcessorCount * 4.0));
        }
    }
    return 0;
}
