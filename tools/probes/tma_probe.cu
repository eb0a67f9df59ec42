// Standalone TMA probe: which descriptor / instruction variants run on this box?
#include "../../dau-convnet_b200/csrc/dau_tma.cuh"
#include <cstdio>
#include <cstdlib>
#include <vector>
using namespace dau::tma;

__global__ void k4(const __grid_constant__ CUtensorMap m, float* out, int c0, int c1, int c2, int c3, int nfloats) {
    extern __shared__ unsigned char dyn[];
    __shared__ uint64_t bar;
    unsigned char* s = dyn + ((128u - (smem_u32(dyn) & 127u)) & 127u);
    if (threadIdx.x == 0) { mbar_init(&bar, 1); fence_barrier_init(); }
    __syncthreads();
    if (threadIdx.x == 0) {
        mbar_arrive_expect_tx(&bar, nfloats * 4);
        tensor4_g2s(s, &m, c0, c1, c2, c3, &bar);
    }
    mbar_wait(&bar, 0);
    for (int i = threadIdx.x; i < nfloats; i += blockDim.x) out[i] = reinterpret_cast<float*>(s)[i];
}
__global__ void k2(const __grid_constant__ CUtensorMap m, float* out, int c0, int c1, int nfloats) {
    extern __shared__ unsigned char dyn[];
    __shared__ uint64_t bar;
    unsigned char* s = dyn + ((128u - (smem_u32(dyn) & 127u)) & 127u);
    if (threadIdx.x == 0) { mbar_init(&bar, 1); fence_barrier_init(); }
    __syncthreads();
    if (threadIdx.x == 0) {
        mbar_arrive_expect_tx(&bar, nfloats * 4);
        asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];\n"
                     ::"r"(smem_u32(s)), "l"(&m), "r"(c0), "r"(c1), "r"(smem_u32(&bar)) : "memory");
    }
    mbar_wait(&bar, 0);
    for (int i = threadIdx.x; i < nfloats; i += blockDim.x) out[i] = reinterpret_cast<float*>(s)[i];
}

int main(int argc, char** argv) {
    const int mode = argc > 1 ? atoi(argv[1]) : 0;
    const size_t n = 1 << 20;
    std::vector<float> h(n);
    for (size_t i = 0; i < n; ++i) h[i] = (float)i;
    float *d, *o;
    cudaMalloc(&d, n * 4); cudaMalloc(&o, 1 << 20);
    cudaMemcpy(d, h.data(), n * 4, cudaMemcpyHostToDevice);
    cudaFuncSetAttribute(k4, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
    cudaFuncSetAttribute(k2, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
    int nf = 0;
    if (mode == 0) {          // 2-D, 64x64, box 16x8
        uint64_t dims[2] = {64, 64}, str[1] = {256}; uint32_t box[2] = {16, 8};
        CUtensorMap m = make_map_f32(d, 2, dims, str, box); nf = 128;
        k2<<<1, 128, 64 * 1024>>>(m, o, 0, 0, nf);
    } else if (mode == 1) {   // 4-D like ours, W=64: dims (32,4,8,6) box (36,4,24,1)
        uint64_t dims[4] = {32, 4, 8, 6}, str[3] = {128, 512, 4096}; uint32_t box[4] = {36, 4, 24, 1};
        CUtensorMap m = make_map_f32(d, 4, dims, str, box); nf = 36 * 4 * 24;
        k4<<<1, 128, 64 * 1024>>>(m, o, -2, 0, -4, 1, nf);
    } else if (mode == 2) {   // same, coordinates inside, box inside the tensor
        uint64_t dims[4] = {32, 4, 8, 6}, str[3] = {128, 512, 4096}; uint32_t box[4] = {16, 4, 8, 1};
        CUtensorMap m = make_map_f32(d, 4, dims, str, box); nf = 16 * 4 * 8;
        k4<<<1, 128, 64 * 1024>>>(m, o, 0, 0, 0, 1, nf);
    } else if (mode == 3) {   // box inner 32 (128 B), otherwise as mode 1
        uint64_t dims[4] = {32, 4, 8, 6}, str[3] = {128, 512, 4096}; uint32_t box[4] = {32, 4, 24, 1};
        CUtensorMap m = make_map_f32(d, 4, dims, str, box); nf = 32 * 4 * 24;
        k4<<<1, 128, 64 * 1024>>>(m, o, -2, 0, -4, 1, nf);
    } else if (mode == 4) {   // as mode 1 but 3-D box rows 8
        uint64_t dims[4] = {32, 4, 8, 6}, str[3] = {128, 512, 4096}; uint32_t box[4] = {36, 4, 8, 1};
        CUtensorMap m = make_map_f32(d, 4, dims, str, box); nf = 36 * 4 * 8;
        k4<<<1, 128, 64 * 1024>>>(m, o, -2, 0, 0, 1, nf);
    }
    else if (mode == 9) {   // generic: d0 d1 d2 d3 b0 b1 b2 b3 c0 c1 c2 c3
        uint64_t dims[4]; uint32_t box[4]; int c[4];
        for (int i = 0; i < 4; ++i) { dims[i] = atoi(argv[2 + i]); box[i] = atoi(argv[6 + i]); c[i] = atoi(argv[10 + i]); }
        uint64_t str[3] = {((dims[0] * 4 + 15) / 16) * 16, 0, 0};
        str[1] = str[0] * dims[1]; str[2] = str[1] * dims[2];
        CUtensorMap m = make_map_f32(d, 4, dims, str, box); nf = box[0] * box[1] * box[2] * box[3];
        k4<<<1, 128, 64 * 1024>>>(m, o, c[0], c[1], c[2], c[3], nf);
    }
    cudaError_t e = cudaDeviceSynchronize();
    std::vector<float> r(nf);
    cudaMemcpy(r.data(), o, nf * 4, cudaMemcpyDeviceToHost);
    printf("mode %d: %s | first values %g %g %g %g %g\n", mode, cudaGetErrorString(e), r[0], r[1], r[2], r[3], r[4]);
    return 0;
}
