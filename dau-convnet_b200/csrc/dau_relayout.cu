// Re-layout of the top error for the gradient kernel (the blurs live in dau_blur_pair.cu).
#include "dau_kernels.cuh"

namespace dau {

namespace {
// top error NCHW -> SGP[n/2][f][y][x%2][x/2][n%2] of the MOSAIC image (row of x/2 padded to an even
// count; gaps and cells beyond the batch are zeros), with the ignore_edge_gradients mask
// (dau_conv_backward_core.hpp:2218-2222) applied on the way.  One thread moves the column pair
// (2j, 2j+1) of one row of both mosaic images of the pair; the float2 stores are coalesced.
__global__ void __launch_bounds__(256) error_relayout_kernel(const float* __restrict__ err, float2* __restrict__ out,
                                                             const __grid_constant__ Mosaic m, int F, int qp,
                                                             int mask_row, int mask_col,
                                                             const float* __restrict__ act, int crop) {
    const int pf = blockIdx.x;                            // (pair, feature)
    const int np = pf / F, f = pf - np * F;
    const int idx = blockIdx.y * 256 + threadIdx.x;       // (row, column pair) of the mosaic plane
    const int vy = idx / qp, j = idx - vy * qp;
    if (vy >= m.Hv) return;
    float2* dst = out + ((size_t)pf * m.Hv + vy) * 2 * qp;
    const int RH = m.H - 2 * crop, RW = m.W - 2 * crop;   // the error tensor sits at (crop, crop) of the full grid
    const size_t rplane = (size_t)RH * RW;
    float v[2][2] = {{0.f, 0.f}, {0.f, 0.f}};             // [column 2j / 2j+1][image of the pair]
    if (m.kx == 1 && m.ky == 1 && (m.W & 1) == 0 && crop == 0) {
        // identity mosaic, even width: one float2 load per image
        const int x = 2 * j;
        if (x < m.W && !(mask_row && vy == m.H - 1)) {
#pragma unroll
            for (int img = 0; img < 2; ++img) {
                const int n = 2 * np + img;
                if (n >= m.N) continue;
                const size_t at = ((size_t)n * F + f) * rplane + (size_t)vy * m.W + x;
                float2 t = __ldg(reinterpret_cast<const float2*>(err + at));
                if (act) {                                // ReLU backward: zero where the activation output was
                    const float2 a2 = __ldg(reinterpret_cast<const float2*>(act + at));
                    t.x = a2.x > 0.f ? t.x : 0.f; t.y = a2.y > 0.f ? t.y : 0.f;
                }
                v[0][img] = t.x;
                v[1][img] = (mask_col && x + 1 == m.W - 1) ? 0.f : t.y;
            }
        }
    } else {
#pragma unroll
        for (int c = 0; c < 2; ++c)
#pragma unroll
            for (int img = 0; img < 2; ++img) {
                const int vx = 2 * j + c, nv = 2 * np + img;
                int rn, ry, rx;
                if (vx < m.Wv && nv < m.Nv && mosaic_to_real(m, nv, vy, vx, &rn, &ry, &rx)) {
                    ry -= crop; rx -= crop;
                    if (ry < 0 || ry >= RH || rx < 0 || rx >= RW) continue;
                    const bool dead = (mask_row && ry == RH - 1) || (mask_col && rx == RW - 1);
                    const size_t at = ((size_t)rn * F + f) * rplane + (size_t)ry * RW + rx;
                    if (!dead && !(act && !(__ldg(act + at) > 0.f))) v[c][img] = __ldg(err + at);
                }
            }
    }
    dst[j] = make_float2(v[0][0], v[0][1]);
    dst[qp + j] = make_float2(v[1][0], v[1][1]);
}
}  // namespace

void launch_error_relayout(const float* err, float* out_sg, const Mosaic& m, int F, int mask_row, int mask_col,
                           cudaStream_t stream, const float* act_mask, int crop) {
    const int qv = (m.Wv + 1) / 2, qp = (qv + 1) / 2 * 2;
    const long long total = (long long)((m.Nv + 1) / 2) * F;
    const int blocks_y = (m.Hv * qp + 255) / 256;
    DAU_CHECK(total < (1LL << 31) && blocks_y <= 65535, 4, "shape too large for the error re-layout grid");
    error_relayout_kernel<<<dim3((unsigned)total, blocks_y), 256, 0, stream>>>(err, reinterpret_cast<float2*>(out_sg), m, F, qp,
                                                                           mask_row, mask_col, act_mask, crop);
    DAU_CUDA_CHECK(cudaGetLastError());
}

}  // namespace dau
