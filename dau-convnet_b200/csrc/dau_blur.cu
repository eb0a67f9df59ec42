// Separable Gaussian / Gaussian-derivative blurs of NCHW fp32 planes for sm_100a.
//
// Replaces caffe_gpu_convolve2 -> convolve2<...> of the reference (dense Kb x Kb "same"
// correlation, filter in __constant__ memory, src/dau_conv/util/convolve.cu:48-131,
// index algebra :107-124) in its three uses: forward blur (base_dau_conv_layer.cu:81-90),
// the 4 derivative blurs of the backward pass (:210-221) and the blur of the top error with
// the rotated filter (:278-288).  out[y,x] = sum_{v,u} in[y+v-c, x+u-c] * fy[v]*fx[u], zero
// outside the image.  Cost per pixel: (NH + NV)*Kb FMAs instead of NOUT*Kb^2.
//
// One CTA = one 32x32 output tile of one plane.  The tile plus a c-pixel halo is staged in
// shared memory with coalesced loads, filtered along x into shared memory (NH row filters),
// then along y; stores are coalesced along x.  HBM bound: reads each plane once (+halo),
// writes NOUT planes once.
#include "dau_kernels.cuh"

namespace dau {

namespace {
constexpr int BT = 32;                       // output tile edge
constexpr int BH = (kMaxBlurTaps - 1) / 2;   // max halo (8)
constexpr int BIN = BT + 2 * BH;             // staged tile edge (48)

struct BlurArgs1 { float fx[kMaxBlurTaps]; float fy[kMaxBlurTaps]; };
struct BlurArgsBank { float h[3][kMaxBlurTaps]; float v[3][kMaxBlurTaps]; };

// MODE 0: one separable filter, one output plane per input plane.
// MODE 1: derivative bank, four output planes [w, mu1, mu2, sigma] per input plane,
//         laid out [plane][4][H][W] (the reference's [N*S, K, H, W], base_dau_conv_layer.cu:210-221).
// (The paired / strip-major layouts of the TMA kernels are written by dau_blur_pair.cu.)
template <int MODE>
__global__ void __launch_bounds__(256)
blur_kernel(const float* __restrict__ in, float* __restrict__ out, int H, int W, int kb,
            const __grid_constant__ BlurArgs1 a1, const __grid_constant__ BlurArgsBank ab,
            int plane0, int N, int Cn, int tiles_x, int tiles_y) {
    constexpr int NH = MODE == 1 ? 3 : 1;
    __shared__ float s_in[BIN][BIN + 1];
    __shared__ float s_h[NH][BIN][BT + 1];
    const int c = kb >> 1;
    const int plane = plane0 + blockIdx.z;
    const int x0 = blockIdx.x * BT, y0 = blockIdx.y * BT;
    const float* src = in + (size_t)plane * H * W;
    const int tid = threadIdx.x;
    const int span = BT + 2 * c;             // rows / cols actually needed
    // stage input tile (zero outside the image)
    for (int i = tid; i < span * span; i += 256) {
        const int r = i / span, q = i - r * span;
        const int gy = y0 + r - c, gx = x0 + q - c;
        float v = 0.f;
        if (gy >= 0 && gy < H && gx >= 0 && gx < W) v = __ldg(src + (size_t)gy * W + gx);
        s_in[r][q] = v;
    }
    __syncthreads();
    // horizontal pass: rows [0, span), output columns [0, BT)
    for (int i = tid; i < span * BT; i += 256) {
        const int r = i / BT, q = i - r * BT;
        float acc[NH];
#pragma unroll
        for (int j = 0; j < NH; ++j) acc[j] = 0.f;
        for (int u = 0; u < kb; ++u) {
            const float v = s_in[r][q + u];
            if (NH == 1) acc[0] = fmaf(v, a1.fx[u], acc[0]);
            else {
#pragma unroll
                for (int j = 0; j < NH; ++j) acc[j] = fmaf(v, ab.h[j][u], acc[j]);
            }
        }
#pragma unroll
        for (int j = 0; j < NH; ++j) s_h[j][r][q] = acc[j];
    }
    __syncthreads();
    // vertical pass
    for (int i = tid; i < BT * BT; i += 256) {
        const int r = i / BT, q = i - r * BT;
        const int gy = y0 + r, gx = x0 + q;
        if (gy >= H || gx >= W) continue;
        if (MODE == 0) {
            float acc = 0.f;
            for (int v = 0; v < kb; ++v) acc = fmaf(s_h[0][r + v][q], a1.fy[v], acc);
            out[(size_t)plane * H * W + (size_t)gy * W + gx] = acc;
        } else {
            float aw = 0.f, am1 = 0.f, am2 = 0.f, asg = 0.f;
            for (int v = 0; v < kb; ++v) {
                const float t0 = s_h[0][r + v][q], t1 = s_h[1][r + v][q], t2 = s_h[2][r + v][q];
                aw = fmaf(t0, ab.v[0][v], aw);
                am1 = fmaf(t1, ab.v[0][v], am1);
                am2 = fmaf(t0, ab.v[1][v], am2);
                asg = fmaf(t2, ab.v[0][v], asg);
                asg = fmaf(t0, ab.v[2][v], asg);
            }
            float* o = out + (size_t)plane * 4 * H * W + (size_t)gy * W + gx;
            const size_t ps = (size_t)H * W;
            o[0] = aw; o[ps] = am1; o[2 * ps] = am2; o[3 * ps] = asg;
        }
    }
}
}  // namespace

static dim3 blur_grid(int planes, int H, int W) {
    return dim3((W + BT - 1) / BT, (H + BT - 1) / BT, planes);
}

void launch_blur_single(const float* in, float* out, int planes, int H, int W, int kb,
                        const float* fx, const float* fy, cudaStream_t stream) {
    BlurArgs1 a1; BlurArgsBank ab;
    memset(&a1, 0, sizeof(a1)); memset(&ab, 0, sizeof(ab));
    for (int i = 0; i < kb; ++i) { a1.fx[i] = fx[i]; a1.fy[i] = fy[i]; }
    // gridDim.z limit is 65535: split the plane range if needed
    for (int p0 = 0; p0 < planes; p0 += 65535) {
        const int np = planes - p0 < 65535 ? planes - p0 : 65535;
        blur_kernel<0><<<blur_grid(np, H, W), 256, 0, stream>>>(
            in + (size_t)p0 * H * W, out + (size_t)p0 * H * W, H, W, kb, a1, ab, 0, 0, 1, 0, 0);
    }
    DAU_CUDA_CHECK(cudaGetLastError());
}

void launch_blur_bank(const float* in, float* out4, int planes, int H, int W,
                      const BlurFactors& f, cudaStream_t stream) {
    BlurArgs1 a1; BlurArgsBank ab;
    memset(&a1, 0, sizeof(a1)); memset(&ab, 0, sizeof(ab));
    for (int i = 0; i < f.kb; ++i) {
        ab.h[0][i] = f.h0[i]; ab.h[1][i] = f.h1[i]; ab.h[2][i] = f.h2[i];
        ab.v[0][i] = f.v0[i]; ab.v[1][i] = f.v1[i]; ab.v[2][i] = f.v2[i];
    }
    for (int p0 = 0; p0 < planes; p0 += 65535) {
        const int np = planes - p0 < 65535 ? planes - p0 : 65535;
        blur_kernel<1><<<blur_grid(np, H, W), 256, 0, stream>>>(
            in + (size_t)p0 * H * W, out4 + (size_t)p0 * 4 * H * W, H, W, f.kb, a1, ab, 0, 0, 1, 0, 0);
    }
    DAU_CUDA_CHECK(cudaGetLastError());
}

namespace {
// top error NCHW -> SGP[n/2][f][y][x%2][x/2][n%2] of the MOSAIC image (row of x/2 padded to an even
// count; gaps and cells beyond the batch are zeros), with the ignore_edge_gradients mask
// (dau_conv_backward_core.hpp:2218-2222) applied on the way.  One thread moves the column pair
// (2j, 2j+1) of both mosaic images of the pair; the float2 stores are coalesced.
__global__ void __launch_bounds__(128) error_relayout_kernel(const float* __restrict__ err, float2* __restrict__ out,
                                                             const __grid_constant__ Mosaic m, int F, int qp,
                                                             int mask_row, int mask_col) {
    const int pf = blockIdx.x;                            // (pair, feature)
    const int np = pf / F, f = pf - np * F;
    const int vy = blockIdx.y;
    float2* dst = out + ((size_t)pf * m.Hv + vy) * 2 * qp;
    const size_t rplane = (size_t)m.H * m.W;
    for (int j = threadIdx.x; j < qp; j += 128) {
        float v[2][2] = {{0.f, 0.f}, {0.f, 0.f}};         // [column 2j / 2j+1][image of the pair]
#pragma unroll
        for (int c = 0; c < 2; ++c)
#pragma unroll
            for (int img = 0; img < 2; ++img) {
                const int vx = 2 * j + c, nv = 2 * np + img;
                int rn, ry, rx;
                if (vx < m.Wv && nv < m.Nv && mosaic_to_real(m, nv, vy, vx, &rn, &ry, &rx)) {
                    const bool dead = (mask_row && ry == m.H - 1) || (mask_col && rx == m.W - 1);
                    if (!dead) v[c][img] = __ldg(err + ((size_t)rn * F + f) * rplane + (size_t)ry * m.W + rx);
                }
            }
        dst[j] = make_float2(v[0][0], v[0][1]);
        dst[qp + j] = make_float2(v[1][0], v[1][1]);
    }
}
}  // namespace

void launch_error_relayout(const float* err, float* out_sg, const Mosaic& m, int F, int mask_row, int mask_col,
                           cudaStream_t stream) {
    const int qv = (m.Wv + 1) / 2, qp = (qv + 1) / 2 * 2;
    const long long total = (long long)((m.Nv + 1) / 2) * F;
    DAU_CHECK(total < (1LL << 31) && m.Hv <= 65535, 4, "shape too large for the error re-layout grid");
    error_relayout_kernel<<<dim3((unsigned)total, m.Hv), 128, 0, stream>>>(err, reinterpret_cast<float2*>(out_sg), m, F, qp,
                                                                           mask_row, mask_col);
    DAU_CUDA_CHECK(cudaGetLastError());
}

}  // namespace dau
