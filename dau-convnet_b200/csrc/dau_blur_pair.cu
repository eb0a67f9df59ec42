// Separable Gaussian / Gaussian-derivative blurs of NCHW fp32 planes for sm_100a.
//
// Replaces caffe_gpu_convolve2 -> convolve2<...> of the reference (dense Kb x Kb "same"
// correlation, filter in __constant__ memory, src/dau_conv/util/convolve.cu:48-131, index
// algebra :107-124) in its three uses: forward blur (base_dau_conv_layer.cu:81-90), the 4
// derivative blurs of the backward pass (:210-221) and the blur of the top error with the
// rotated filter (:278-288).  out[y,x] = sum_{v,u} in[y+v-c, x+u-c] * fy[v]*fx[u], zero outside
// the image.  Cost per pixel: (NH + NV)*Kb FMAs instead of NOUT*Kb^2.  Built for HBM speed:
//
//  * one CTA = one 64 x 32 output tile of one channel of an image PAIR: both images are filtered
//    one after the other through the same shared-memory buffers and their results leave the SM
//    together, because both consumer layouts interleave the two images of a pair
//        PG[n/2][c][y][x%4][x/4][n%2]                    (forward / dL/dx gather, dau_gather_tma.cu)
//        XS[item][s][k][r<4][lane<32][c<2][img<2]        (parameter gradients, dau_grad_tma.cu)
//    so every store is a full float2 / float4 and a warp's stores are contiguous;
//  * a thread filters 4 consecutive pixels: the row pass reads 2*CQ+1 float4 from shared memory
//    for 4 outputs, the column pass one float4 per tap (0.4 LDS wavefronts per pixel instead
//    of 14 scalar loads), with the row filters pre-shifted on the host so that all register
//    indices are static.
#include "dau_kernels.cuh"

namespace dau {

namespace {
__device__ __forceinline__ void cp_async4_zfill(float* dst, const float* src, bool ok) {
    const unsigned d = (unsigned)__cvta_generic_to_shared(dst);
    const int sz = ok ? 4 : 0;
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;\n" ::"r"(d), "l"(src), "r"(sz) : "memory");
}
__device__ __forceinline__ void cp_async16_zfill(float* dst, const float* src, bool ok) {
    const unsigned d = (unsigned)__cvta_generic_to_shared(dst);
    const int sz = ok ? 16 : 0;
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(d), "l"(src), "r"(sz) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N) : "memory"); }

constexpr int PT_W = 64, PT_H = 32, PT_THREADS = 256;
constexpr int PT_MAXTAPS = 17;

struct PairBlurArgs {
    float fx[3][PT_MAXTAPS];     // row filters, shifted: tap u' applies to tile column 4g + u' - 4*CQ
    float fy[3][PT_MAXTAPS];     // column filters, tap v applies to row y + v - c
    int kb, H, W, N, Cn, z0;     // H, W, N: mosaic extents
    Mosaic m;                    // mosaic <-> real image mapping (identity for wide images)
    int tiles_x, tiles_y;        // XS: 64-column / 4-row item grid
    int qp, qv;                  // PG: padded / valid float2-columns per plane
};

// MODE 2: one filter -> PG.   MODE 3: derivative bank (h0,h1,h2 / v0,v1,v2) -> XS.
template <int MODE, int CQ, bool IDENT>
__global__ void __launch_bounds__(PT_THREADS)
blur_pair_kernel(const float* __restrict__ in, float* __restrict__ out, const __grid_constant__ PairBlurArgs a) {
    constexpr int NH = MODE == 3 ? 3 : 1, NOUT = MODE == 3 ? 4 : 1;
    constexpr int PADX = 4 * CQ, NT = 8 * CQ + 1, SW = PT_W + 2 * PADX, NV = 2 * CQ + 1;
    extern __shared__ float4 smem4[];
    const int c = a.kb >> 1, rows = PT_H + 2 * c;
    float* s_in0 = reinterpret_cast<float*>(smem4);         // [2 images][rows][SW]
    float* s_h = s_in0 + 2 * rows * SW;                      // [NH][rows][PT_W]
    const int tid = threadIdx.x;
    const int x0 = blockIdx.x * PT_W, y0 = blockIdx.y * PT_H;
    const int pc = a.z0 + blockIdx.z;                        // (pair, channel)
    const int np = pc / a.Cn, cn = pc - np * a.Cn;
    const int H = a.H, W = a.W;
    constexpr bool ident = IDENT;
    const size_t rplane = (size_t)a.m.H * a.m.W;

    // ---- stage the haloed tiles of BOTH images with cp.async (zero fill outside the image): the
    //      second image's HBM latency hides behind the first image's filtering --------------------
    const bool vec16 = ident && (W & 3) == 0 && (reinterpret_cast<size_t>(in) & 15) == 0;
#pragma unroll
    for (int img = 0; img < 2; ++img) {
        const int n = 2 * np + img;                          // mosaic image
        const bool live = n < a.N;
        float* dst = s_in0 + img * rows * SW;
        if (vec16) {
            const float* src = in + ((size_t)n * a.Cn + cn) * rplane;
            for (int i = tid; i < rows * (SW / 4); i += PT_THREADS) {
                const int r = i / (SW / 4), q = (i - r * (SW / 4)) * 4;
                const int gy = y0 + r - c, gx = x0 + q - PADX;       // gx % 4 == 0: all in or all out
                const bool ok = live && gy >= 0 && gy < H && gx >= 0 && gx < W;
                cp_async16_zfill(dst + r * SW + q, ok ? src + (size_t)gy * W + gx : in, ok);
            }
        } else {
            for (int i = tid; i < rows * SW; i += PT_THREADS) {
                const int r = i / SW, q = i - r * SW;
                const int gy = y0 + r - c, gx = x0 + q - PADX;
                const float* src = in;
                bool ok = live && gy >= 0 && gy < H && gx >= 0 && gx < W;
                if (ok) {
                    if (ident) {
                        src = in + ((size_t)n * a.Cn + cn) * rplane + (size_t)gy * W + gx;
                    } else {
                        int rn, ry, rx;
                        ok = mosaic_to_real(a.m, n, gy, gx, &rn, &ry, &rx);
                        if (ok) src = in + ((size_t)rn * a.Cn + cn) * rplane + (size_t)ry * a.m.W + rx;
                    }
                }
                cp_async4_zfill(dst + i, src, ok);
            }
        }
        cp_async_commit();
    }

    float4 res[2][2][NOUT];
#pragma unroll
    for (int img = 0; img < 2; ++img) {
        const int n = 2 * np + img;
        const bool live = n < a.N;
        const float* s_in = s_in0 + img * rows * SW;
        if (img == 0) cp_async_wait<1>(); else cp_async_wait<0>();
        __syncthreads();
        // ---- row pass: 4 outputs per task --------------------------------------------------------
        for (int t = tid; t < rows * (PT_W / 4); t += PT_THREADS) {
            const int r = t >> 4, g = t & 15;
            float xin[4 * NV];
            const float4* sp = reinterpret_cast<const float4*>(s_in + r * SW + 4 * g);
#pragma unroll
            for (int j = 0; j < NV; ++j) {
                const float4 v = sp[j];
                xin[4 * j] = v.x; xin[4 * j + 1] = v.y; xin[4 * j + 2] = v.z; xin[4 * j + 3] = v.w;
            }
#pragma unroll
            for (int j = 0; j < NH; ++j) {
                float4 h = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
                for (int u = 0; u < NT; ++u) {
                    const float f = a.fx[j][u];
                    h.x = fmaf(xin[u], f, h.x); h.y = fmaf(xin[u + 1], f, h.y);
                    h.z = fmaf(xin[u + 2], f, h.z); h.w = fmaf(xin[u + 3], f, h.w);
                }
                *reinterpret_cast<float4*>(s_h + (j * rows + r) * PT_W + 4 * g) = h;
            }
        }
        __syncthreads();
        // ---- column pass: 2 tasks of 4 outputs per thread ---------------------------------------
#pragma unroll
        for (int tt = 0; tt < 2; ++tt) {
            const int t = tid + tt * PT_THREADS;
            const int y = t >> 4, g = t & 15;
            const float4* hp = reinterpret_cast<const float4*>(s_h + y * PT_W + 4 * g);
            const int hs = rows * PT_W / 4;                  // float4 stride between the NH planes
            float4 o[NOUT];
#pragma unroll
            for (int k = 0; k < NOUT; ++k) o[k] = make_float4(0.f, 0.f, 0.f, 0.f);
            for (int v = 0; v < a.kb; ++v) {
                const float4 t0 = hp[v * (PT_W / 4)];
                if (MODE == 3) {
                    const float4 t1 = hp[hs + v * (PT_W / 4)], t2 = hp[2 * hs + v * (PT_W / 4)];
                    const float v0 = a.fy[0][v], v1 = a.fy[1][v], v2 = a.fy[2][v];
                    // X_w = v0[h0]   X_mu1 = v0[h1]   X_mu2 = v1[h0]   X_sigma = v0[h2] + v2[h0]
                    o[0].x = fmaf(t0.x, v0, o[0].x); o[0].y = fmaf(t0.y, v0, o[0].y);
                    o[0].z = fmaf(t0.z, v0, o[0].z); o[0].w = fmaf(t0.w, v0, o[0].w);
                    o[1 % NOUT].x = fmaf(t1.x, v0, o[1 % NOUT].x); o[1 % NOUT].y = fmaf(t1.y, v0, o[1 % NOUT].y);
                    o[1 % NOUT].z = fmaf(t1.z, v0, o[1 % NOUT].z); o[1 % NOUT].w = fmaf(t1.w, v0, o[1 % NOUT].w);
                    o[2 % NOUT].x = fmaf(t0.x, v1, o[2 % NOUT].x); o[2 % NOUT].y = fmaf(t0.y, v1, o[2 % NOUT].y);
                    o[2 % NOUT].z = fmaf(t0.z, v1, o[2 % NOUT].z); o[2 % NOUT].w = fmaf(t0.w, v1, o[2 % NOUT].w);
                    o[3 % NOUT].x = fmaf(t2.x, v0, o[3 % NOUT].x); o[3 % NOUT].y = fmaf(t2.y, v0, o[3 % NOUT].y);
                    o[3 % NOUT].z = fmaf(t2.z, v0, o[3 % NOUT].z); o[3 % NOUT].w = fmaf(t2.w, v0, o[3 % NOUT].w);
                    o[3 % NOUT].x = fmaf(t0.x, v2, o[3 % NOUT].x); o[3 % NOUT].y = fmaf(t0.y, v2, o[3 % NOUT].y);
                    o[3 % NOUT].z = fmaf(t0.z, v2, o[3 % NOUT].z); o[3 % NOUT].w = fmaf(t0.w, v2, o[3 % NOUT].w);
                } else {
                    const float v0 = a.fy[0][v];
                    o[0].x = fmaf(t0.x, v0, o[0].x); o[0].y = fmaf(t0.y, v0, o[0].y);
                    o[0].z = fmaf(t0.z, v0, o[0].z); o[0].w = fmaf(t0.w, v0, o[0].w);
                }
            }
            // zero what lies outside the image (pad columns / rows of the consumer layouts)
            const int gy = y0 + y, gx = x0 + 4 * g;
            const bool rowok = live && gy < H;
            bool ok[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                ok[i] = rowok && gx + i < W;
                if (ok[i] && !ident) {                       // blur tails inside the gaps are not image
                    int rn, ry, rx;
                    ok[i] = mosaic_to_real(a.m, n, gy, gx + i, &rn, &ry, &rx);
                }
            }
#pragma unroll
            for (int k = 0; k < NOUT; ++k) {
                if (!ok[0]) o[k].x = 0.f;
                if (!ok[1]) o[k].y = 0.f;
                if (!ok[2]) o[k].z = 0.f;
                if (!ok[3]) o[k].w = 0.f;
                res[img][tt][k] = o[k];
            }
        }
        __syncthreads();                                     // the next image re-uses both buffers
    }

    // ---- both images leave together --------------------------------------------------------------
#pragma unroll
    for (int tt = 0; tt < 2; ++tt) {
        const int t = tid + tt * PT_THREADS;
        const int y = t >> 4, g = t & 15;
        const int gy = y0 + y, gx = x0 + 4 * g;
        if (MODE == 2) {
            const int q = gx >> 2;
            if (gy >= H || q >= a.qv) continue;
            float2* o = reinterpret_cast<float2*>(out) + (((size_t)pc * H + gy) * 4) * a.qp + q;
            const float4 r0 = res[0][tt][0], r1 = res[1][tt][0];
            o[0] = make_float2(r0.x, r1.x);
            o[a.qp] = make_float2(r0.y, r1.y);
            o[2 * a.qp] = make_float2(r0.z, r1.z);
            o[3 * a.qp] = make_float2(r0.w, r1.w);
        } else {
            if (gy >= a.tiles_y * 4 || gx >= a.tiles_x * 64) continue;
            const size_t item = ((size_t)np * a.tiles_y + (gy >> 2)) * a.tiles_x + (gx >> 6);
            float4* o = reinterpret_cast<float4*>(out + ((item * a.Cn + cn) * 16 + (gy & 3)) * 128 + (gx & 63) * 2);
#pragma unroll
            for (int k = 0; k < NOUT; ++k) {
                const float4 r0 = res[0][tt][k], r1 = res[1][tt][k];
                o[k * 128] = make_float4(r0.x, r1.x, r0.y, r1.y);          // k stride: 4 rows * 128 floats
                o[k * 128 + 1] = make_float4(r0.z, r1.z, r0.w, r1.w);
            }
        }
    }
}

int cq_of(int kb) { const int cq = ((kb >> 1) + 3) / 4; return cq < 1 ? 1 : cq; }

void fill_row_filter(float* dst, const float* f, int kb, int cq) {
    // tap u' <-> tile column offset u' - 4*cq; the filter's tap u sits at offset u - c
    const int c = kb >> 1, shift = 4 * cq - c;
    for (int i = 0; i < PT_MAXTAPS; ++i) dst[i] = 0.f;
    for (int u = 0; u < kb; ++u) dst[u + shift] = f[u];
}

template <int MODE>
void launch_pair(const float* in, float* out, PairBlurArgs& a, int pairs_times_c, cudaStream_t stream) {
    const int c = a.kb >> 1, cq = cq_of(a.kb);
    const int rows = PT_H + 2 * c, sw = PT_W + 8 * cq, nh = MODE == 3 ? 3 : 1;
    const size_t smem = (size_t)(2 * rows * sw + nh * rows * PT_W) * sizeof(float);
    static bool attr_set = false;
    if (!attr_set) {
        DAU_CUDA_CHECK(cudaFuncSetAttribute(blur_pair_kernel<MODE, 1, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
        DAU_CUDA_CHECK(cudaFuncSetAttribute(blur_pair_kernel<MODE, 2, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
        DAU_CUDA_CHECK(cudaFuncSetAttribute(blur_pair_kernel<MODE, 1, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
        DAU_CUDA_CHECK(cudaFuncSetAttribute(blur_pair_kernel<MODE, 2, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
        attr_set = true;
    }
    const int gx = MODE == 3 ? a.tiles_x : (a.W + PT_W - 1) / PT_W;
    const int gy = MODE == 3 ? (a.tiles_y * 4 + PT_H - 1) / PT_H : (a.H + PT_H - 1) / PT_H;
    for (int z0 = 0; z0 < pairs_times_c; z0 += 65535) {
        const int nz = pairs_times_c - z0 < 65535 ? pairs_times_c - z0 : 65535;
        a.z0 = z0;
        const bool ident = a.m.kx == 1 && a.m.ky == 1;
        const dim3 grid(gx, gy, nz);
        if (cq == 1 && ident) blur_pair_kernel<MODE, 1, true><<<grid, PT_THREADS, smem, stream>>>(in, out, a);
        else if (cq == 1) blur_pair_kernel<MODE, 1, false><<<grid, PT_THREADS, smem, stream>>>(in, out, a);
        else if (ident) blur_pair_kernel<MODE, 2, true><<<grid, PT_THREADS, smem, stream>>>(in, out, a);
        else blur_pair_kernel<MODE, 2, false><<<grid, PT_THREADS, smem, stream>>>(in, out, a);
    }
    DAU_CUDA_CHECK(cudaGetLastError());
}
}  // namespace

void launch_blur_single_paired(const float* in, float* out_paired, const Mosaic& m, int Cn, int kb,
                               const float* fx, const float* fy, cudaStream_t stream) {
    const int N = m.Nv, H = m.Hv, W = m.Wv;
    DAU_CHECK(m.gap >= (kb >> 1) || (m.kx == 1 && m.ky == 1), 4,
              "blur wider than the mosaic gap: call dau_set_sigma_hint before dau_reshape");
    DAU_CHECK(kb >= 1 && kb <= PT_MAXTAPS && (kb & 1), 1, "blur kernel size must be odd and <= 17");
    PairBlurArgs a; memset(&a, 0, sizeof(a));
    fill_row_filter(a.fx[0], fx, kb, cq_of(kb));
    for (int i = 0; i < kb; ++i) a.fy[0][i] = fy[i];
    a.kb = kb; a.H = H; a.W = W; a.N = N; a.Cn = Cn; a.m = m;
    a.qv = (W + 3) / 4; a.qp = (a.qv + 1) / 2 * 2;
    launch_pair<2>(in, out_paired, a, ((N + 1) / 2) * Cn, stream);
}

void launch_blur_bank_strips(const float* in, float* out_strips, const Mosaic& m, int S,
                             const BlurFactors& f, cudaStream_t stream) {
    const int N = m.Nv, H = m.Hv, W = m.Wv;
    DAU_CHECK(m.gap >= (f.kb >> 1) || (m.kx == 1 && m.ky == 1), 4,
              "blur wider than the mosaic gap: call dau_set_sigma_hint before dau_reshape");
    DAU_CHECK(f.kb >= 1 && f.kb <= PT_MAXTAPS && (f.kb & 1), 1, "blur kernel size must be odd and <= 17");
    PairBlurArgs a; memset(&a, 0, sizeof(a));
    const int cq = cq_of(f.kb);
    fill_row_filter(a.fx[0], f.h0, f.kb, cq);
    fill_row_filter(a.fx[1], f.h1, f.kb, cq);
    fill_row_filter(a.fx[2], f.h2, f.kb, cq);
    for (int i = 0; i < f.kb; ++i) { a.fy[0][i] = f.v0[i]; a.fy[1][i] = f.v1[i]; a.fy[2][i] = f.v2[i]; }
    a.kb = f.kb; a.H = H; a.W = W; a.N = N; a.Cn = S; a.m = m;
    a.tiles_x = (W + 63) / 64; a.tiles_y = (H + 3) / 4;
    launch_pair<3>(in, out_strips, a, ((N + 1) / 2) * S, stream);
}

}  // namespace dau
