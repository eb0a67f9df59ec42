// Separable Gaussian / Gaussian-derivative blurs of NCHW fp32 planes for sm_100a.
//
// Replaces caffe_gpu_convolve2 -> convolve2<...> of the reference (dense Kb x Kb "same"
// correlation, filter in __constant__ memory, src/dau_conv/util/convolve.cu:48-131, index
// algebra :107-124) in its three uses: forward blur (base_dau_conv_layer.cu:81-90), the 4
// derivative blurs of the backward pass (:210-221) and the blur of the top error with the
// rotated filter (:278-288).  out[y,x] = sum_{v,u} in[y+v-c, x+u-c] * fy[v]*fx[u], zero outside
// the image.  Cost per pixel: (NH + NV)*Kb FMAs instead of NOUT*Kb^2.  Built for HBM speed:
//
//  * one CTA = one 64 x TH output tile of one channel of an image PAIR.  Both consumer layouts
//    interleave the two images of a pair
//        PG[n/2][c][y][x%4][x/4][n%2]                    (forward / dL/dx gather, dau_gather_tma.cu)
//        XS[item][s][k][r<4][lane<32][c<2][img<2]        (parameter gradients, dau_grad_tma.cu)
//    so the two images travel through the kernel as the two halves of fma.rn.f32x2 operands:
//    the row pass turns the two haloed input tiles into (img0, img1) float2 planes in shared
//    memory, the column pass reads them with LDS.128 (2 pixels x 2 images), and every result
//    leaves as a full float2 / float4 with a warp's stores contiguous;
//  * the haloed tiles arrive by TMA (one cp.async.bulk.tensor.3d per image, zero fill outside the
//    plane) when the planes are not a mosaic and rows are 16-byte multiples; mosaics and odd
//    widths are staged element-wise with cp.async through the mosaic -> real mapping;
//  * the filter half-width C is a template parameter: both passes are fully unrolled, taps come
//    from the kernel parameter bank as (f, f) pairs, no zero taps are multiplied;
//  * the column pass is register-blocked over RY output rows (2 for the bank, 4 for one
//    filter), so a shared-memory row is read once for RY results.
#include "dau_kernels.cuh"
#include "dau_tma.cuh"

namespace dau {

namespace {
using namespace tma;

__device__ __forceinline__ void cp_async4_zfill(float* dst, const float* src, bool ok) {
    const unsigned d = (unsigned)__cvta_generic_to_shared(dst);
    const int sz = ok ? 4 : 0;
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;\n" ::"r"(d), "l"(src), "r"(sz) : "memory");
}
__device__ __forceinline__ void cp_async_commit_wait_all() {
    asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;\n" ::: "memory");
}
// 3-D tiled tensor load global -> shared (dst 128-byte aligned); coordinates innermost first
__device__ __forceinline__ void tensor3_g2s(void* dst, const CUtensorMap* map, int c0, int c1, int c2, uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];\n"
        ::"r"(smem_u32(dst)), "l"(map), "r"(c0), "r"(c1), "r"(c2), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ float2 lo(const float4& v) { return make_float2(v.x, v.y); }
__device__ __forceinline__ float2 hi(const float4& v) { return make_float2(v.z, v.w); }

constexpr int PT_W = 64, PT_THREADS = 256;
constexpr int PT_MAXTAPS = 17;

struct PairBlurArgs {
    float2 fx[3][PT_MAXTAPS];    // row filters as (f, f): tap u applies to column x + u - c
    float2 fy[3][PT_MAXTAPS];    // column filters as (f, f): tap v applies to row y + v - c
    int H, W, N, Cn, z0;         // H, W, N: mosaic extents
    Mosaic m;                    // mosaic <-> real image mapping (identity for wide images)
    int tiles_x;                 // XS: 64-column tiles
    int sq, stail;               // XS: full 4-row strips per cell row, rows of the tail strip (StripGeom)
    long long tail_base;         // XS: float offset of the tail blocks
    int qp, qv;                  // PG: padded / valid float2-columns per plane
    // MODE 2 extras (the top error on its way into the backward pass):
    const float* mask;           // same shape as `in`: in[i] is taken as 0 where mask[i] <= 0 (ReLU backward)
    float2* sgp;                 // also emit the (masked) input as SGP[n/2][c][y][x%2][x/2][n%2] for the gradient kernel
    int sgp_qp;                  // padded float2-columns per SGP plane
    int edge_row, edge_col;      // ignore_edge_gradients: SGP gets zeros in the last real row / column
    int crop;                    // `in` / `mask` are [.., m.H-2crop, m.W-2crop] planes centred in the full grid
};

template <int MODE, int C>
struct TileGeom {
    static constexpr int TH = MODE == 3 ? 16 : 32;           // output rows of a CTA
    static constexpr int RY = MODE == 3 ? 2 : 4;             // output rows of a column-pass task
    static constexpr int NH = MODE == 3 ? 3 : 1, NOUT = MODE == 3 ? 4 : 1;
    static constexpr int CQ = (C + 3) / 4, PADX = 4 * CQ, SW = PT_W + 2 * PADX, NV = 2 * CQ + 1;
    static constexpr int ROWS = TH + 2 * C, KB = 2 * C + 1;
    static constexpr int IMG_FLOATS = (ROWS * SW + 31) / 32 * 32;     // a 128-byte multiple
    static constexpr int PLANE4 = ROWS * 32;                 // float4 per (img0, img1) plane
    static constexpr size_t SMEM = 128 + (size_t)2 * IMG_FLOATS * 4 + (size_t)NH * PLANE4 * 16 + 128;
    static constexpr size_t SMEM_MASK = (size_t)2 * IMG_FLOATS * 4;   // extra bytes when a mask is staged
};

// MODE 2: one filter -> PG.   MODE 3: derivative bank (h0,h1,h2 / v0,v1,v2) -> XS.
template <int MODE, int C, bool TMA>
__global__ void __launch_bounds__(PT_THREADS, 4)
blur_pair_kernel(const __grid_constant__ CUtensorMap in_map, const __grid_constant__ CUtensorMap mask_map,
                 const float* __restrict__ in, float* __restrict__ out, const __grid_constant__ PairBlurArgs a) {
    using G = TileGeom<MODE, C>;
    constexpr int TH = G::TH, RY = G::RY, NH = G::NH, NOUT = G::NOUT, PADX = G::PADX, SW = G::SW, NV = G::NV;
    constexpr int ROWS = G::ROWS, KB = G::KB, IMG_FLOATS = G::IMG_FLOATS, PLANE4 = G::PLANE4;
    extern __shared__ unsigned char smem_dyn[];
    unsigned char* smem_raw = smem_dyn + ((128u - (smem_u32(smem_dyn) & 127u)) & 127u);
    float* s_in = reinterpret_cast<float*>(smem_raw);                 // [2 images][ROWS][SW]
    float4* s_h = reinterpret_cast<float4*>(s_in + 2 * IMG_FLOATS);   // [NH][ROWS][2 halves][16] (img0,img1) pairs
    uint64_t* bar = reinterpret_cast<uint64_t*>(s_h + NH * PLANE4);
    float* s_mask = reinterpret_cast<float*>(bar + 16);              // [2 images][ROWS][SW], only with a.mask
    const bool masked = MODE == 2 && a.mask != nullptr;
    const int tid = threadIdx.x;
    const int x0 = blockIdx.x * PT_W, y0 = blockIdx.y * TH;
    const int pc = a.z0 + blockIdx.z;                        // (pair, channel)
    const int np = pc / a.Cn, cn = pc - np * a.Cn;
    const int H = a.H, W = a.W;
    const bool ident = a.m.kx == 1 && a.m.ky == 1;
    const bool live1 = 2 * np + 1 < a.N;

    // ---- mosaic maps of the tile's rows and columns: (cell index << 16 | pixel inside the cell), or
    //      -1 for gap pixels and for what lies outside the mosaic.  Integer divisions happen here,
    //      ROWS + SW times per CTA, and not once per staged element ----------------------------------
    __shared__ int s_rowmap[ROWS], s_colmap[SW];
    for (int i = tid; i < ROWS + SW; i += PT_THREADS) {
        const bool row = i < ROWS;
        const int v = row ? y0 + i - C : x0 + (i - ROWS) - PADX;
        const int ext = row ? H : W, pitch = row ? a.m.py : a.m.px, size = row ? a.m.H : a.m.W, cells = row ? a.m.ky : a.m.kx;
        int code = -1;
        if (v >= 0 && v < ext) {
            const int j = v / pitch, l = v - j * pitch;
            if (l < size && j < cells) code = (j << 16) | l;
        }
        if (row) s_rowmap[i] = code; else s_colmap[i - ROWS] = code;
    }
    if (!TMA) __syncthreads();                               // (the TMA branch synchronises below)

    // ---- stage the haloed tiles of both images -----------------------------------------------------
    if (TMA) {
        if (tid == 0) {
            mbar_init(bar, 1);
            fence_barrier_init();
            mbar_arrive_expect_tx(bar, (live1 ? 2u : 1u) * (masked ? 2u : 1u) * ROWS * SW * 4u);
            tensor3_g2s(s_in, &in_map, x0 - PADX, y0 - C, (2 * np) * a.Cn + cn, bar);
            if (live1) tensor3_g2s(s_in + IMG_FLOATS, &in_map, x0 - PADX, y0 - C, (2 * np + 1) * a.Cn + cn, bar);
            if (masked) {
                tensor3_g2s(s_mask, &mask_map, x0 - PADX, y0 - C, (2 * np) * a.Cn + cn, bar);
                if (live1) tensor3_g2s(s_mask + IMG_FLOATS, &mask_map, x0 - PADX, y0 - C, (2 * np + 1) * a.Cn + cn, bar);
            }
        }
        if (!live1)
            for (int i = tid; i < ROWS * SW; i += PT_THREADS) {
                s_in[IMG_FLOATS + i] = 0.f;
                if (masked) s_mask[IMG_FLOATS + i] = 0.f;
            }
        __syncthreads();                                     // barrier initialised (and zeros visible)
        mbar_wait(bar, 0);
    } else {
        const int crop = a.crop, RH = a.m.H - 2 * crop, RW = a.m.W - 2 * crop;
        const size_t rplane = (size_t)RH * RW;
#pragma unroll
        for (int img = 0; img < 2; ++img) {
            const int n = 2 * np + img;                      // mosaic image
            float* dst = s_in + img * IMG_FLOATS;
            for (int i = tid; i < ROWS * SW; i += PT_THREADS) {
                const int r = i / SW, q = i - r * SW;
                const int rm = s_rowmap[r], cm = s_colmap[q];
                const int nr = (n * a.m.ky + (rm >> 16)) * a.m.kx + (cm >> 16);     // real image
                const int ry = (rm & 0xffff) - crop, rx = (cm & 0xffff) - crop;
                const bool ok = (rm | cm) >= 0 && nr < a.m.N && (unsigned)ry < (unsigned)RH && (unsigned)rx < (unsigned)RW;
                const size_t src = ok ? ((size_t)nr * a.Cn + cn) * rplane + (size_t)ry * RW + rx : 0;
                cp_async4_zfill(dst + i, in + src, ok);
                if (masked) cp_async4_zfill(s_mask + img * IMG_FLOATS + i, a.mask + src, ok);
            }
        }
        cp_async_commit_wait_all();
        __syncthreads();
    }

    if (MODE == 2) {
        // ---- ReLU backward fused into the load: the error is zero where the activation output was --
        if (masked) {
            float4* e4 = reinterpret_cast<float4*>(s_in);
            const float4* m4 = reinterpret_cast<const float4*>(s_mask);
            for (int i = tid; i < 2 * IMG_FLOATS / 4; i += PT_THREADS) {
                float4 e = e4[i];
                const float4 mk = m4[i];
                e.x = mk.x > 0.f ? e.x : 0.f; e.y = mk.y > 0.f ? e.y : 0.f;
                e.z = mk.z > 0.f ? e.z : 0.f; e.w = mk.w > 0.f ? e.w : 0.f;
                e4[i] = e;
            }
            __syncthreads();
        }
        // ---- the un-blurred tile in the gradient kernel's layout (what error_relayout_kernel writes) -
        if (a.sgp != nullptr) {
            const bool edges = a.edge_row || a.edge_col;
            for (int t = tid; t < TH * 16; t += PT_THREADS) {
                const int y = t >> 4, g = t & 15;
                const int gy = y0 + y, j = (x0 >> 1) + 2 * g;           // float2-column of pixel x0 + 4g
                if (gy >= H || j >= a.sgp_qp) continue;
                float4 v0 = *reinterpret_cast<const float4*>(s_in + (y + C) * SW + PADX + 4 * g);
                float4 v1 = *reinterpret_cast<const float4*>(s_in + IMG_FLOATS + (y + C) * SW + PADX + 4 * g);
                if (edges) {
                    float* p0 = &v0.x; float* p1 = &v1.x;
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        int rn, ry, rx;
                        mosaic_to_real(a.m, 2 * np, gy, x0 + 4 * g + i, &rn, &ry, &rx);   // (ry, rx) do not depend on the image
                        if ((a.edge_row && ry == a.m.H - 1 - a.crop) || (a.edge_col && rx == a.m.W - 1 - a.crop)) p0[i] = p1[i] = 0.f;
                    }
                }
                float4* dst = reinterpret_cast<float4*>(a.sgp + ((size_t)pc * H + gy) * 2 * a.sgp_qp + j);
                dst[0] = make_float4(v0.x, v1.x, v0.z, v1.z);                 // even columns
                dst[a.sgp_qp / 2] = make_float4(v0.y, v1.y, v0.w, v1.w);     // odd columns
            }
        }
    }

    // ---- row pass: 4 pixels of both images per task, results stored as (img0, img1) pairs.  Scalar
    //      FMAs: the two images arrive in separate registers, and pairing them for f32x2 costs more
    //      moves than it saves ------------------------------------------------------------------------
    for (int t = tid; t < ROWS * 16; t += PT_THREADS) {
        const int r = t >> 4, g = t & 15;
        const float4* sp0 = reinterpret_cast<const float4*>(s_in + r * SW + 4 * g);
        const float4* sp1 = reinterpret_cast<const float4*>(s_in + IMG_FLOATS + r * SW + 4 * g);
        float x0v[4 * NV], x1v[4 * NV];
#pragma unroll
        for (int j = 0; j < NV; ++j) {
            const float4 v0 = sp0[j], v1 = sp1[j];
            x0v[4 * j] = v0.x; x0v[4 * j + 1] = v0.y; x0v[4 * j + 2] = v0.z; x0v[4 * j + 3] = v0.w;
            x1v[4 * j] = v1.x; x1v[4 * j + 1] = v1.y; x1v[4 * j + 2] = v1.z; x1v[4 * j + 3] = v1.w;
        }
#pragma unroll
        for (int j = 0; j < NH; ++j) {
            float h0[4], h1[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                h0[i] = x0v[PADX - C + i] * a.fx[j][0].x;
                h1[i] = x1v[PADX - C + i] * a.fx[j][0].x;
            }
#pragma unroll
            for (int u = 1; u < KB; ++u) {
                const float f = a.fx[j][u].x;
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    h0[i] = fmaf(x0v[u + PADX - C + i], f, h0[i]);
                    h1[i] = fmaf(x1v[u + PADX - C + i], f, h1[i]);
                }
            }
            float4* hp = s_h + j * PLANE4 + r * 32 + g;
            hp[0] = make_float4(h0[0], h1[0], h0[1], h1[1]);
            hp[16] = make_float4(h0[2], h1[2], h0[3], h1[3]);
        }
    }
    __syncthreads();

    // ---- column pass: a warp = RY output rows x 64 columns x 2 images; lane = (half p, group g) ----
    const int warp = tid >> 5, lane = tid & 31, p = lane >> 4, g = lane & 15;
    const bool interior = ident && live1 && x0 + PT_W <= W && y0 + TH <= H;
    for (int yb = warp; yb < TH / RY; yb += PT_THREADS / 32) {
        float2 acc[RY][NOUT][2];
#pragma unroll
        for (int o = 0; o < RY; ++o)
#pragma unroll
            for (int k = 0; k < NOUT; ++k) acc[o][k][0] = acc[o][k][1] = make_float2(0.f, 0.f);
        const float4* hp = s_h + (yb * RY) * 32 + p * 16 + g;
#pragma unroll
        for (int i = 0; i < KB + RY - 1; ++i) {              // input row yb*RY + i of the haloed tile
            const float4 t0 = hp[i * 32];
            float4 t1 = t0, t2 = t0;
            if (MODE == 3) { t1 = hp[PLANE4 + i * 32]; t2 = hp[2 * PLANE4 + i * 32]; }
#pragma unroll
            for (int o = 0; o < RY; ++o) {
                const int v = i - o;                         // tap of output row o
                if (v < 0 || v >= KB) continue;
                const float2 v0 = a.fy[0][v];
                acc[o][0][0] = __ffma2_rn(lo(t0), v0, acc[o][0][0]);
                acc[o][0][1] = __ffma2_rn(hi(t0), v0, acc[o][0][1]);
                if (MODE == 3) {
                    // X_w = v0[h0]   X_mu1 = v0[h1]   X_mu2 = v1[h0]   X_sigma = v0[h2] + v2[h0]
                    const float2 v1 = a.fy[1][v], v2 = a.fy[2][v];
                    acc[o][1 % NOUT][0] = __ffma2_rn(lo(t1), v0, acc[o][1 % NOUT][0]);
                    acc[o][1 % NOUT][1] = __ffma2_rn(hi(t1), v0, acc[o][1 % NOUT][1]);
                    acc[o][2 % NOUT][0] = __ffma2_rn(lo(t0), v1, acc[o][2 % NOUT][0]);
                    acc[o][2 % NOUT][1] = __ffma2_rn(hi(t0), v1, acc[o][2 % NOUT][1]);
                    acc[o][3 % NOUT][0] = __ffma2_rn(lo(t2), v0, acc[o][3 % NOUT][0]);
                    acc[o][3 % NOUT][1] = __ffma2_rn(hi(t2), v0, acc[o][3 % NOUT][1]);
                    acc[o][3 % NOUT][0] = __ffma2_rn(lo(t0), v2, acc[o][3 % NOUT][0]);
                    acc[o][3 % NOUT][1] = __ffma2_rn(hi(t0), v2, acc[o][3 % NOUT][1]);
                }
            }
        }
#pragma unroll
        for (int o = 0; o < RY; ++o) {
            const int gy = y0 + yb * RY + o;
            if (!interior) {
                // zero what lies outside the image: pad columns / rows of the consumer layouts, the
                // second image of an odd batch, blur tails inside the mosaic gaps
                const int rm = s_rowmap[yb * RY + o + C];
#pragma unroll
                for (int i = 0; i < 2; ++i) {
                    const int cm = s_colmap[4 * g + 2 * p + i + PADX];
                    const int cell = (rm >> 16) * a.m.kx + (cm >> 16), per = a.m.kx * a.m.ky;
                    const bool px = (rm | cm) >= 0;
                    const bool ok0 = px && 2 * np * per + cell < a.m.N;
                    const bool ok1 = px && (2 * np + 1) * per + cell < a.m.N;
#pragma unroll
                    for (int k = 0; k < NOUT; ++k) {
                        if (!ok0) acc[o][k][i].x = 0.f;
                        if (!ok1) acc[o][k][i].y = 0.f;
                    }
                }
            }
            if (MODE == 2) {
                const int q = (x0 >> 2) + g;
                if (gy >= H || q >= a.qv) continue;
                float2* o2 = reinterpret_cast<float2*>(out) + (((size_t)pc * H + gy) * 4 + 2 * p) * a.qp + q;
                o2[0] = acc[o][0][0];
                o2[a.qp] = acc[o][0][1];
            } else {
                // strip-major Xk: the row belongs to a full 4-row strip or to the tail strip of its cell
                // row; gap rows (and rows beyond the mosaic) belong to no strip and are not written
                const int rm = s_rowmap[yb * RY + o + C];
                if (rm < 0) continue;
                const int jy = rm >> 16, ly = rm & 0xffff, t = ly >> 2, r = ly & 3;
                const size_t cellrow = (size_t)np * a.m.ky + jy;
                size_t base; int rows;
                if (t < a.sq) {
                    rows = 4;
                    base = (((cellrow * a.sq + t) * a.tiles_x + blockIdx.x) * a.Cn + cn) * (size_t)(16 * 128);
                } else {
                    rows = a.stail;
                    base = (size_t)a.tail_base + ((cellrow * a.tiles_x + blockIdx.x) * a.Cn + cn) * (size_t)(4 * rows * 128);
                }
                float4* o4 = reinterpret_cast<float4*>(out + base + r * 128) + 2 * g + p;
#pragma unroll
                for (int k = 0; k < NOUT; ++k)               // k stride: `rows` rows of 128 floats
                    o4[k * rows * 32] = make_float4(acc[o][k][0].x, acc[o][k][0].y, acc[o][k][1].x, acc[o][k][1].y);
            }
        }
    }
}

template <int MODE, int C>
void launch_c(const float* in, float* out, PairBlurArgs& a, int pairs_times_c, cudaStream_t stream) {
    using G = TileGeom<MODE, C>;
    const bool ident = a.m.kx == 1 && a.m.ky == 1;
    // TMA needs 16-byte row pitch and base; everything else goes through the element-wise path
    const bool use_tma = ident && a.crop == 0 && (a.W & 3) == 0 && (reinterpret_cast<size_t>(in) & 15) == 0 &&
                         (reinterpret_cast<size_t>(a.mask) & 15) == 0;
    const size_t smem = G::SMEM + (a.mask ? G::SMEM_MASK : 0);
    static PerDeviceOnce once;                 // the attribute is per (kernel, device)
    if (once.need()) {
        DAU_CUDA_CHECK(cudaFuncSetAttribute(blur_pair_kernel<MODE, C, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(G::SMEM + G::SMEM_MASK)));
        DAU_CUDA_CHECK(cudaFuncSetAttribute(blur_pair_kernel<MODE, C, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(G::SMEM + G::SMEM_MASK)));
        once.done();
    }
    CUtensorMap map, mmap;
    memset(&map, 0, sizeof(map));
    memset(&mmap, 0, sizeof(mmap));
    if (use_tma) {
        const uint64_t dims[3] = {(uint64_t)a.W, (uint64_t)a.H, (uint64_t)a.N * a.Cn};
        const uint64_t strides[2] = {(uint64_t)a.W * 4, (uint64_t)a.H * a.W * 4};
        const uint32_t box[3] = {(uint32_t)G::SW, (uint32_t)G::ROWS, 1};
        map = make_map_f32(in, 3, dims, strides, box);
        if (a.mask) mmap = make_map_f32(a.mask, 3, dims, strides, box);
    }
    const int gx = MODE == 3 ? a.tiles_x : (a.W + PT_W - 1) / PT_W;
    const int gy = (a.H + G::TH - 1) / G::TH;
    for (int z0 = 0; z0 < pairs_times_c; z0 += 65535) {
        const int nz = pairs_times_c - z0 < 65535 ? pairs_times_c - z0 : 65535;
        a.z0 = z0;
        const dim3 grid(gx, gy, nz);
        if (use_tma) blur_pair_kernel<MODE, C, true><<<grid, PT_THREADS, smem, stream>>>(map, mmap, in, out, a);
        else blur_pair_kernel<MODE, C, false><<<grid, PT_THREADS, smem, stream>>>(map, mmap, in, out, a);
    }
    DAU_CUDA_CHECK(cudaGetLastError());
}

template <int MODE>
void launch_pair(const float* in, float* out, PairBlurArgs& a, int kb, int pairs_times_c, cudaStream_t stream) {
    switch (kb >> 1) {
        case 1: launch_c<MODE, 1>(in, out, a, pairs_times_c, stream); break;
        case 2: launch_c<MODE, 2>(in, out, a, pairs_times_c, stream); break;
        case 3: launch_c<MODE, 3>(in, out, a, pairs_times_c, stream); break;
        case 4: launch_c<MODE, 4>(in, out, a, pairs_times_c, stream); break;
        case 5: launch_c<MODE, 5>(in, out, a, pairs_times_c, stream); break;
        case 6: launch_c<MODE, 6>(in, out, a, pairs_times_c, stream); break;
        case 7: launch_c<MODE, 7>(in, out, a, pairs_times_c, stream); break;
        case 8: launch_c<MODE, 8>(in, out, a, pairs_times_c, stream); break;
        default: DAU_CHECK(false, 1, "blur kernel size must be odd, >= 3 and <= 17");
    }
}

void fill_pairs(float2* dst, const float* f, int kb) {
    for (int i = 0; i < PT_MAXTAPS; ++i) dst[i] = make_float2(0.f, 0.f);
    for (int i = 0; i < kb; ++i) dst[i] = make_float2(f[i], f[i]);
}
}  // namespace

void launch_blur_single_paired(const float* in, float* out_paired, const Mosaic& m, int Cn, int kb,
                               const float* fx, const float* fy, cudaStream_t stream, const ErrorSideArgs* extra) {
    const int N = m.Nv, H = m.Hv, W = m.Wv;
    DAU_CHECK(m.gap >= (kb >> 1) || (m.kx == 1 && m.ky == 1), 4,
              "blur wider than the mosaic gap: call dau_set_sigma_hint before dau_reshape");
    DAU_CHECK(kb >= 3 && kb <= PT_MAXTAPS && (kb & 1), 1, "blur kernel size must be odd, >= 3 and <= 17");
    PairBlurArgs a; memset(&a, 0, sizeof(a));
    fill_pairs(a.fx[0], fx, kb);
    fill_pairs(a.fy[0], fy, kb);
    a.H = H; a.W = W; a.N = N; a.Cn = Cn; a.m = m;
    a.qv = (W + 3) / 4; a.qp = (a.qv + 1) / 2 * 2;
    if (extra) {
        a.mask = extra->mask; a.sgp = reinterpret_cast<float2*>(extra->sgp);
        a.sgp_qp = ((W + 1) / 2 + 1) / 2 * 2;
        a.edge_row = extra->edge_row; a.edge_col = extra->edge_col; a.crop = extra->crop;
    }
    launch_pair<2>(in, out_paired, a, kb, ((N + 1) / 2) * Cn, stream);
}

void launch_blur_bank_strips(const float* in, float* out_strips, const Mosaic& m, int S,
                             const BlurFactors& f, cudaStream_t stream) {
    const int N = m.Nv, H = m.Hv, W = m.Wv;
    DAU_CHECK(m.gap >= (f.kb >> 1) || (m.kx == 1 && m.ky == 1), 4,
              "blur wider than the mosaic gap: call dau_set_sigma_hint before dau_reshape");
    DAU_CHECK(f.kb >= 3 && f.kb <= PT_MAXTAPS && (f.kb & 1), 1, "blur kernel size must be odd, >= 3 and <= 17");
    PairBlurArgs a; memset(&a, 0, sizeof(a));
    fill_pairs(a.fx[0], f.h0, f.kb); fill_pairs(a.fx[1], f.h1, f.kb); fill_pairs(a.fx[2], f.h2, f.kb);
    fill_pairs(a.fy[0], f.v0, f.kb); fill_pairs(a.fy[1], f.v1, f.kb); fill_pairs(a.fy[2], f.v2, f.kb);
    a.H = H; a.W = W; a.N = N; a.Cn = S; a.m = m;
    const StripGeom sg = strip_geometry(m, S);
    a.tiles_x = sg.tiles_x; a.sq = sg.q; a.stail = sg.tail; a.tail_base = (long long)sg.tail_base_floats;
    launch_pair<3>(in, out_strips, a, f.kb, ((N + 1) / 2) * S, stream);
}

}  // namespace dau
