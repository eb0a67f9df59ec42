// Displaced bilinear gather-accumulate for sm_100a: the DAU forward pass and, with the
// transposed / negated unit table, the error back-propagation dL/dx.
//
//   out[n,co,y,x] = sum_{ci,g} sum_{dy,dx in {0,1}} wt[ci,co,g][dy][dx] *
//                   in0[n,ci, y + offy + dy, x + offx + dx]        (in0 = in, zero outside)
//
// Replaces DAUConv_forward_pipeline_kernel + interleave_input_data_kernel of the reference
// (include/dau_conv/dau_conv_impl/dau_conv_forward_core.hpp:804-1605, :1607-1732) and their
// dispatch ladder (src/dau_conv/dau_conv_impl/dau_conv_forward.cpp:129-264).  Differences by
// design: no patch copy of the input in HBM (the haloed tile is staged straight from NCHW
// with zero-filling cp.async), no atomics and no pre-zeroed output (one CTA owns an output
// tile for ALL (ci,g) and stores it once), shapes are not restricted to F%16==0 / even G.
//
// Work decomposition
//   CTA  = NW warps; tile = IMGS images x TH x TW output pixels x NW output channels.
//   warp = one output channel co of the chunk; its 32 lanes are LX x LY x LN over
//          (x, y, image); each lane owns an R x C strip (accumulators in registers).
//   loop = input channels ci, double-buffered: cp.async of stage ci+1 (haloed tile of the
//          LN images + the NW*G unit records of (ci, chunk)) overlaps the math of stage ci.
//   unit = warp-uniform record {4 tap weights, integer offset}.  A lane reads the
//          (R+1) x (C+1) window once and uses every word for up to 4 FMAs (both dx taps
//          of two neighbouring columns, both dy taps of two neighbouring rows).
//
// Shared-memory layout of a tile row: columns are de-interleaved modulo C,
//   word(y, x) = y*RS + (x % C)*QS + x / C,
// so the C+1 window columns of the 32 lanes, which sit C apart in x, land on consecutive
// words whatever the (data-dependent) offset is: every LDS is one conflict-free wavefront.
// RS and the image stride carry a few padding words chosen at compile time so that the
// LY / LN lane groups fall on disjoint banks.
//
// Roofline: FP32 FMA pipe (4 FMA per unit-pixel), co-limited by shared-memory wavefronts:
// (R+1)(C+1) LDS per 4RC FMA per lane = 45 : 128 for R=8, C=4.
#include "dau_geom.cuh"

namespace dau {

namespace {

using namespace geom;

struct GatherParams {
    const float* in; const UnitRec* table; const float* bias; float* out;
    int N, Cin, Cout, G, H, W;
    int tiles_x;
    int stage_words;   // TILE_WORDS + NW*G*8
};

template <class GEO>
__global__ void __launch_bounds__(GEO::NW * 32, 2) dau_gather_kernel(const GatherParams p) {
    constexpr int R = GEO::R, C = GEO::C, LX = GEO::LX, LY = GEO::LY, LN = GEO::LN, NW = GEO::NW;
    constexpr int TW = GEO::TW, TH = GEO::TH, HX = GEO::HX, HY = GEO::HY;
    constexpr int TWH = GEO::TWH, THH = GEO::THH, QS = GEO::QS, RS = GEO::RS, IS = GEO::IS;
    constexpr int LOGC = ilog2(C);
    extern __shared__ __align__(16) float smem[];

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int lx = lane % LX, ly = (lane / LX) % LY, ln = lane / (LX * LY);
    const int tyi = blockIdx.x / p.tiles_x, txi = blockIdx.x - tyi * p.tiles_x;
    const int x0 = txi * TW, y0 = tyi * TH;
    const int n0 = blockIdx.z * LN;
    const int co0 = blockIdx.y * NW;
    const int co = co0 + warp;
    const bool co_valid = co < p.Cout;
    const int G = p.G, H = p.H, W = p.W;
    const int nrec_chunks = min(NW, p.Cout - co0) * G * 2;     // 16-byte chunks of unit records

    auto issue_stage = [&](int ci, float* buf) {
#pragma unroll
        for (int img = 0; img < LN; ++img) {
            const int n = n0 + img;
            const float* src = p.in + ((size_t)min(n, p.N - 1) * p.Cin + ci) * (size_t)H * W;
            for (int y = warp; y < THH; y += NW) {
                const int gy = y0 - HY + y;
                const bool rowok = n < p.N && gy >= 0 && gy < H;
                float* drow = buf + img * IS + y * RS;
                const float* srow = src + (size_t)(rowok ? gy : 0) * W;
#pragma unroll
                for (int xb = 0; xb < TWH; xb += 32) {
                    const int x = xb + lane;
                    if (x < TWH) {
                        const int gx = x0 - HX + x;
                        const bool ok = rowok && gx >= 0 && gx < W;
                        cp_async4_zfill(drow + (x & (C - 1)) * QS + (x >> LOGC), srow + (ok ? gx : 0), ok);
                    }
                }
            }
        }
        const float* rsrc = reinterpret_cast<const float*>(p.table + ((size_t)ci * p.Cout + co0) * G);
        float* rdst = buf + GEO::TILE_WORDS;
        for (int i = tid; i < nrec_chunks; i += NW * 32) cp_async16(rdst + i * 4, rsrc + i * 4);
    };

    float acc[R][C];
#pragma unroll
    for (int r = 0; r < R; ++r)
#pragma unroll
        for (int c = 0; c < C; ++c) acc[r][c] = 0.f;

    issue_stage(0, smem);
    cp_async_commit();
    cp_async_wait_all();
    __syncthreads();

    const int lane_base = ln * IS + ly * R * RS + lx;
    for (int ci = 0; ci < p.Cin; ++ci) {
        float* buf = smem + (ci & 1) * p.stage_words;
        if (ci + 1 < p.Cin) {
            issue_stage(ci + 1, smem + ((ci + 1) & 1) * p.stage_words);
            cp_async_commit();
        }
        if (co_valid) {
            const UnitRec* recs = reinterpret_cast<const UnitRec*>(buf + GEO::TILE_WORDS) + warp * G;
            const float* tb = buf + lane_base;
            for (int g = 0; g < G; ++g) {
                const float4 wv = *reinterpret_cast<const float4*>(&recs[g]);
                const int2 off = *reinterpret_cast<const int2*>(&recs[g].offx_h);
                const int b = off.x & (C - 1);
                const float* base = tb + off.y * RS + (off.x >> LOGC);
                const float* colp[C + 1];
#pragma unroll
                for (int c = 0; c <= C; ++c) {
                    const int m = b + c;
                    colp[c] = base + (m & (C - 1)) * QS + (m >> LOGC);
                }
#pragma unroll
                for (int r = 0; r <= R; ++r) {
                    float xv[C + 1];
#pragma unroll
                    for (int c = 0; c <= C; ++c) xv[c] = colp[c][r * RS];
                    if (r < R) {
#pragma unroll
                        for (int c = 0; c < C; ++c) {
                            acc[r][c] = fmaf(wv.x, xv[c], acc[r][c]);
                            acc[r][c] = fmaf(wv.y, xv[c + 1], acc[r][c]);
                        }
                    }
                    if (r > 0) {
#pragma unroll
                        for (int c = 0; c < C; ++c) {
                            acc[r - 1][c] = fmaf(wv.z, xv[c], acc[r - 1][c]);
                            acc[r - 1][c] = fmaf(wv.w, xv[c + 1], acc[r - 1][c]);
                        }
                    }
                }
            }
        }
        cp_async_wait_all();
        __syncthreads();
    }

    // epilogue: one store per output element, optional bias (forward_gpu_bias of the reference)
    const int n = n0 + ln;
    if (co_valid && n < p.N) {
        const float bv = p.bias ? __ldg(p.bias + co) : 0.f;
        float* obase = p.out + ((size_t)n * p.Cout + co) * (size_t)H * W;
        const int xs = x0 + lx * C;
        const bool vec = (C == 4) && ((W & 3) == 0) && (xs + C <= W);
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const int y = y0 + ly * R + r;
            if (y >= H) continue;
            float* orow = obase + (size_t)y * W;
            if (vec) {
                *reinterpret_cast<float4*>(orow + xs) =
                    make_float4(acc[r][0] + bv, acc[r][1 % C] + bv, acc[r][2 % C] + bv, acc[r][3 % C] + bv);
            } else {
#pragma unroll
                for (int c = 0; c < C; ++c)
                    if (xs + c < W) orow[xs + c] = acc[r][c] + bv;
            }
        }
    }
}

// ---- instantiated shapes -------------------------------------------------------------------
// variant = tile shape (x halo class).  R=8, C=4 everywhere; lanes re-split for small images.
template <int LX, int LY, int LN, int HALO> using Geo = GatherGeom<8, 4, LX, LY, LN, 8, HALO>;

struct VariantInfo { int tile_w, tile_h, imgs, halo; };
constexpr int kNumShapes = 4;
constexpr int kHalos[4] = {4, 8, 16, 32};

template <class GEO>
void launch_one(const GatherPlan& plan, const GatherParams& p, cudaStream_t stream) {
    static bool attr_set = false;
    if (!attr_set) {
        DAU_CUDA_CHECK(cudaFuncSetAttribute(dau_gather_kernel<GEO>,
                                            cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        attr_set = true;
    }
    dim3 grid(plan.grid_x, plan.grid_y, plan.grid_z);
    dau_gather_kernel<GEO><<<grid, GEO::NW * 32, plan.smem_bytes, stream>>>(p);
}

template <class GEO>
void fill_plan(GatherPlan& pl) {
    pl.tile_w = GEO::TW; pl.tile_h = GEO::TH; pl.imgs = GEO::LN;
    pl.halo_x = GEO::HX; pl.halo_y = GEO::HY;
    pl.row_stride = GEO::RS; pl.img_stride = GEO::IS;
    pl.stage_words = GEO::TILE_WORDS;     // unit records added by the caller
    pl.cout_per_cta = GEO::NW;
}

#define DAU_FOR_SHAPE(HALO, shape, STMT)                                         \
    switch (shape) {                                                             \
        case 0: { using GEO = Geo<16, 2, 1, HALO>; STMT; } break;                \
        case 1: { using GEO = Geo<8, 4, 1, HALO>; STMT; } break;                 \
        case 2: { using GEO = Geo<4, 2, 4, HALO>; STMT; } break;                 \
        default: { using GEO = Geo<2, 1, 16, HALO>; STMT; } break;               \
    }
#define DAU_FOR_VARIANT(variant, STMT)                                           \
    switch ((variant) / kNumShapes) {                                            \
        case 0: DAU_FOR_SHAPE(4, (variant) % kNumShapes, STMT) break;            \
        case 1: DAU_FOR_SHAPE(8, (variant) % kNumShapes, STMT) break;            \
        case 2: DAU_FOR_SHAPE(16, (variant) % kNumShapes, STMT) break;           \
        default: DAU_FOR_SHAPE(32, (variant) % kNumShapes, STMT) break;          \
    }

}  // namespace

GatherPlan plan_gather(int N, int Cin, int Cout, int G, int H, int W, int max_off) {
    DAU_CHECK(max_off >= 1 && max_off <= 32, 4, "kernel_size/2 must be in [1,32]");
    int hclass = 0;
    while (kHalos[hclass] < max_off) ++hclass;
    GatherPlan best; memset(&best, 0, sizeof(best));
    double best_cost = 1e300;
    for (int shape = 0; shape < kNumShapes; ++shape) {
        GatherPlan pl; memset(&pl, 0, sizeof(pl));
        pl.kind = 0; pl.nstage = 2;
        pl.variant = hclass * kNumShapes + shape;
        DAU_FOR_VARIANT(pl.variant, fill_plan<GEO>(pl));
        const long tx = (W + pl.tile_w - 1) / pl.tile_w, ty = (H + pl.tile_h - 1) / pl.tile_h;
        const long ng = (N + pl.imgs - 1) / pl.imgs;
        // padded pixels actually computed; ties go to the wider tile (fewer, larger CTAs)
        const double cost = (double)tx * ty * ng * pl.tile_w * pl.tile_h * pl.imgs * (1.0 + 1e-3 * shape);
        pl.stage_words += pl.cout_per_cta * G * 8;
        pl.smem_bytes = (size_t)2 * pl.stage_words * sizeof(float);
        if (pl.smem_bytes > 200 * 1024) continue;
        pl.grid_x = (int)(tx * ty); pl.grid_y = (Cout + pl.cout_per_cta - 1) / pl.cout_per_cta;
        pl.grid_z = (int)ng;
        if (cost < best_cost) { best_cost = cost; best = pl; }
    }
    DAU_CHECK(best_cost < 1e300, 4, "no gather tile shape fits shared memory (G too large)");
    (void)Cin;
    return best;
}

void launch_gather(const GatherPlan& plan, const float* in, const UnitRec* table, const float* bias,
                   float* out, int N, int Cin, int Cout, int G, int H, int W, cudaStream_t stream) {
    GatherParams p;
    p.in = in; p.table = table; p.bias = bias; p.out = out;
    p.N = N; p.Cin = Cin; p.Cout = Cout; p.G = G; p.H = H; p.W = W;
    p.tiles_x = (W + plan.tile_w - 1) / plan.tile_w;
    p.stage_words = plan.stage_words;
    DAU_CHECK(plan.grid_z <= 65535 && plan.grid_y <= 65535, 4, "batch / channel count too large for one launch");
    DAU_FOR_VARIANT(plan.variant, launch_one<GEO>(plan, p, stream));
    DAU_CUDA_CHECK(cudaGetLastError());
}

}  // namespace dau
