// Per-unit parameter gradients of the DAU layer for sm_100a.
//
//   R[k,s,g,f] = sum_{n,y,x} E[n,f,y,x] * sum_{dy,dx} b[dy][dx] * Xk[n,s, y+oy+dy, x+ox+dx]
//   k in {w, mu1, mu2, sigma}: Xk = X blurred with the k-th derivative filter (zero outside).
//
// Replaces DAUConv_bwd_multi_pipeline_kernel + interleave_input_data_kernel +
// interleave_error_data_kernel of the reference
// (include/dau_conv/dau_conv_impl/dau_conv_backward_core.hpp:1017-1820, :2248-2380, :2118-2246)
// and the K-grouping dispatch (src/dau_conv/dau_conv_impl/dau_conv_backward.cpp:173-320).
//
// Like the reference (:846-894, comment :2055-2057) the bilinear weights are moved onto the
// error so the four k share one interpolation, but the roles of the two operands are swapped
// with respect to it: substituting p' = p + o + tap,
//   R[k,s,g,f] = sum_{n,p'} Xk[n,s,p'] * Ebar_{sgf}[n,p'],
//   Ebar[p']   = sum_{dy,dx} b[dy][dx] * E0[n,f, p' - o - (dy,dx)]       (E0 zero outside)
// so the operand that is read ONCE per (n,s,pixel) -- the four Xk values -- stays in
// registers for all G*FC units of the pass, and the operand that is gathered at a
// data-dependent displacement -- E -- comes from a haloed shared-memory tile where each
// word feeds up to four FMAs (same window trick as dau_gather.cu).  Per unit and lane:
// (R+1)(C+1)=25 LDS for 8*R*C=128 FMA, against 4 LDS.32 per 8 FMA in the reference scheme.
//
// Work decomposition
//   CTA  = NW warps = NW input channels s (one per warp) x FC features (unit records of the
//          CTA are fixed for its lifetime and live in shared memory).
//   item = (image group, spatial tile); a CTA walks its split of the items, double-buffered:
//          cp.async of item i+1 (error tiles of the FC features + each warp's 4 Xk strips)
//          overlaps the math of item i.
//   lane = R x C strip of p'; accumulators acc[unit][k] live in registers for the whole walk
//          and are reduced across the warp once at the end -> one plain store per output into
//          partial[split]; no atomics, deterministic.
#include "dau_geom.cuh"

namespace dau {

namespace {

using namespace geom;

constexpr int UPW = 16;     // units per warp and pass (accumulators: 4*UPW registers)

struct GradParams {
    const float* xk; const float* err; const UnitRec* table; float* partial;
    int N, S, F, G, H, W;
    int FC, GC, n_fchunks, n_gchunks;
    int tiles_x, tiles, items, items_per_split;
    int feat_stride;         // words between the error tiles of two features in a stage
    int stage_words;         // FC*feat_stride + NW*XK_WORDS
    int mask_last_row, mask_last_col;
};

template <class GEO>
__global__ void __launch_bounds__(GEO::NW * 32, 1) dau_grad_kernel(const GradParams p) {
    constexpr int R = GEO::R, C = GEO::C, LX = GEO::LX, LY = GEO::LY, LN = GEO::LN, NW = GEO::NW;
    constexpr int TW = GEO::TW, TH = GEO::TH, HX = GEO::HX, HY = GEO::HY;
    constexpr int TWH = GEO::TWH, THH = GEO::THH, QS = GEO::QS, RS = GEO::RS, IS = GEO::IS;
    constexpr int LOGC = ilog2(C);
    constexpr int XK_WORDS = 4 * R * 32 * C;          // per warp: [k][r][lane][c]
    static_assert(C == 4, "Xk staging uses 16-byte lanes");
    extern __shared__ __align__(16) float smem[];
    __shared__ UnitRec s_units[NW][UPW];

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int lx = lane % LX, ly = (lane / LX) % LY, ln = lane / (LX * LY);
    const int s = blockIdx.x * NW + warp;
    const bool s_valid = s < p.S;
    const int fchunk = blockIdx.y / p.n_gchunks, gchunk = blockIdx.y - fchunk * p.n_gchunks;
    const int f0 = fchunk * p.FC, g0 = gchunk * p.GC;
    const int H = p.H, W = p.W, G = p.G, F = p.F;
    const int nfeat = min(p.FC, F - f0);
    const int ngc = min(p.GC, G - g0);
    const int nunits = nfeat * ngc;                   // <= UPW, warp-uniform

    // unit records of this warp: u = fi*ngc + gi ; pad0 <- word offset of feature fi's tile
    if (s_valid && lane < nunits) {
        const int fi = lane / ngc, gi = lane - fi * ngc;
        UnitRec r = p.table[((size_t)s * F + f0 + fi) * G + g0 + gi];
        r.pad0 = fi * p.feat_stride;
        s_units[warp][lane] = r;
    }

    const bool vec_ok = (W & 3) == 0;
    auto issue_item = [&](int item, float* buf) {
        const int ng = item / p.tiles, t = item - ng * p.tiles;
        const int tyi = t / p.tiles_x, txi = t - tyi * p.tiles_x;
        const int x0 = txi * TW, y0 = tyi * TH, n0 = ng * LN;
        // error tiles: FC features x LN images, haloed, zero outside, edge mask applied
        for (int fi = 0; fi < nfeat; ++fi) {
#pragma unroll
            for (int img = 0; img < LN; ++img) {
                const int n = n0 + img;
                const float* src = p.err + ((size_t)min(n, p.N - 1) * F + f0 + fi) * (size_t)H * W;
                float* dimg = buf + fi * p.feat_stride + img * IS;
                for (int y = warp; y < THH; y += NW) {
                    const int gy = y0 - HY + y;
                    const bool rowok = n < p.N && gy >= 0 && gy < H && !(p.mask_last_row && gy == H - 1);
                    float* drow = dimg + y * RS;
                    const float* srow = src + (size_t)(rowok ? gy : 0) * W;
#pragma unroll
                    for (int xb = 0; xb < TWH; xb += 32) {
                        const int x = xb + lane;
                        if (x < TWH) {
                            const int gx = x0 - HX + x;
                            const bool ok = rowok && gx >= 0 && gx < W && !(p.mask_last_col && gx == W - 1);
                            cp_async4_zfill(drow + (x & (C - 1)) * QS + (x >> LOGC), srow + (ok ? gx : 0), ok);
                        }
                    }
                }
            }
        }
        // this warp's Xk strips: [k][r][lane][c]
        if (s_valid) {
            float* xdst = buf + p.FC * p.feat_stride + warp * XK_WORDS + lane * C;
            const int n = n0 + ln;
            const int xs = x0 + lx * C, ys = y0 + ly * R;
            const float* src = p.xk + ((size_t)min(n, p.N - 1) * p.S + s) * 4 * (size_t)H * W;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    const int y = ys + r;
                    const bool rowok = n < p.N && y < H;
                    const float* srow = src + ((size_t)k * H + (rowok ? y : 0)) * W;
                    float* d = xdst + (k * R + r) * 32 * C;
                    if (vec_ok) {
                        const bool ok = rowok && xs < W;
                        const unsigned da = (unsigned)__cvta_generic_to_shared(d);
                        const int sz = ok ? 16 : 0;
                        asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(da),
                                     "l"(srow + (ok ? xs : 0)), "r"(sz) : "memory");
                    } else {
#pragma unroll
                        for (int c = 0; c < C; ++c) {
                            const bool ok = rowok && xs + c < W;
                            cp_async4_zfill(d + c, srow + (ok ? xs + c : 0), ok);
                        }
                    }
                }
            }
        }
    };

    float acc[UPW][4];
#pragma unroll
    for (int u = 0; u < UPW; ++u)
#pragma unroll
        for (int k = 0; k < 4; ++k) acc[u][k] = 0.f;

    const int item_begin = blockIdx.z * p.items_per_split;
    const int item_end = min(p.items, item_begin + p.items_per_split);
    if (item_begin < item_end) {
        issue_item(item_begin, smem);
        cp_async_commit();
    }
    cp_async_wait_all();
    __syncthreads();

    const int lane_base = ln * IS + ly * R * RS + lx;
    for (int item = item_begin; item < item_end; ++item) {
        float* buf = smem + ((item - item_begin) & 1) * p.stage_words;
        if (item + 1 < item_end) {
            issue_item(item + 1, smem + ((item + 1 - item_begin) & 1) * p.stage_words);
            cp_async_commit();
        }
        if (s_valid) {
            float xk[4][R][C];
            const float* xsrc = buf + p.FC * p.feat_stride + warp * XK_WORDS + lane * C;
#pragma unroll
            for (int k = 0; k < 4; ++k)
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    const float4 v = *reinterpret_cast<const float4*>(xsrc + (k * R + r) * 32 * C);
                    xk[k][r][0] = v.x; xk[k][r][1] = v.y; xk[k][r][2] = v.z; xk[k][r][3] = v.w;
                }
            const float* tb = buf + lane_base;
#pragma unroll
            for (int u = 0; u < UPW; ++u) {
                if (u < nunits) {
                    const UnitRec* rec = &s_units[warp][u];
                    const float4 wv = *reinterpret_cast<const float4*>(rec);
                    const int4 off = *reinterpret_cast<const int4*>(&rec->offx_h);   // offx, offy, feat offset
                    const int b = off.x & (C - 1);
                    const float* base = tb + off.z + off.y * RS + (off.x >> LOGC);
                    const float* colp[C + 1];
#pragma unroll
                    for (int c = 0; c <= C; ++c) {
                        const int m = b + c;
                        colp[c] = base + (m & (C - 1)) * QS + (m >> LOGC);
                    }
                    float prev[C + 1];
#pragma unroll
                    for (int c = 0; c <= C; ++c) prev[c] = colp[c][0];
#pragma unroll
                    for (int r = 1; r <= R; ++r) {
                        float cur[C + 1];
#pragma unroll
                        for (int c = 0; c <= C; ++c) cur[c] = colp[c][r * RS];
#pragma unroll
                        for (int c = 0; c < C; ++c) {
                            float e = wv.x * prev[c];
                            e = fmaf(wv.y, prev[c + 1], e);
                            e = fmaf(wv.z, cur[c], e);
                            e = fmaf(wv.w, cur[c + 1], e);
#pragma unroll
                            for (int k = 0; k < 4; ++k) acc[u][k] = fmaf(e, xk[k][r - 1][c], acc[u][k]);
                        }
#pragma unroll
                        for (int c = 0; c <= C; ++c) prev[c] = cur[c];
                    }
                }
            }
        }
        cp_async_wait_all();
        __syncthreads();
    }

    // one warp reduction per accumulator, one store per output element of this split
    if (s_valid) {
        const size_t sgf = (size_t)p.S * G * F;
        float* pbase = p.partial + (size_t)blockIdx.z * 4 * sgf;
#pragma unroll
        for (int u = 0; u < UPW; ++u) {
            if (u < nunits) {
                const int fi = u / ngc, gi = u - fi * ngc;
                const size_t idx = ((size_t)s * G + g0 + gi) * F + f0 + fi;
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    float v = acc[u][k];
#pragma unroll
                    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
                    if (lane == 0) pbase[(size_t)k * sgf + idx] = v;
                }
            }
        }
    }
}

template <int LX, int LY, int LN, int HALO> using GGeo = GatherGeom<4, 4, LX, LY, LN, 8, HALO>;
constexpr int kNumShapes = 4;
constexpr int kHalos[4] = {4, 8, 16, 32};

#define DAU_GFOR_SHAPE(HALO, shape, STMT)                                        \
    switch (shape) {                                                             \
        case 0: { using GEO = GGeo<16, 2, 1, HALO>; STMT; } break;               \
        case 1: { using GEO = GGeo<8, 4, 1, HALO>; STMT; } break;                \
        case 2: { using GEO = GGeo<4, 4, 2, HALO>; STMT; } break;                \
        default: { using GEO = GGeo<2, 2, 8, HALO>; STMT; } break;               \
    }
#define DAU_GFOR_VARIANT(variant, STMT)                                          \
    switch ((variant) / kNumShapes) {                                            \
        case 0: DAU_GFOR_SHAPE(4, (variant) % kNumShapes, STMT) break;           \
        case 1: DAU_GFOR_SHAPE(8, (variant) % kNumShapes, STMT) break;           \
        case 2: DAU_GFOR_SHAPE(16, (variant) % kNumShapes, STMT) break;          \
        default: DAU_GFOR_SHAPE(32, (variant) % kNumShapes, STMT) break;         \
    }

template <class GEO>
void fill_plan(GradPlan& pl) {
    pl.tile_w = GEO::TW; pl.tile_h = GEO::TH; pl.imgs = GEO::LN;
    pl.halo_x = GEO::HX; pl.halo_y = GEO::HY;
    pl.row_stride = GEO::RS; pl.img_stride = GEO::IS;
    pl.feat_stride = GEO::TILE_WORDS;
    pl.smem_bytes = (size_t)GEO::NW * 4 * GEO::R * 32 * GEO::C;   // Xk words per stage (temporarily)
}

template <class GEO>
void launch_one(const GradPlan& plan, const GradParams& p, cudaStream_t stream) {
    static bool attr_set = false;
    if (!attr_set) {
        DAU_CUDA_CHECK(cudaFuncSetAttribute(dau_grad_kernel<GEO>,
                                            cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
        attr_set = true;
    }
    dim3 grid(plan.grid_x, plan.grid_y, plan.grid_z);
    dau_grad_kernel<GEO><<<grid, GEO::NW * 32, plan.smem_bytes, stream>>>(p);
}

}  // namespace

GradPlan plan_grad(int N, int S, int F, int G, int H, int W, int max_off, int num_sms) {
    DAU_CHECK(max_off >= 1 && max_off <= 32, 4, "kernel_size/2 must be in [1,32]");
    int hclass = 0;
    while (kHalos[hclass] < max_off) ++hclass;
    const size_t budget = 200 * 1024;
    GradPlan best; memset(&best, 0, sizeof(best));
    double best_cost = 1e300;
    for (int shape = 0; shape < kNumShapes; ++shape) {
        GradPlan pl; memset(&pl, 0, sizeof(pl));
        pl.kind = 0;
        pl.variant = hclass * kNumShapes + shape;
        DAU_GFOR_VARIANT(pl.variant, fill_plan<GEO>(pl));
        const size_t xk_words = pl.smem_bytes;
        pl.g_chunk = G < UPW ? G : UPW;
        int fc = UPW / pl.g_chunk;
        if (fc > F) fc = F;
        while (fc > 1 && 2 * ((size_t)fc * pl.feat_stride + xk_words) * 4 > budget) --fc;
        if (2 * ((size_t)fc * pl.feat_stride + xk_words) * 4 > budget) continue;
        pl.feats_per_cta = fc;
        pl.units_per_warp = fc * pl.g_chunk;
        pl.smem_bytes = 2 * ((size_t)fc * pl.feat_stride + xk_words) * 4;
        const long tx = (W + pl.tile_w - 1) / pl.tile_w, ty = (H + pl.tile_h - 1) / pl.tile_h;
        const long ng = (N + pl.imgs - 1) / pl.imgs;
        const double cost = (double)tx * ty * ng * pl.tile_w * pl.tile_h * pl.imgs * (1.0 + 1e-3 * shape) *
                            (16.0 / pl.units_per_warp);
        if (cost < best_cost) { best_cost = cost; best = pl; }
    }
    DAU_CHECK(best_cost < 1e300, 4, "no gradient tile shape fits shared memory");
    GradPlan& pl = best;
    const long tx = (W + pl.tile_w - 1) / pl.tile_w, ty = (H + pl.tile_h - 1) / pl.tile_h;
    const long items = tx * ty * ((N + pl.imgs - 1) / pl.imgs);
    const int n_fchunks = (F + pl.feats_per_cta - 1) / pl.feats_per_cta;
    const int n_gchunks = (G + pl.g_chunk - 1) / pl.g_chunk;
    pl.grid_x = (S + 7) / 8;
    pl.grid_y = n_fchunks * n_gchunks;
    // split the (image,tile) walk so that the grid fills whole waves of the SMs (1 CTA / SM)
    const long base = (long)pl.grid_x * pl.grid_y;
    int best_split = 1; double best_eff = -1.0;
    const int max_split = (int)(items < 64 ? items : 64);
    for (int sp = 1; sp <= max_split; ++sp) {
        const long per = (items + sp - 1) / sp;
        const long ctas = base * sp;
        const long waves = (ctas + num_sms - 1) / num_sms;
        // efficiency = useful item-slots / scheduled item-slots, small penalty per extra split
        const double eff = (double)(base * items) / ((double)waves * num_sms * per) - 0.002 * sp;
        if (eff > best_eff + 1e-9) { best_eff = eff; best_split = sp; }
    }
    pl.n_splits = best_split;
    pl.grid_z = best_split;
    pl.partial_floats = (size_t)best_split * 4 * S * G * F;
    return pl;
}

void launch_grad(const GradPlan& plan, const float* xk, const float* err, const UnitRec* table,
                 float* partial, int N, int S, int F, int G, int H, int W,
                 int mask_last_row, int mask_last_col, cudaStream_t stream) {
    GradParams p;
    p.xk = xk; p.err = err; p.table = table; p.partial = partial;
    p.N = N; p.S = S; p.F = F; p.G = G; p.H = H; p.W = W;
    p.FC = plan.feats_per_cta; p.GC = plan.g_chunk;
    p.n_fchunks = (F + p.FC - 1) / p.FC; p.n_gchunks = (G + p.GC - 1) / p.GC;
    p.tiles_x = (W + plan.tile_w - 1) / plan.tile_w;
    p.tiles = p.tiles_x * ((H + plan.tile_h - 1) / plan.tile_h);
    p.items = p.tiles * ((N + plan.imgs - 1) / plan.imgs);
    p.items_per_split = (p.items + plan.n_splits - 1) / plan.n_splits;
    p.feat_stride = plan.feat_stride;
    p.stage_words = (int)(plan.smem_bytes / 8);
    p.mask_last_row = mask_last_row; p.mask_last_col = mask_last_col;
    DAU_CHECK(plan.grid_y <= 65535 && plan.grid_z <= 65535, 4, "too many feature chunks for one launch");
    DAU_GFOR_VARIANT(plan.variant, launch_one<GEO>(plan, p, stream));
    DAU_CUDA_CHECK(cudaGetLastError());
}

}  // namespace dau
