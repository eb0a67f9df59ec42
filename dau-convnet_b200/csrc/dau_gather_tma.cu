// Displaced bilinear gather-accumulate for sm_100a: the DAU forward pass and, with the
// transposed / negated unit table, the error back-propagation dL/dx.
//
//   out[n,co,y,x] = sum_{ci,g} sum_{dy,dx in {0,1}} wt[ci,co,g][dy][dx] *
//                   in0[n,ci, y + offy + dy, x + offx + dx]        (in0 = in, zero outside)
//
// Replaces DAUConv_forward_pipeline_kernel + interleave_input_data_kernel of the reference
// (include/dau_conv/dau_conv_impl/dau_conv_forward_core.hpp:804-1605, :1607-1732) and their
// dispatch ladder (src/dau_conv/dau_conv_impl/dau_conv_forward.cpp:129-264).  Differences by
// design: no patch copy of the input in HBM, no atomics and no pre-zeroed output (one CTA owns
// an output tile for ALL (ci,g) and stores it once), shapes are not restricted to F%16==0 /
// even G.
//
//  * the producing blur kernel writes its output in the "paired gather layout"
//        PG[n/2][c][y][x%4][x/4][n%2]           (n = mosaic image, see Mosaic in dau_kernels.cuh)
//    so that (a) a haloed tile of an image PAIR is one 4-D TMA box
//    (cp.async.bulk.tensor, zero fill outside the image = the reference's zero padding of the
//    blurred input, dau_conv_forward_core.hpp:1707-1710), landing in shared memory already
//    column-de-interleaved, word(y, x) = y*ROW + (x % 4)*QS + x / 4, so the 5 window columns of
//    16 lanes that sit 4 apart in x are consecutive float2 whatever the (data-dependent) offset
//    is: every LDS.64 is two conflict-free wavefronts; and (b) one LDS.64 fetches the same
//    pixel of two images;
//  * a dedicated producer warp streams input channels through an NSTAGE ring of
//    {tile, unit records} guarded by full/empty mbarriers; the 8 consumer warps never execute a
//    load instruction for global memory and never meet at a CTA barrier;
//  * unit = warp-uniform record {4 tap weights, integer offset}.  A lane reads the 9 x 5 window
//    once and uses every word for up to 4 FMAs (both dx taps of two neighbouring columns, both
//    dy taps of two neighbouring rows);
//  * the math is fma.rn.f32x2 (FFMA2): both images of the pair share the unit's weights, so
//    one instruction issues two FMAs -- the kernel is bound by shared-memory wavefronts
//    (85 LDS.64 per 256 FFMA2 per lane), not by instruction issue.
//
// CTA = 8 consumer warps (8 output channels) + a producer warpgroup; tile = 32 x 64 pixels x 2
// images; lane = (lx 0..15, ly 0..1) owning a 16 x 4 strip of both images (128 accumulators).
#include "dau_kernels.cuh"
#include "dau_tma.cuh"

#include <cstdlib>

namespace dau {

namespace {

using namespace tma;

// lane = (lx 0..15, ly 0..1) owns an R x 4 strip of both images.  Two ways to place the two strips of a tile:
//  * stacked (ly_stride == R): the tile is 2R consecutive mosaic rows; R (16, 15 or 14) is picked per shape so
//    that the 2R-row tiles waste the fewest rows (28-row maps: R = 14);
//  * cell rows (ly_stride == the mosaic's row pitch, R == the image height, 13..16): ly = 0 / 1 own the rows of
//    two consecutive CELL rows of the mosaic and nobody computes the gap rows between them (13 x 13 maps:
//    26 rows of work per tile instead of 30)
constexpr int C = 4, LX = 16, LY = 2, NW = 8;
// CTA = 3 warpgroups: warpgroup 0 holds the producer lane and gives its registers away
// (setmaxnreg.dec), warpgroups 1-2 are the 8 consumer warps and grow to CONS_REGS each.
constexpr int NTHREADS = (NW + 4) * 32, PROD_REGS = 40, CONS_REGS = 232;
constexpr int TW = LX * C;

template <int HALO, int R>
struct TGeom {
    static constexpr int TH = LY * R;                     // rows a tile computes (not its extent in cell-row mode)
    // the innermost TMA start coordinate must be 16-byte aligned: 2*(x0-HX)/4 floats -> HX % 8 == 0
    static constexpr int HX = (HALO + 7) / 8 * 8, HY = HALO;
    static constexpr int QS = (TW + 2 * HX) / 4;          // float2 per column plane
    static constexpr int ROW = 4 * QS;                    // float2 per tile row
};

struct TGatherParams {
    const UnitRec* table; const float* bias; float* out;
    int N, Cin, Cout, G, H, W;
    int tiles_x, nstage;
    int ly_stride;         // mosaic rows between the strips of ly = 0 and ly = 1: R (stacked) or the cell pitch
    int tile_pitch_y;      // mosaic rows between consecutive tiles: 2R (stacked) or two cell pitches
    int tile_bytes;        // (ly_stride + R + 2 halo) haloed rows of the input tile
    int stage_bytes;       // tile + records, multiple of 128
    int debug;             // development switches, only with -DDAU_DEV_SWITCHES (dau_common.hpp)
    int relu;              // epilogue: store max(0, y + bias)  
    int gx, gy;            // items are numbered (pair * gy + channel group) * gx + tile
    int full_items, split_k;   // items >= full_items are cut into split_k parts of the input-channel loop
    float2* part;          // partial strips of those parts: [slot][warp][r][c][lane]
    int crop;              // `out` is [.., H-2crop, W-2crop]: the centre of the same-size result (pad < (ks-1)/2)
    Mosaic m;              // N, H, W above are the mosaic extents; `out` is the real NCHW tensor
};

// Mosaic maps of a tile's rows and columns for the store: (cell << 16 | pixel inside the cell), -1
// for gap pixels -- the integer divisions happen once per CTA, not once per output.
template <int TH>
__device__ __forceinline__ void build_store_maps(const TGatherParams& p, int x0, int y0, int tid, int* s_rowmap,
                                                 int* s_colmap) {
    if (tid < TH + TW) {
        const bool row = tid < TH;
        constexpr int R = TH / LY;                       // row tid = strip ly = tid / R, row r = tid % R of it
        const int v = row ? y0 + (tid / R) * p.ly_stride + tid % R : x0 + tid - TH;
        const int ext = row ? p.H : p.W, pitch = row ? p.m.py : p.m.px, size = row ? p.m.H : p.m.W,
                  cells = row ? p.m.ky : p.m.kx;
        int code = -1;
        if (v < ext) {
            const int j = v / pitch, l = v - j * pitch;
            if (l < size && j < cells) code = (j << 16) | l;
        }
        if (row) s_rowmap[tid] = code; else s_colmap[tid - TH] = code;
    }
}

// Epilogue: bias, activation, and the store of a lane's 16 x 4 strip of both images of the pair
// into the caller's NCHW tensor (float4 per row when the planes allow it).
template <int R>
__device__ __forceinline__ void store_strip(const TGatherParams& p, const float2 (&acc)[R][C], int np, int co, int lx,
                                            int ly, int x0, int y0, const int* s_rowmap, const int* s_colmap) {
    const float bv = p.bias ? __ldg(p.bias + co) : 0.f;
    const int xs = x0 + lx * C;
    const int H = p.H, W = p.W;
    const bool ident = p.m.kx == 1 && p.m.ky == 1;
    const int crop = p.crop, RH = p.m.H - 2 * crop, RW = p.m.W - 2 * crop;
    const bool vec = ident && crop == 0 && ((W & 3) == 0) && (xs + C <= W);
    const size_t rplane = (size_t)RH * RW;
#pragma unroll
    for (int img = 0; img < 2; ++img) {
        const int n = 2 * np + img;
        if (n >= p.N) break;
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const int y = y0 + ly * p.ly_stride + r;
            if (y >= H) continue;
            float v[C];
#pragma unroll
            for (int c = 0; c < C; ++c) {
                v[c] = (img == 0 ? acc[r][c].x : acc[r][c].y) + bv;
                if (p.relu) v[c] = fmaxf(v[c], 0.f);
            }
            if (vec) {
                float* orow = p.out + ((size_t)n * p.Cout + co) * rplane + (size_t)y * W;
                *reinterpret_cast<float4*>(orow + xs) = make_float4(v[0], v[1], v[2], v[3]);
            } else {
                // mosaic: gap pixels and cells beyond the batch are computed but never stored;
                // crop: neither is the border of the same-size result
                const int rm = s_rowmap[ly * R + r];
#pragma unroll
                for (int c = 0; c < C; ++c) {
                    const int cm = s_colmap[lx * C + c];
                    const int rn = (n * p.m.ky + (rm >> 16)) * p.m.kx + (cm >> 16);
                    const int ry = (rm & 0xffff) - crop, rx = (cm & 0xffff) - crop;
                    if ((rm | cm) >= 0 && rn < p.m.N && (unsigned)ry < (unsigned)RH && (unsigned)rx < (unsigned)RW)
                        p.out[((size_t)rn * p.Cout + co) * rplane + (size_t)ry * RW + rx] = v[c];
                }
            }
        }
    }
}

template <int HALO, int R>
__global__ void __launch_bounds__(NTHREADS, 1)
dau_gather_tma_kernel(const __grid_constant__ CUtensorMap in_map, const TGatherParams p) {
    using GEO = TGeom<HALO, R>;
    constexpr int TH = GEO::TH;
    constexpr int QS = GEO::QS, ROW = GEO::ROW;
    const int TILE_BYTES = p.tile_bytes;
    extern __shared__ unsigned char smem_dyn[];
    __shared__ uint64_t full_bar[8], empty_bar[8];
    // TMA destinations must be 128-byte aligned; the dynamic segment carries 128 spare bytes
    unsigned char* smem_raw = smem_dyn + ((128u - (smem_u32(smem_dyn) & 127u)) & 127u);

    const int tid = threadIdx.x, lane = tid & 31;
    const int warp = (tid >> 5) - 4;               // consumer warp 0..7; negative: producer warpgroup
    // item = (tile, group of 8 output channels, image pair).  The items of the last, partial wave are
    // cut along the input channels into split_k CTAs each (their partial strips are summed in a fixed
    // order by gather_tail_kernel), so that the tail costs 1/split_k of a wave instead of a whole one.
    int item = blockIdx.x, slot = -1, ci_begin = 0, ci_end = p.Cin;
    if (item >= p.full_items) {
        slot = item - p.full_items;
        const int part = slot % p.split_k;
        item = p.full_items + slot / p.split_k;
        ci_begin = (int)((long long)part * p.Cin / p.split_k);
        ci_end = (int)((long long)(part + 1) * p.Cin / p.split_k);
    }
    const int tile_i = item % p.gx, cg = (item / p.gx) % p.gy;
    const int np = item / (p.gx * p.gy);       // image pair
    const int tyi = tile_i / p.tiles_x, txi = tile_i - tyi * p.tiles_x;
    const int x0 = txi * TW, y0 = tyi * p.tile_pitch_y;
    const int co0 = cg * NW;
    const int nrec = min(NW, p.Cout - co0) * p.G;
    const int nstage = p.nstage;

    if (tid == 0) {
        for (int s = 0; s < nstage; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], NW); }
        fence_barrier_init();
    }
    __shared__ int s_rowmap[TH], s_colmap[TW];
    build_store_maps<TH>(p, x0, y0, tid, s_rowmap, s_colmap);
    __syncthreads();

    if (warp < 0) {
        // ---- producer warpgroup: one lane streams the input channels ----------------------------
        asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;\n" ::"n"(PROD_REGS));
        if (warp == -4 && lane == 0) {
            if (!(DAU_DEV_BITS(p.debug) & 1)) prefetch_map(&in_map);
            const uint32_t tx_bytes = ((DAU_DEV_BITS(p.debug) & 2) ? 0 : TILE_BYTES) + ((DAU_DEV_BITS(p.debug) & 4) ? 0 : nrec * (uint32_t)sizeof(UnitRec));
            for (int ci = ci_begin; ci < ci_end; ++ci) {
                const int st = (ci - ci_begin) % nstage, use = (ci - ci_begin) / nstage;
                if (use > 0) mbar_wait_backoff(&empty_bar[st], (use - 1) & 1);
                unsigned char* dst = smem_raw + (size_t)st * p.stage_bytes;
                mbar_arrive_expect_tx(&full_bar[st], tx_bytes);
                if (!(DAU_DEV_BITS(p.debug) & 2)) tensor4_g2s(dst, &in_map, 2 * ((x0 - GEO::HX) / 4), 0, y0 - GEO::HY, np * p.Cin + ci, &full_bar[st]);
                if (!(DAU_DEV_BITS(p.debug) & 4)) bulk_g2s(dst + TILE_BYTES, p.table + ((size_t)ci * p.Cout + co0) * p.G,
                         nrec * (uint32_t)sizeof(UnitRec), &full_bar[st]);
            }
        }
        return;
    }

    // ---- consumer warps ---------------------------------------------------------------------------
    asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;\n" ::"n"(CONS_REGS));
    const int lx = lane & (LX - 1), ly = lane >> 4;
    const int co = co0 + warp;
    const bool co_valid = co < p.Cout;
    float2 acc[R][C];
#pragma unroll
    for (int r = 0; r < R; ++r)
#pragma unroll
        for (int c = 0; c < C; ++c) acc[r][c] = make_float2(0.f, 0.f);

    const int lane_base = ly * p.ly_stride * ROW + lx;
    for (int ci = ci_begin; ci < ci_end; ++ci) {
        const int st = (ci - ci_begin) % nstage, use = (ci - ci_begin) / nstage;
        mbar_wait(&full_bar[st], use & 1);
        const unsigned char* sbase = smem_raw + (size_t)st * p.stage_bytes;
        if (co_valid) {
            const float2* tile = reinterpret_cast<const float2*>(sbase) + lane_base;
            const UnitRec* recs = reinterpret_cast<const UnitRec*>(sbase + TILE_BYTES) + warp * p.G;
            for (int g = 0; g < p.G; ++g) {
                const float4 wv = *reinterpret_cast<const float4*>(&recs[g]);
                const int2 off = *reinterpret_cast<const int2*>(&recs[g].offx_h);
                const float2 w00 = make_float2(wv.x, wv.x), w01 = make_float2(wv.y, wv.y);
                const float2 w10 = make_float2(wv.z, wv.z), w11 = make_float2(wv.w, wv.w);
                const int b = off.x & 3;
                const float2* base = tile + off.y * ROW + (off.x >> 2);
                const float2* colp[C + 1];
#pragma unroll
                for (int c = 0; c <= C; ++c) {
                    const int m = b + c;
                    colp[c] = base + (m & 3) * QS + (m >> 2);
                }
#pragma unroll
                for (int r = 0; r <= R; ++r) {
                    float2 xv[C + 1];
#pragma unroll
                    for (int c = 0; c <= C; ++c) xv[c] = colp[c][r * ROW];
                    if (r < R) {
#pragma unroll
                        for (int c = 0; c < C; ++c) {
                            acc[r][c] = __ffma2_rn(w00, xv[c], acc[r][c]);
                            acc[r][c] = __ffma2_rn(w01, xv[c + 1], acc[r][c]);
                        }
                    }
                    if (r > 0) {
#pragma unroll
                        for (int c = 0; c < C; ++c) {
                            acc[r - 1][c] = __ffma2_rn(w10, xv[c], acc[r - 1][c]);
                            acc[r - 1][c] = __ffma2_rn(w11, xv[c + 1], acc[r - 1][c]);
                        }
                    }
                }
            }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty_bar[st]);
    }

    if (!co_valid) return;
    if (slot >= 0) {
        // part of a split item: the raw strip goes to the workspace, coalesced over the lanes
        float2* dst = p.part + ((size_t)slot * NW + warp) * (R * C * 32) + lane;
#pragma unroll
        for (int r = 0; r < R; ++r)
#pragma unroll
            for (int c = 0; c < C; ++c) dst[(r * C + c) * 32] = acc[r][c];
        return;
    }
    store_strip<R>(p, acc, np, co, lx, ly, x0, y0, s_rowmap, s_colmap);
}

// One row (r) of a lane's strip, both images: the epilogue of store_strip for the kernel that sums split items.
template <int R>
__device__ __forceinline__ void store_row(const TGatherParams& p, const float2 (&acc)[C], int np, int co, int lx, int ly,
                                          int r, int x0, int y0, const int* s_rowmap, const int* s_colmap) {
    const float bv = p.bias ? __ldg(p.bias + co) : 0.f;
    const int xs = x0 + lx * C;
    const int H = p.H, W = p.W;
    const bool ident = p.m.kx == 1 && p.m.ky == 1;
    const int crop = p.crop, RH = p.m.H - 2 * crop, RW = p.m.W - 2 * crop;
    const bool vec = ident && crop == 0 && ((W & 3) == 0) && (xs + C <= W);
    const size_t rplane = (size_t)RH * RW;
    const int y = y0 + ly * p.ly_stride + r;
    if (y >= H) return;
#pragma unroll
    for (int img = 0; img < 2; ++img) {
        const int n = 2 * np + img;
        if (n >= p.N) break;
        float v[C];
#pragma unroll
        for (int c = 0; c < C; ++c) {
            v[c] = (img == 0 ? acc[c].x : acc[c].y) + bv;
            if (p.relu) v[c] = fmaxf(v[c], 0.f);
        }
        if (vec) {
            float* orow = p.out + ((size_t)n * p.Cout + co) * rplane + (size_t)y * W;
            *reinterpret_cast<float4*>(orow + xs) = make_float4(v[0], v[1], v[2], v[3]);
        } else {
            const int rm = s_rowmap[ly * R + r];
#pragma unroll
            for (int c = 0; c < C; ++c) {
                const int cm = s_colmap[lx * C + c];
                const int rn = (n * p.m.ky + (rm >> 16)) * p.m.kx + (cm >> 16);
                const int ry = (rm & 0xffff) - crop, rx = (cm & 0xffff) - crop;
                if ((rm | cm) >= 0 && rn < p.m.N && (unsigned)ry < (unsigned)RH && (unsigned)rx < (unsigned)RW)
                    p.out[((size_t)rn * p.Cout + co) * rplane + (size_t)ry * RW + rx] = v[c];
            }
        }
    }
}

// Sums the split_k partial strips of one tail item in part order (deterministic) and stores them like any other.
// grid = (tail items, output channels of the group), block = R warps: thread (lane, r) owns row r of lane's strip, so
// the serial work of a thread is split_k x 4 coalesced loads -- on the latency-bound shapes (BASELINE config C1, the
// reference's speed-test shape) the first version of this kernel, one thread per whole strip, took longer than the
// gather it finishes (39 us against 15 us on C1).
template <int R>
__global__ void __launch_bounds__(32 * R) gather_tail_kernel(const TGatherParams p) {
    constexpr int TH = LY * R;
    const int item = p.full_items + blockIdx.x;
    const int tile_i = item % p.gx, cg = (item / p.gx) % p.gy, np = item / (p.gx * p.gy);
    const int tyi = tile_i / p.tiles_x, txi = tile_i - tyi * p.tiles_x;
    const int x0 = txi * TW, y0 = tyi * p.tile_pitch_y;
    const int tid = threadIdx.x, lane = tid & 31, r = tid >> 5, warp = blockIdx.y;
    __shared__ int s_rowmap[TH], s_colmap[TW];
    build_store_maps<TH>(p, x0, y0, tid, s_rowmap, s_colmap);
    __syncthreads();
    const int co = cg * NW + warp;
    if (co >= p.Cout) return;
    float2 acc[C];
#pragma unroll
    for (int c = 0; c < C; ++c) acc[c] = make_float2(0.f, 0.f);
    for (int part = 0; part < p.split_k; ++part) {
        const float2* src = p.part + ((size_t)(blockIdx.x * p.split_k + part) * NW + warp) * (R * C * 32) + lane;
#pragma unroll
        for (int c = 0; c < C; ++c) {
            const float2 v = src[(r * C + c) * 32];
            acc[c].x += v.x; acc[c].y += v.y;
        }
    }
    store_row<R>(p, acc, np, co, lane & (LX - 1), lane >> 4, r, x0, y0, s_rowmap, s_colmap);
}

constexpr int kHalos[4] = {4, 8, 16, 32};

template <int HALO, int R>
void launch_t(const GatherPlan& plan, const CUtensorMap& map, const TGatherParams& p, cudaStream_t stream) {
    static PerDeviceOnce once;                 // the attribute is per (kernel, device)
    if (once.need()) {
        DAU_CUDA_CHECK(cudaFuncSetAttribute(dau_gather_tma_kernel<HALO, R>,
                                            cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
        once.done();
    }
    const long long items = (long long)plan.grid_x * plan.grid_y * plan.grid_z;
    const long long ctas = plan.full_items + (items - plan.full_items) * plan.split_k;
    dau_gather_tma_kernel<HALO, R><<<(unsigned)ctas, NTHREADS, plan.smem_bytes, stream>>>(map, p);
    if (plan.full_items < items)
        gather_tail_kernel<R><<<dim3((unsigned)(items - plan.full_items), NW), 32 * R, 0, stream>>>(p);
}

template <int HALO, int R>
void fill_t(GatherPlan& pl) {
    using GEO = TGeom<HALO, R>;
    pl.halo_x = GEO::HX; pl.halo_y = GEO::HY;
    pl.row_stride = GEO::ROW * 2; pl.img_stride = 0;
}

}  // namespace

size_t paired_layout_floats(int N, int Cn, int H, int W) {
    const size_t qv = (W + 3) / 4, qp = (qv + 1) / 2 * 2;
    return (size_t)((N + 1) / 2) * Cn * H * 4 * qp * 2;
}

bool plan_gather_tma(const Mosaic& m, int Cin, int Cout, int G, int max_off, int num_sms, GatherPlan* out) {
    const int N = m.Nv, H = m.Hv, W = m.Wv;
    int hclass = 0;
    while (kHalos[hclass] < max_off) ++hclass;
    GatherPlan pl; memset(&pl, 0, sizeof(pl));
    pl.kind = 1; pl.variant = hclass;
    // strip rows R.  Cell-row mode when the mosaic stacks cells of 13..16 rows: R = the image height, the two
    // strips of a tile sit one cell pitch apart.  Else stacked strips: the 2R-row tiling that computes the
    // fewest rows (ties: the taller strip).
    int R = 16;
    const bool cell_rows = m.ky >= 2 && m.H >= 13 && m.H <= 16;
    if (cell_rows) {
        R = m.H;
    } else {
        long best = -1;
        for (int r : {16, 15, 14}) {
            const long rows = (long)((H + 2 * r - 1) / (2 * r)) * 2 * r;
            if (best < 0 || rows < best) { best = rows; R = r; }
        }
    }
    const int TH = LY * R;
    pl.tile_w = TW; pl.tile_h = TH; pl.imgs = 2; pl.cout_per_cta = NW;
    pl.ly_stride = cell_rows ? m.py : R;
    pl.tile_pitch_y = cell_rows ? 2 * m.py : TH;
#define DAU_FILL(HALO)                                              \
    switch (R) {                                                    \
        case 16: fill_t<HALO, 16>(pl); break;                       \
        case 15: fill_t<HALO, 15>(pl); break;                       \
        case 14: fill_t<HALO, 14>(pl); break;                       \
        default: fill_t<HALO, 13>(pl); break;                       \
    }
    switch (hclass) {
        case 0: DAU_FILL(4); break;
        case 1: DAU_FILL(8); break;
        case 2: DAU_FILL(16); break;
        default: DAU_FILL(32); break;
    }
#undef DAU_FILL
    pl.tile_rows = pl.ly_stride + R + 2 * pl.halo_y;
    const size_t tile_bytes = (size_t)pl.tile_rows * (pl.row_stride / 2) * 8;      // rows x ROW float2
    pl.tile_bytes = (int)tile_bytes;
    const size_t stage = (tile_bytes + (size_t)NW * G * sizeof(UnitRec) + 127) / 128 * 128;
    int nstage = (int)((200 * 1024 - 128) / stage);
    if (nstage > 4) nstage = 4;
    if (nstage < 2) return false;
    pl.stage_words = (int)(stage / 4);
    pl.nstage = nstage;
    pl.smem_bytes = stage * nstage + 128;
    const int tiles_y = cell_rows ? (m.ky + 1) / 2 : (H + TH - 1) / TH;
    pl.grid_x = ((W + TW - 1) / TW) * tiles_y;
    pl.grid_y = (Cout + NW - 1) / NW;
    pl.grid_z = (N + 1) / 2;
    // One CTA per SM: `items` equal CTAs take ceil(items / SMs) rounds.  The items of the last,
    // partial round are cut along the input channels into split_k CTAs whose partial strips are
    // summed afterwards: the tail then costs ceil(tail * k / SMs) / k rounds instead of one.
    const long long items = (long long)pl.grid_x * pl.grid_y * pl.grid_z;
    if (items >= (1LL << 30)) return false;
    const long long full = items / num_sms * num_sms, tail = items - full;
    pl.full_items = (int)items; pl.split_k = 1;
    pl.rounds = (double)((items + num_sms - 1) / num_sms);
    if (tail > 0) {
        int best_k = 1; double best_t = 1.0;
        for (int k = 2; k <= 8 && Cin / k >= 4 && tail * k <= 512; ++k) {
            const double t = (double)((tail * k + num_sms - 1) / num_sms) / k + 0.02;   // + the summing kernel
            if (t < best_t - 1e-9) { best_t = t; best_k = k; }
        }
        if (best_k > 1) {
            pl.full_items = (int)full; pl.split_k = best_k;
            pl.rounds = (double)(full / num_sms) + best_t;
        }
    }
    pl.part_bytes = (size_t)(items - pl.full_items) * pl.split_k * NW * R * C * 32 * sizeof(float2);
    *out = pl;
    return true;
}

void launch_gather_tma(const GatherPlan& plan, const float* in_paired, const UnitRec* table, const float* bias,
                       float* out, const Mosaic& m, int Cin, int Cout, int G, cudaStream_t stream, int relu, int crop,
                       float* part_buf) {
    const int N = m.Nv, H = m.Hv, W = m.Wv;
    // tensor map of the paired gather layout: dims (q*2+i, p, y, plane)
    const uint64_t qv = (W + 3) / 4, qp = (qv + 1) / 2 * 2;
    const uint64_t dims[4] = {2 * qv, 4, (uint64_t)H, (uint64_t)((N + 1) / 2) * Cin};
    const uint64_t strides[3] = {qp * 8, 4 * qp * 8, (uint64_t)H * 4 * qp * 8};
    const int halo_q2 = 2 * ((TW + 2 * plan.halo_x) / 4);
    const uint32_t box[4] = {(uint32_t)halo_q2, 4, (uint32_t)plan.tile_rows, 1};
    const CUtensorMap map = make_map_f32(in_paired, 4, dims, strides, box);
    TGatherParams p;
    p.table = table; p.bias = bias; p.out = out;
    p.N = N; p.Cin = Cin; p.Cout = Cout; p.G = G; p.H = H; p.W = W;
    p.tiles_x = (W + TW - 1) / TW; p.nstage = plan.nstage;
    p.stage_bytes = plan.stage_words * 4;
    p.ly_stride = plan.ly_stride; p.tile_pitch_y = plan.tile_pitch_y; p.tile_bytes = plan.tile_bytes;
    p.m = m; p.relu = relu; p.crop = crop;
    p.gx = plan.grid_x; p.gy = plan.grid_y; p.full_items = plan.full_items; p.split_k = plan.split_k;
    p.part = reinterpret_cast<float2*>(part_buf);
    DAU_CHECK(plan.part_bytes == 0 || part_buf != nullptr, 3, "gather plan splits its tail but no partial buffer was given");
    p.debug = 0;
#ifdef DAU_DEV_SWITCHES
    p.debug = getenv("DAU_DEBUG") ? atoi(getenv("DAU_DEBUG")) : 0;
#endif
#define DAU_LAUNCH(HALO)                                                     \
    switch (plan.tile_h / LY) {                                              \
        case 16: launch_t<HALO, 16>(plan, map, p, stream); break;            \
        case 15: launch_t<HALO, 15>(plan, map, p, stream); break;            \
        case 14: launch_t<HALO, 14>(plan, map, p, stream); break;            \
        default: launch_t<HALO, 13>(plan, map, p, stream); break;            \
    }
    switch (plan.variant) {
        case 0: DAU_LAUNCH(4); break;
        case 1: DAU_LAUNCH(8); break;
        case 2: DAU_LAUNCH(16); break;
        default: DAU_LAUNCH(32); break;
    }
#undef DAU_LAUNCH
    DAU_CUDA_CHECK(cudaGetLastError());
}

}  // namespace dau
