// Per-unit parameter gradients of the DAU layer for sm_100a (TMA + packed FFMA2).
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
// registers for all units of the pass, and the operand that is gathered at a data-dependent
// displacement -- E -- comes from a haloed shared-memory tile where each word feeds up to
// four FMAs.  Accumulators acc[unit][k] live in registers for the CTA's whole walk over its
// items and are reduced across the warp once at the end -> one plain store per output into
// partial[split]; no atomics, deterministic.
//
// How operands move and how the FMAs are packed:
//
//  * work item = (image PAIR, row strip, 64-column tile).  Row strips never contain a mosaic gap row:
//    every cell row of the mosaic is cut into H/4 full strips of 4 rows and one tail strip of H%4
//    rows (StripGeom in dau_kernels.cuh); the kernel walks its full items with the 4-row code and
//    then its tail items with a copy of the same code instantiated for the tail height, so a 13-row
//    map costs 13 rows of work and not the 16 of four 4-row tiles (or the 32 of eight mosaic tiles
//    for 26 rows, as before).  Everything is laid out so that the two images of a pair are the two
//    halves of one fma.rn.f32x2 (SASS FFMA2) operand: they share the unit's offsets and bilinear
//    weights, so no register shuffling is ever needed to form pairs;
//  * the derivative-blur kernel writes Xk "strip-major":
//        XS[item][s][k][r<4][lane<32][c<2][img<2]
//    so that everything the 8 consumer warps (= 8 input channels) of a CTA need for one item is
//    ONE contiguous 64 KB block fetched by ONE cp.async.bulk, and a lane reads its (k,r) strip
//    row with one LDS.128;
//  * the top error is re-laid out once (edge mask applied) as
//        SGP[n/2][f][y][x%2][x/2][n%2]
//    so that the haloed tile of one (image pair, feature) is one 4-D TMA box, already column
//    de-interleaved; 32 lanes x 2 columns cover the 64-px tile and every LDS.64 of the window is
//    two conflict-free wavefronts whatever the unit's offset;
//  * a producer lane streams the error tiles of the items through a 2-stage ring and Xk through
//    ONE buffer (the consumers move Xk to registers first thing, so the next item's copy
//    overlaps the math), all guarded by full/empty mbarriers; the consumer warps execute no
//    global loads and meet no CTA barrier.
//
// Per unit and lane (strip = 4 rows x 2 columns x 2 images): 15 LDS.64 feed 64 FFMA2
//   e2(r,c)  = w00*P[r][c] + w01*P[r][c+1] + w10*P[r+1][c] + w11*P[r+1][c+1]     (pairs over images)
//   acc2[k] += e2(r,c) * xk2[k][r][c]            (image partial sums added at the very end)
// A scalar FFMA whose three source registers all change issues at ~0.6 / clk / sub-partition on
// sm_100a and a packed one with two changing pairs at ~0.4 (= 0.8 FMA-equivalents), see
// tools/probes/fma_probe.cu; the packed form also halves the issue slots the math needs.
#include "dau_kernels.cuh"
#include "dau_tma.cuh"

#include <cstdlib>

namespace dau {

namespace {

using namespace tma;

constexpr int R = 4, C = 2, TW = 64, TH = 4;
// Two shapes of the CTA, selected by the template parameter UPW = units per warp and pass (= accumulator slots
// acc[UPW][4] float2 per lane):
//   UPW 12: 8 consumer warps (2 per SM sub-partition) of 232 registers       -- the shipped shape
//   UPW  8: 12 consumer warps (3 per sub-partition) of 160 registers          -- more warps to hide latency, fewer
//           units per pass (Xk re-read from shared memory once per 8 units); measured, see DESIGN.md
// Warpgroup 0 holds the producer lane and gives its registers away (setmaxnreg.dec); the consumer warpgroups
// grow to CONS_REGS each (setmaxnreg.inc).
template <int UPW>
struct Cfg {
    static constexpr int NW = UPW == 8 ? 12 : 8;
    static constexpr int NTHREADS = (NW + 4) * 32;
    static constexpr int PROD_REGS = UPW == 8 ? 24 : 40;
    static constexpr int CONS_REGS = UPW == 8 ? 160 : 232;
};
constexpr int XK_ROW_BYTES = 32 * C * 2 * 4;            // one (k, r) row of a warp's strip: 512 B
constexpr int XK_WARP_BYTES = 4 * R * XK_ROW_BYTES;     // 8 KB: [k][r][lane][c][img] of a full strip

template <int HALO>
struct GGeomT {
    // innermost TMA start coordinate (x0-HX) floats must be a multiple of 4 floats
    static constexpr int HX = (HALO + 3) / 4 * 4, HY = HALO;
    static constexpr int QS = (TW + 2 * HX) / 2;       // float2 per column plane
    static constexpr int THH = TH + 2 * HY;
    static constexpr int ROW = 2 * QS;                 // float2 per tile row
    static constexpr int FEAT_BYTES = THH * ROW * 8;
};

// CTA-local copy of a unit record with the four tap weights duplicated into FFMA2 operands
struct __align__(16) GUnit {
    float2 w00, w01, w10, w11;
    int offx, offy, tile_off, pad;
};
static_assert(sizeof(GUnit) == 48, "GUnit must be 48 bytes");

struct TGradParams {
    const float* xs; const UnitRec* table; float* partial;
    int N, S, F, G;
    int FC, GC, n_gchunks;
    int tiles_x, ky, py, q;            // strip geometry (StripGeom)
    long long full_items, tail_items;
    long long full_per_split, tail_per_split;
    long long tail_base_floats;
    int feat_bytes_al;     // FEAT_BYTES rounded up to 128
    int stage_bytes;
    int debug;             // development switches, only with -DDAU_DEV_SWITCHES (dau_common.hpp)
};

// (pair, first mosaic row, column tile) of a full (tail == false) or tail item
__host__ __device__ __forceinline__ void item_origin(const TGradParams& p, long long item, bool tail, int* np, int* y0, int* txi) {
    const int tx = (int)(item % p.tiles_x);
    long long v = item / p.tiles_x;
    int t = p.q;
    if (!tail) { t = (int)(v % p.q); v /= p.q; }
    const int jy = (int)(v % p.ky);
    *np = (int)(v / p.ky); *y0 = jy * p.py + 4 * t; *txi = tx;
}

// One item of a consumer warp: its Xk strip (RR rows) to registers, then the walk over the units.
template <int HALO, int RR, int UPW>
__device__ __forceinline__ void consume_item(const unsigned char* smem_raw, const TGradParams& p, const GUnit* units,
                                             float2 (&acc)[UPW][4], int warp, int lane, bool s_valid, int kk,
                                             uint64_t* xfull_bar, uint64_t* xempty_bar, uint64_t* full_bar,
                                             uint64_t* empty_bar) {
    using GEO = GGeomT<HALO>;
    constexpr int QS = GEO::QS, ROW = GEO::ROW;
    const int st = kk & 1, use = kk >> 1;
    float2 xk[4][RR][C];
    mbar_wait(xfull_bar, kk & 1);
    if (s_valid) {
        const float4* xsrc = reinterpret_cast<const float4*>(smem_raw + 2 * (size_t)p.stage_bytes +
                                                             warp * (4 * RR * XK_ROW_BYTES)) + lane;
#pragma unroll
        for (int k = 0; k < 4; ++k)
#pragma unroll
            for (int r = 0; r < RR; ++r) {
                const float4 v = xsrc[(k * RR + r) * 32];
                xk[k][r][0] = make_float2(v.x, v.y);
                xk[k][r][1] = make_float2(v.z, v.w);
            }
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(xempty_bar);               // the buffer may be refilled while we compute
    mbar_wait(&full_bar[st], use & 1);
    const unsigned char* sbase = smem_raw + (size_t)st * p.stage_bytes;
    if (s_valid && !(DAU_DEV_BITS(p.debug) & 8)) {
        const float2* tb = reinterpret_cast<const float2*>(sbase) + lane;
#pragma unroll
        for (int u = 0; u < UPW; ++u) {
            const GUnit* rec = &units[u];
            const float4 wa = *reinterpret_cast<const float4*>(&rec->w00);   // w00 w00 w01 w01
            const float4 wb = *reinterpret_cast<const float4*>(&rec->w10);   // w10 w10 w11 w11
            const int4 off = *reinterpret_cast<const int4*>(&rec->offx);     // offx, offy, tile offset
            const float2 w00 = make_float2(wa.x, wa.y), w01 = make_float2(wa.z, wa.w);
            const float2 w10 = make_float2(wb.x, wb.y), w11 = make_float2(wb.z, wb.w);
            const int b = off.x & 1;
            const float2* base = tb + off.z + off.y * ROW + (off.x >> 1);
            const float2* colp[C + 1];
#pragma unroll
            for (int c = 0; c <= C; ++c) {
                const int m = b + c;
                colp[c] = base + (m & 1) * QS + (m >> 1);
            }
            float2 prev[C + 1];
#pragma unroll
            for (int c = 0; c <= C; ++c) prev[c] = colp[c][0];
#pragma unroll
            for (int r = 1; r <= RR; ++r) {
                float2 cur[C + 1];
#pragma unroll
                for (int c = 0; c <= C; ++c) cur[c] = colp[c][r * ROW];
#pragma unroll
                for (int c = 0; c < C; ++c) {
                    float2 e = __fmul2_rn(w00, prev[c]);
                    e = __ffma2_rn(w01, prev[c + 1], e);
                    e = __ffma2_rn(w10, cur[c], e);
                    e = __ffma2_rn(w11, cur[c + 1], e);
#pragma unroll
                    for (int k = 0; k < 4; ++k) acc[u][k] = __ffma2_rn(e, xk[k][r - 1][c], acc[u][k]);
                }
#pragma unroll
                for (int c = 0; c <= C; ++c) prev[c] = cur[c];
            }
        }
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(&empty_bar[st]);
}

template <int HALO, int TAIL, int UPW>
__global__ void __launch_bounds__(Cfg<UPW>::NTHREADS, 1)
dau_grad_tma_kernel(const __grid_constant__ CUtensorMap err_map, const TGradParams p) {
    using GEO = GGeomT<HALO>;
    constexpr int NW = Cfg<UPW>::NW, PROD_REGS = Cfg<UPW>::PROD_REGS, CONS_REGS = Cfg<UPW>::CONS_REGS;
    extern __shared__ unsigned char smem_dyn[];
    __shared__ uint64_t full_bar[2], empty_bar[2];      // error-tile ring
    __shared__ uint64_t xfull_bar, xempty_bar;           // the single Xk buffer
    __shared__ GUnit s_units[NW][UPW];               // (UPW: template parameter)
    unsigned char* smem_raw = smem_dyn + ((128u - (smem_u32(smem_dyn) & 127u)) & 127u);

    const int tid = threadIdx.x, lane = tid & 31;
    const int warp = (tid >> 5) - 4;                   // consumer warp 0..7; negative: producer warpgroup
    // f-chunks vary fastest across the grid: CTAs that run together share the Xk block in L2
    const int fchunk = blockIdx.x / p.n_gchunks, gchunk = blockIdx.x - fchunk * p.n_gchunks;
    const int s0 = blockIdx.y * NW;
    const int f0 = fchunk * p.FC, g0 = gchunk * p.GC;
    const int F = p.F, G = p.G;
    const int nfeat = min(p.FC, F - f0), ngc = min(p.GC, G - g0);
    const int nunits = nfeat * ngc;
    const int nsv = min(NW, p.S - s0);                 // valid input channels of this CTA
    // this CTA's share of the reduction axis: a range of full items, then a range of tail items
    const long long fb = min(p.full_items, (long long)blockIdx.z * p.full_per_split);
    const long long fe = min(p.full_items, fb + p.full_per_split);
    const long long tb_ = TAIL ? min(p.tail_items, (long long)blockIdx.z * p.tail_per_split) : 0;
    const long long te = TAIL ? min(p.tail_items, tb_ + p.tail_per_split) : 0;
    const int nfull = (int)(fe - fb), nsteps = nfull + (int)(te - tb_);

    if (tid == 0) {
        for (int s = 0; s < 2; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], NW); }
        mbar_init(&xfull_bar, 1); mbar_init(&xempty_bar, NW);
        fence_barrier_init();
    }
    __syncthreads();                                   // barriers initialised: the producer starts copying NOW

    if (warp < 0) {
        asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;\n" ::"n"(PROD_REGS));
        if (warp == -4 && lane == 0) {
            prefetch_map(&err_map);
            // Error tiles go through a 2-stage ring.  Xk needs ONE buffer only: the consumers move it
            // to registers at the start of an item, so the next item's copy overlaps the math.
            const uint32_t tx_e = (DAU_DEV_BITS(p.debug) & 2) ? 0 : nfeat * GEO::FEAT_BYTES;
            unsigned char* xdst = smem_raw + 2 * (size_t)p.stage_bytes;
            for (int k = 0; k < nsteps; ++k) {
                const bool tail = k >= nfull;
                const long long item = tail ? tb_ + (k - nfull) : fb + k;
                const int st = k & 1, use = k >> 1;
                int n, y0, txi;                                       // n = image pair
                item_origin(p, item, tail, &n, &y0, &txi);
                const uint32_t xbytes = (tail ? TAIL : R) * 4 * XK_ROW_BYTES;      // per input channel
                const float* xsrc = tail ? p.xs + p.tail_base_floats + ((size_t)item * p.S + s0) * (xbytes / 4)
                                         : p.xs + ((size_t)item * p.S + s0) * (XK_WARP_BYTES / 4);
                if (k > 0) mbar_wait_backoff(&xempty_bar, (k - 1) & 1);
                const uint32_t tx_x = (DAU_DEV_BITS(p.debug) & 4) ? 0 : nsv * xbytes;
                mbar_arrive_expect_tx(&xfull_bar, tx_x);
                if (!(DAU_DEV_BITS(p.debug) & 4)) bulk_g2s(xdst, xsrc, nsv * xbytes, &xfull_bar);
                if (use > 0) mbar_wait_backoff(&empty_bar[st], (use - 1) & 1);
                unsigned char* dst = smem_raw + (size_t)st * p.stage_bytes;
                mbar_arrive_expect_tx(&full_bar[st], tx_e);
                for (int fi = 0; fi < nfeat && !(DAU_DEV_BITS(p.debug) & 2); ++fi)
                    tensor4_g2s(dst + fi * p.feat_bytes_al, &err_map, txi * TW - GEO::HX, 0,
                                y0 - GEO::HY, n * F + f0 + fi, &full_bar[st]);
            }
        }
        return;
    }

    // ---- consumer warps: the unit records of this CTA, fetched while the first copies are in flight ----
    asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;\n" ::"n"(CONS_REGS));
    if (lane < UPW) {
        // slots beyond the chunk's units are dummies (zero weights, in-tile offsets): the unit loop
        // is branch-free so that one unit's prologue overlaps the previous unit's math
        GUnit gu;
        gu.w00 = gu.w01 = gu.w10 = gu.w11 = make_float2(0.f, 0.f);
        gu.offx = GEO::HX; gu.offy = GEO::HY; gu.tile_off = 0; gu.pad = 0;
        if (s0 + warp < p.S && lane < nunits) {
            const int fi = lane / ngc, gi = lane - fi * ngc;
            const UnitRec r = p.table[((size_t)(s0 + warp) * F + f0 + fi) * G + g0 + gi];
            gu.w00 = make_float2(r.w00, r.w00); gu.w01 = make_float2(r.w01, r.w01);
            gu.w10 = make_float2(r.w10, r.w10); gu.w11 = make_float2(r.w11, r.w11);
            gu.offx = r.offx_h; gu.offy = r.offy_h;
            gu.tile_off = fi * (p.feat_bytes_al / 8);  // float2 offset of the feature's error tile
        }
        s_units[warp][lane] = gu;                      // a warp reads only its own row: a warp-level sync is enough
    }
    __syncwarp();

    // warp = input channel s0 + warp
    const bool s_valid = warp < nsv;
    float2 acc[UPW][4];                                // {image 0, image 1} partial sums
#pragma unroll
    for (int u = 0; u < UPW; ++u)
#pragma unroll
        for (int k = 0; k < 4; ++k) acc[u][k] = make_float2(0.f, 0.f);

    for (int kk = 0; kk < nfull; ++kk)
        consume_item<HALO, R, UPW>(smem_raw, p, s_units[warp], acc, warp, lane, s_valid, kk, &xfull_bar, &xempty_bar,
                              full_bar, empty_bar);
    if (TAIL > 0) {
        for (int kk = nfull; kk < nsteps; ++kk)
            consume_item<HALO, (TAIL > 0 ? TAIL : 1), UPW>(smem_raw, p, s_units[warp], acc, warp, lane, s_valid, kk,
                                                      &xfull_bar, &xempty_bar, full_bar, empty_bar);
    }

    if (s_valid) {
        const int s = s0 + warp;
        const size_t sgf = (size_t)p.S * G * F;
        float* pbase = p.partial + (size_t)blockIdx.z * 4 * sgf;
#pragma unroll
        for (int u = 0; u < UPW; ++u) {
            if (u < nunits) {
                const int fi = u / ngc, gi = u - fi * ngc;
                const size_t idx = ((size_t)s * G + g0 + gi) * F + f0 + fi;
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    float v = acc[u][k].x + acc[u][k].y;
#pragma unroll
                    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
                    if (lane == 0) pbase[(size_t)k * sgf + idx] = v;
                }
            }
        }
    }
}

constexpr int kHalos[4] = {4, 8, 16, 32};

template <int HALO, int TAIL, int UPW>
void launch_t(const GradPlan& plan, const CUtensorMap& map, const TGradParams& p, cudaStream_t stream) {
    static PerDeviceOnce once;                 // the attribute is per (kernel, device)
    if (once.need()) {
        DAU_CUDA_CHECK(cudaFuncSetAttribute(dau_grad_tma_kernel<HALO, TAIL, UPW>,
                                            cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));   // + up to 6.2 KB static <= 227 KB
        once.done();
    }
    dim3 grid(plan.grid_x, plan.grid_y, plan.grid_z);
    dau_grad_tma_kernel<HALO, TAIL, UPW><<<grid, Cfg<UPW>::NTHREADS, plan.smem_bytes, stream>>>(map, p);
}

template <int HALO, int UPW>
void launch_u(const GradPlan& plan, const CUtensorMap& map, const TGradParams& p, cudaStream_t stream) {
    switch (plan.tail) {
        case 0: launch_t<HALO, 0, UPW>(plan, map, p, stream); break;
        case 1: launch_t<HALO, 1, UPW>(plan, map, p, stream); break;
        case 2: launch_t<HALO, 2, UPW>(plan, map, p, stream); break;
        default: launch_t<HALO, 3, UPW>(plan, map, p, stream); break;
    }
}

template <int HALO>
void launch_h(const GradPlan& plan, const CUtensorMap& map, const TGradParams& p, cudaStream_t stream) {
#ifdef DAU_DEV_SWITCHES
    if (plan.upw == 8) { launch_u<HALO, 8>(plan, map, p, stream); return; }
    if (plan.upw == 16) { launch_u<HALO, 16>(plan, map, p, stream); return; }
#endif
    launch_u<HALO, 12>(plan, map, p, stream);
}

template <int HALO>
void fill_t(GradPlan& pl) {
    using GEO = GGeomT<HALO>;
    pl.halo_x = GEO::HX; pl.halo_y = GEO::HY;
    pl.row_stride = GEO::ROW; pl.img_stride = 0;
    pl.feat_stride = (GEO::FEAT_BYTES + 127) / 128 * 128 / 4;    // words
}

}  // namespace

// ---- layouts -------------------------------------------------------------------------------------
StripGeom strip_geometry(const Mosaic& m, int S) {
    StripGeom g;
    g.tiles_x = (m.Wv + TW - 1) / TW;
    g.ky = m.ky; g.py = m.py; g.H = m.H;
    g.q = m.H / R; g.tail = m.H % R;
    g.pairs = (m.Nv + 1) / 2;
    const long long cell_rows = (long long)g.pairs * g.ky * g.tiles_x;
    g.full_items = cell_rows * g.q;
    g.tail_items = g.tail ? cell_rows : 0;
    g.tail_base_floats = (size_t)g.full_items * S * (XK_WARP_BYTES / 4);
    g.total_floats = g.tail_base_floats + (size_t)g.tail_items * S * (g.tail * 4 * XK_ROW_BYTES / 4);
    return g;
}
// Host view of the kernel's own item -> (image pair, first mosaic row, rows, column tile) map (test hook)
void strip_item_origin(const Mosaic& m, long long item, bool tail, int* pair, int* y0, int* rows, int* tile_x) {
    const StripGeom g = strip_geometry(m, 1);
    TGradParams p;
    memset(&p, 0, sizeof(p));
    p.tiles_x = g.tiles_x; p.ky = g.ky; p.py = g.py; p.q = g.q;
    item_origin(p, item, tail, pair, y0, tile_x);
    *rows = tail ? g.tail : R;
}

size_t error_layout_floats(int N, int F, int H, int W) {
    const size_t qv = (W + 1) / 2, qp = (qv + 1) / 2 * 2;
    return (size_t)((N + 1) / 2) * F * H * 2 * qp * 2;
}

bool plan_grad_tma(const Mosaic& m, int S, int F, int G, int max_off, int num_sms, GradPlan* out) {
    int hclass = 0;
    while (kHalos[hclass] < max_off) ++hclass;
    GradPlan pl; memset(&pl, 0, sizeof(pl));
    pl.kind = 1; pl.variant = hclass;
    pl.tile_w = TW; pl.tile_h = TH; pl.imgs = 2;
    switch (hclass) {
        case 0: fill_t<4>(pl); break;
        case 1: fill_t<8>(pl); break;
        case 2: fill_t<16>(pl); break;
        default: fill_t<32>(pl); break;
    }
    // Units per warp and pass: 12 (8 consumer warps of 232 registers).  Measured alternatives, both slower and
    // therefore only in development builds: 16 (same warps, 128 accumulator registers: C2
    // 12.21 -> 12.81 ms, L4 6.05 -> 6.50 ms) and 8 (12 consumer warps of 160 registers).
    int UPW = 12;
#ifdef DAU_DEV_SWITCHES
    if (getenv("DAU_GRAD_UPW")) { const int u = atoi(getenv("DAU_GRAD_UPW")); UPW = (u == 16 || u == 8) ? u : 12; }
#endif
    const int NW = UPW == 8 ? 12 : 8;
    const size_t budget = 220 * 1024 - 128, xk_bytes = (size_t)NW * XK_WARP_BYTES;
    pl.upw = UPW;
    pl.g_chunk = G < UPW ? G : UPW;
    int fc = UPW / pl.g_chunk;
    if (fc > F) fc = F;
    while (fc > 1 && 2 * (size_t)fc * pl.feat_stride * 4 + xk_bytes > budget) --fc;
    if (2 * (size_t)fc * pl.feat_stride * 4 + xk_bytes > budget) return false;
    pl.feats_per_cta = fc;
    pl.units_per_warp = fc * pl.g_chunk;
    const size_t stage = (size_t)fc * pl.feat_stride * 4;     // error tiles of one ring stage
    pl.smem_bytes = 2 * stage + xk_bytes + 128;
    const StripGeom sg = strip_geometry(m, S);
    pl.tail = sg.tail;
    const int n_fchunks = (F + fc - 1) / fc, n_gchunks = (G + pl.g_chunk - 1) / pl.g_chunk;
    pl.grid_x = n_fchunks * n_gchunks;
    pl.grid_y = (S + NW - 1) / NW;
    if (pl.grid_y > 65535) return false;
    // CTAs along the reduction axis: each takes ceil(full / sp) full and ceil(tail / sp) tail items; the cost of
    // a split is counted in strip rows, the score is the fraction of the SM-rounds spent on real rows
    const long base = (long)pl.grid_x * pl.grid_y;
    const double work = (double)sg.full_items * R + (double)sg.tail_items * sg.tail;
    const long long most = sg.full_items > sg.tail_items ? sg.full_items : sg.tail_items;
    int best_split = 1; double best_eff = -1.0;
    const int max_split = (int)(most < 64 ? (most < 1 ? 1 : most) : 64);
    for (int sp = 1; sp <= max_split; ++sp) {
        const double per = (double)((sg.full_items + sp - 1) / sp) * R + (double)((sg.tail_items + sp - 1) / sp) * sg.tail;
        const long waves = (base * sp + num_sms - 1) / num_sms;
        const double eff = (double)base * work / ((double)waves * num_sms * per) - 0.002 * sp;
        if (eff > best_eff + 1e-9) { best_eff = eff; best_split = sp; }
    }
    pl.n_splits = best_split; pl.grid_z = best_split;
    pl.partial_floats = (size_t)best_split * 4 * S * G * F;
    *out = pl;
    return true;
}

void launch_grad_tma(const GradPlan& plan, const float* xk_strips, const float* err_sg, const UnitRec* table,
                     float* partial, const Mosaic& m, int S, int F, int G, cudaStream_t stream) {
    const int N = m.Nv, H = m.Hv, W = m.Wv;
    // tensor map of SGP[n/2][f][y][x%2][x/2][n%2]: dims (q*2+img, plane, y, (pair, f))
    const uint64_t qv = (W + 1) / 2, qp = (qv + 1) / 2 * 2;
    const uint64_t dims[4] = {2 * qv, 2, (uint64_t)H, (uint64_t)((N + 1) / 2) * F};
    const uint64_t strides[3] = {qp * 8, 2 * qp * 8, (uint64_t)H * 2 * qp * 8};
    const uint32_t box[4] = {(uint32_t)(TW + 2 * plan.halo_x), 2, (uint32_t)(TH + 2 * plan.halo_y), 1};
    const CUtensorMap map = make_map_f32(err_sg, 4, dims, strides, box);
    const StripGeom sg = strip_geometry(m, S);
    TGradParams p;
    p.xs = xk_strips; p.table = table; p.partial = partial;
    p.N = N; p.S = S; p.F = F; p.G = G;
    p.FC = plan.feats_per_cta; p.GC = plan.g_chunk; p.n_gchunks = (G + p.GC - 1) / p.GC;
    p.tiles_x = sg.tiles_x; p.ky = sg.ky; p.py = sg.py; p.q = sg.q;
    p.full_items = sg.full_items; p.tail_items = sg.tail_items;
    p.full_per_split = (sg.full_items + plan.n_splits - 1) / plan.n_splits;
    p.tail_per_split = (sg.tail_items + plan.n_splits - 1) / plan.n_splits;
    p.tail_base_floats = (long long)sg.tail_base_floats;
    p.feat_bytes_al = plan.feat_stride * 4;
    p.stage_bytes = plan.feats_per_cta * plan.feat_stride * 4;
    p.debug = 0;
#ifdef DAU_DEV_SWITCHES
    p.debug = getenv("DAU_DEBUG_GRAD") ? atoi(getenv("DAU_DEBUG_GRAD")) : 0;
#endif
    DAU_CHECK(plan.tail == sg.tail, 3, "gradient plan was made for another shape");
    switch (plan.variant) {
        case 0: launch_h<4>(plan, map, p, stream); break;
        case 1: launch_h<8>(plan, map, p, stream); break;
        case 2: launch_h<16>(plan, map, p, stream); break;
        default: launch_h<32>(plan, map, p, stream); break;
    }
    DAU_CUDA_CHECK(cudaGetLastError());
}

}  // namespace dau
