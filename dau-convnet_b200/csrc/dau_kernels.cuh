// Device-side contracts shared by the DAU kernels and their host launchers.
#pragma once
#include "dau_common.hpp"

#include <cstring>

namespace dau {

// ---- unit table -----------------------------------------------------------------------
// One record per DAU unit and per use (forward / dL/dx / parameter gradients), 32 bytes so
// a warp fetches it with two broadcast LDS.128.  A record describes a displaced 2x2-tap
// gather:   out[y][x] (+)= sum_{dy,dx in {0,1}} wt[dy][dx] * in[y + offy + dy][x + offx + dx]
// with (offx, offy) stored relative to the top-left corner of the haloed tile, i.e.
// offx_h = offx + halo_x >= 0, offy_h = offy + halo_y >= 0.
struct __align__(16) UnitRec {
    float w00, w01, w10, w11;   // wt[dy][dx]
    int offx_h, offy_h;         // integer displacement + halo
    int pad0, pad1;
};
static_assert(sizeof(UnitRec) == 32, "UnitRec must be 32 bytes");

enum UnitTableMode {
    UNITS_FORWARD = 0,   // table[s][f][g]: off=floor(mu),    wt = w * bilinear(frac)
    UNITS_BACKPROP = 1,  // table[f][s][g]: off=floor(-mu),   wt = w * bilinear(frac(-mu))
    UNITS_GRADIENT = 2   // table[s][f][g]: off=-floor(mu)-1, wt = bilinear flipped, no w
};

struct UnitPrepArgs {
    const float* w; const float* mu1; const float* mu2;   // [1,S,G,F]
    int S, G, F;
    int g_active;            // G - number_units_ignore
    int halo_x, halo_y;      // halo of the forward / back-propagation tables (x halo is rounded up by the kernel)
    int ghalo_x, ghalo_y;    // halo of the gradient table (its kernel may round differently)
    int max_off;             // kernel_size/2: integer offsets are clamped to [-max_off, max_off-1]
    int use_interpolation;
    float mu_center = 0.f;   // kernel_size/2 when offsets_already_centered == 0, else 0
};

void launch_unit_prep(const UnitPrepArgs& a, UnitRec* fwd, UnitRec* bwd, UnitRec* grad,
                      int* nan_flag, cudaStream_t stream);

// in-place parameter clipping (unit_bounds_guard, base_dau_conv_layer.cu:33-49)
void launch_clip_params(float* mu1, float* mu2, float* sigma, int n, float mu_lo, float mu_hi,
                        float sigma_lo, cudaStream_t stream);

// ---- image mosaic -------------------------------------------------------------------------
// The TMA kernels work on 64-pixel-wide tiles.  Narrow images (AlexNet's 27x27 / 13x13 maps, the
// 28x28 stress layer) are therefore packed kx x ky to a "mosaic" image, separated by `gap` zero
// pixels (gap >= the largest displacement and >= the blur half-width), so that neither the blur
// nor a displaced gather of one image can reach its neighbour: the zero gap IS the zero padding
// the reference applies around every image.  Only the layout producers (blur, error re-layout)
// and the gather's store know about it; wide images use the identity mosaic (kx = ky = 1).
struct Mosaic {
    int kx, ky;          // real images per mosaic row / column
    int px, py;          // pitch: image extent + gap
    int W, H;            // real image extent
    int Wv, Hv;          // mosaic extent = k * pitch - gap
    int N, Nv;           // real images, mosaic images
    int gap;
};
Mosaic plan_mosaic(int N, int H, int W, int gap);

// mosaic pixel -> real pixel; false for gap pixels and for cells beyond the batch
__host__ __device__ inline bool mosaic_to_real(const Mosaic& m, int nv, int vy, int vx, int* n, int* y, int* x) {
    const int jy = vy / m.py, jx = vx / m.px;
    const int ly = vy - jy * m.py, lx = vx - jx * m.px;
    const int nr = (nv * m.ky + jy) * m.kx + jx;
    *n = nr; *y = ly; *x = lx;
    return ly < m.H && lx < m.W && jy < m.ky && jx < m.kx && nr < m.N;
}

// ---- displaced gather (forward and dL/dx) -------------------------------------------------
struct GatherPlan {            // chosen once per shape by plan_gather_tma()
    int kind;                  // 1: TMA + FFMA2 kernel on the paired gather layout (dau_gather_tma.cu)
    int nstage;                // pipeline depth
    int variant;               // index into the instantiated tile shapes
    int tile_w, tile_h, imgs;  // output tile of one CTA: imgs x tile_h x tile_w
    int ly_stride, tile_pitch_y;   // rows between the two strips of a tile / between tiles (see dau_gather_tma.cu)
    int tile_rows, tile_bytes;     // haloed input tile in shared memory
    int halo_x, halo_y;
    int row_stride, img_stride;// shared-memory strides in words (bank-conflict padded)
    int stage_words;           // words per pipeline stage (tile + unit records)
    int cout_per_cta;
    size_t smem_bytes;
    int grid_x, grid_y, grid_z;// items: tiles, groups of cout_per_cta output channels, image pairs
    int full_items, split_k;   // items >= full_items (the last partial wave) run as split_k CTAs each
    size_t part_bytes;         // workspace for their partial strips
    double rounds;             // CTA rounds on this GPU (1 CTA per SM), for the host-step model
};
// out[n,co,y,x] = sum_{ci,g} gather(in[n,ci], table[ci][co][g]) (+ bias[co])

// TMA variant: `in_paired` is the blurred input in the paired gather layout
//   PG[n/2][c][y][x%4][x/4][n%2]   (x/4 padded to an even count; written by launch_blur_single_paired)
// (N, H, W below are the MOSAIC extents m.Nv, m.Hv, m.Wv; `in` / `out` of the producers and of
// the gather's store are the caller's real NCHW tensors)
size_t paired_layout_floats(int N, int Cn, int H, int W);
bool plan_gather_tma(const Mosaic& m, int Cin, int Cout, int G, int max_off, int num_sms, GatherPlan* out);
void launch_gather_tma(const GatherPlan& p, const float* in_paired, const UnitRec* table, const float* bias,
                       float* out, const Mosaic& m, int Cin, int Cout, int G, cudaStream_t stream, int relu = 0,
                       int crop = 0, float* part_buf = nullptr);
// Extras of the top-error blur (the prologue fused into the error load)
struct ErrorSideArgs {
    const float* mask = nullptr;   // activation output, same shape as the error: error := 0 where mask <= 0
    float* sgp = nullptr;          // also write the (masked) un-blurred error in the SGP layout of the
                                   // gradient kernel, i.e. what launch_error_relayout produces
    int edge_row = 0, edge_col = 0;// ignore_edge_gradients mask of the SGP copy
    int crop = 0;                  // the error (and mask) tensors are [.., H-2crop, W-2crop] and sit at offset
                                   // (crop, crop) of the full grid, zeros around (base_dau_conv_layer.cu:258-266)
};
void launch_blur_single_paired(const float* in, float* out_paired, const Mosaic& m, int Cn, int kb,
                               const float* fx, const float* fy, cudaStream_t stream,
                               const ErrorSideArgs* extra = nullptr);

// ---- parameter gradients -----------------------------------------------------------------
struct GradPlan {
    int kind;                  // 1: TMA + FFMA2 kernel (dau_grad_tma.cu)
    int variant;
    int tile_w, tile_h, imgs;
    int halo_x, halo_y;
    int row_stride, feat_stride, img_stride;
    int feats_per_cta;         // FC: features whose error tiles share one stage
    int units_per_warp;        // FC * G  (<= 16)
    int g_chunk;               // units of one (s,f) handled per pass (G or 16)
    int n_splits;              // CTAs along the (image,tile) reduction axis
    int tail;                  // rows of the tail strips (H % 4): selects the kernel instantiation
    int upw;                   // accumulator slots per warp (12 or 16): selects the kernel instantiation
    size_t smem_bytes;
    int grid_x, grid_y, grid_z;
    size_t partial_floats;     // n_splits * 4 * S*G*F
};
// partial[split][k][s][g][f] = sum over the split's (n,y,x) of  xk[n,s,k] * gather(err[n,f])

// Row strips of the gradient kernel.  Every cell row of the mosaic (m.ky of them; 1 for wide images) is
// cut into q = H / 4 full strips of 4 rows and, when H % 4 != 0, ONE tail strip of tail = H % 4 rows;
// the gap rows between cells belong to no strip, so no lane of the gradient kernel ever sits on a gap
// ROW (13-row maps: 13 of 13 rows are work instead of 26 of the 32 rows of eight 4-row mosaic tiles).
// A work item is one strip x one 64-column tile x one image pair.  XS, the strip-major Xk layout the
// derivative-blur kernel writes and the gradient kernel copies with one cp.async.bulk per item:
//     full items: XS[item][s][k<4][r<4][lane<32][c<2][img<2],  item = ((pair*ky + jy)*q + t)*tiles_x + tx
//     tail items: XS[tail_base + ...][titem][s][k<4][r<tail][lane][c][img], titem = (pair*ky + jy)*tiles_x + tx
struct StripGeom {
    int tiles_x, ky, py, H, q, tail, pairs;
    long long full_items, tail_items;
    size_t tail_base_floats;         // where the tail blocks start in XS
    size_t total_floats;
};
StripGeom strip_geometry(const Mosaic& m, int S);
void strip_item_origin(const Mosaic& m, long long item, bool tail, int* pair, int* y0, int* rows, int* tile_x);

// TMA variant: Xk in the strip-major layout XS (launch_blur_bank_strips), the error in the de-interleaved
// pair layout SGP[n/2][f][y][x%2][x/2][n%2] with the edge mask applied (launch_error_relayout)
size_t error_layout_floats(int N, int F, int H, int W);
bool plan_grad_tma(const Mosaic& m, int S, int F, int G, int max_off, int num_sms, GradPlan* out);
void launch_grad_tma(const GradPlan& p, const float* xk_strips, const float* err_sg, const UnitRec* table,
                     float* partial, const Mosaic& m, int S, int F, int G, cudaStream_t stream);
void launch_blur_bank_strips(const float* in, float* out_strips, const Mosaic& m, int S,
                             const BlurFactors& f, cudaStream_t stream);
void launch_error_relayout(const float* err, float* out_sg, const Mosaic& m, int F, int mask_row, int mask_col,
                           cudaStream_t stream, const float* act_mask = nullptr, int crop = 0);

struct GradFinalizeArgs {
    const float* partial; int n_splits;
    const float* w;
    float* dw; float* dmu1; float* dmu2; float* dsigma;   // any may be NULL
    int S, G, F, g_active;
    float mu_lr;
};
void launch_grad_finalize(const GradFinalizeArgs& a, cudaStream_t stream);

// bias gradient: dbias[f] += sum_{n,y,x} err[n,f,y,x]; `partial` = bias_grad_partial_floats(F) floats of
// workspace (two-stage deterministic sum, no atomics)
size_t bias_grad_partial_floats(int F);
void launch_bias_grad(const float* err, const float* act, float* dbias, float* partial, int N, int F, int HW,
                      cudaStream_t stream);

// the same with the partials in a static device array of this library (no workspace, nothing allocated);
// F <= static_bias_grad_max_features(); the CURRENT device is used; calls that overlap on one GPU race
int static_bias_grad_max_features();
void launch_bias_grad_static(const float* err, const float* act, float* dbias, int N, int F, int HW,
                             cudaStream_t stream);

// FP32 FMA peak probe (register-resident FFMA chains on every SM)
void launch_fma_probe(float* sink, int iters, int blocks, cudaStream_t stream);

}  // namespace dau
