// Host orchestration of the B200 DAU layer (see dau_layer.hpp).  Follows the order of work of
// BaseDAUConvLayer<T>::Forward_gpu / Backward_gpu (reference
// src/dau_conv/base_dau_conv_layer.cu:16-127, 131-363) with the re-layout kernels, memsets,
// side streams, cuBLAS calls and host syncs of the reference removed:
//   forward : [clip] -> blur(X) -> unit table -> gather                         (3-4 launches)
//   backward: blur bank(X) -> unit tables -> gradient kernel -> finalize
//             blur(E, rotated) -> gather with the transposed table  [-> bias]   (6-7 launches)
#include "dau_layer.hpp"

#include <algorithm>
#include <cmath>
#include <cstring>

namespace dau {

namespace {
size_t align_up(size_t v, size_t a = 256) { return (v + a - 1) / a * a; }

// ignore_edge_gradients: zero the last error column/row when the extent is a multiple of the
// reference's patch size (dau_conv_test.py:110-136, base_dau_conv_layer.cpp:763-771,
// dau_conv_backward_core.hpp:2218-2222).
bool edge_rule(int n) {
    const int patches[4] = {64, 32, 16, 8};
    for (int p : patches)
        if (n >= p) return n % p == 0;
    return false;
}

}  // namespace

// Packs narrow images kx x ky to mosaic images so that the 64-pixel-wide tiles of the TMA kernels are
// mostly image and not padding; scores every (kx, ky) by the fraction of computed pixels that
// are real ones (tiles are 64 wide; 32 rows for the gathers, 4 for the gradients) and by how
// evenly the batch divides into pairs of mosaics.
Mosaic plan_mosaic(int N, int H, int W, int gap) {
    Mosaic best; memset(&best, 0, sizeof(best));
    double best_score = -1.0;
    for (int ky = 1; ky <= 16; ++ky) {
        const int Hv = ky * (H + gap) - gap;
        if (ky > 1 && Hv > 256) break;
        for (int kx = 1; kx <= 16; ++kx) {
            const int Wv = kx * (W + gap) - gap;
            if (kx > 1 && Wv > 512) break;
            const int cells = kx * ky;
            if (cells > 1 && 2 * cells > N) continue;          // never pad the batch by more than a pair
            const int Nv = (N + cells - 1) / cells, pairs = (Nv + 1) / 2;
            const double ex = (double)kx * W / (64.0 * ((Wv + 63) / 64));
            // half of the flops run on the gathers' tiles (32, 30 or 28 consecutive rows, whichever wastes the
            // fewest, or -- cells of 13..16 rows stacked in y -- pairs of cell rows: plan_gather_tma), half on
            // the gradient kernel's row strips, which skip the gap rows altogether (strip_geometry)
            int grows = 1 << 30;
            for (int th : {32, 30, 28}) grows = std::min(grows, (Hv + th - 1) / th * th);
            if (ky >= 2 && H >= 13 && H <= 16) grows = (ky + 1) / 2 * 2 * H;
            const double ey = 0.5 * (double)ky * H / grows + 0.5;
            const double eb = (double)N / ((double)pairs * 2 * cells);
            const double score = ex * ey * eb - 1e-3 * cells;
            if (score > best_score + 1e-12) {
                best_score = score;
                best.kx = kx; best.ky = ky; best.px = W + gap; best.py = H + gap;
                best.W = W; best.H = H; best.Wv = Wv; best.Hv = Hv; best.N = N; best.Nv = Nv; best.gap = gap;
            }
        }
    }
    return best;
}

namespace {
const char* kFwdNames[] = {"clip_params", "blur_input", "unit_table", "gather_forward"};
const char* kBwdNames[] = {"blur_bank", "unit_tables", "grad_units", "grad_finalize",
                           "blur_error", "gather_backprop", "bias_grad"};
}  // namespace

DAUConvLayerB200::DAUConvLayerB200(const dau_settings_t& s, int device) : st_(s), device_(device) {
    // LayerSetUp checks (base_dau_conv_layer.cpp:29-187)
    DAU_CHECK(s.num_output > 0, 1, "num_output must be positive");
    DAU_CHECK(s.number_units_x > 0 && s.number_units_y > 0, 1, "number_units must be positive");
    DAU_CHECK(s.stride == 1, 4, "DAUConvolution current supports only stride == 1");   // .cpp:68-69
    DAU_CHECK(s.kernel_size >= 3 && (s.kernel_size & 1) == 1, 1, "kernel_size must be odd and >= 3");
    DAU_CHECK(s.kernel_size / 2 <= 32, 4, "kernel_size above 65 is not supported (dau_conv_op.cpp:250-253)");
    // pad < (ks-1)/2: the output is the centre [H-2crop, W-2crop] of the same-size result
    // (height_out_ = H + 2*pad - ks + 1, base_dau_conv_layer.cpp:20-22; border = H/2 - H_out/2, :544-545)
    crop_ = (s.kernel_size - 1) / 2 - s.pad;
    DAU_CHECK(s.pad >= 0 && crop_ >= 0, 4, "pad must be in [0, (kernel_size-1)/2] (output never larger than the input)");
    G_ = s.number_units_x * s.number_units_y;
    DAU_CHECK(s.number_units_ignore >= 0 && s.number_units_ignore < G_ + 1, 1, "number_units_ignore out of range");
    F_ = s.num_output;
    max_off_ = s.kernel_size / 2;
    if (device_ < 0) DAU_CUDA_CHECK(cudaGetDevice(&device_));   // negative = the caller's current device
    DeviceGuard guard(device_);
    cudaDeviceProp prop;
    DAU_CUDA_CHECK(cudaGetDeviceProperties(&prop, device_));
    DAU_CHECK(prop.major >= 10, 4, "this library is built for sm_100a (B200) only");
    num_sms_ = prop.multiProcessorCount;
    // sticky "a mu was NaN" flag, read and cleared by take_nan_flag()
    DAU_CUDA_CHECK(cudaMalloc(&nan_flag_, sizeof(int)));
    DAU_CUDA_CHECK(cudaMemset(nan_flag_, 0, sizeof(int)));
}

bool DAUConvLayerB200::take_nan_flag(cudaStream_t stream) {
    int v = 0;
    DeviceGuard guard(device_);
    DAU_CUDA_CHECK(cudaMemcpyAsync(&v, nan_flag_, sizeof(int), cudaMemcpyDeviceToHost, stream));
    DAU_CUDA_CHECK(cudaMemsetAsync(nan_flag_, 0, sizeof(int), stream));
    DAU_CUDA_CHECK(cudaStreamSynchronize(stream));
    return v != 0;
}

DAUConvLayerB200::~DAUConvLayerB200() {
    if (nan_flag_) cudaFree(nan_flag_);
    for (int w = 0; w < 2; ++w)
        for (auto& sl : slots_[w]) { cudaEventDestroy(sl.beg); cudaEventDestroy(sl.end); }
}

void DAUConvLayerB200::Reshape(int N, int S, int H, int W, size_t* fwd_ws, size_t* bwd_ws) {
    DAU_CHECK(N > 0 && S > 0 && H > 0 && W > 0, 1, "bottom shape must be a positive 4-D NCHW shape");
    DAU_CHECK(H > 2 * crop_ && W > 2 * crop_, 1, "input smaller than kernel_size - 2*pad: empty output");
    N_ = N; S_ = S; H_ = H; W_ = W;
    const size_t units = (size_t)S * G_ * F_;
    // TMA + FFMA2 kernels on the (mosaic) pair layouts
    const float sig = sigma_hint_ > 0.f ? sigma_hint_ : 0.5f;      // the op's default sigma (dau_conv.py:296-297)
    // displacement bucket: the whole kernel_size/2 unless the caller bounded |mu| (set_offset_hint):
    // the reference's op picks its kernel from max|mu| on every call (dau_conv_op.cpp:222-256,
    // buckets 9 / 17 / 33 / 65 = displacement < 4 / 8 / 16 / 32)
    eff_off_ = max_off_;
    if (offset_hint_ >= 0.f) {
        for (int c : {4, 8, 16, 32})
            if (offset_hint_ < (float)c && c < eff_off_) { eff_off_ = c; break; }
    }
    const int gap = std::max(eff_off_, blur_kernel_size(sig) / 2);
    mosaic_ = plan_mosaic(N, H, W, gap);
    const Mosaic& m = mosaic_;
    DAU_CHECK(plan_gather_tma(m, S, F_, G_, eff_off_, num_sms_, &fwd_plan_), 4, "no forward gather plan for this shape");
    DAU_CHECK(plan_gather_tma(m, F_, S, G_, eff_off_, num_sms_, &bwd_plan_), 4, "no back-propagation gather plan for this shape");
    DAU_CHECK(plan_grad_tma(m, S, F_, G_, eff_off_, num_sms_, &grad_plan_), 4, "no gradient plan for this shape");
    // forward workspace: blurred input | unit table | partial strips of the gather's split tail
    size_t o = 0;
    const size_t blur_in = paired_layout_floats(m.Nv, S, m.Hv, m.Wv);
    const size_t blur_out = paired_layout_floats(m.Nv, F_, m.Hv, m.Wv);
    off_fwd_blur_ = o;  o = align_up(o + blur_in * sizeof(float));
    off_fwd_table_ = o; o = align_up(o + units * sizeof(UnitRec));
    off_fwd_part_ = o;  o = align_up(o + fwd_plan_.part_bytes);
    fwd_ws_ = o;
    // backward workspace: 4 derivative blurs | blurred error | two unit tables | split partials | gather tail strips
    o = 0;
    const size_t xk_floats = strip_geometry(m, S).total_floats;
    const size_t esg_floats = error_layout_floats(m.Nv, F_, m.Hv, m.Wv);
    off_bwd_xk_ = o;      o = align_up(o + xk_floats * sizeof(float));
    off_bwd_esg_ = o;     o = align_up(o + esg_floats * sizeof(float));
    off_bwd_eb_ = o;      o = align_up(o + blur_out * sizeof(float));
    off_bwd_table_ = o;   o = align_up(o + units * sizeof(UnitRec));
    off_bwd_gtable_ = o;  o = align_up(o + units * sizeof(UnitRec));
    off_bwd_partial_ = o; o = align_up(o + grad_plan_.partial_floats * sizeof(float));
    off_bwd_part_ = o;    o = align_up(o + bwd_plan_.part_bytes);
    off_bwd_bias_ = o;    o = align_up(o + (st_.bias_term ? bias_grad_partial_floats(F_) * sizeof(float) : 0));
    bwd_ws_ = o;
    shaped_ = true;
    if (fwd_ws) *fwd_ws = fwd_ws_;
    if (bwd_ws) *bwd_ws = bwd_ws_;
}

const BlurFactors& DAUConvLayerB200::factors_for(const float* sigma_dev, cudaStream_t stream) {
    float sigma = sigma_hint_;
    if (!(sigma > 0.f)) {
        // the reference reads sigma[0] with a blocking copy in LayerSetUp
        // (base_dau_conv_layer.hpp:266-275); same here unless a hint was given
        DAU_CHECK(sigma_dev != nullptr, 1, "sigma pointer is NULL");
        DAU_CUDA_CHECK(cudaMemcpyAsync(&sigma, sigma_dev, sizeof(float), cudaMemcpyDeviceToHost, stream));
        DAU_CUDA_CHECK(cudaStreamSynchronize(stream));
    }
    DAU_CHECK(sigma > 0.f && std::isfinite(sigma), 1, "sigma must be a positive finite number");
    if (sigma != cached_sigma_) {
        FilterOptions opt;
        opt.unit_normalization = st_.unit_normalization != 0;
        opt.square_unit_normalization = st_.square_unit_normalization != 0;
        opt.single_dimension_kernel = st_.single_dimension_kernel != 0;
        opt.forbid_positive_dim1 = st_.forbid_positive_dim1 != 0;
        compute_blur_factors(sigma, opt, &factors_);
        cached_sigma_ = sigma;
    }
    return factors_;
}

bool DAUConvLayerB200::edge_mask_rule(int extent) { return edge_rule(extent); }

void DAUConvLayerB200::debug_unit_tables(const float* w, const float* mu1, const float* mu2, UnitRec* fwd,
                                         UnitRec* bwd, UnitRec* grad, int* halos4, cudaStream_t stream) {
    DAU_CHECK(shaped_, 3, "dau_debug_unit_tables called before dau_reshape");
    DAU_CHECK(w && mu1 && mu2, 1, "NULL tensor pointer");
    DeviceGuard guard(device_);
    UnitPrepArgs ua;
    ua.w = w; ua.mu1 = mu1; ua.mu2 = mu2; ua.S = S_; ua.G = G_; ua.F = F_;
    ua.g_active = G_ - st_.number_units_ignore;
    ua.max_off = eff_off_; ua.use_interpolation = st_.use_interpolation; ua.mu_center = mu_center();
    // exactly the two launches of Forward_gpu / Backward_gpu
    if (fwd) {
        ua.halo_x = fwd_plan_.halo_x; ua.halo_y = fwd_plan_.halo_y; ua.ghalo_x = ua.halo_x; ua.ghalo_y = ua.halo_y;
        launch_unit_prep(ua, fwd, nullptr, nullptr, nan_flag_, stream);
    }
    if (bwd || grad) {
        ua.halo_x = bwd_plan_.halo_x; ua.halo_y = bwd_plan_.halo_y;
        ua.ghalo_x = grad_plan_.halo_x; ua.ghalo_y = grad_plan_.halo_y;
        launch_unit_prep(ua, nullptr, bwd, grad, nan_flag_, stream);
    }
    if (halos4) {
        halos4[0] = fwd_plan_.halo_x; halos4[1] = fwd_plan_.halo_y;
        halos4[2] = grad_plan_.halo_x; halos4[3] = grad_plan_.halo_y;
    }
}

void DAUConvLayerB200::set_activation(int act) {
    DAU_CHECK(act == DAU_ACT_NONE || act == DAU_ACT_RELU, 1, "activation must be DAU_ACT_NONE or DAU_ACT_RELU");
    activation_ = act;
}

void DAUConvLayerB200::enable_timing(bool e) {
    timing_ = e;
    if (e && slots_[0].empty()) {
        for (int w = 0; w < 2; ++w) {
            const int n = w == 0 ? 4 : 7;
            for (int i = 0; i < n; ++i) {
                TimingSlot sl; sl.name = w == 0 ? kFwdNames[i] : kBwdNames[i]; sl.used = false;
                DAU_CUDA_CHECK(cudaEventCreate(&sl.beg));
                DAU_CUDA_CHECK(cudaEventCreate(&sl.end));
                slots_[w].push_back(sl);
            }
        }
    }
}

const char* DAUConvLayerB200::timing_name(int which, int idx) {
    if (which == 0 && idx >= 0 && idx < 4) return kFwdNames[idx];
    if (which == 1 && idx >= 0 && idx < 7) return kBwdNames[idx];
    return nullptr;
}

int DAUConvLayerB200::get_timing(int which, float* ms, int cap) {
    if (which < 0 || which > 1 || slots_[which].empty()) return 0;
    int n = 0;
    for (auto& sl : slots_[which]) {
        if (n >= cap) break;
        float t = 0.f;
        if (sl.used) { cudaEventSynchronize(sl.end); cudaEventElapsedTime(&t, sl.beg, sl.end); }
        ms[n++] = t;
    }
    return n;
}

void DAUConvLayerB200::tic(int which, int idx, cudaStream_t s) {
    ++launches_;
    if (timing_) { slots_[which][idx].used = true; cudaEventRecord(slots_[which][idx].beg, s); }
}
void DAUConvLayerB200::toc(int which, int idx, cudaStream_t s) {
    if (timing_) cudaEventRecord(slots_[which][idx].end, s);
}

void DAUConvLayerB200::Forward_gpu(const float* bottom, float* w, float* mu1, float* mu2, float* sigma,
                                   const float* bias, float* top, void* ws, size_t ws_bytes,
                                   cudaStream_t stream) {
    DAU_CHECK(shaped_, 3, "Forward_gpu called before Reshape");
    DAU_CHECK(bottom && w && mu1 && mu2 && top, 1, "NULL tensor pointer");
    DAU_CHECK(ws != nullptr && ws_bytes >= fwd_ws_, 3, "forward workspace missing or too small");
    DAU_CHECK(!st_.bias_term || bias, 1, "bias_term is set but bias pointer is NULL");
    DeviceGuard guard(device_);
    if (timing_) for (auto& sl : slots_[0]) sl.used = false;
    char* base = static_cast<char*>(ws);
    float* xb = reinterpret_cast<float*>(base + off_fwd_blur_);
    UnitRec* table = reinterpret_cast<UnitRec*>(base + off_fwd_table_);
    const int units = S_ * G_ * F_;

    if (st_.unit_bounds_guard) {                      // base_dau_conv_layer.cu:33-49
        tic(0, 0, stream);
        // centred offsets live in [-ks/2 + bound, ks/2 - bound], corner-based ones in [bound, ks-1-bound]
        const float lim = (float)max_off_ - st_.component_border_bound, c = mu_center();
        launch_clip_params(mu1, mu2, sigma, units, c - lim, c + lim, st_.sigma_lower_bound, stream);
        toc(0, 0, stream);
    }
    const BlurFactors& f = factors_for(sigma, stream);

    tic(0, 1, stream);
    launch_blur_single_paired(bottom, xb, mosaic_, S_, f.kb, f.fwd_x, f.fwd_y, stream);
    toc(0, 1, stream);

    tic(0, 2, stream);
    UnitPrepArgs ua;
    ua.w = w; ua.mu1 = mu1; ua.mu2 = mu2; ua.S = S_; ua.G = G_; ua.F = F_;
    ua.g_active = G_ - st_.number_units_ignore;
    ua.halo_x = fwd_plan_.halo_x; ua.halo_y = fwd_plan_.halo_y; ua.max_off = eff_off_;
    ua.ghalo_x = ua.halo_x; ua.ghalo_y = ua.halo_y;
    ua.use_interpolation = st_.use_interpolation; ua.mu_center = mu_center();
    launch_unit_prep(ua, table, nullptr, nullptr, nan_flag_, stream);
    toc(0, 2, stream);

    tic(0, 3, stream);
    launch_gather_tma(fwd_plan_, xb, table, st_.bias_term ? bias : nullptr, top, mosaic_, S_, F_, G_, stream,
                      activation_ == DAU_ACT_RELU ? 1 : 0, crop_, reinterpret_cast<float*>(base + off_fwd_part_));
    if (fwd_plan_.split_k > 1) ++launches_;           // the kernel that sums the split tail
    toc(0, 3, stream);
}

void DAUConvLayerB200::Backward_gpu(const float* bottom, const float* top_error, const float* w,
                                    const float* mu1, const float* mu2, const float* sigma,
                                    float* bottom_error, float* dw, float* dmu1, float* dmu2,
                                    float* dsigma, float* dbias, unsigned flags, void* ws,
                                    size_t ws_bytes, cudaStream_t stream, const float* top_data) {
    DAU_CHECK(shaped_, 3, "Backward_gpu called before Reshape");
    DAU_CHECK(bottom && top_error && w && mu1 && mu2, 1, "NULL tensor pointer");
    DAU_CHECK(activation_ == DAU_ACT_NONE || top_data, 1,
              "an activation is fused into this layer: the backward pass needs top_data (dau_backward_act)");
    // ReLU backward fused into every load of the top error (row f3): error := 0 where top_data <= 0
    const float* act = activation_ == DAU_ACT_RELU ? top_data : nullptr;
    DAU_CHECK(ws != nullptr && ws_bytes >= bwd_ws_, 3, "backward workspace missing or too small");
    DeviceGuard guard(device_);
    if (timing_) for (auto& sl : slots_[1]) sl.used = false;
    char* base = static_cast<char*>(ws);
    float* xk = reinterpret_cast<float*>(base + off_bwd_xk_);
    float* eb = reinterpret_cast<float*>(base + off_bwd_eb_);
    float* esg = reinterpret_cast<float*>(base + off_bwd_esg_);
    UnitRec* btable = reinterpret_cast<UnitRec*>(base + off_bwd_table_);
    UnitRec* gtable = reinterpret_cast<UnitRec*>(base + off_bwd_gtable_);
    float* partial = reinterpret_cast<float*>(base + off_bwd_partial_);
    const BlurFactors& f = factors_for(sigma, stream);

    const bool want_params = (flags & (DAU_GRAD_W | DAU_GRAD_MU1 | DAU_GRAD_MU2 | DAU_GRAD_SIGMA)) != 0;
    const bool want_input = (flags & DAU_GRAD_INPUT) != 0;
    if (want_input) DAU_CHECK(bottom_error != nullptr, 1, "bottom_error is NULL but DAU_GRAD_INPUT is set");

    UnitPrepArgs ua;
    ua.w = w; ua.mu1 = mu1; ua.mu2 = mu2; ua.S = S_; ua.G = G_; ua.F = F_;
    ua.g_active = G_ - st_.number_units_ignore;
    ua.max_off = eff_off_; ua.use_interpolation = st_.use_interpolation; ua.mu_center = mu_center();

    if (want_params) {
        tic(1, 0, stream);
        launch_blur_bank_strips(bottom, xk, mosaic_, S_, f, stream);
        toc(1, 0, stream);
    }
    {
        // both tables in one launch
        tic(1, 1, stream);
        ua.halo_x = bwd_plan_.halo_x; ua.halo_y = bwd_plan_.halo_y;
        ua.ghalo_x = grad_plan_.halo_x; ua.ghalo_y = grad_plan_.halo_y;
        launch_unit_prep(ua, nullptr, want_input ? btable : nullptr, want_params ? gtable : nullptr, nan_flag_, stream);
        toc(1, 1, stream);
    }
    const int mrow = st_.ignore_edge_gradients && edge_rule(H_out()) ? 1 : 0;
    const int mcol = st_.ignore_edge_gradients && edge_rule(W_out()) ? 1 : 0;
    if (want_input) {
        // blur the error with the rotated filter (base_dau_conv_layer.cu:278-288); the same kernel
        // hands the un-blurred tile to the gradient kernel in its own layout, so the error is read
        // from HBM once
        tic(1, 4, stream);
        ErrorSideArgs ex;
        ex.mask = act; ex.sgp = want_params ? esg : nullptr; ex.edge_row = mrow; ex.edge_col = mcol; ex.crop = crop_;
        launch_blur_single_paired(top_error, eb, mosaic_, F_, f.kb, f.err_x, f.err_y, stream, &ex);
        toc(1, 4, stream);
    }
    if (want_params) {
        tic(1, 2, stream);
        if (!want_input) launch_error_relayout(top_error, esg, mosaic_, F_, mrow, mcol, stream, act, crop_);
        launch_grad_tma(grad_plan_, xk, esg, gtable, partial, mosaic_, S_, F_, G_, stream);
        toc(1, 2, stream);
        tic(1, 3, stream);
        GradFinalizeArgs fa;
        fa.partial = partial; fa.n_splits = grad_plan_.n_splits; fa.w = w;
        fa.dw = (flags & DAU_GRAD_W) ? dw : nullptr;
        fa.dmu1 = (flags & DAU_GRAD_MU1) ? dmu1 : nullptr;
        fa.dmu2 = (flags & DAU_GRAD_MU2) ? dmu2 : nullptr;
        fa.dsigma = ((flags & DAU_GRAD_SIGMA) && st_.learn_sigma) ? dsigma : nullptr;
        fa.S = S_; fa.G = G_; fa.F = F_; fa.g_active = G_ - st_.number_units_ignore;
        fa.mu_lr = st_.mu_learning_rate_factor;
        launch_grad_finalize(fa, stream);
        toc(1, 3, stream);
    }
    if (want_input) {
        // gather with -mu and the parameters read transposed (base_dau_conv_layer.cu:290-325)
        tic(1, 5, stream);
        launch_gather_tma(bwd_plan_, eb, btable, nullptr, bottom_error, mosaic_, F_, S_, G_, stream, 0, 0,
                          reinterpret_cast<float*>(base + off_bwd_part_));
        if (bwd_plan_.split_k > 1) ++launches_;
        toc(1, 5, stream);
    }
    if ((flags & DAU_GRAD_BIAS) && st_.bias_term && dbias) {
        tic(1, 6, stream);
        launch_bias_grad(top_error, act, dbias, reinterpret_cast<float*>(base + off_bwd_bias_), N_, F_,
                         H_out() * W_out(), stream);
        ++launches_;                                      // the second stage
        toc(1, 6, stream);
    }
}

}  // namespace dau
