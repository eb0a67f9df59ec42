// DAUConvLayerB200: host-side orchestration of the B200 DAU layer.
//
// Keeps the contract of BaseDAUConvLayer<float> (reference
// include/dau_conv/base_dau_conv_layer.hpp:177-225): LayerSetUp -> Reshape -> Forward_gpu /
// Backward_gpu, the enable_* / set_* toggles and "the caller owns every buffer"
// (:229-264, :413-415).  The virtual buffer accessors of the reference (param_w(),
// temp_interm_buffer(), allocate_workspace_mem() ...) become explicit pointer arguments plus
// ONE caller-provided workspace whose size Reshape reports.
#pragma once
#include "dau_kernels.cuh"
#include "../../include/dau_conv_b200.h"

#include <string>
#include <vector>

namespace dau {

struct TimingSlot { const char* name; cudaEvent_t beg, end; bool used; };

class DAUConvLayerB200 {
public:
    DAUConvLayerB200(const dau_settings_t& settings, int device);
    ~DAUConvLayerB200();

    // (LayerSetUp happens in the constructor: settings validation, G, halo.)
    void Reshape(int N, int S, int H, int W, size_t* fwd_ws, size_t* bwd_ws);

    void Forward_gpu(const float* bottom, float* w, float* mu1, float* mu2, float* sigma,
                     const float* bias, float* top, void* ws, size_t ws_bytes, cudaStream_t stream);

    void Backward_gpu(const float* bottom, const float* top_error, const float* w, const float* mu1,
                      const float* mu2, const float* sigma, float* bottom_error, float* dw,
                      float* dmu1, float* dmu2, float* dsigma, float* dbias, unsigned flags,
                      void* ws, size_t ws_bytes, cudaStream_t stream, const float* top_data = nullptr);

    // Fused epilogue / prologue (the activation the Python layer applies
    // right after the op, dau_conv.py:500-526): with DAU_ACT_RELU Forward_gpu stores
    // max(0, y + bias) and Backward_gpu, given the forward output as top_data, takes the top error
    // as zero wherever that output was not positive.
    void set_activation(int act);
    int activation() const { return activation_; }

    // sigma known to the caller: skips the blocking device read of sigma[0] in every call (the
    // reference reads it once in LayerSetUp, base_dau_conv_layer.hpp:266-275).  Reshape sizes the
    // mosaic gap from it, so set it BEFORE Reshape when sigma is larger than the op default 0.5.
    void set_sigma_hint(float v) { sigma_hint_ = v; }
    // max|mu - centre| known to the caller (set BEFORE Reshape; negative = unknown): the halos, the
    // mosaic gap and the clamp of the integer offsets shrink to the displacement bucket (< 4, 8, 16,
    // 32) that holds it, the per-call choice the reference's op makes from max|mu| (dau_conv_op.cpp:222-256)
    void set_offset_hint(float v) { offset_hint_ = v; }
    // true if any mu was NaN in a forward / backward call since the last query (such units are
    // treated as w = 0); blocks on `stream`.  The TF op fails the call instead (dau_conv_op.cpp:258-262)
    // because it reads max|mu| back on every call anyway; here the read is the caller's choice.
    bool take_nan_flag(cudaStream_t stream);
    // Test hook (dau_debug_unit_tables): the three unit tables the kernels would read for these parameters,
    // written to caller-provided device buffers of S*G*F records (any may be NULL); halos4 = {gather
    // halo x, y, gradient halo x, y} are the amounts added to the integer offsets in the records.
    void debug_unit_tables(const float* w, const float* mu1, const float* mu2, UnitRec* fwd, UnitRec* bwd,
                           UnitRec* grad, int* halos4, cudaStream_t stream);
    // ignore_edge_gradients rule for one output extent (dau_conv_test.py:110-136)
    static bool edge_mask_rule(int extent);
    void enable_timing(bool e);
    int  get_timing(int which, float* ms, int cap);
    static const char* timing_name(int which, int idx);
    long long launches() const { return launches_; }

    int N() const { return N_; } int S() const { return S_; } int F() const { return F_; }
    int G() const { return G_; } int H() const { return H_; } int W() const { return W_; }
    // top / top_error are [N, F, H_out, W_out]
    int H_out() const { return H_ - 2 * crop_; } int W_out() const { return W_ - 2 * crop_; }
    int device() const { return device_; }
    const dau_settings_t& settings() const { return st_; }
    const GatherPlan& fwd_plan() const { return fwd_plan_; }
    const GatherPlan& bwd_plan() const { return bwd_plan_; }
    const GradPlan& grad_plan() const { return grad_plan_; }
    int num_sms() const { return num_sms_; }
    const Mosaic& mosaic() const { return mosaic_; }

private:
    const BlurFactors& factors_for(const float* sigma_dev, cudaStream_t stream);
    // offsets_already_centered == 0: mu counts from the kernel's corner (the Caffe layer's convention)
    float mu_center() const { return st_.offsets_already_centered ? 0.f : (float)(st_.kernel_size / 2); }
    void tic(int which, int idx, cudaStream_t s);
    void toc(int which, int idx, cudaStream_t s);

    dau_settings_t st_;
    int device_, num_sms_;
    int N_ = 0, S_ = 0, H_ = 0, W_ = 0, F_ = 0, G_ = 0, max_off_ = 0, eff_off_ = 0;
    float offset_hint_ = -1.f;
    bool shaped_ = false;
    GatherPlan fwd_plan_{}, bwd_plan_{};
    GradPlan grad_plan_{};
    Mosaic mosaic_{};          // how narrow images are packed for the TMA kernels (identity when wide)
    size_t fwd_ws_ = 0, bwd_ws_ = 0;
    // workspace carving (byte offsets)
    size_t off_fwd_blur_ = 0, off_fwd_table_ = 0, off_fwd_part_ = 0, off_bwd_part_ = 0;
    size_t off_bwd_xk_ = 0, off_bwd_esg_ = 0, off_bwd_eb_ = 0, off_bwd_table_ = 0, off_bwd_gtable_ = 0,
           off_bwd_partial_ = 0, off_bwd_bias_ = 0;
    int activation_ = 0;
    int* nan_flag_ = nullptr;  // device int, owned
    int crop_ = 0;             // (kernel_size-1)/2 - pad
    float sigma_hint_ = 0.f;
    float cached_sigma_ = -1.f;
    BlurFactors factors_{};
    bool timing_ = false;
    std::vector<TimingSlot> slots_[2];
    long long launches_ = 0;
};

}  // namespace dau
