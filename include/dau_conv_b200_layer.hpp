/*
 * dau_conv_b200_layer.hpp -- header-only C++ face of libdau_conv_b200.so for hosts that are
 * written against the reference's BaseDAUConvLayer<float> (Caffe-style layers, the TensorFlow
 * op's DAUConvLayerTensorflowGPU).  It needs nothing but dau_conv_b200.h: no CUDA, cuBLAS or
 * torch headers (streams travel as void*).
 *
 * Reference interface kept (paths relative to the reference checkout):
 *   DAUConvSettings                 include/dau_conv/base_dau_conv_layer.hpp:109-130   (same fields)
 *   BaseDAUConvLayer ctor flags     include/dau_conv/base_dau_conv_layer.hpp:177-184
 *   LayerSetUp / Reshape            include/dau_conv/base_dau_conv_layer.hpp:190-198
 *   Forward_gpu / Backward_gpu      include/dau_conv/base_dau_conv_layer.hpp:205-208
 *   enable_* / set_* toggles        include/dau_conv/base_dau_conv_layer.hpp:210-225
 *   param_* / param_*_grad /
 *   allocate_workspace_mem virtuals include/dau_conv/base_dau_conv_layer.hpp:229-264
 *   DAUException                    include/dau_conv/util/common.hpp:38-64
 *
 * What is NOT here, on purpose: the temp_* buffer virtuals (temp_bwd_gradients, temp_interm_buffer,
 * temp_param_buffer, temp_col_buffer, temp_bias_multiplier :247-261) -- the library needs ONE
 * workspace, asked from allocate_workspace_mem(); the kernel_compute / kernel_param / kernel_output
 * objects and the initializer of LayerSetUp (:190-195) -- filters are derived inside the library and
 * parameter initialisation stays with the host framework; Forward_cpu / Backward_cpu (there is no
 * CPU path behind this library).  A subclass written for the reference keeps its param_* overrides
 * and its workspace allocator and drops the rest.
 */
#ifndef DAU_CONV_B200_LAYER_HPP_
#define DAU_CONV_B200_LAYER_HPP_

#include <algorithm>
#include <stdexcept>
#include <string>
#include <vector>

#include "dau_conv_b200.h"

namespace dau_b200 {

/* common.hpp:38-64: every failed check of the reference throws this */
class DAUException : public std::runtime_error {
 public:
    DAUException(dau_status_t status, const std::string& what)
        : std::runtime_error(what), status_(status) {}
    dau_status_t status() const { return status_; }
 private:
    dau_status_t status_;
};

/* base_dau_conv_layer.hpp:109-130, field for field */
struct DAUConvSettings {
    int num_output = 0;
    std::vector<int> number_units;          /* {units_x} or {units_x, units_y} (:46-48 of the .cpp) */
    bool bias_term = false;
    int pad = 4, kernel_size = 9, stride = 1;
    bool unit_normalization = true, square_unit_normalization = false;
    int mean_iteration_step = 1, sigma_iteration_step = 1;
    float component_border_bound = 0.01f, sigma_lower_bound = 0.3f;
    int merge_iteration_step = 0;
    float merge_threshold = 1.0f;
    bool offsets_already_centered = true;
    bool use_interpolation = true;
};

class BaseDAUConvLayer {
 public:
    /* same flags, same defaults as base_dau_conv_layer.hpp:177-184; the cuBLAS handle the reference
     * borrows there is not needed (accepted and ignored so that call sites compile unchanged) */
    explicit BaseDAUConvLayer(void* /*cublas_handle*/ = nullptr, bool ignore_edge_gradients = false,
                              bool offsets_already_centered = true, bool dynamic_kernel_size = true)
        : ignore_edge_gradients_(ignore_edge_gradients),
          offsets_already_centered_(offsets_already_centered),
          dynamic_kernel_size_(dynamic_kernel_size) {}

    virtual ~BaseDAUConvLayer() {
        if (handle_) dau_destroy(handle_);
    }
    BaseDAUConvLayer(const BaseDAUConvLayer&) = delete;
    BaseDAUConvLayer& operator=(const BaseDAUConvLayer&) = delete;

    /* base_dau_conv_layer.cpp:29-187.  number_units_ignore / mu_learning_rate_factor are the two
     * extras the TF subclass threads through (dau_conv_layer_tensorflow.hpp LayerSetUp overload,
     * dau_conv_grad_op.cpp:296-303).  The handle is (re)built lazily in Reshape, so the toggles
     * below may be called before or after LayerSetUp, as the TF op does (dau_conv_op.cpp:270-290). */
    virtual void LayerSetUp(const DAUConvSettings& settings, const std::vector<int>& bottom_shape,
                            int number_units_ignore = 0, bool in_train = true,
                            float mu_learning_rate_factor = 1.0f, int device = -1) {
        check_shape(bottom_shape, "bottom_shape");
        if (settings.number_units.empty() || settings.number_units.size() > 2)
            throw DAUException(DAU_ERR_INVALID_ARGUMENT, "number_units must hold one or two values");
        settings_ = settings;
        number_units_ignore_ = number_units_ignore;
        mu_learning_rate_factor_ = mu_learning_rate_factor;
        device_ = device;
        in_train_ = in_train;
        dirty_ = true;
    }

    /* base_dau_conv_layer.cpp:190-394: returns the top shape, sizes and asks for the workspace */
    virtual std::vector<int> Reshape(const std::vector<int>& bottom_shape,
                                     const std::vector<int>& /*top_shape*/) {
        check_shape(bottom_shape, "bottom_shape");
        if (dirty_) rebuild();
        if (bottom_shape != bottom_shape_ || !reshaped_) {
            size_t fwd = 0, bwd = 0;
            check(dau_reshape(handle_, bottom_shape[0], bottom_shape[1], bottom_shape[2],
                              bottom_shape[3], &fwd, &bwd));
            /* the reference sizes one buffer for max(forward, backward) (:302-321) */
            size_t need = enabled_bwd_op_ && in_train_ ? std::max(fwd, bwd) : fwd;
            if (need > workspace_bytes_ || workspace_ == nullptr) {
                if (workspace_) deallocate_workspace_mem();
                workspace_ = allocate_workspace_mem(need);
                workspace_bytes_ = need;
                if (workspace_ == nullptr && need > 0)
                    throw DAUException(DAU_ERR_NOT_READY, "allocate_workspace_mem returned NULL");
            }
            bottom_shape_ = bottom_shape;
            reshaped_ = true;
        }
        int ho = 0, wo = 0;
        check(dau_get_output_size(handle_, &ho, &wo));
        return {bottom_shape[0], settings_.num_output, ho, wo};
    }

    /* base_dau_conv_layer.cu:16-127 */
    virtual void Forward_gpu(const float* bottom_data, const std::vector<int>& bottom_shape,
                             float* top_data, const std::vector<int>& top_shape) {
        if (!enabled_fwd_op_) return;
        require_ready(bottom_shape);
        check_top(top_shape);
        check(dau_forward(handle_, bottom_data, param_w(), param_mu1(), param_mu2(), param_sigma(),
                          settings_.bias_term ? param_bias() : nullptr, top_data,
                          workspace_, workspace_bytes_, stream_));
    }

    /* base_dau_conv_layer.cu:131-363; params_propagate_down = [w, mu1, mu2, sigma, bias] (:198-206).
     * top_data is unused by the reference (:293 passes NULL) and by this call. */
    virtual void Backward_gpu(const float* /*top_data*/, const float* top_error,
                              const std::vector<int>& top_shape, bool propagate_down,
                              const float* bottom_data, float* bottom_error,
                              const std::vector<int>& bottom_shape,
                              const std::vector<bool>& params_propagate_down) {
        if (!enabled_bwd_op_) return;
        require_ready(bottom_shape);
        check_top(top_shape);
        if (params_propagate_down.size() < 4)
            throw DAUException(DAU_ERR_INVALID_ARGUMENT, "params_propagate_down needs [w,mu1,mu2,sigma(,bias)]");
        unsigned flags = propagate_down ? (unsigned)DAU_GRAD_INPUT : 0u;
        /* the reference runs its gradient kernel for all k or not at all, keyed on flag[0] (:198) */
        if (params_propagate_down[0]) flags |= DAU_GRAD_W;
        if (params_propagate_down[0] && params_propagate_down[1]) flags |= DAU_GRAD_MU1;
        if (params_propagate_down[0] && params_propagate_down[2]) flags |= DAU_GRAD_MU2;
        if (params_propagate_down[0] && params_propagate_down[3] && learn_sigma_) flags |= DAU_GRAD_SIGMA;
        if (settings_.bias_term && params_propagate_down.size() > 4 && params_propagate_down[4])
            flags |= DAU_GRAD_BIAS;
        check(dau_backward(handle_, bottom_data, top_error, param_w(), param_mu1(), param_mu2(),
                           param_sigma(), propagate_down ? bottom_error : nullptr,
                           param_w_grad(), param_mu1_grad(), param_mu2_grad(), param_sigma_grad(),
                           (flags & DAU_GRAD_BIAS) ? param_bias_grad() : nullptr, flags,
                           workspace_, workspace_bytes_, stream_));
    }

    /* toggles, base_dau_conv_layer.hpp:210-225 */
    void enable_forward(bool enable) { enabled_fwd_op_ = enable; }
    void enable_backward(bool enable) { enabled_bwd_op_ = enable; }
    void enable_memalloc_info(bool) {}                       /* the reference only prints with it */
    void enable_unit_bounds_guard(bool enable) { set(unit_bounds_guard_, enable); }
    void enable_sigma_learning(bool enable) { set(learn_sigma_, enable); }
    void set_single_dimensional_kernel(bool v) { set(single_dimension_kernel_, v); }
    void set_forbid_positive_dim1(bool v) { set(forbid_positive_dim1_, v); }
    void set_default_cuda_stream(void* cuda_stream) { stream_ = cuda_stream; }

    /* Extras of this library (no counterpart in the reference class): */
    /* sigma is one number per layer (:150); telling it spares the 4-byte read-back of get_sigma_val (:266-275) */
    void set_sigma_hint(float sigma) { sigma_hint_ = sigma; if (handle_) check(dau_set_sigma_hint(handle_, sigma)); }
    /* the bound the TF op reads back with caffe_gpu_amax on every call (dau_conv_op.cpp:222-256) */
    /* honoured like the reference's own bucket choice: only with centred offsets and dynamic_kernel_size (base_dau_conv_layer.cu:100) */
    void set_max_offset_hint(float max_abs_mu) {
        offset_hint_ = (offsets_already_centered_ && dynamic_kernel_size_) ? max_abs_mu : -1.0f;
        reshaped_ = false;
        if (handle_) check(dau_set_offset_hint(handle_, offset_hint_));
    }
    /* dau_conv_op.cpp:258-262: true when a NaN offset was seen since the last query */
    bool take_nan_flag() { int f = 0; if (handle_) check(dau_take_nan_flag(handle_, &f, stream_)); return f != 0; }

    dau_handle_t* handle() { return handle_; }
    size_t workspace_bytes() const { return workspace_bytes_; }

 protected:
    /* learnable parameters [1, S, G, F] on the device (base_dau_conv_layer.hpp:233-238) */
    virtual float* param_w() = 0;
    virtual float* param_mu1() = 0;
    virtual float* param_mu2() = 0;
    virtual float* param_sigma() = 0;
    virtual float* param_bias() = 0;
    /* gradient buffers, accumulated into (:241-245) */
    virtual float* param_w_grad() = 0;
    virtual float* param_mu1_grad() = 0;
    virtual float* param_mu2_grad() = 0;
    virtual float* param_sigma_grad() = 0;
    virtual float* param_bias_grad() = 0;
    /* the host framework owns the workspace (:263-264, :413-415) */
    virtual void* allocate_workspace_mem(size_t bytes) = 0;
    virtual void deallocate_workspace_mem() = 0;

    static void check(dau_status_t st) {
        if (st != DAU_OK) throw DAUException(st, dau_last_error());
    }

 private:
    void set(bool& field, bool v) { if (field != v) { field = v; dirty_ = true; } }

    static void check_shape(const std::vector<int>& s, const char* what) {
        if (s.size() != 4)   /* base_dau_conv_layer.cpp:33-35: only 4-D NCHW blobs */
            throw DAUException(DAU_ERR_INVALID_ARGUMENT, std::string(what) + " must be [N, C, H, W]");
    }
    void require_ready(const std::vector<int>& bottom_shape) {
        if (!handle_ || !reshaped_ || dirty_ || bottom_shape != bottom_shape_)
            throw DAUException(DAU_ERR_NOT_READY, "call LayerSetUp and Reshape with this bottom shape first");
    }
    void check_top(const std::vector<int>& top_shape) {
        int ho = 0, wo = 0;
        check(dau_get_output_size(handle_, &ho, &wo));
        if (top_shape.size() != 4 || top_shape[0] != bottom_shape_[0] || top_shape[1] != settings_.num_output ||
            top_shape[2] != ho || top_shape[3] != wo)
            throw DAUException(DAU_ERR_INVALID_ARGUMENT, "top_shape differs from what Reshape returned");
    }
    void rebuild() {
        dau_settings_t st;
        dau_settings_default(&st);
        st.num_output = settings_.num_output;
        st.number_units_x = settings_.number_units[0];
        st.number_units_y = settings_.number_units.size() > 1 ? settings_.number_units[1] : 1;
        st.number_units_ignore = number_units_ignore_;
        st.kernel_size = settings_.kernel_size;
        st.pad = settings_.pad;
        st.stride = settings_.stride;
        st.bias_term = settings_.bias_term;
        st.unit_normalization = settings_.unit_normalization;
        st.square_unit_normalization = settings_.square_unit_normalization;
        st.use_interpolation = settings_.use_interpolation;
        st.single_dimension_kernel = single_dimension_kernel_;
        st.forbid_positive_dim1 = forbid_positive_dim1_;
        st.ignore_edge_gradients = ignore_edge_gradients_;
        st.unit_bounds_guard = unit_bounds_guard_;
        st.offsets_already_centered = offsets_already_centered_ && settings_.offsets_already_centered;
        st.learn_sigma = learn_sigma_;
        st.component_border_bound = settings_.component_border_bound;
        st.sigma_lower_bound = settings_.sigma_lower_bound;
        st.mu_learning_rate_factor = mu_learning_rate_factor_;
        st.mean_iteration_step = settings_.mean_iteration_step;
        st.sigma_iteration_step = settings_.sigma_iteration_step;
        st.merge_iteration_step = settings_.merge_iteration_step;
        st.merge_threshold = settings_.merge_threshold;
        if (handle_) { dau_destroy(handle_); handle_ = nullptr; }
        check(dau_create(&st, device_, &handle_));
        if (sigma_hint_ > 0) check(dau_set_sigma_hint(handle_, sigma_hint_));
        if (offset_hint_ >= 0) check(dau_set_offset_hint(handle_, offset_hint_));
        dirty_ = false;
        reshaped_ = false;
    }

    DAUConvSettings settings_;
    bool ignore_edge_gradients_, offsets_already_centered_, dynamic_kernel_size_;
    bool enabled_fwd_op_ = true, enabled_bwd_op_ = true;
    bool unit_bounds_guard_ = true;            /* reference default (:180) */
    bool learn_sigma_ = true, single_dimension_kernel_ = false, forbid_positive_dim1_ = false;
    bool in_train_ = true, dirty_ = true, reshaped_ = false;
    int number_units_ignore_ = 0, device_ = -1;
    float mu_learning_rate_factor_ = 1.0f, sigma_hint_ = -1.0f, offset_hint_ = -1.0f;
    dau_handle_t* handle_ = nullptr;
    void* stream_ = nullptr;
    void* workspace_ = nullptr;
    size_t workspace_bytes_ = 0;
    std::vector<int> bottom_shape_;
};

}  // namespace dau_b200
#endif /* DAU_CONV_B200_LAYER_HPP_ */
