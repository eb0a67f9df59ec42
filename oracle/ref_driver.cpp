// ref_driver.cpp -- TEST INFRASTRUCTURE.  A plain-cudaMalloc host for the reference's own
// BaseDAUConvLayer<float> (the only in-tree host is the TensorFlow one, which cannot be built here).
// It follows what DAUConvOp / DAUConvGradOp do around the layer (plugins/tensorflow/src/
// dau_conv_op.cpp:264-311, dau_conv_grad_op.cpp:251-303) and the buffer shapes of
// DAUConvLayerTensorflowGPU / DAUKernel*TF (plugins/tensorflow/src/dau_conv_layer_tensorflow.cpp:
// 189-276, 347-383, 401-507, 530-556), with cudaMalloc in place of allocate_temp.
// Linked with the reference's objects into oracle/_ref/libdau_ref_cuda.so by oracle/Makefile.refcuda;
// only files under tests/ load it.  Nothing here is this repo's product path.
#include <cmath>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "dau_conv/base_dau_conv_layer.hpp"

using namespace DAUConvNet;

namespace {

struct DBuf {                       // owning device buffer
    void* p = nullptr; size_t bytes = 0;
    void need(size_t b) {
        if (b <= bytes && p) return;
        if (p) cudaFree(p);
        CUDA_CHECK(cudaMalloc(&p, b ? b : 4));
        bytes = b;
    }
    float* f() { return (float*)p; }
    ~DBuf() { if (p) cudaFree(p); }
};

class PlainParams : public BaseDAUKernelParams<float> {
 public:
    void reshape(int s, int f, int g) override {
        size_t b = sizeof(float) * s * f * g;
        w_.need(b); mu1_.need(b); mu2_.need(b);
        if (!sigma_ext_) sigma_.need(b);
    }
    float* weight() override { return w_.f(); }
    float* mu1() override { return mu1_.f(); }
    float* mu2() override { return mu2_.f(); }
    float* sigma() override { return sigma_ext_ ? sigma_ext_ : sigma_.f(); }
    float* sigma_ext_ = nullptr;    // the layer's own sigma tensor (dau_conv_layer_tensorflow.cpp:181-185)
 private:
    DBuf w_, mu1_, mu2_, sigma_;
};

class PlainOutput : public BaseDAUKernelOutput<float> {
 public:
    void reshape(int s, int f, int g, int kh, int kw) override {
        size_t n = (size_t)s * f * g * kh * kw;
        weight_.need(sizeof(float) * n); d_error_.need(sizeof(float) * n); d_params_.need(sizeof(float) * 4 * n);
    }
    float* weight() override { return weight_.f(); }
    float* d_error() override { return d_error_.f(); }
    float* d_params() override { return d_params_.f(); }
 private:
    DBuf weight_, d_error_, d_params_;
};

class PlainCompute : public BaseDAUKernelCompute<float> {
 public:
    void reshape(int s, int f, int g, int kh, int kw) override {
        this->num_in_channels = s; this->num_out_channels = f; this->num_gauss = g;
        this->kernel_h = kh; this->kernel_w = kw;
        size_t n = (size_t)s * f * g;
        for (int i = 0; i < 5; ++i) kernels_[i].need(sizeof(float) * n * kh * kw);
        for (int i = 0; i < 7; ++i) params_[i].need(sizeof(float) * n);
        std::vector<int> idx(n + 1);
        for (size_t i = 0; i <= n; ++i) idx[i] = (int)(kh * kw * i);
        index_.need(sizeof(int) * (n + 1));
        CUDA_CHECK(cudaMemcpy(index_.p, idx.data(), sizeof(int) * (n + 1), cudaMemcpyHostToDevice));
    }
    float* param_temp(Param_IDX i) override { return params_[i].f(); }
    float* kernels_temp(Kernel_IDX i) override { return kernels_[i].f(); }
    int* precomp_index() override { return (int*)index_.p; }
 private:
    DBuf kernels_[5], params_[7], index_;
};

class NullInit : public BaseDAUComponentInitializer<float> {
 public:
    void InitializeParameters(const DAUConvSettings&, float*, float*, float*, float*, bool, int, int, int, int, int,
                              int, int) const override {}
};

class PlainLayer : public BaseDAUConvLayer<float> {
 public:
    PlainLayer(cublasHandle_t h, bool ignore_edge_gradients)
        : BaseDAUConvLayer<float>(h, ignore_edge_gradients, /*offsets_already_centered=*/true) {}
    ~PlainLayer() override {}

    float *w = nullptr, *mu1 = nullptr, *mu2 = nullptr, *sigma = nullptr;
    float *dw = nullptr, *dmu1 = nullptr, *dmu2 = nullptr, *dsigma = nullptr;
    size_t workspace_bytes = 0;

    void SetUp(const DAUConvSettings& st, PlainCompute* kc, PlainParams* kp, PlainOutput* ko,
               const std::vector<int>& bottom_shape, int units_ignore) {
        BaseDAUConvLayer<float>::LayerSetUp(st, NullInit(), kc, kp, ko, bottom_shape, true);
        this->num_units_ignore = units_ignore;          // dau_conv_layer_tensorflow.cpp:176-179
    }
    std::vector<int> Reshape(const std::vector<int>& bottom_shape, const std::vector<int>& top_shape) override {
        std::vector<int> top = BaseDAUConvLayer<float>::Reshape(bottom_shape, top_shape);
        const size_t mh = std::max(this->height_out_, this->height_), mw = std::max(this->width_out_, this->width_);
        const size_t P = (size_t)this->conv_in_channels_ * this->units_per_channel * this->conv_out_channels_;
        col_.need(sizeof(float) * this->aggregation.kernel_h_ * this->aggregation.kernel_w_ * this->height_ * this->width_);
        size_t ch = std::max(this->conv_out_channels_, this->conv_in_channels_ * this->NUM_K);
        interm_.need(sizeof(float) * this->batch_num_ * ch * mh * mw);
        bwd_gradients_.need(sizeof(float) * this->NUM_K * P);
        tmp_param_.need(sizeof(float) * 2 * P);
        bias_mult_.need(sizeof(float) * this->height_out_ * this->width_out_);
        return top;
    }
    void set_max_kernel_size(int kw, int kh) { this->max_kernel_w_ = kw; this->max_kernel_h_ = kh; }
    cudaStream_t stream() { return this->stream_; }

 protected:
    bool is_data_on_gpu() override { return true; }
    void reshape_params(const std::vector<int>&) override {}
    float* param_w() override { return w; }
    float* param_mu1() override { return mu1; }
    float* param_mu2() override { return mu2; }
    float* param_sigma() override { return sigma; }
    float* param_bias() override { return nullptr; }
    float* param_w_grad() override { return dw; }
    float* param_mu1_grad() override { return dmu1; }
    float* param_mu2_grad() override { return dmu2; }
    float* param_sigma_grad() override { return dsigma; }
    float* param_bias_grad() override { return nullptr; }
    float* temp_bwd_gradients() override { return bwd_gradients_.f(); }
    float* temp_interm_buffer() override { return interm_.f(); }
    float* temp_param_buffer() override { return tmp_param_.f(); }
    float* temp_col_buffer() override { return col_.f(); }
    float* temp_bias_multiplier() override { return bias_mult_.f(); }
    void* allocate_workspace_mem(size_t bytes) override { workspace_.need(bytes); workspace_bytes = bytes; return workspace_.p; }
    void deallocate_workspace_mem() override {}
 private:
    DBuf col_, interm_, bwd_gradients_, tmp_param_, bias_mult_, workspace_;
};

}  // namespace

// One DAUConv + DAUConvGrad pair on device tensors, `iters` times, timed with CUDA events on the layer's
// stream (layer objects built once, outside the timed region: kinder to the reference than its TF op,
// which rebuilds the layer on every call).  Returns 0, or -1 with the exception text in errbuf.
extern "C" int dau_ref_cuda_run(int N, int S, int F, int Gx, int Gy, int H, int W, int kernel_size,
                                int ignore_edge_gradients, int units_ignore, float mu_learning_rate_factor,
                                const float* x, const float* top_error, float* w, float* mu1, float* mu2, float* sigma,
                                float* y, float* dx, float* dw, float* dmu1, float* dmu2, float* dsigma,
                                int iters, float* fwd_ms, float* bwd_ms, size_t* workspace_bytes,
                                char* errbuf, int errlen, int opt_flags) {
    // opt_flags: 1 use_interpolation = false, 2 single_dim_kernel, 4 forbid_positive_dim1 (TF op attrs of those names)
    cublasHandle_t cublas = nullptr;
    try {
        cudaStream_t stream;
        CUDA_CHECK(cudaStreamCreate(&stream));
        CUBLAS_CHECK(cublasCreate(&cublas));
        CUBLAS_CHECK(cublasSetStream(cublas, stream));
        DAUConvSettings st;
        st.num_output = F; st.number_units = {Gx, Gy}; st.bias_term = false;
        st.kernel_size = kernel_size; st.pad = (kernel_size - 1) / 2; st.stride = 1;
        st.unit_normalization = true; st.square_unit_normalization = false;
        st.mean_iteration_step = 1; st.sigma_iteration_step = 1;
        st.component_border_bound = 0.01f; st.sigma_lower_bound = 0.3f;
        st.merge_iteration_step = 0; st.merge_threshold = 1;
        st.offsets_already_centered = true; st.use_interpolation = !(opt_flags & 1);
        const std::vector<int> bottom{N, S, H, W}, top{N, F, H, W};
        const size_t P = (size_t)S * Gx * Gy * F;
        {
            PlainCompute kc; PlainParams kp; PlainOutput ko;
            kp.sigma_ext_ = sigma;
            PlainLayer layer(cublas, ignore_edge_gradients != 0);
            layer.w = w; layer.mu1 = mu1; layer.mu2 = mu2; layer.sigma = sigma;
            layer.dw = dw; layer.dmu1 = dmu1; layer.dmu2 = dmu2; layer.dsigma = dsigma;
            layer.set_default_cuda_stream(stream);
            layer.enable_forward(true);
            layer.enable_backward(true);
            layer.enable_memalloc_info(false);
            layer.enable_unit_bounds_guard(false);
            layer.enable_sigma_learning(true);
            layer.set_single_dimensional_kernel((opt_flags & 2) != 0);      // dau_conv_op.cpp:287-289
            layer.set_forbid_positive_dim1((opt_flags & 4) != 0);
            layer.SetUp(st, &kc, &kp, &ko, bottom, units_ignore);
            std::vector<int> got = layer.Reshape(bottom, top);
            if (got != top) throw DAUException("ref_driver: unexpected top shape");
            layer.set_max_kernel_size(kernel_size, kernel_size);
            if (workspace_bytes) *workspace_bytes = layer.workspace_bytes;

            cudaEvent_t e0, e1, e2;
            CUDA_CHECK(cudaEventCreate(&e0)); CUDA_CHECK(cudaEventCreate(&e1)); CUDA_CHECK(cudaEventCreate(&e2));
            float tf = 0, tb = 0;
            for (int it = 0; it < iters; ++it) {
                // DAUConvGradOp zeroes the four outputs it accumulates into (dau_conv_grad_op.cpp:202-205)
                CUDA_CHECK(cudaMemsetAsync(dw, 0, sizeof(float) * P, stream));
                CUDA_CHECK(cudaMemsetAsync(dmu1, 0, sizeof(float) * P, stream));
                CUDA_CHECK(cudaMemsetAsync(dmu2, 0, sizeof(float) * P, stream));
                CUDA_CHECK(cudaMemsetAsync(dsigma, 0, sizeof(float) * P, stream));
                CUDA_CHECK(cudaDeviceSynchronize());
                CUDA_CHECK(cudaEventRecord(e0, stream));
                layer.Forward_gpu(x, bottom, y, top);
                // the layer also works on four side streams of its own (base_dau_conv_layer.cpp:168-171): wait for
                // the whole device before each stamp so that nothing escapes the timed spans
                CUDA_CHECK(cudaDeviceSynchronize());
                CUDA_CHECK(cudaEventRecord(e1, stream));
                layer.Backward_gpu(nullptr, top_error, top, true, x, dx, bottom,
                                   {true, true, !(opt_flags & 2), true, false});   // dau_conv_grad_op.cpp:291-292
                if (mu_learning_rate_factor != 1.0f) {
                    caffe_gpu_scal<float>((int)P, mu_learning_rate_factor, dmu1, cublas);
                    caffe_gpu_scal<float>((int)P, mu_learning_rate_factor, dmu2, cublas);
                }
                CUDA_CHECK(cudaDeviceSynchronize());
                CUDA_CHECK(cudaEventRecord(e2, stream));
                CUDA_CHECK(cudaDeviceSynchronize());
                float a = 0, b = 0;
                CUDA_CHECK(cudaEventElapsedTime(&a, e0, e1)); CUDA_CHECK(cudaEventElapsedTime(&b, e1, e2));
                if (it == 0 && iters > 1) continue;      // first pass pays the lazy filter setup
                tf += a; tb += b;
            }
            const int counted = iters > 1 ? iters - 1 : 1;
            if (fwd_ms) *fwd_ms = tf / counted;
            if (bwd_ms) *bwd_ms = tb / counted;
            cudaEventDestroy(e0); cudaEventDestroy(e1); cudaEventDestroy(e2);
        }
        cublasDestroy(cublas);
        cudaStreamDestroy(stream);
        return 0;
    } catch (const std::exception& ex) {
        if (errbuf && errlen > 0) { std::strncpy(errbuf, ex.what(), errlen - 1); errbuf[errlen - 1] = 0; }
        return -1;
    }
}

// DAUKernelCompute alone (row a1): the six Kb x Kb filters the reference derives from sigma for one unit
// (w = 1, mu = 0), exactly as update_prefiltering_kernels drives it (base_dau_conv_layer.cpp:415-431,
// :120-126, :165-183).  out_host = [blur, d_w, d_mu1, d_mu2, d_sigma, d_error] (the order of this repo's
// dau_get_kernels).  flags: 1 single_dimension_kernel, 2 forbid_positive_dim1, 4 no unit normalisation,
// 8 square unit normalisation.
extern "C" int dau_ref_cuda_kernels(float sigma, int flags, int* kb_out, float* out_host, int capacity_floats,
                                    char* errbuf, int errlen) {
    cublasHandle_t cublas = nullptr;
    try {
        const int kb = 2 * (int)ceil(5 * sigma) + 1;
        if (kb_out) *kb_out = kb;
        if (capacity_floats < 6 * kb * kb) throw DAUException("ref_driver: output buffer too small");
        CUBLAS_CHECK(cublasCreate(&cublas));
        {
            PlainCompute kc; PlainParams kp; PlainOutput ko;
            kc.setup(!(flags & 4), (flags & 8) != 0, 0.3f, 0, /*offsets_already_centered=*/true,
                     (flags & 1) != 0, (flags & 2) != 0);
            kp.reshape(1, 1, 1);
            const float one = 1.f, zero = 0.f;
            CUDA_CHECK(cudaMemcpy(kp.weight(), &one, 4, cudaMemcpyHostToDevice));
            CUDA_CHECK(cudaMemcpy(kp.mu1(), &zero, 4, cudaMemcpyHostToDevice));
            CUDA_CHECK(cudaMemcpy(kp.mu2(), &zero, 4, cudaMemcpyHostToDevice));
            CUDA_CHECK(cudaMemcpy(kp.sigma(), &sigma, 4, cudaMemcpyHostToDevice));
            kc.reshape(1, 1, 1, kb, kb);
            ko.reshape(1, 1, 1, kb, kb);
            kc.get_kernels(kp, ko, /*enable_unit_bounds_guard=*/false, cublas, 0);
            CUDA_CHECK(cudaDeviceSynchronize());
            const size_t n = (size_t)kb * kb;
            CUDA_CHECK(cudaMemcpy(out_host, ko.weight(), 4 * n, cudaMemcpyDeviceToHost));
            CUDA_CHECK(cudaMemcpy(out_host + n, ko.d_params(), 4 * 4 * n, cudaMemcpyDeviceToHost));
            CUDA_CHECK(cudaMemcpy(out_host + 5 * n, ko.d_error(), 4 * n, cudaMemcpyDeviceToHost));
        }
        cublasDestroy(cublas);
        return 0;
    } catch (const std::exception& ex) {
        if (errbuf && errlen > 0) { std::strncpy(errbuf, ex.what(), errlen - 1); errbuf[errlen - 1] = 0; }
        return -1;
    }
}
