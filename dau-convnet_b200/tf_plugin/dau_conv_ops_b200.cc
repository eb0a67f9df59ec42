// TensorFlow registration of the ops `DAUConv` and `DAUConvGrad` on top of libdau_conv_b200.so.
//
// Same op names, inputs, outputs and attributes (names and defaults) as the reference's
// plugins/tensorflow/src/dau_conv_op.cpp:22-48 and dau_conv_grad_op.cpp:18-49, so that the reference's
// Python layer (plugins/tensorflow/dau_conv/dau_conv.py, _dau_conv_grad_op.py) loads this library in place
// of its own dau_conv_op.so without a change.
//
// STATUS: UNVERIFIED.  TensorFlow is not installed in the authoring image or on the GPU box, so this file
// has never been compiled; it is compile-gated (DAU_B200_WITH_TENSORFLOW) and not part of
// csrc/Makefile.  The verified bindings of the same C ABI are the ctypes mirror (dau_conv/functional.py)
// and the C++ face (include/dau_conv_b200_layer.hpp).  Build, where TensorFlow exists:
//   g++ -std=c++17 -shared -fPIC -DDAU_B200_WITH_TENSORFLOW dau_conv_ops_b200.cc -o dau_conv_op.so \
//       -I../../include $(python -c "import tensorflow as tf; print(' '.join(tf.sysconfig.get_compile_flags()))") \
//       -L../dau_conv -ldau_conv_b200 $(python -c "import tensorflow as tf; print(' '.join(tf.sysconfig.get_link_flags()))")
#ifdef DAU_B200_WITH_TENSORFLOW

#include <algorithm>
#include <cmath>
#include <mutex>
#include <vector>

#include "tensorflow/core/framework/op.h"
#include "tensorflow/core/framework/op_kernel.h"
#include "tensorflow/core/framework/shape_inference.h"
#define EIGEN_USE_GPU
#include "tensorflow/core/util/gpu_kernel_helper.h"

#include "dau_conv_b200.h"

namespace dau_b200_tf {
using namespace tensorflow;  // NOLINT

#define DAU_COMMON_ATTRS(border_default)                                                           \
    .Attr("number_units_x : int = 2").Attr("number_units_y : int = 2")                             \
    .Attr("number_units_ignore : int = 0").Attr("num_output : int = 64")                           \
    .Attr("kernel_size: int = 9").Attr("pad: int = 4").Attr("stride: int = 1")                     \
    .Attr("unit_normalization: bool = true").Attr("square_unit_normalization: bool = false")       \
    .Attr("mean_iteration_step: int = 1").Attr("sigma_iteration_step: int = 1")                    \
    .Attr("component_border_bound: float = " border_default).Attr("sigma_lower_bound: float = 0.3") \
    .Attr("merge_iteration_step: int = 0").Attr("merge_threshold: int = 1")                        \
    .Attr("unit_testing: bool = false").Attr("mu_learning_rate_factor: float = 1.0")               \
    .Attr("single_dim_kernel: bool = false").Attr("forbid_positive_dim1: bool = false")            \
    .Attr("use_interpolation: bool = true")

REGISTER_OP("DAUConv")
    .Input("input: float").Input("weights: float").Input("mu1: float").Input("mu2: float").Input("sigma: float")
    .Output("output: float")
    DAU_COMMON_ATTRS("1")
    .SetShapeFn([](shape_inference::InferenceContext* c) {
        // rank-4 everything, parameters end in num_output, output = input with dim 1 replaced (dau_conv_op.cpp:49-84)
        shape_inference::ShapeHandle in, p;
        TF_RETURN_IF_ERROR(c->WithRank(c->input(0), 4, &in));
        int num_output = 0;
        TF_RETURN_IF_ERROR(c->GetAttr("num_output", &num_output));
        shape_inference::DimensionHandle f;
        for (int i = 1; i <= 4; ++i) {
            TF_RETURN_IF_ERROR(c->WithRank(c->input(i), 4, &p));
            TF_RETURN_IF_ERROR(c->WithValue(c->Dim(p, 3), num_output, &f));
        }
        shape_inference::ShapeHandle out;
        TF_RETURN_IF_ERROR(c->ReplaceDim(in, 1, f, &out));
        c->set_output(0, out);
        return OkStatus();
    });

REGISTER_OP("DAUConvGrad")
    .Input("grad: float32").Input("input: float32").Input("weights: float32")
    .Input("mu1: float32").Input("mu2: float32").Input("sigma: float32")
    .Output("grad_input: float32").Output("grad_weights: float32").Output("grad_mu1: float32")
    .Output("grad_mu2: float32").Output("grad_sigma: float32")
    DAU_COMMON_ATTRS("0");

// Attributes -> dau_settings_t, one handle per op instance, re-planned only when the input shape or the
// displacement bucket changes (the reference builds a whole layer object on every Compute, dau_conv_op.cpp:267).
class DAUOpBase : public OpKernel {
 public:
    explicit DAUOpBase(OpKernelConstruction* ctx) : OpKernel(ctx) {
        dau_settings_default(&st_);
        int merge_threshold = 1;
        bool b = false;
        OP_REQUIRES_OK(ctx, ctx->GetAttr("number_units_x", &st_.number_units_x));
        OP_REQUIRES_OK(ctx, ctx->GetAttr("number_units_y", &st_.number_units_y));
        OP_REQUIRES_OK(ctx, ctx->GetAttr("number_units_ignore", &st_.number_units_ignore));
        OP_REQUIRES_OK(ctx, ctx->GetAttr("num_output", &st_.num_output));
        OP_REQUIRES_OK(ctx, ctx->GetAttr("kernel_size", &st_.kernel_size));
        OP_REQUIRES_OK(ctx, ctx->GetAttr("stride", &st_.stride));
        OP_REQUIRES_OK(ctx, ctx->GetAttr("unit_normalization", &b)); st_.unit_normalization = b;
        OP_REQUIRES_OK(ctx, ctx->GetAttr("square_unit_normalization", &b)); st_.square_unit_normalization = b;
        OP_REQUIRES_OK(ctx, ctx->GetAttr("mean_iteration_step", &st_.mean_iteration_step));
        OP_REQUIRES_OK(ctx, ctx->GetAttr("sigma_iteration_step", &st_.sigma_iteration_step));
        OP_REQUIRES_OK(ctx, ctx->GetAttr("component_border_bound", &st_.component_border_bound));
        OP_REQUIRES_OK(ctx, ctx->GetAttr("sigma_lower_bound", &st_.sigma_lower_bound));
        OP_REQUIRES_OK(ctx, ctx->GetAttr("merge_iteration_step", &st_.merge_iteration_step));
        OP_REQUIRES_OK(ctx, ctx->GetAttr("merge_threshold", &merge_threshold)); st_.merge_threshold = merge_threshold;
        OP_REQUIRES_OK(ctx, ctx->GetAttr("unit_testing", &b)); st_.ignore_edge_gradients = b;   // dau_conv_grad_op.cpp:253
        OP_REQUIRES_OK(ctx, ctx->GetAttr("mu_learning_rate_factor", &st_.mu_learning_rate_factor));
        OP_REQUIRES_OK(ctx, ctx->GetAttr("single_dim_kernel", &b)); st_.single_dimension_kernel = b;
        OP_REQUIRES_OK(ctx, ctx->GetAttr("forbid_positive_dim1", &b)); st_.forbid_positive_dim1 = b;
        OP_REQUIRES_OK(ctx, ctx->GetAttr("use_interpolation", &b)); st_.use_interpolation = b;
        st_.pad = (st_.kernel_size - 1) / 2;      // the op always runs "same" (dau_conv_op.cpp:256)
        st_.unit_bounds_guard = 0;                // inputs are immutable; clipping lives in Python (:283-285)
        st_.offsets_already_centered = 1;
        st_.learn_sigma = 1;                      // dau_conv_grad_op.cpp:123
    }
    ~DAUOpBase() override { if (h_) dau_destroy(h_); }

 protected:
    // Validates the five tensors (dau_conv_op.cpp:150-208) and (re)plans the handle for this shape.  The caller
    // holds mu_ from here until its launch and NaN read are enqueued: two concurrent Compute() calls on one kernel
    // instance with different shapes must not re-plan the shared handle under a running call.
    Status Prepare(OpKernelContext* ctx, const Tensor& input, const Tensor& w, const Tensor& mu1, const Tensor& mu2,
                   const Tensor& sigma, size_t* fwd, size_t* bwd) {
        if (input.dims() != 4) return errors::InvalidArgument("DAUConv: input must be [N, C, H, W]");
        for (const Tensor* p : {&w, &mu1, &mu2, &sigma}) {
            if (p->dims() != 4 || p->dim_size(0) != 1 || p->dim_size(1) != input.dim_size(1) ||
                p->dim_size(2) != st_.number_units_x * st_.number_units_y || p->dim_size(3) != st_.num_output)
                return errors::InvalidArgument("DAUConv: parameters must be [1, S, G, F]");
        }
        if (!h_ && dau_create(&st_, /*device=*/-1, &h_) != DAU_OK) return errors::Internal(dau_last_error());
        std::vector<int64_t> shape{input.dim_size(0), input.dim_size(1), input.dim_size(2), input.dim_size(3)};
        if (shape != shape_) {
            if (dau_reshape(h_, (int)shape[0], (int)shape[1], (int)shape[2], (int)shape[3], &fwd_, &bwd_) != DAU_OK)
                return errors::Internal(dau_last_error());
            shape_ = shape;
        }
        *fwd = fwd_; *bwd = bwd_;
        return OkStatus();
    }
    // The reference fails the call on a NaN offset (dau_conv_op.cpp:258-262); here NaN units are dropped
    // inside the kernels and the sticky flag is read back after the launch.
    Status CheckNan(cudaStream_t stream) {
        int found = 0;
        if (dau_take_nan_flag(h_, &found, stream) != DAU_OK) return errors::Internal(dau_last_error());
        if (found) return errors::FailedPrecondition("DAUConvOp ERROR: got NaN value in offset (mu1,mu2) variable");
        return OkStatus();
    }
    static float* ptr(const Tensor& t) { return const_cast<float*>(t.flat<float>().data()); }

    dau_settings_t st_;
    dau_handle_t* h_ = nullptr;
    std::mutex mu_;
    std::vector<int64_t> shape_;
    size_t fwd_ = 0, bwd_ = 0;
};

class DAUConvOpB200 : public DAUOpBase {
 public:
    using DAUOpBase::DAUOpBase;
    void Compute(OpKernelContext* ctx) override {
        const Tensor &input = ctx->input(0), &w = ctx->input(1), &mu1 = ctx->input(2), &mu2 = ctx->input(3),
                     &sigma = ctx->input(4);
        size_t fwd = 0, bwd = 0;
        std::lock_guard<std::mutex> lock(mu_);          // held across planning, the launch and the flag read
        OP_REQUIRES_OK(ctx, Prepare(ctx, input, w, mu1, mu2, sigma, &fwd, &bwd));
        Tensor* out = nullptr;
        OP_REQUIRES_OK(ctx, ctx->allocate_output(0, TensorShape({input.dim_size(0), st_.num_output, input.dim_size(2),
                                                                 input.dim_size(3)}), &out));
        Tensor ws;
        OP_REQUIRES_OK(ctx, ctx->allocate_temp(DT_UINT8, TensorShape({(int64_t)std::max<size_t>(fwd, 256)}), &ws));
        cudaStream_t stream = ctx->eigen_device<Eigen::GpuDevice>().stream();
        OP_REQUIRES(ctx, dau_forward(h_, input.flat<float>().data(), ptr(w), ptr(mu1), ptr(mu2), ptr(sigma), nullptr,
                                     out->flat<float>().data(), ws.flat<uint8>().data(), fwd, stream) == DAU_OK,
                    errors::Internal(dau_last_error()));
        OP_REQUIRES_OK(ctx, CheckNan(stream));
    }
};

class DAUConvGradOpB200 : public DAUOpBase {
 public:
    using DAUOpBase::DAUOpBase;
    void Compute(OpKernelContext* ctx) override {
        const Tensor &grad = ctx->input(0), &input = ctx->input(1), &w = ctx->input(2), &mu1 = ctx->input(3),
                     &mu2 = ctx->input(4), &sigma = ctx->input(5);
        size_t fwd = 0, bwd = 0;
        std::lock_guard<std::mutex> lock(mu_);          // held across planning, the launch and the flag read
        OP_REQUIRES_OK(ctx, Prepare(ctx, input, w, mu1, mu2, sigma, &fwd, &bwd));
        Tensor *gi = nullptr, *gw = nullptr, *gm1 = nullptr, *gm2 = nullptr, *gs = nullptr;
        OP_REQUIRES_OK(ctx, ctx->allocate_output(0, input.shape(), &gi));
        OP_REQUIRES_OK(ctx, ctx->allocate_output(1, w.shape(), &gw));
        OP_REQUIRES_OK(ctx, ctx->allocate_output(2, mu1.shape(), &gm1));
        OP_REQUIRES_OK(ctx, ctx->allocate_output(3, mu2.shape(), &gm2));
        OP_REQUIRES_OK(ctx, ctx->allocate_output(4, sigma.shape(), &gs));
        Tensor ws;
        OP_REQUIRES_OK(ctx, ctx->allocate_temp(DT_UINT8, TensorShape({(int64_t)std::max<size_t>(bwd, 256)}), &ws));
        cudaStream_t stream = ctx->eigen_device<Eigen::GpuDevice>().stream();
        // the layer accumulates into the parameter gradients; the op owns zeroing them (dau_conv_grad_op.cpp:202-205)
        for (Tensor* t : {gw, gm1, gm2, gs})
            cudaMemsetAsync(t->flat<float>().data(), 0, sizeof(float) * t->NumElements(), stream);
        // which gradients: dau_conv_grad_op.cpp:291-292 ({w, mu1, mu2 unless single_dim_kernel, sigma, no bias})
        unsigned flags = DAU_GRAD_INPUT | DAU_GRAD_W | DAU_GRAD_MU1 | DAU_GRAD_SIGMA |
                         (st_.single_dimension_kernel ? 0u : (unsigned)DAU_GRAD_MU2);
        OP_REQUIRES(ctx, dau_backward(h_, input.flat<float>().data(), grad.flat<float>().data(), ptr(w), ptr(mu1),
                                      ptr(mu2), ptr(sigma), gi->flat<float>().data(), gw->flat<float>().data(),
                                      gm1->flat<float>().data(), gm2->flat<float>().data(), gs->flat<float>().data(),
                                      nullptr, flags, ws.flat<uint8>().data(), bwd, stream) == DAU_OK,
                    errors::Internal(dau_last_error()));
        // mu_learning_rate_factor (:296-303) is applied inside dau_backward
        OP_REQUIRES_OK(ctx, CheckNan(stream));
    }
};

// GPU / float only, as the reference (dau_conv_op.cpp:339-347)
REGISTER_KERNEL_BUILDER(Name("DAUConv").Device(DEVICE_GPU), DAUConvOpB200);
REGISTER_KERNEL_BUILDER(Name("DAUConvGrad").Device(DEVICE_GPU), DAUConvGradOpB200);

}  // namespace dau_b200_tf
#endif  // DAU_B200_WITH_TENSORFLOW
