/*
 * dau_conv_b200.h -- C ABI of the B200-native DAU convolution layer.
 *
 * This is the drop-in boundary for the ONE hot path of skokec/DAU-ConvNet:
 * BaseDAUConvLayer<float>::Forward_gpu / Backward_gpu and the TensorFlow ops
 * DAUConv / DAUConvGrad built on top of them.  Plain pointers and sizes only; every
 * pointer named *_dev is DEVICE memory owned by the caller, every call is ordered on
 * the caller's CUDA stream (passed as void*; NULL = legacy default stream).  dau_forward /
 * dau_backward allocate nothing behind the caller's back except 4 bytes per handle (the NaN flag);
 * the *_host conveniences are the exception and say so: dau_forward_host / dau_backward_host
 * cudaMalloc and free their device copies inside the call, and dau_train_step_host keeps a device
 * pool per handle (two staging buffers per tensor plus the workspace of the largest chunk, sized on
 * first use; growing it synchronises the device once) until dau_destroy.  Every entry point leaves
 * the calling thread's current CUDA device as it found it.
 *
 * Reference interfaces replaced (paths relative to the reference checkout):
 *   dau_settings_t      <- DAUConvSettings                   include/dau_conv/base_dau_conv_layer.hpp:109-130
 *                          + ctor flags / toggles            include/dau_conv/base_dau_conv_layer.hpp:177-184, 208-225
 *                          + TF op attrs                     plugins/tensorflow/src/dau_conv_op.cpp:29-48
 *   dau_create          <- BaseDAUConvLayer ctor+LayerSetUp  src/dau_conv/base_dau_conv_layer.cpp:29-187
 *   dau_reshape         <- Reshape / workspace sizing        src/dau_conv/base_dau_conv_layer.cpp:190-394
 *   dau_forward         <- Forward_gpu                       src/dau_conv/base_dau_conv_layer.cu:16-127
 *   dau_backward        <- Backward_gpu (+ TF grad scaling)  src/dau_conv/base_dau_conv_layer.cu:131-363,
 *                                                            plugins/tensorflow/src/dau_conv_grad_op.cpp:296-303
 *   dau_get_kernels     <- BaseDAUKernelCompute::get_kernels src/dau_conv/base_dau_conv_layer.cu:537-710
 *   dau_status_t/_last_error <- DAUException                 include/dau_conv/util/common.hpp:38-64, 102-106
 *
 * Layouts are the reference's: activations NCHW fp32, parameters [1,S,G,F] fp32 with F
 * fastest (base_dau_conv_layer.hpp:233); mu1 is the x (width) offset, mu2 the y (height)
 * offset (base_dau_conv_layer.cpp:574-578).
 */
#ifndef DAU_CONV_B200_H_
#define DAU_CONV_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DAU_B200_ABI_VERSION 1

typedef enum dau_status_t {
    DAU_OK = 0,
    DAU_ERR_INVALID_ARGUMENT = 1, /* bad shape / setting (reference: DAU_CHECK failure) */
    DAU_ERR_CUDA = 2,             /* a CUDA call or launch failed (reference: CUDA_CHECK) */
    DAU_ERR_NOT_READY = 3,        /* forward/backward before reshape, workspace too small */
    DAU_ERR_UNSUPPORTED = 4,      /* e.g. stride != 1, |mu| bucket > 32 (dau_conv_op.cpp:250-253) */
    DAU_ERR_NAN_OFFSET = 5        /* NaN in mu1/mu2 (dau_conv_op.cpp:258-262, FAILED_PRECONDITION) */
} dau_status_t;

/* Mirrors DAUConvSettings + the layer toggles.  Zero-initialise, then dau_settings_default(). */
typedef struct dau_settings_t {
    int32_t num_output;            /* F */
    int32_t number_units_x;        /* G = x*y (DAUConvSettings::number_units) */
    int32_t number_units_y;
    int32_t number_units_ignore;   /* trailing dummy units whose gradients are zeroed */
    int32_t kernel_size;           /* max_kernel_size: |mu| < kernel_size/2; halo of the gather */
    int32_t pad;                   /* only pad == (kernel_size-1)/2 (TF op, dau_conv_op.cpp:256) */
    int32_t stride;                /* must be 1 (base_dau_conv_layer.cpp:68-69) */
    int32_t bias_term;             /* 0/1: adds bias[F] in forward, reduces dbias in backward */
    int32_t unit_normalization;    /* default 1 */
    int32_t square_unit_normalization;
    int32_t use_interpolation;     /* default 1; 0 = nearest-lower tap only */
    int32_t single_dimension_kernel; /* DAUConv1d: only the centre row of the blur survives */
    int32_t forbid_positive_dim1;
    int32_t ignore_edge_gradients; /* TF attr unit_testing */
    int32_t unit_bounds_guard;     /* clip mu/sigma IN PLACE before use (Forward_gpu :33-49) */
    int32_t offsets_already_centered; /* default 1 (TF op); 0: mu counts from the kernel corner, centre = mu - kernel_size/2 (Caffe layer, base_dau_conv_layer.cpp:574-575) */
    int32_t learn_sigma;           /* 0 => sigma gradient skipped (last_k_optional) */
    float   component_border_bound;
    float   sigma_lower_bound;
    float   mu_learning_rate_factor; /* applied to dmu1,dmu2 by dau_backward (TF grad op) */
    /* parsed-but-unused reference fields, accepted for signature parity */
    int32_t mean_iteration_step, sigma_iteration_step, merge_iteration_step;
    float   merge_threshold;
} dau_settings_t;

typedef struct dau_handle_s dau_handle_t;

/* flags for dau_backward: which gradients to produce (params_propagate_down + propagate_down) */
enum {
    DAU_GRAD_INPUT = 1, DAU_GRAD_W = 2, DAU_GRAD_MU1 = 4, DAU_GRAD_MU2 = 8,
    DAU_GRAD_SIGMA = 16, DAU_GRAD_BIAS = 32, DAU_GRAD_ALL = 31
};

const char* dau_last_error(void);          /* thread-local message of the last failure */
int         dau_abi_version(void);
void        dau_settings_default(dau_settings_t* s);

/* device: CUDA ordinal; negative = the calling thread's current device */
dau_status_t dau_create(const dau_settings_t* settings, int device, dau_handle_t** out);
void         dau_destroy(dau_handle_t* h);

/* Shape-dependent setup.  Output shape is [N, F, H, W].  Reports the workspace the caller
 * must lend to forward and backward (bytes, 256-byte aligned pointer expected). */
dau_status_t dau_reshape(dau_handle_t* h, int N, int S, int H, int W,
                         size_t* fwd_workspace_bytes, size_t* bwd_workspace_bytes);

/* Optional: tell the layer sigma's value so that forward/backward need no 4-byte D2H read
 * of sigma_dev[0] (the reference reads it every LayerSetUp, base_dau_conv_layer.hpp:266-275).
 * value <= 0 re-enables the read. */
dau_status_t dau_set_sigma_hint(dau_handle_t* h, float value);

/* Optional, BEFORE dau_reshape: an upper bound of |mu - centre| over all units.  The halos, the
 * mosaic gap and the clamp of the integer offsets then follow the displacement bucket (< 4, 8, 16
 * or 32 pixels) that holds the bound instead of kernel_size/2 -- the choice the reference's TF op
 * makes on every call from max|mu| (dau_conv_op.cpp:222-256: kernel 9 / 17 / 33 / 65), here without
 * the device read.  Offsets beyond the bucket are clamped.  A negative value removes the bound. */
dau_status_t dau_set_offset_hint(dau_handle_t* h, float max_abs_offset);

/* Output size after dau_reshape: H_out = H + 2*pad - kernel_size + 1 (base_dau_conv_layer.cpp:20-22).
 * pad == (kernel_size-1)/2 (what the TF op always passes) gives H_out == H; a smaller pad gives the
 * centre of the same-size result, and the backward pass takes top_error[N,F,H_out,W_out] as if
 * zero-padded back to H x W (base_dau_conv_layer.cu:258-266).  pad > (kernel_size-1)/2 is refused. */
dau_status_t dau_get_output_size(dau_handle_t* h, int* h_out, int* w_out);

/* top[N,F,H_out,W_out] = DAU(bottom[N,S,H,W]); top is overwritten. bias_dev may be NULL. */
dau_status_t dau_forward(dau_handle_t* h, const float* bottom_dev,
                         float* w_dev, float* mu1_dev, float* mu2_dev, float* sigma_dev,
                         const float* bias_dev, float* top_dev,
                         void* workspace_dev, size_t workspace_bytes, void* stream);

/* top_error is [N,F,H_out,W_out]; bottom_error[N,S,H,W] is overwritten; d{w,mu1,mu2,sigma}[1,S,G,F] (and dbias[F]) are
 * ACCUMULATED INTO (caller pre-zeroes, dau_conv_grad_op.cpp:202-205).  dmu1/dmu2/dsigma
 * carry the xw factor, dmu1/dmu2 also mu_learning_rate_factor, NaNs in dmu are zeroed. */
dau_status_t dau_backward(dau_handle_t* h, const float* bottom_dev, const float* top_error_dev,
                          const float* w_dev, const float* mu1_dev, const float* mu2_dev,
                          const float* sigma_dev, float* bottom_error_dev,
                          float* dw_dev, float* dmu1_dev, float* dmu2_dev, float* dsigma_dev,
                          float* dbias_dev, unsigned flags,
                          void* workspace_dev, size_t workspace_bytes, void* stream);

/* *found = 1 if any mu1 / mu2 was NaN in a dau_forward / dau_backward call on this handle since the
 * last query (such units contribute nothing), then clears the flag.  Blocks on `stream`.  The TF op
 * fails such a call with FAILED_PRECONDITION (dau_conv_op.cpp:258-262) out of the max|mu| it reads
 * back on every call; this library reads nothing back unless asked. */
dau_status_t dau_take_nan_flag(dau_handle_t* h, int* found, void* stream);

/* Fused epilogue / prologue (the steps the reference's Python layer runs right after the op:
 * bias add and activation, plugins/tensorflow/dau_conv/dau_conv.py:500-526; ReLU is what every
 * reference model uses).  With DAU_ACT_RELU dau_forward stores max(0, y + bias) and
 * dau_backward_act -- dau_backward plus the forward OUTPUT top_data[N,F,H,W] -- takes the top
 * error as zero wherever top_data <= 0 while it loads it (blur, gradient and bias-gradient
 * kernels alike), so neither the activation nor its backward mask cost a pass over HBM.
 * With DAU_ACT_NONE dau_backward_act ignores top_data and equals dau_backward. */
enum { DAU_ACT_NONE = 0, DAU_ACT_RELU = 1 };
dau_status_t dau_set_activation(dau_handle_t* h, int activation);
dau_status_t dau_backward_act(dau_handle_t* h, const float* bottom_dev, const float* top_data_dev,
                              const float* top_error_dev,
                              const float* w_dev, const float* mu1_dev, const float* mu2_dev,
                              const float* sigma_dev, float* bottom_error_dev,
                              float* dw_dev, float* dmu1_dev, float* dmu2_dev, float* dsigma_dev,
                              float* dbias_dev, unsigned flags,
                              void* workspace_dev, size_t workspace_bytes, void* stream);

/* DAUKernelCompute for one unit (w=1, mu=0): six Kb x Kb HOST arrays, row-major, axis 0 = y.
 * Order: blur, d_w, d_mu1, d_mu2, d_sigma, d_error.  *kb_out = 2*ceil(5*sigma)+1. */
dau_status_t dau_get_kernels(const dau_settings_t* settings, float sigma, int* kb_out,
                             float* out6_host, size_t out_capacity_floats);

/* Host-buffer convenience (what a non-CUDA host binds): copies H2D, runs, copies back. */
dau_status_t dau_forward_host(dau_handle_t* h, int N, int S, int H, int W, const float* bottom,
                              const float* w, const float* mu1, const float* mu2, const float* sigma,
                              const float* bias, float* top);
dau_status_t dau_backward_host(dau_handle_t* h, int N, int S, int H, int W, const float* bottom,
                               const float* top_error, const float* w, const float* mu1,
                               const float* mu2, const float* sigma, float* bottom_error,
                               float* dw, float* dmu1, float* dmu2, float* dsigma, float* dbias,
                               unsigned flags);

/* One training step (dau_forward + dau_backward, all gradients) on a batch that lives in HOST
 * memory, cut into chunks of at most `chunk_images` whose PCIe copies overlap the kernels of the
 * neighbouring chunks (three internal streams, double-buffered staging owned by the handle; the
 * chunk sizes -- short at both ends, where nothing can hide a copy -- are searched per shape
 * against a model of the streams and of the kernels' last-wave fill).
 * x, top_error, y, dx: host pointers, NCHW fp32 (pinned memory for real overlap; y / dx may be
 * NULL).  Parameters and parameter gradients: DEVICE pointers [1,S,G,F]; the gradients are
 * accumulated into, as in dau_backward.  Stream-ordered with respect to `stream`, asynchronous
 * to the host.  sigma_value: the layer's sigma (it is one number, base_dau_conv_layer.hpp:150).
 * Replaces, for a host outside a CUDA framework, the sequence  H2D -> DAUConv -> DAUConvGrad ->
 * D2H  around plugins/tensorflow/src/dau_conv_op.cpp:150-324 / dau_conv_grad_op.cpp:115-318. */
dau_status_t dau_train_step_host(dau_handle_t* h, int N, int S, int H, int W, int chunk_images,
                                 float sigma_value, const float* x_host, const float* top_error_host,
                                 float* w_dev, float* mu1_dev, float* mu2_dev, float* sigma_dev,
                                 float* y_host, float* dx_host, float* dw_dev, float* dmu1_dev,
                                 float* dmu2_dev, float* dsigma_dev, void* stream);

/* Host-side planning helpers (no device needed; used by the CPU test-suite).
 * dau_plan_mosaic: how a batch of N narrow H x W images is packed into mosaic images for the
 * 64-pixel-wide tiles of the kernels; out10 = {kx, ky, pitch_x, pitch_y, W, H, Wv, Hv, Nv, gap}.
 * dau_mosaic_to_real: mosaic pixel (nv, vy, vx) -> real pixel; returns 1 and fills n/y/x for
 * image pixels, 0 for gap pixels and for cells beyond the batch. */
dau_status_t dau_plan_mosaic(int N, int H, int W, int gap, int* out10);
/* dau_plan_strips: the work items of the parameter-gradient kernel for that mosaic -- row strips of 4 rows and one
 * tail strip of H % 4 rows per cell row, never a gap row (reference: the 8-row warp tiles of its patches,
 * dau_conv_backward_core.hpp:1017-1100, have no counterpart of this).  counts2 = {full items, tail items};
 * items_out (may be NULL) receives {image pair, first mosaic row, rows, 64-column tile} per item, full items first. */
dau_status_t dau_plan_strips(int N, int H, int W, int gap, long long* counts2, int* items_out, long long capacity);
int dau_mosaic_to_real(const int* mosaic10, int N, int nv, int vy, int vx, int* n, int* y, int* x);

/* Allocation-free bias gradient.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  dbias[f] += sum_{n,y,x} top_error[n,f,y,x], and with top_data != NULL only where
 * top_data[n,f,y,x] > 0 (the fused ReLU backward) -- exactly what dau_backward / dau_backward_act compute for
 * DAU_GRAD_BIAS (same two kernels, same summation order, same bits), but the per-slice partial sums live in a
 * file-scope `static __device__` array of libdau_conv_b200.so instead of the caller's workspace:
 *   - limit: F <= 4096 (DAU_STATIC_BIAS_MAX_F); a larger F is refused with DAU_ERR_UNSUPPORTED and a message;
 *   - cost: 16 x 4096 x 4 B = 256 KB of device memory on every GPU that loads the library, until the process exits;
 *   - the call allocates and frees nothing (no cudaMalloc / cudaFree, no allocating library routine); its first
 *     call on a device looks the array's address up once (cudaGetSymbolAddress) -- make that call outside a capture;
 *   - every call on a GPU shares the one array: calls that OVERLAP on the same GPU (two streams, two threads) race;
 *     stream-ordered calls are safe.  Nothing in the array has to start at zero: a call writes every entry it reads.
 * device: CUDA ordinal, negative = the calling thread's current device (left as found).  All pointers are DEVICE
 * memory owned by the caller; dbias[F] is accumulated into.  No handle, no workspace. */
#define DAU_STATIC_BIAS_MAX_F 4096
dau_status_t dau_bias_grad_static(int device, const float* top_error_dev, const float* top_data_dev,
                                  float* dbias_dev, int N, int F, int H_out, int W_out, void* stream);

/* ignore_edge_gradients rule for one output extent: 1 if the last row / column of the top error is
 * zeroed (extent a multiple of the reference's patch size 64 / 32 / 16 / 8, the largest that is
 * <= extent; plugins/tensorflow/tests/dau_conv_test.py:110-136, base_dau_conv_layer.cpp:763-771). */
int dau_edge_mask_rule(int extent);

/* Test hook (integer-only model of the index math, SURVEY section 8(d) parity gate): the unit tables
 * the kernels read for these DEVICE parameters [1,S,G,F], written into caller-provided DEVICE buffers
 * of S*G*F 32-byte records {float w00, w01, w10, w11; int32 offx + halo_x, offy + halo_y, 0, 0} (any of
 * the three may be NULL).  Orders: forward [s][f][g] (offset floor(mu), dau_conv_forward_core.hpp:2025-2028,
 * weights w * bilinear(frac), :2155-2213), back-propagation [f][s][g] (offset floor(-mu)), gradient
 * [s][f][g] (offset -floor(mu)-1, bilinear weights flipped, no w).  halos4_host = {halo_x, halo_y of the
 * first two tables, halo_x, halo_y of the gradient table}.  After dau_reshape. */
dau_status_t dau_debug_unit_tables(dau_handle_t* h, const float* w_dev, const float* mu1_dev,
                                   const float* mu2_dev, void* fwd_out_dev, void* bwd_out_dev,
                                   void* grad_out_dev, int* halos4_host, void* stream);

/* Measurement helpers used by bench.py (not part of the reference contract). */
dau_status_t dau_measure_fma_peak(int device, int iters, double* tflops_out, double* ms_out);
/* Per-kernel device times (ms) of the last forward / backward on this handle, measured with
 * CUDA events on the caller's stream when profiling is enabled. names: see dau_layer.cpp. */
dau_status_t dau_enable_timing(dau_handle_t* h, int enable);
dau_status_t dau_get_timing(dau_handle_t* h, int which, float* ms_out, int capacity, int* count_out);
const char*  dau_timing_name(int which, int index);
/* number of kernels launched by this handle since creation (for bench.py's gpu_launches) */
long long    dau_launch_count(const dau_handle_t* h);

#ifdef __cplusplus
}
#endif
#endif /* DAU_CONV_B200_H_ */
