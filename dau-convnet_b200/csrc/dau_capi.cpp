// extern "C" boundary of libdau_conv_b200.so (include/dau_conv_b200.h).  Exceptions of the
// C++ layer (DAUException, the reference's error convention) become status codes plus a
// thread-local message.
#include "dau_host_step.hpp"
#include "dau_layer.hpp"

#include <cstdio>
#include <cstring>
#include <new>

using dau::DAUConvLayerB200;

struct dau_handle_s { DAUConvLayerB200* layer; dau::HostPipeline* pipe; };

namespace {
thread_local char g_err[1024] = "";

dau_status_t fail(int status, const char* msg) {
    std::snprintf(g_err, sizeof(g_err), "%s", msg);
    return (dau_status_t)status;
}

template <class Fn>
dau_status_t guarded(Fn&& fn) {
    try {
        fn();
        return DAU_OK;
    } catch (const dau::DAUException& e) {
        return fail(e.status, e.what());
    } catch (const std::bad_alloc&) {
        return fail(DAU_ERR_CUDA, "host allocation failed");
    } catch (const std::exception& e) {
        return fail(DAU_ERR_CUDA, e.what());
    }
}

dau::FilterOptions filter_options(const dau_settings_t* s) {
    dau::FilterOptions o;
    o.unit_normalization = s->unit_normalization != 0;
    o.square_unit_normalization = s->square_unit_normalization != 0;
    o.single_dimension_kernel = s->single_dimension_kernel != 0;
    o.forbid_positive_dim1 = s->forbid_positive_dim1 != 0;
    return o;
}

struct DevBuf {   // RAII device buffer for the *_host convenience calls
    void* p = nullptr;
    explicit DevBuf(size_t bytes) { if (bytes) DAU_CUDA_CHECK(cudaMalloc(&p, bytes)); }
    ~DevBuf() { if (p) cudaFree(p); }
    float* f() { return static_cast<float*>(p); }
};
void h2d(DevBuf& d, const float* src, size_t n) {
    if (src && n) DAU_CUDA_CHECK(cudaMemcpy(d.p, src, n * sizeof(float), cudaMemcpyHostToDevice));
}
void d2h(float* dst, DevBuf& s, size_t n) {
    if (dst && n) DAU_CUDA_CHECK(cudaMemcpy(dst, s.p, n * sizeof(float), cudaMemcpyDeviceToHost));
}
}  // namespace

extern "C" {

const char* dau_last_error(void) { return g_err; }
int dau_abi_version(void) { return DAU_B200_ABI_VERSION; }

void dau_settings_default(dau_settings_t* s) {
    if (!s) return;
    std::memset(s, 0, sizeof(*s));
    // defaults of the TF op attrs (plugins/tensorflow/src/dau_conv_op.cpp:29-48)
    s->num_output = 64; s->number_units_x = 2; s->number_units_y = 2; s->number_units_ignore = 0;
    s->kernel_size = 9; s->pad = 4; s->stride = 1; s->bias_term = 0;
    s->unit_normalization = 1; s->square_unit_normalization = 0; s->use_interpolation = 1;
    s->single_dimension_kernel = 0; s->forbid_positive_dim1 = 0; s->ignore_edge_gradients = 0;
    s->unit_bounds_guard = 0;          // the TF op disables it (dau_conv_op.cpp:283)
    s->offsets_already_centered = 1; s->learn_sigma = 1;
    s->component_border_bound = 1.f; s->sigma_lower_bound = 0.3f; s->mu_learning_rate_factor = 1.f;
    s->mean_iteration_step = 1; s->sigma_iteration_step = 1; s->merge_iteration_step = 0;
    s->merge_threshold = 1.f;
}

dau_status_t dau_create(const dau_settings_t* settings, int device, dau_handle_t** out) {
    if (!settings || !out) return fail(DAU_ERR_INVALID_ARGUMENT, "dau_create: NULL argument");
    *out = nullptr;
    return guarded([&] {
        DAUConvLayerB200* l = new DAUConvLayerB200(*settings, device);
        dau_handle_t* h = new dau_handle_t;
        h->layer = l;
        h->pipe = nullptr;
        *out = h;
    });
}

void dau_destroy(dau_handle_t* h) {
    if (!h) return;
    delete h->pipe;
    delete h->layer;
    delete h;
}

dau_status_t dau_reshape(dau_handle_t* h, int N, int S, int H, int W, size_t* fwd_ws, size_t* bwd_ws) {
    if (!h) return fail(DAU_ERR_INVALID_ARGUMENT, "dau_reshape: NULL handle");
    return guarded([&] { h->layer->Reshape(N, S, H, W, fwd_ws, bwd_ws); });
}

dau_status_t dau_set_sigma_hint(dau_handle_t* h, float value) {
    if (!h) return fail(DAU_ERR_INVALID_ARGUMENT, "dau_set_sigma_hint: NULL handle");
    h->layer->set_sigma_hint(value);
    return DAU_OK;
}

dau_status_t dau_set_offset_hint(dau_handle_t* h, float max_abs_offset) {
    if (!h) return fail(DAU_ERR_INVALID_ARGUMENT, "dau_set_offset_hint: NULL handle");
    h->layer->set_offset_hint(max_abs_offset);
    return DAU_OK;
}

dau_status_t dau_forward(dau_handle_t* h, const float* bottom, float* w, float* mu1, float* mu2,
                         float* sigma, const float* bias, float* top, void* ws, size_t ws_bytes,
                         void* stream) {
    if (!h) return fail(DAU_ERR_INVALID_ARGUMENT, "dau_forward: NULL handle");
    return guarded([&] {
        h->layer->Forward_gpu(bottom, w, mu1, mu2, sigma, bias, top, ws, ws_bytes, (cudaStream_t)stream);
    });
}

dau_status_t dau_backward(dau_handle_t* h, const float* bottom, const float* top_error, const float* w,
                          const float* mu1, const float* mu2, const float* sigma, float* bottom_error,
                          float* dw, float* dmu1, float* dmu2, float* dsigma, float* dbias,
                          unsigned flags, void* ws, size_t ws_bytes, void* stream) {
    if (!h) return fail(DAU_ERR_INVALID_ARGUMENT, "dau_backward: NULL handle");
    return guarded([&] {
        h->layer->Backward_gpu(bottom, top_error, w, mu1, mu2, sigma, bottom_error, dw, dmu1, dmu2,
                               dsigma, dbias, flags, ws, ws_bytes, (cudaStream_t)stream);
    });
}

dau_status_t dau_get_output_size(dau_handle_t* h, int* h_out, int* w_out) {
    if (!h || !h_out || !w_out) return fail(DAU_ERR_INVALID_ARGUMENT, "dau_get_output_size: NULL argument");
    return guarded([&] {
        DAU_CHECK(h->layer->H() > 0, 3, "dau_get_output_size called before dau_reshape");
        *h_out = h->layer->H_out(); *w_out = h->layer->W_out();
    });
}

dau_status_t dau_take_nan_flag(dau_handle_t* h, int* found, void* stream) {
    if (!h || !found) return fail(DAU_ERR_INVALID_ARGUMENT, "dau_take_nan_flag: NULL argument");
    return guarded([&] { *found = h->layer->take_nan_flag((cudaStream_t)stream) ? 1 : 0; });
}

dau_status_t dau_set_activation(dau_handle_t* h, int activation) {
    if (!h) return fail(DAU_ERR_INVALID_ARGUMENT, "dau_set_activation: NULL handle");
    return guarded([&] { h->layer->set_activation(activation); });
}

dau_status_t dau_backward_act(dau_handle_t* h, const float* bottom, const float* top_data, const float* top_error,
                              const float* w, const float* mu1, const float* mu2, const float* sigma,
                              float* bottom_error, float* dw, float* dmu1, float* dmu2, float* dsigma,
                              float* dbias, unsigned flags, void* ws, size_t ws_bytes, void* stream) {
    if (!h) return fail(DAU_ERR_INVALID_ARGUMENT, "dau_backward_act: NULL handle");
    return guarded([&] {
        h->layer->Backward_gpu(bottom, top_error, w, mu1, mu2, sigma, bottom_error, dw, dmu1, dmu2,
                               dsigma, dbias, flags, ws, ws_bytes, (cudaStream_t)stream, top_data);
    });
}

dau_status_t dau_get_kernels(const dau_settings_t* settings, float sigma, int* kb_out, float* out6,
                             size_t cap) {
    if (!settings || !kb_out || !out6) return fail(DAU_ERR_INVALID_ARGUMENT, "dau_get_kernels: NULL argument");
    return guarded([&] {
        DAU_CHECK(sigma > 0.f, 1, "sigma must be positive");
        const int kb = dau::blur_kernel_size(sigma);
        DAU_CHECK(kb > 1 && kb <= dau::kMaxBlurTaps, 4, "blur support 2*ceil(5*sigma)+1 must be in (1,17]");
        DAU_CHECK(cap >= (size_t)6 * kb * kb, 1, "output buffer too small for 6*Kb*Kb floats");
        dau::compute_dense_kernels(sigma, filter_options(settings), kb_out, out6);
    });
}

dau_status_t dau_forward_host(dau_handle_t* h, int N, int S, int H, int W, const float* bottom,
                              const float* w, const float* mu1, const float* mu2, const float* sigma,
                              const float* bias, float* top) {
    if (!h) return fail(DAU_ERR_INVALID_ARGUMENT, "dau_forward_host: NULL handle");
    return guarded([&] {
        DAUConvLayerB200* l = h->layer;
        dau::DeviceGuard guard(l->device());
        size_t fws = 0, bws = 0;
        l->Reshape(N, S, H, W, &fws, &bws);
        const size_t nin = (size_t)N * S * H * W, nout = (size_t)N * l->F() * l->H_out() * l->W_out();
        const size_t np = (size_t)S * l->G() * l->F();
        DevBuf dx(nin * 4), dy(nout * 4), dwt(np * 4), dm1(np * 4), dm2(np * 4), dsg(np * 4),
            db(bias ? (size_t)l->F() * 4 : 0), dws(fws);
        h2d(dx, bottom, nin); h2d(dwt, w, np); h2d(dm1, mu1, np); h2d(dm2, mu2, np); h2d(dsg, sigma, np);
        if (bias) h2d(db, bias, l->F());
        l->Forward_gpu(dx.f(), dwt.f(), dm1.f(), dm2.f(), dsg.f(), bias ? db.f() : nullptr, dy.f(),
                       dws.p, fws, nullptr);
        DAU_CUDA_CHECK(cudaDeviceSynchronize());
        d2h(top, dy, nout);
    });
}

dau_status_t dau_backward_host(dau_handle_t* h, int N, int S, int H, int W, const float* bottom,
                               const float* top_error, const float* w, const float* mu1,
                               const float* mu2, const float* sigma, float* bottom_error, float* dw,
                               float* dmu1, float* dmu2, float* dsigma, float* dbias, unsigned flags) {
    if (!h) return fail(DAU_ERR_INVALID_ARGUMENT, "dau_backward_host: NULL handle");
    return guarded([&] {
        DAUConvLayerB200* l = h->layer;
        dau::DeviceGuard guard(l->device());
        size_t fws = 0, bws = 0;
        l->Reshape(N, S, H, W, &fws, &bws);
        const size_t nin = (size_t)N * S * H * W, nout = (size_t)N * l->F() * l->H_out() * l->W_out();
        const size_t np = (size_t)S * l->G() * l->F();
        DevBuf dx(nin * 4), de(nout * 4), dwt(np * 4), dm1(np * 4), dm2(np * 4), dsg(np * 4),
            gdx(nin * 4), gw(np * 4), gm1(np * 4), gm2(np * 4), gsg(np * 4), gb((size_t)l->F() * 4), dws(bws);
        h2d(dx, bottom, nin); h2d(de, top_error, nout);
        h2d(dwt, w, np); h2d(dm1, mu1, np); h2d(dm2, mu2, np); h2d(dsg, sigma, np);
        // accumulate-into semantics: start from the caller's values
        h2d(gw, dw, np); h2d(gm1, dmu1, np); h2d(gm2, dmu2, np); h2d(gsg, dsigma, np);
        if (dbias) h2d(gb, dbias, l->F());
        l->Backward_gpu(dx.f(), de.f(), dwt.f(), dm1.f(), dm2.f(), dsg.f(), gdx.f(), gw.f(), gm1.f(),
                        gm2.f(), gsg.f(), dbias ? gb.f() : nullptr, flags, dws.p, bws, nullptr);
        DAU_CUDA_CHECK(cudaDeviceSynchronize());
        if (flags & DAU_GRAD_INPUT) d2h(bottom_error, gdx, nin);
        d2h(dw, gw, np); d2h(dmu1, gm1, np); d2h(dmu2, gm2, np); d2h(dsigma, gsg, np);
        if (dbias) d2h(dbias, gb, l->F());
    });
}

dau_status_t dau_train_step_host(dau_handle_t* h, int N, int S, int H, int W, int chunk_images,
                                 float sigma_value, const float* x_host, const float* top_error_host,
                                 float* w, float* mu1, float* mu2, float* sigma, float* y_host,
                                 float* dx_host, float* dw, float* dmu1, float* dmu2, float* dsigma,
                                 void* stream) {
    if (!h) return fail(DAU_ERR_INVALID_ARGUMENT, "dau_train_step_host: NULL handle");
    return guarded([&] {
        DAU_CHECK(!h->layer->settings().bias_term, 4, "dau_train_step_host: bias_term is not supported");
        if (!h->pipe) h->pipe = new dau::HostPipeline(h->layer->settings(), h->layer->device());
        h->pipe->step(N, S, H, W, chunk_images, sigma_value, x_host, top_error_host, w, mu1, mu2, sigma,
                      y_host, dx_host, dw, dmu1, dmu2, dsigma, (cudaStream_t)stream);
    });
}

dau_status_t dau_plan_mosaic(int N, int H, int W, int gap, int* out10) {
    if (!out10) return fail(DAU_ERR_INVALID_ARGUMENT, "dau_plan_mosaic: NULL output");
    return guarded([&] {
        DAU_CHECK(N > 0 && H > 0 && W > 0 && gap >= 0, 1, "dau_plan_mosaic: sizes must be positive");
        const dau::Mosaic m = dau::plan_mosaic(N, H, W, gap);
        const int v[10] = {m.kx, m.ky, m.px, m.py, m.W, m.H, m.Wv, m.Hv, m.Nv, m.gap};
        std::memcpy(out10, v, sizeof(v));
    });
}

dau_status_t dau_plan_strips(int N, int H, int W, int gap, long long* counts2, int* items_out, long long capacity) {
    if (!counts2) return fail(DAU_ERR_INVALID_ARGUMENT, "dau_plan_strips: NULL output");
    return guarded([&] {
        DAU_CHECK(N > 0 && H > 0 && W > 0 && gap >= 0, 1, "dau_plan_strips: sizes must be positive");
        const dau::Mosaic m = dau::plan_mosaic(N, H, W, gap);
        const dau::StripGeom g = dau::strip_geometry(m, 1);
        counts2[0] = g.full_items; counts2[1] = g.tail_items;
        if (!items_out) return;
        DAU_CHECK(capacity >= g.full_items + g.tail_items, 1, "dau_plan_strips: items_out too small");
        long long k = 0;
        for (int tail = 0; tail < 2; ++tail)
            for (long long i = 0; i < (tail ? g.tail_items : g.full_items); ++i, ++k)
                dau::strip_item_origin(m, i, tail != 0, &items_out[4 * k], &items_out[4 * k + 1], &items_out[4 * k + 2],
                                       &items_out[4 * k + 3]);
    });
}

int dau_mosaic_to_real(const int* q, int N, int nv, int vy, int vx, int* n, int* y, int* x) {
    if (!q || !n || !y || !x) return 0;
    dau::Mosaic m;
    m.kx = q[0]; m.ky = q[1]; m.px = q[2]; m.py = q[3]; m.W = q[4]; m.H = q[5]; m.Wv = q[6]; m.Hv = q[7];
    m.Nv = q[8]; m.gap = q[9]; m.N = N;
    if (nv < 0 || vy < 0 || vx < 0 || vy >= m.Hv || vx >= m.Wv || nv >= m.Nv) return 0;
    return dau::mosaic_to_real(m, nv, vy, vx, n, y, x) ? 1 : 0;
}

dau_status_t dau_bias_grad_static(int device, const float* top_error, const float* top_data, float* dbias,
                                  int N, int F, int H_out, int W_out, void* stream) {
    return guarded([&] {
        dau::DeviceGuard guard(device);                 // negative: the caller's current device
        DAU_CHECK(H_out > 0 && W_out > 0, 1, "dau_bias_grad_static: N, F, H_out and W_out must be positive");
        dau::launch_bias_grad_static(top_error, top_data, dbias, N, F, H_out * W_out, (cudaStream_t)stream);
    });
}

int dau_edge_mask_rule(int extent) { return DAUConvLayerB200::edge_mask_rule(extent) ? 1 : 0; }

dau_status_t dau_debug_unit_tables(dau_handle_t* h, const float* w, const float* mu1, const float* mu2,
                                   void* fwd_out, void* bwd_out, void* grad_out, int* halos4, void* stream) {
    if (!h) return fail(DAU_ERR_INVALID_ARGUMENT, "dau_debug_unit_tables: NULL handle");
    return guarded([&] {
        h->layer->debug_unit_tables(w, mu1, mu2, static_cast<dau::UnitRec*>(fwd_out),
                                    static_cast<dau::UnitRec*>(bwd_out), static_cast<dau::UnitRec*>(grad_out),
                                    halos4, (cudaStream_t)stream);
    });
}

dau_status_t dau_measure_fma_peak(int device, int iters, double* tflops_out, double* ms_out) {
    return guarded([&] {
        dau::DeviceGuard guard(device);
        cudaDeviceProp prop;
        DAU_CUDA_CHECK(cudaGetDeviceProperties(&prop, device));
        const int blocks = prop.multiProcessorCount * 8;
        DevBuf sink(256);
        cudaEvent_t a, b;
        DAU_CUDA_CHECK(cudaEventCreate(&a));
        DAU_CUDA_CHECK(cudaEventCreate(&b));
        dau::launch_fma_probe(sink.f(), iters / 8 + 1, blocks, nullptr);   // warm-up
        DAU_CUDA_CHECK(cudaEventRecord(a, nullptr));
        dau::launch_fma_probe(sink.f(), iters, blocks, nullptr);
        DAU_CUDA_CHECK(cudaEventRecord(b, nullptr));
        DAU_CUDA_CHECK(cudaEventSynchronize(b));
        float ms = 0.f;
        DAU_CUDA_CHECK(cudaEventElapsedTime(&ms, a, b));
        cudaEventDestroy(a); cudaEventDestroy(b);
        const double flops = (double)blocks * 256.0 * (double)iters * 64.0 * 2.0;
        if (tflops_out) *tflops_out = flops / (ms * 1e-3) / 1e12;
        if (ms_out) *ms_out = ms;
    });
}

dau_status_t dau_enable_timing(dau_handle_t* h, int enable) {
    if (!h) return fail(DAU_ERR_INVALID_ARGUMENT, "dau_enable_timing: NULL handle");
    return guarded([&] { h->layer->enable_timing(enable != 0); });
}

dau_status_t dau_get_timing(dau_handle_t* h, int which, float* ms_out, int capacity, int* count_out) {
    if (!h || !ms_out || !count_out) return fail(DAU_ERR_INVALID_ARGUMENT, "dau_get_timing: NULL argument");
    *count_out = h->layer->get_timing(which, ms_out, capacity);
    return DAU_OK;
}

const char* dau_timing_name(int which, int index) { return DAUConvLayerB200::timing_name(which, index); }

long long dau_launch_count(const dau_handle_t* h) {
    return h ? h->layer->launches() + (h->pipe ? h->pipe->launches() : 0) : 0;
}

}  // extern "C"
