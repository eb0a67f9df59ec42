// Host-buffer training step: forward + backward of one DAU layer on a batch that lives in HOST
// memory, pipelined over chunks of images so that the PCIe copies hide behind the kernels.
//
// The reference's op receives device tensors from its framework and ends every call with a
// device-wide synchronisation (plugins/tensorflow/src/dau_conv_op.cpp:311); a caller outside
// such a framework pays  H2D(x, top_error) -> forward -> backward -> D2H(y, dL/dx)  in sequence.
// Here the batch is cut into chunks that flow through three streams
//     copy-in  : x_c, e_c  host -> device            (double-buffered)
//     compute  : Forward_gpu(x_c) ; Backward_gpu(x_c, e_c)   -- parameter gradients ACCUMULATE
//                into the caller's buffers, exactly the contract of Backward_gpu
//                (base_dau_conv_layer.cu:340-343), so chunking changes nothing but the order of
//                the fp32 additions
//     copy-out : y_c, dx_c device -> host
// ordered by events only; the call itself is asynchronous with respect to the host and ordered
// after / before the caller's stream like any other stream-ordered call.
#include "dau_host_step.hpp"

#include <algorithm>
#include <vector>

namespace dau {

HostPipeline::HostPipeline(const dau_settings_t& st, int device) : st_(st), device_(device) {
    DAU_CUDA_CHECK(cudaSetDevice(device));
    DAU_CUDA_CHECK(cudaStreamCreateWithFlags(&s_in_, cudaStreamNonBlocking));
    DAU_CUDA_CHECK(cudaStreamCreateWithFlags(&s_comp_, cudaStreamNonBlocking));
    DAU_CUDA_CHECK(cudaStreamCreateWithFlags(&s_out_, cudaStreamNonBlocking));
    for (int b = 0; b < 2; ++b) {
        DAU_CUDA_CHECK(cudaEventCreateWithFlags(&in_done_[b], cudaEventDisableTiming));
        DAU_CUDA_CHECK(cudaEventCreateWithFlags(&comp_done_[b], cudaEventDisableTiming));
        DAU_CUDA_CHECK(cudaEventCreateWithFlags(&out_done_[b], cudaEventDisableTiming));
    }
    DAU_CUDA_CHECK(cudaEventCreateWithFlags(&fork_, cudaEventDisableTiming));
}

HostPipeline::~HostPipeline() {
    cudaSetDevice(device_);
    cudaDeviceSynchronize();
    for (auto& kv : layers_) delete kv.second;
    if (pool_) cudaFree(pool_);
    for (int b = 0; b < 2; ++b) {
        cudaEventDestroy(in_done_[b]); cudaEventDestroy(comp_done_[b]); cudaEventDestroy(out_done_[b]);
    }
    cudaEventDestroy(fork_);
    cudaStreamDestroy(s_in_); cudaStreamDestroy(s_comp_); cudaStreamDestroy(s_out_);
}

DAUConvLayerB200* HostPipeline::layer_for(int n, int S, int H, int W, float sigma_hint, size_t* fws, size_t* bws) {
    DAUConvLayerB200*& l = layers_[n];
    if (!l) l = new DAUConvLayerB200(st_, device_);
    l->set_sigma_hint(sigma_hint);
    l->Reshape(n, S, H, W, fws, bws);
    return l;
}

long long HostPipeline::launches() const {
    long long t = 0;
    for (auto& kv : layers_) t += kv.second->launches();
    return t;
}

void HostPipeline::step(int N, int S, int H, int W, int chunk, float sigma_hint, const float* x,
                        const float* top_error, float* w, float* mu1, float* mu2, float* sigma, float* y,
                        float* dx, float* dw, float* dmu1, float* dmu2, float* dsigma, cudaStream_t stream) {
    DAU_CHECK(N > 0 && chunk > 0, 1, "batch and chunk must be positive");
    DAU_CHECK(x && top_error && w && mu1 && mu2 && dw && dmu1 && dmu2 && dsigma, 1, "NULL tensor pointer");
    DAU_CUDA_CHECK(cudaSetDevice(device_));
    chunk = std::min(chunk, N);
    const int F = st_.num_output;
    const int crop = (st_.kernel_size - 1) / 2 - st_.pad;             // top is [N, F, H-2crop, W-2crop]
    const size_t in_img = (size_t)S * H * W, out_img = (size_t)F * (H - 2 * crop) * (W - 2 * crop);
    auto up = [](size_t v) { return (v + 255) / 256 * 256; };

    // chunk schedule: short chunks at both ends (the first copy-in and the last copy-out are the
    // only transfers nothing can hide), full chunks in between
    std::vector<int> sizes;
    {
        const int q = chunk / 4, h = chunk / 2;
        if (q >= 1 && N >= 2 * (q + h) + chunk) {
            int mid = N - 2 * (q + h);
            sizes.push_back(q); sizes.push_back(h);
            while (mid > 0) { const int n = std::min(chunk, mid); sizes.push_back(n); mid -= n; }
            sizes.push_back(h); sizes.push_back(q);
        } else {
            for (int left = N; left > 0; left -= chunk) sizes.push_back(std::min(chunk, left));
        }
    }
    // plan every chunk shape and size the pool
    size_t fws = 0, bws = 0, ws_bytes = 0;
    for (int n : sizes) {
        layer_for(n, S, H, W, sigma_hint, &fws, &bws);
        ws_bytes = std::max(ws_bytes, std::max(fws, bws));
    }
    const size_t xin = up(chunk * in_img * 4), eout = up(chunk * out_img * 4);
    const size_t need = 2 * (2 * xin + 2 * eout) + up(ws_bytes);
    if (need > pool_bytes_) {
        DAU_CUDA_CHECK(cudaDeviceSynchronize());
        if (pool_) cudaFree(pool_);
        pool_ = nullptr; pool_bytes_ = 0;
        DAU_CUDA_CHECK(cudaMalloc(&pool_, need));
        pool_bytes_ = need;
    }
    char* base = static_cast<char*>(pool_);
    float *xd[2], *ed[2], *yd[2], *dxd[2];
    for (int b = 0; b < 2; ++b) {
        xd[b] = reinterpret_cast<float*>(base);  base += xin;
        dxd[b] = reinterpret_cast<float*>(base); base += xin;
        ed[b] = reinterpret_cast<float*>(base);  base += eout;
        yd[b] = reinterpret_cast<float*>(base);  base += eout;
    }
    void* ws = base;

    // the three internal streams start after everything already queued on the caller's stream
    DAU_CUDA_CHECK(cudaEventRecord(fork_, stream));
    DAU_CUDA_CHECK(cudaStreamWaitEvent(s_in_, fork_, 0));
    DAU_CUDA_CHECK(cudaStreamWaitEvent(s_comp_, fork_, 0));
    DAU_CUDA_CHECK(cudaStreamWaitEvent(s_out_, fork_, 0));

    const unsigned flags = DAU_GRAD_W | DAU_GRAD_MU1 | DAU_GRAD_MU2 | DAU_GRAD_SIGMA | (dx ? DAU_GRAD_INPUT : 0u);
    int i = 0;
    size_t n0 = 0;
    for (size_t ci = 0; ci < sizes.size(); n0 += sizes[ci], ++ci, ++i) {
        const int n = sizes[ci], b = i & 1;
        DAUConvLayerB200* l = layer_for(n, S, H, W, sigma_hint, &fws, &bws);
        // copy-in: the buffers of chunk i-2 must have been consumed
        if (i >= 2) DAU_CUDA_CHECK(cudaStreamWaitEvent(s_in_, comp_done_[b], 0));
        DAU_CUDA_CHECK(cudaMemcpyAsync(xd[b], x + n0 * in_img, n * in_img * 4, cudaMemcpyHostToDevice, s_in_));
        DAU_CUDA_CHECK(cudaMemcpyAsync(ed[b], top_error + n0 * out_img, n * out_img * 4, cudaMemcpyHostToDevice, s_in_));
        DAU_CUDA_CHECK(cudaEventRecord(in_done_[b], s_in_));
        // compute: needs the inputs, and y/dx buffers of chunk i-2 copied out
        DAU_CUDA_CHECK(cudaStreamWaitEvent(s_comp_, in_done_[b], 0));
        if (i >= 2) DAU_CUDA_CHECK(cudaStreamWaitEvent(s_comp_, out_done_[b], 0));
        l->Forward_gpu(xd[b], w, mu1, mu2, sigma, nullptr, yd[b], ws, ws_bytes, s_comp_);
        l->Backward_gpu(xd[b], ed[b], w, mu1, mu2, sigma, dx ? dxd[b] : nullptr, dw, dmu1, dmu2, dsigma,
                        nullptr, flags, ws, ws_bytes, s_comp_);
        DAU_CUDA_CHECK(cudaEventRecord(comp_done_[b], s_comp_));
        // copy-out
        DAU_CUDA_CHECK(cudaStreamWaitEvent(s_out_, comp_done_[b], 0));
        if (y) DAU_CUDA_CHECK(cudaMemcpyAsync(y + n0 * out_img, yd[b], n * out_img * 4, cudaMemcpyDeviceToHost, s_out_));
        if (dx) DAU_CUDA_CHECK(cudaMemcpyAsync(dx + n0 * in_img, dxd[b], n * in_img * 4, cudaMemcpyDeviceToHost, s_out_));
        DAU_CUDA_CHECK(cudaEventRecord(out_done_[b], s_out_));
    }
    // join: the caller's stream continues after the last copy-out (which follows all compute)
    const int last = (i - 1) & 1;
    DAU_CUDA_CHECK(cudaStreamWaitEvent(stream, out_done_[last], 0));
    if (i >= 2) DAU_CUDA_CHECK(cudaStreamWaitEvent(stream, out_done_[last ^ 1], 0));
}

}  // namespace dau
