// HostPipeline: chunked, copy/compute-overlapped forward+backward on host-resident batches
// (see dau_host_step.cpp).  Owned by a dau_handle_t; created on first use.
#pragma once
#include "dau_layer.hpp"

#include <map>
#include <vector>

namespace dau {

class HostPipeline {
public:
    HostPipeline(const dau_settings_t& st, int device);
    ~HostPipeline();

    // x, top_error, y, dx: HOST pointers (pinned memory for real overlap), NCHW fp32.
    // w, mu1, mu2, sigma, dw, dmu1, dmu2, dsigma: DEVICE pointers [1,S,G,F]; the parameter
    // gradients are accumulated into.  y / dx may be NULL (not copied back / not computed).
    void step(int N, int S, int H, int W, int chunk, float sigma_hint, const float* x, const float* top_error,
              float* w, float* mu1, float* mu2, float* sigma, float* y, float* dx, float* dw, float* dmu1,
              float* dmu2, float* dsigma, cudaStream_t stream);
    long long launches() const;

private:
    DAUConvLayerB200* layer_for(int n, int S, int H, int W, float sigma_hint, size_t* fws, size_t* bws);
    const std::vector<int>& schedule(int N, int S, int H, int W, int cap, float sigma_hint, bool want_dx);

    dau_settings_t st_;
    int device_;
    std::map<int, DAUConvLayerB200*> layers_;      // one planned layer per chunk size
    std::map<std::vector<int>, std::vector<int>> schedules_;   // (N,S,H,W,cap,dx) -> chunk sizes
    cudaStream_t s_in_ = nullptr, s_comp_ = nullptr, s_out_ = nullptr;
    cudaEvent_t in_done_[2], comp_done_[2], out_done_[2], x_done_[2], fwd_done_[2], fork_;
    void* pool_ = nullptr;                          // staging buffers + layer workspace
    size_t pool_bytes_ = 0;
};

}  // namespace dau
