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
#include <cstdio>
#include <cstdlib>
#include <vector>

namespace dau {

HostPipeline::HostPipeline(const dau_settings_t& st, int device) : st_(st), device_(device) {
    DeviceGuard guard(device);
    DAU_CUDA_CHECK(cudaStreamCreateWithFlags(&s_in_, cudaStreamNonBlocking));
    DAU_CUDA_CHECK(cudaStreamCreateWithFlags(&s_comp_, cudaStreamNonBlocking));
    DAU_CUDA_CHECK(cudaStreamCreateWithFlags(&s_out_, cudaStreamNonBlocking));
    for (int b = 0; b < 2; ++b) {
        DAU_CUDA_CHECK(cudaEventCreateWithFlags(&in_done_[b], cudaEventDisableTiming));
        DAU_CUDA_CHECK(cudaEventCreateWithFlags(&comp_done_[b], cudaEventDisableTiming));
        DAU_CUDA_CHECK(cudaEventCreateWithFlags(&out_done_[b], cudaEventDisableTiming));
        DAU_CUDA_CHECK(cudaEventCreateWithFlags(&x_done_[b], cudaEventDisableTiming));
        DAU_CUDA_CHECK(cudaEventCreateWithFlags(&fwd_done_[b], cudaEventDisableTiming));
    }
    DAU_CUDA_CHECK(cudaEventCreateWithFlags(&fork_, cudaEventDisableTiming));
}

HostPipeline::~HostPipeline() {
    int prev = -1;
    cudaGetDevice(&prev);
    cudaSetDevice(device_);
    cudaDeviceSynchronize();
    for (auto& kv : layers_) delete kv.second;
    if (pool_) cudaFree(pool_);
    for (int b = 0; b < 2; ++b) {
        cudaEventDestroy(in_done_[b]); cudaEventDestroy(comp_done_[b]); cudaEventDestroy(out_done_[b]);
        cudaEventDestroy(x_done_[b]); cudaEventDestroy(fwd_done_[b]);
    }
    cudaEventDestroy(fork_);
    cudaStreamDestroy(s_in_); cudaStreamDestroy(s_comp_); cudaStreamDestroy(s_out_);
    if (prev >= 0 && prev != device_) cudaSetDevice(prev);
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

// Chunk schedule.  Three things pull against each other: the first copy-in and the last copy-out
// are the only transfers nothing can hide (wants short chunks at both ends), a chunk's copy-in must
// fit under the previous chunk's kernels and its copy-out under the next one's (wants a geometric
// ramp, the kernels being ~2.5x slower per image than the link), and a short chunk fills the last
// wave of the gather grids badly (wants sizes whose CTA counts are near-multiples of the SM count).
// The schedule is therefore searched: symmetric ramps [a, b, c, middle..., c, b, a] are run through
// a model of the three streams (double-buffered staging exactly as step() orders it) with the wave
// efficiency read from the kernels' own launch plans, and the fastest is kept per shape.
const std::vector<int>& HostPipeline::schedule(int N, int S, int H, int W, int cap, float sigma_hint, bool want_dx) {
    const std::vector<int> key = {N, S, H, W, cap, want_dx ? 1 : 0, (int)(sigma_hint * 1000.f)};
    auto it = schedules_.find(key);
    if (it != schedules_.end()) return it->second;

    const int F = st_.num_output, G = st_.number_units_x * st_.number_units_y;
    const int crop = (st_.kernel_size - 1) / 2 - st_.pad;
    const double link = 50e9;                                   // PCIe 5 x16, both directions busy (tools/pcie_probe.py)
    const double t_x = 4.0 * S * H * W / link, t_e = 4.0 * F * (H - 2 * crop) * (W - 2 * crop) / link;
    const double t_comp = 32.0 * F * S * G * H * W / 45e12;     // per image at the C2 rate
    const double t_fixed = 60e-6;                               // ~10 launches per chunk
    struct Slow { double fwd, bwd; };                           // time per image / t_comp, partial waves included
    std::map<int, Slow> slow;
    auto slowdown = [&](int n) {
        auto f = slow.find(n);
        if (f != slow.end()) return f->second;
        size_t fws, bws;
        DAUConvLayerB200* l = layer_for(n, S, H, W, sigma_hint, &fws, &bws);
        auto inv_eff = [&](long long ctas) {
            const long long sms = l->num_sms();
            return (double)((ctas + sms - 1) / sms * sms) / (double)ctas;
        };
        auto inv_eff_g = [&](const GatherPlan& g) {            // the gathers split their last wave
            return g.rounds * l->num_sms() / ((double)g.grid_x * g.grid_y * g.grid_z);
        };
        const GatherPlan& fp = l->fwd_plan(); const GatherPlan& bp = l->bwd_plan(); const GradPlan& gp = l->grad_plan();
        Slow v;
        v.fwd = 0.24 * inv_eff_g(fp);
        v.bwd = (want_dx ? 0.24 * inv_eff_g(bp) : 0.0) +
                0.52 * inv_eff((long long)gp.grid_x * gp.grid_y * gp.grid_z);
        return slow[n] = v;
    };
    auto simulate = [&](const std::vector<int>& sz) {          // mirrors the event graph of step()
        double in_end = 0, comp_end = 0, out_end = 0;
        std::vector<double> cd(sz.size()), od(sz.size());
        for (size_t k = 0; k < sz.size(); ++k) {
            const int n = sz[k];
            const Slow sl = slowdown(n);
            double t = in_end;
            if (k >= 2) t = std::max(t, cd[k - 2]);
            const double x_end = t + n * t_x;
            in_end = x_end + n * t_e;
            t = std::max(x_end, comp_end);
            if (k >= 2) t = std::max(t, od[k - 2]);
            const double f_end = t + n * t_comp * sl.fwd + 0.3 * t_fixed;
            const double y_end = std::max(f_end, out_end) + n * t_e;
            comp_end = cd[k] = std::max(f_end, in_end) + n * t_comp * sl.bwd + 0.7 * t_fixed;
            out_end = od[k] = std::max(comp_end, y_end) + (want_dx ? n * t_x : 0.0);
        }
        return out_end;
    };
    auto build = [&](int a, int b, int c) {
        std::vector<int> sz;
        const int ramp = a + b + c, mid = N - 2 * ramp;
        if (mid < 0) return sz;
        for (int v : {a, b, c}) if (v > 0) sz.push_back(v);
        if (mid > 0) {
            const int k = (mid + cap - 1) / cap;
            for (int i = 0, left = mid; i < k; ++i) {
                int n = (left / (k - i) + 1) / 2 * 2;           // even chunks: whole image pairs
                n = std::min(std::min(std::max(n, 1), left), cap);
                sz.push_back(n); left -= n;
            }
        }
        for (int v : {c, b, a}) if (v > 0) sz.push_back(v);
        return sz;
    };
    std::vector<int> best;
    for (int left = N; left > 0; left -= cap) best.push_back(std::min(cap, left));       // no ramp
    double best_t = simulate(best);
    const int top = std::min(cap, N / 2);
    for (int a = 2; a <= top; a += 2)
        for (int b = 0; b <= top; b += 2) {
            if (b != 0 && b <= a) continue;
            for (int c = 0; c <= top; c += 2) {
                if (c != 0 && (b == 0 || c <= b)) continue;
                const std::vector<int> sz = build(a, b, c);
                if (sz.empty()) continue;
                const double t = simulate(sz);
                if (t < best_t) { best_t = t; best = sz; }
            }
        }
    if (getenv("DAU_DEBUG_SCHEDULE")) {
        fprintf(stderr, "[dau] host-step schedule for N=%d cap=%d (model %.2f ms):", N, cap, best_t * 1e3);
        for (int n : best) fprintf(stderr, " %d", n);
        fprintf(stderr, "\n");
    }
    return schedules_[key] = best;
}

void HostPipeline::step(int N, int S, int H, int W, int chunk, float sigma_hint, const float* x,
                        const float* top_error, float* w, float* mu1, float* mu2, float* sigma, float* y,
                        float* dx, float* dw, float* dmu1, float* dmu2, float* dsigma, cudaStream_t stream) {
    DAU_CHECK(N > 0 && chunk > 0, 1, "batch and chunk must be positive");
    DAU_CHECK(x && top_error && w && mu1 && mu2 && dw && dmu1 && dmu2 && dsigma, 1, "NULL tensor pointer");
    DeviceGuard guard(device_);
    chunk = std::min(chunk, N);
    const int F = st_.num_output;
    const int crop = (st_.kernel_size - 1) / 2 - st_.pad;             // top is [N, F, H-2crop, W-2crop]
    const size_t in_img = (size_t)S * H * W, out_img = (size_t)F * (H - 2 * crop) * (W - 2 * crop);
    auto up = [](size_t v) { return (v + 255) / 256 * 256; };

    const std::vector<int>& sizes = schedule(N, S, H, W, chunk, sigma_hint, dx != nullptr);
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
        // copy-in: the buffers of chunk i-2 must have been consumed.  x first: the forward pass
        // starts as soon as it is there, the error follows while the forward kernels run
        if (i >= 2) DAU_CUDA_CHECK(cudaStreamWaitEvent(s_in_, comp_done_[b], 0));
        DAU_CUDA_CHECK(cudaMemcpyAsync(xd[b], x + n0 * in_img, n * in_img * 4, cudaMemcpyHostToDevice, s_in_));
        DAU_CUDA_CHECK(cudaEventRecord(x_done_[b], s_in_));
        DAU_CUDA_CHECK(cudaMemcpyAsync(ed[b], top_error + n0 * out_img, n * out_img * 4, cudaMemcpyHostToDevice, s_in_));
        DAU_CUDA_CHECK(cudaEventRecord(in_done_[b], s_in_));
        // forward: needs x, and the y/dx buffers of chunk i-2 copied out
        DAU_CUDA_CHECK(cudaStreamWaitEvent(s_comp_, x_done_[b], 0));
        if (i >= 2) DAU_CUDA_CHECK(cudaStreamWaitEvent(s_comp_, out_done_[b], 0));
        l->Forward_gpu(xd[b], w, mu1, mu2, sigma, nullptr, yd[b], ws, ws_bytes, s_comp_);
        DAU_CUDA_CHECK(cudaEventRecord(fwd_done_[b], s_comp_));
        // y leaves while the backward pass runs
        DAU_CUDA_CHECK(cudaStreamWaitEvent(s_out_, fwd_done_[b], 0));
        if (y) DAU_CUDA_CHECK(cudaMemcpyAsync(y + n0 * out_img, yd[b], n * out_img * 4, cudaMemcpyDeviceToHost, s_out_));
        // backward: needs the error
        DAU_CUDA_CHECK(cudaStreamWaitEvent(s_comp_, in_done_[b], 0));
        l->Backward_gpu(xd[b], ed[b], w, mu1, mu2, sigma, dx ? dxd[b] : nullptr, dw, dmu1, dmu2, dsigma,
                        nullptr, flags, ws, ws_bytes, s_comp_);
        DAU_CUDA_CHECK(cudaEventRecord(comp_done_[b], s_comp_));
        DAU_CUDA_CHECK(cudaStreamWaitEvent(s_out_, comp_done_[b], 0));
        if (dx) DAU_CUDA_CHECK(cudaMemcpyAsync(dx + n0 * in_img, dxd[b], n * in_img * 4, cudaMemcpyDeviceToHost, s_out_));
        DAU_CUDA_CHECK(cudaEventRecord(out_done_[b], s_out_));
    }
    // join: the caller's stream continues after the last copy-out (which follows all compute)
    const int last = (i - 1) & 1;
    DAU_CUDA_CHECK(cudaStreamWaitEvent(stream, out_done_[last], 0));
    if (i >= 2) DAU_CUDA_CHECK(cudaStreamWaitEvent(stream, out_done_[last ^ 1], 0));
}

}  // namespace dau
