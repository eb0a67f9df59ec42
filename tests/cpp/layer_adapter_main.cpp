// Test driver for include/dau_conv_b200_layer.hpp: a host written against the reference's
// BaseDAUConvLayer<float> (a subclass that lends device buffers through the param_* virtuals,
// base_dau_conv_layer.hpp:229-264, the way DAUConvLayerTensorflowGPU does) runs
// LayerSetUp -> Reshape -> Forward_gpu -> Backward_gpu and dumps the results for
// tests/test_cpp_adapter.py to compare against the oracle.  Test infrastructure, not product.
//
// usage: layer_adapter_main <in.bin> <out.bin>
//   in : int32 {N,S,F,Gx,Gy,H,W,ks,pad,ignore_edge,bias_term,learn_sigma}, float mu_lr,
//        then x[N,S,H,W], err[N,F,Ho,Wo], w, mu1, mu2, sigma [S*G*F] each, bias[F]
//   out: int32 {Ho,Wo}, y, dx, dw, dmu1, dmu2, dsigma, dbias
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>
#include <vector>

#include "dau_conv_b200_layer.hpp"

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { std::fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e_)); std::exit(3); } } while (0)

namespace {
struct DevBuf {
    float* p = nullptr; size_t n = 0;
    void alloc(size_t count) { n = count; CK(cudaMalloc(&p, (count ? count : 1) * sizeof(float))); CK(cudaMemset(p, 0, (count ? count : 1) * sizeof(float))); }
    void put(const std::vector<float>& h) { alloc(h.size()); if (n) CK(cudaMemcpy(p, h.data(), n * sizeof(float), cudaMemcpyHostToDevice)); }
    std::vector<float> get() const { std::vector<float> h(n); if (n) CK(cudaMemcpy(h.data(), p, n * sizeof(float), cudaMemcpyDeviceToHost)); return h; }
    ~DevBuf() { cudaFree(p); }
};

class PlainCudaDAULayer : public dau_b200::BaseDAUConvLayer {
 public:
    using BaseDAUConvLayer::BaseDAUConvLayer;
    DevBuf w, mu1, mu2, sigma, bias, dw, dmu1, dmu2, dsigma, dbias;
    int allocations = 0;
    ~PlainCudaDAULayer() override { deallocate_workspace_mem(); }
 protected:
    float* param_w() override { return w.p; }
    float* param_mu1() override { return mu1.p; }
    float* param_mu2() override { return mu2.p; }
    float* param_sigma() override { return sigma.p; }
    float* param_bias() override { return bias.p; }
    float* param_w_grad() override { return dw.p; }
    float* param_mu1_grad() override { return dmu1.p; }
    float* param_mu2_grad() override { return dmu2.p; }
    float* param_sigma_grad() override { return dsigma.p; }
    float* param_bias_grad() override { return dbias.p; }
    void* allocate_workspace_mem(size_t bytes) override { ++allocations; CK(cudaMalloc(&ws_, bytes ? bytes : 256)); return ws_; }
    void deallocate_workspace_mem() override { if (ws_) cudaFree(ws_); ws_ = nullptr; }
 private:
    void* ws_ = nullptr;
};

std::vector<float> read_floats(FILE* f, size_t n) {
    std::vector<float> v(n);
    if (n && std::fread(v.data(), sizeof(float), n, f) != n) { std::fprintf(stderr, "short input file\n"); std::exit(2); }
    return v;
}
void write_floats(FILE* f, const std::vector<float>& v) { if (!v.empty()) std::fwrite(v.data(), sizeof(float), v.size(), f); }
}  // namespace

int main(int argc, char** argv) {
    if (argc != 3) { std::fprintf(stderr, "usage: %s in.bin out.bin\n", argv[0]); return 2; }
    FILE* fi = std::fopen(argv[1], "rb");
    if (!fi) { std::perror(argv[1]); return 2; }
    int hdr[12]; float mu_lr = 1.f;
    if (std::fread(hdr, sizeof(int), 12, fi) != 12 || std::fread(&mu_lr, sizeof(float), 1, fi) != 1) return 2;
    const int N = hdr[0], S = hdr[1], F = hdr[2], Gx = hdr[3], Gy = hdr[4], H = hdr[5], W = hdr[6];
    const int ks = hdr[7], pad = hdr[8], ignore_edge = hdr[9], bias_term = hdr[10], learn_sigma = hdr[11];
    const size_t P = (size_t)S * Gx * Gy * F;

    try {
        PlainCudaDAULayer layer(nullptr, ignore_edge != 0, /*offsets_already_centered=*/true, /*dynamic_kernel_size=*/true);
        dau_b200::DAUConvSettings st;
        st.num_output = F; st.number_units = {Gx, Gy}; st.kernel_size = ks; st.pad = pad; st.bias_term = bias_term != 0;
        // the order the TF op uses (dau_conv_op.cpp:266-290): stream, toggles, LayerSetUp, Reshape, Forward_gpu
        cudaStream_t stream; CK(cudaStreamCreate(&stream));
        layer.set_default_cuda_stream(stream);
        layer.enable_unit_bounds_guard(false);
        layer.enable_sigma_learning(learn_sigma != 0);
        const std::vector<int> bottom_shape{N, S, H, W};
        layer.LayerSetUp(st, bottom_shape, /*number_units_ignore=*/0, /*in_train=*/true, mu_lr);
        std::vector<int> top_shape = layer.Reshape(bottom_shape, {});
        const int Ho = top_shape[2], Wo = top_shape[3];
        if (top_shape[0] != N || top_shape[1] != F) { std::fprintf(stderr, "bad top shape\n"); return 4; }

        DevBuf x, err, y, dx;
        x.put(read_floats(fi, (size_t)N * S * H * W));
        err.put(read_floats(fi, (size_t)N * F * Ho * Wo));
        layer.w.put(read_floats(fi, P)); layer.mu1.put(read_floats(fi, P)); layer.mu2.put(read_floats(fi, P));
        layer.sigma.put(read_floats(fi, P)); layer.bias.put(read_floats(fi, F));
        std::fclose(fi);
        layer.dw.alloc(P); layer.dmu1.alloc(P); layer.dmu2.alloc(P); layer.dsigma.alloc(P); layer.dbias.alloc(F);
        y.alloc((size_t)N * F * Ho * Wo); dx.alloc((size_t)N * S * H * W);

        // calling before Reshape with another shape must throw, not corrupt memory
        bool threw = false;
        try { layer.Forward_gpu(x.p, {N + 1, S, H, W}, y.p, top_shape); } catch (const dau_b200::DAUException&) { threw = true; }
        if (!threw) { std::fprintf(stderr, "shape mismatch was not refused\n"); return 4; }

        layer.Forward_gpu(x.p, bottom_shape, y.p, top_shape);
        // a second Reshape with the same shape must not ask for the workspace again (base_dau_conv_layer.cpp:200-205)
        layer.Reshape(bottom_shape, top_shape);
        if (layer.allocations != 1) { std::fprintf(stderr, "workspace re-allocated for an unchanged shape\n"); return 4; }
        layer.Backward_gpu(nullptr, err.p, top_shape, /*propagate_down=*/true, x.p, dx.p, bottom_shape,
                           {true, true, true, true, bias_term != 0});
        CK(cudaStreamSynchronize(stream));

        FILE* fo = std::fopen(argv[2], "wb");
        if (!fo) { std::perror(argv[2]); return 2; }
        int oh[2] = {Ho, Wo};
        std::fwrite(oh, sizeof(int), 2, fo);
        write_floats(fo, y.get()); write_floats(fo, dx.get()); write_floats(fo, layer.dw.get());
        write_floats(fo, layer.dmu1.get()); write_floats(fo, layer.dmu2.get()); write_floats(fo, layer.dsigma.get());
        write_floats(fo, layer.dbias.get());
        std::fclose(fo);
        CK(cudaStreamDestroy(stream));
    } catch (const dau_b200::DAUException& ex) {
        std::fprintf(stderr, "DAUException(%d): %s\n", (int)ex.status(), ex.what());
        return 5;
    }
    return 0;
}
