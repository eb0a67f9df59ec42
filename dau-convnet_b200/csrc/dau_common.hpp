// Shared host-side declarations of the B200 DAU layer (no CUDA types in this header
// beyond cudaStream_t so plain C++ translation units can include it).
#pragma once
#include <cuda_runtime.h>
#include <atomic>
#include <cstddef>
#include <cstdint>
#include <stdexcept>
#include <string>

namespace dau {

// Error convention of the reference: DAU_CHECK / CUDA_CHECK throw DAUException
// (include/dau_conv/util/common.hpp:38-64,102-106).  The C ABI converts to status codes.
struct DAUException : public std::runtime_error {
    int status;
    DAUException(int status_, const std::string& what) : std::runtime_error(what), status(status_) {}
};

#define DAU_CHECK(cond, status, msg)                                                        \
    do { if (!(cond)) throw ::dau::DAUException((status), std::string(msg) +                \
            " [" #cond "] at " __FILE__ ":" + std::to_string(__LINE__)); } while (0)

#define DAU_CUDA_CHECK(expr)                                                                \
    do { cudaError_t e__ = (expr); if (e__ != cudaSuccess) throw ::dau::DAUException(2,     \
            std::string("CUDA error: ") + cudaGetErrorString(e__) + " in " #expr " at "     \
            __FILE__ ":" + std::to_string(__LINE__)); } while (0)

// Kernel attributes (cudaFuncAttributeMaxDynamicSharedMemorySize) belong to a (kernel, DEVICE) pair:
// one of these per kernel instantiation remembers on which devices the attribute has been set, so a
// process that drives handles on several GPUs (the reference's model: one layer object per device,
// base_dau_conv_layer.hpp:177-225) opts in on each of them.  Lock-free; two host threads may both set
// the attribute the first time, which is harmless.
struct PerDeviceOnce {
    std::atomic<unsigned long long> mask[4] = {};       // up to 256 device ordinals
    int dev = 0;
    bool need() {
        if (cudaGetDevice(&dev) != cudaSuccess) return true;
        if (dev < 0 || dev >= 256) return true;
        return !((mask[dev >> 6].load(std::memory_order_acquire) >> (dev & 63)) & 1ull);
    }
    void done() {
        if (dev >= 0 && dev < 256) mask[dev >> 6].fetch_or(1ull << (dev & 63), std::memory_order_release);
    }
};

// Restores the calling thread's current device when a C-ABI entry point returns: the library switches
// to the handle's device for its launches and must not leave the caller (torch, a Caffe-style host)
// on another GPU than it was on.
struct DeviceGuard {
    int prev = -1;
    explicit DeviceGuard(int device) {
        cudaGetDevice(&prev);
        if (device >= 0 && device != prev) {
            cudaError_t e = cudaSetDevice(device);
            if (e != cudaSuccess) throw DAUException(2, std::string("CUDA error: ") + cudaGetErrorString(e) + " in cudaSetDevice");
        } else {
            prev = -1;
        }
    }
    ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
    DeviceGuard(const DeviceGuard&) = delete;
    DeviceGuard& operator=(const DeviceGuard&) = delete;
};

// Development switches (skip the TMA loads / the math inside the hot kernels to time their parts) exist
// only in builds made with -DDAU_DEV_SWITCHES; the shipped library has no environment variable that
// changes results.
#ifdef DAU_DEV_SWITCHES
#define DAU_DEV_BITS(x) (x)
#else
#define DAU_DEV_BITS(x) 0
#endif

constexpr int kMaxBlurTaps = 17;   // reference limit: Kb <= 17, sigma <= 1.6 (convolve.cu:40,230-246)

// Separable factors of the six DAUKernelCompute filters (see dau_filters.cpp).
struct BlurFactors {
    int kb;                       // 2*ceil(5*sigma)+1
    // forward blur  G = fx (x) fy                (normalisation folded into fx)
    float fwd_x[kMaxBlurTaps], fwd_y[kMaxBlurTaps];
    // error blur = 180-degree rotation of G     (base_dau_conv_layer.cu:699-704)
    float err_x[kMaxBlurTaps], err_y[kMaxBlurTaps];
    // derivative bank: horizontal passes h0,h1,h2 ; vertical passes v0,v1,v2
    //   X_w   = v0[h0]            X_mu1 = v0[h1]
    //   X_mu2 = v1[h0]            X_sig = v0[h2] + v2[h0]
    // with every normalisation / mean-correction constant folded into the factors.
    float h0[kMaxBlurTaps], h1[kMaxBlurTaps], h2[kMaxBlurTaps];
    float v0[kMaxBlurTaps], v1[kMaxBlurTaps], v2[kMaxBlurTaps];
};

struct FilterOptions {
    bool unit_normalization = true;
    bool square_unit_normalization = false;
    bool single_dimension_kernel = false;
    bool forbid_positive_dim1 = false;
};

int  blur_kernel_size(float sigma);
void compute_blur_factors(float sigma, const FilterOptions& opt, BlurFactors* out);
// Dense Kb x Kb filters (blur, d_w, d_mu1, d_mu2, d_sigma, d_error), row-major, axis 0 = y.
void compute_dense_kernels(float sigma, const FilterOptions& opt, int* kb, float* out6);

}  // namespace dau
