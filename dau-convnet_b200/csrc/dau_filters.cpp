// DAUKernelCompute for the B200 layer: the Gaussian and its three derivative filters,
// computed on the host in double and handed to the blur kernels as SEPARABLE factors.
//
// What it replaces: BaseDAUKernelCompute<T>::get_kernels and its eight small kernels
// (reference src/dau_conv/base_dau_conv_layer.cu:387-448, 537-710), which the reference
// runs with S=G=F=1, w=1, mu=(0,0) (base_dau_conv_layer.cpp:179-183, 232-235) and ~20 tiny
// launches.  The math is a few hundred flops, so it is done here once per sigma value.
//
// Separability (not used by the reference, whose blur is a dense Kb x Kb correlation,
// util/convolve.cu:45-131): g(dx,dy) = gx(dx)*gy(dy); the single_dimension_kernel and
// forbid_positive_dim1 masks (.cu:432-438) are themselves separable, the normalisation and
// the mean-correction terms (.cu:669-683) are scalars.  d_sigma is a sum of two outer
// products, every other filter is one.
#include "dau_common.hpp"

#include <cmath>
#include <cstring>
#include <vector>

namespace dau {

int blur_kernel_size(float sigma) {
    // Kb = 2*ceil(5*sigma)+1 with the product taken in fp32, as the reference's
    // Dtype=float arithmetic does (base_dau_conv_layer.cpp:146-147): 5*0.8f == 4.0f -> Kb = 9
    const float prod = 5.0f * sigma;
    return 2 * (int)std::ceil(prod) + 1;
}

namespace {
struct Factors1D {
    int kb, c;
    std::vector<double> gx, gy;   // masked 1-D Gaussians
    double norm, s_mu1, s_mu2, s_sig, inv_s2, inv_s3;
};

Factors1D build(float sigma_f, const FilterOptions& opt) {
    Factors1D f;
    const double sigma = (double)sigma_f;
    DAU_CHECK(sigma > 0.0, 1, "sigma must be positive");
    f.kb = blur_kernel_size(sigma_f);
    DAU_CHECK(f.kb > 1 && f.kb <= kMaxBlurTaps, 4,
              "blur support 2*ceil(5*sigma)+1 must be in (1,17] (sigma <= 1.6)");
    f.c = f.kb / 2;
    f.inv_s2 = 1.0 / (sigma * sigma);
    f.inv_s3 = 1.0 / (sigma * sigma * sigma);
    f.gx.resize(f.kb);
    f.gy.resize(f.kb);
    for (int i = 0; i < f.kb; ++i) {
        const double d = (double)(i - f.c);
        const double g = std::exp(-d * d / (2.0 * sigma * sigma));
        f.gx[i] = (opt.forbid_positive_dim1 && d > 0) ? 0.0 : g;       // .cu:436-438
        f.gy[i] = (opt.single_dimension_kernel && i != f.c) ? 0.0 : g; // .cu:432-434
    }
    double sx = 0, sy = 0, sxx = 0, syy = 0;         // sum g, sum g^2
    double mx = 0, my = 0, qx = 0, qy = 0;           // sum d*g, sum d^2*g
    double mxx = 0, myy = 0, qxx = 0, qyy = 0;       // same with g^2
    for (int i = 0; i < f.kb; ++i) {
        const double d = (double)(i - f.c);
        sx += f.gx[i];            sy += f.gy[i];
        sxx += f.gx[i] * f.gx[i]; syy += f.gy[i] * f.gy[i];
        mx += d * f.gx[i];        my += d * f.gy[i];
        qx += d * d * f.gx[i];    qy += d * d * f.gy[i];
        mxx += d * f.gx[i] * f.gx[i];     myy += d * f.gy[i] * f.gy[i];
        qxx += d * d * f.gx[i] * f.gx[i]; qyy += d * d * f.gy[i] * f.gy[i];
    }
    if (!opt.unit_normalization) {                    // .cu:612-616, 642-644
        f.norm = 1.0; f.s_mu1 = f.s_mu2 = f.s_sig = 0.0;
    } else if (opt.square_unit_normalization) {       // .cu:618-634, 646-651
        f.norm = 1.0 / (sxx * syy);
        f.s_mu1 = 2.0 * f.inv_s2 * mxx * syy;
        f.s_mu2 = 2.0 * f.inv_s2 * sxx * myy;
        f.s_sig = 2.0 * f.inv_s3 * (qxx * syy + sxx * qyy);
    } else {                                          // .cu:635-640, 652-655
        f.norm = 1.0 / (sx * sy);
        f.s_mu1 = f.inv_s2 * mx * sy;
        f.s_mu2 = f.inv_s2 * sx * my;
        f.s_sig = f.inv_s3 * (qx * sy + sx * qy);
    }
    if (std::fabs(f.s_mu1) <= 1e-10) f.s_mu1 = 0.0;   // .cu:661-662
    if (std::fabs(f.s_mu2) <= 1e-10) f.s_mu2 = 0.0;
    return f;
}
}  // namespace

void compute_blur_factors(float sigma, const FilterOptions& opt, BlurFactors* out) {
    const Factors1D f = build(sigma, opt);
    std::memset(out, 0, sizeof(*out));
    out->kb = f.kb;
    for (int i = 0; i < f.kb; ++i) {
        const double d = (double)(i - f.c);
        out->fwd_x[i] = (float)(f.norm * f.gx[i]);
        out->fwd_y[i] = (float)f.gy[i];
        out->h0[i] = (float)(f.norm * f.gx[i]);
        out->h1[i] = (float)(f.norm * (d * f.inv_s2 - f.s_mu1 * f.norm) * f.gx[i]);
        out->h2[i] = (float)(f.norm * (d * d * f.inv_s3 - f.s_sig * f.norm) * f.gx[i]);
        out->v0[i] = (float)f.gy[i];
        out->v1[i] = (float)((d * f.inv_s2 - f.s_mu2 * f.norm) * f.gy[i]);
        out->v2[i] = (float)(d * d * f.inv_s3 * f.gy[i]);
    }
    for (int i = 0; i < f.kb; ++i) {                  // 180-degree rotation (.cu:699-704)
        out->err_x[i] = out->fwd_x[f.kb - 1 - i];
        out->err_y[i] = out->fwd_y[f.kb - 1 - i];
    }
}

void compute_dense_kernels(float sigma, const FilterOptions& opt, int* kb_out, float* out6) {
    const Factors1D f = build(sigma, opt);
    const int kb = f.kb, n = kb * kb;
    *kb_out = kb;
    for (int v = 0; v < kb; ++v)
        for (int u = 0; u < kb; ++u) {
            const double dx = (double)(u - f.c), dy = (double)(v - f.c);
            const double g = f.gx[u] * f.gy[v];
            const double dw = g * f.norm;                                      // .cu:669
            const double dm1 = f.norm * dx * f.inv_s2 * g - f.s_mu1 * f.norm * dw;   // .cu:676-683
            const double dm2 = f.norm * dy * f.inv_s2 * g - f.s_mu2 * f.norm * dw;
            const double dsg = f.norm * (dx * dx + dy * dy) * f.inv_s3 * g - f.s_sig * f.norm * dw;
            const int i = v * kb + u;
            out6[0 * n + i] = (float)dw;      // blur (w=1, S=G=F=1; .cu:688-693)
            out6[1 * n + i] = (float)dw;
            out6[2 * n + i] = (float)dm1;
            out6[3 * n + i] = (float)dm2;
            out6[4 * n + i] = (float)dsg;
            out6[5 * n + (kb - 1 - v) * kb + (kb - 1 - u)] = (float)dw;   // mirrored (.cu:699-704)
        }
}

}  // namespace dau
