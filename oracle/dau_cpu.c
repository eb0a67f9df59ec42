/*
 * dau_cpu.c -- plain-C (OpenMP) restatement of the reference's CPU path for the DAU layer.
 * TEST INFRASTRUCTURE ONLY: it is the checker for tests/ and the timed `cpu_baseline` /
 * `--impl reference` leg of bench.py.  Nothing under dau-convnet_b200/ links or calls it.
 *
 * Parity status: PINNED through oracle/dau_oracle.py -- tests/test_oracle_c.py checks this file
 * against the NumPy oracle, which is itself pinned to tests/golden/ (outputs of the
 * reference's DAUConvPython, plugins/tensorflow/tests/dau_conv_test.py:13-295).
 *
 * Why a restatement and not oracle/_ref: the reference's own Forward_cpu/Backward_cpu cannot
 * serve: cpu_sum_elementwise_skip multiplies into a zeroed buffer
 * (src/dau_conv/base_dau_conv_layer.cpp:498-501, 550-551) so the CPU forward returns zeros,
 * and Forward_cpu needs CUDA for its filters (base_dau_conv_layer.cu:537-710).
 *
 * Structure followed (paths relative to the reference checkout):
 *   filters ............ BaseDAUKernelCompute::get_kernels   src/dau_conv/base_dau_conv_layer.cu:387-448, 537-710
 *   blur ............... Forward_cpu im2col+sgemm blur        src/dau_conv/base_dau_conv_layer.cpp:680-691
 *                        (here: direct "same" correlation, util/convolve.cu:107-124 semantics)
 *   shift-and-sum ...... offset_and_sum_opencv (ROI axpy)     src/dau_conv/base_dau_conv_layer.cpp:513-636
 *                        with `dst += alpha*src` instead of the defective `dst = dst*src*alpha`
 *   shift-and-dot ...... offset_and_dot_opencv (ROI dot)      src/dau_conv/base_dau_conv_layer.cpp:714-887
 *   edge masks ......... ignore_edge_gradients                src/dau_conv/base_dau_conv_layer.cpp:754-791
 *   dL/dx .............. blur(E, mirrored) -> gather(-mu, FGS) src/dau_conv/base_dau_conv_layer.cpp:930-989
 *   x w, accumulate .... Backward_cpu epilogue                src/dau_conv/base_dau_conv_layer.cpp:1077-1099
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define MAXKB 17

int dau_cpu_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/* bench.py's reference arm: torch.distributed.run exports OMP_NUM_THREADS=1 to its workers */
void dau_cpu_set_num_threads(int n) {
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}

int dau_cpu_kernel_size(float sigma) { return 2 * (int)ceilf(5.0f * sigma) + 1; }

/* six Kb x Kb double filters: blur, d_w, d_mu1, d_mu2, d_sigma, d_error (axis 0 = y) */
int dau_cpu_filters(float sigma_f, int unit_norm, int square_norm, int single_dim, int forbid_pos,
                    int* kb_out, double* out6) {
    const double sigma = (double)sigma_f;
    const int kb = dau_cpu_kernel_size(sigma_f), c = kb / 2, n = kb * kb;
    if (kb <= 1 || kb > MAXKB) return 1;
    double g[MAXKB * MAXKB], m1[MAXKB * MAXKB], m2[MAXKB * MAXKB], sg[MAXKB * MAXKB];
    double sum_g = 0, sum_g2 = 0, s1 = 0, s2 = 0, s3 = 0;
    for (int v = 0; v < kb; ++v)
        for (int u = 0; u < kb; ++u) {
            const double dx = u - c, dy = v - c, d2 = dx * dx + dy * dy;
            double val = exp(-d2 / (2 * sigma * sigma));
            if (single_dim && v != c) val = 0;             /* .cu:432-434 */
            if (forbid_pos && dx > 0) val = 0;             /* .cu:436-438 */
            const int i = v * kb + u;
            g[i] = val;
            m1[i] = dx / (sigma * sigma) * val;
            m2[i] = dy / (sigma * sigma) * val;
            sg[i] = d2 / (sigma * sigma * sigma) * val;
            sum_g += val; sum_g2 += val * val;
            if (square_norm) { s1 += 2 * m1[i] * val; s2 += 2 * m2[i] * val; s3 += 2 * sg[i] * val; }
            else { s1 += m1[i]; s2 += m2[i]; s3 += sg[i]; }
        }
    double norm = 1.0;
    if (!unit_norm) { s1 = s2 = s3 = 0; }                  /* .cu:612-616 */
    else norm = square_norm ? 1.0 / sum_g2 : 1.0 / sum_g;  /* .cu:618-655 */
    if (fabs(s1) <= 1e-10) s1 = 0;                         /* .cu:661-662 */
    if (fabs(s2) <= 1e-10) s2 = 0;
    for (int i = 0; i < n; ++i) {
        const double dw = g[i] * norm;                     /* .cu:669 */
        out6[0 * n + i] = dw;
        out6[1 * n + i] = dw;
        out6[2 * n + i] = norm * m1[i] - s1 * norm * dw;   /* .cu:676-683 */
        out6[3 * n + i] = norm * m2[i] - s2 * norm * dw;
        out6[4 * n + i] = norm * sg[i] - s3 * norm * dw;
    }
    for (int v = 0; v < kb; ++v)                           /* .cu:699-704 mirror */
        for (int u = 0; u < kb; ++u) out6[5 * n + (kb - 1 - v) * kb + (kb - 1 - u)] = out6[v * kb + u];
    *kb_out = kb;
    return 0;
}

/* zero-padded "same" correlation of one H x W plane (util/convolve.cu:107-124) */
static void blur_plane(const float* in, float* out, int H, int W, const double* filt, int kb) {
    const int c = kb / 2;
    float ff[MAXKB * MAXKB];
    for (int i = 0; i < kb * kb; ++i) ff[i] = (float)filt[i];
    memset(out, 0, sizeof(float) * (size_t)H * W);
    for (int v = 0; v < kb; ++v)
        for (int u = 0; u < kb; ++u) {
            const float f = ff[v * kb + u];
            if (f == 0.f) continue;
            const int dy = v - c, dx = u - c;
            const int ylo = dy < 0 ? -dy : 0, yhi = dy > 0 ? H - dy : H;
            const int xlo = dx < 0 ? -dx : 0, xhi = dx > 0 ? W - dx : W;
            for (int y = ylo; y < yhi; ++y) {
                const float* s = in + (size_t)(y + dy) * W + dx;
                float* d = out + (size_t)y * W;
                for (int x = xlo; x < xhi; ++x) d[x] += f * s[x];
            }
        }
}

/* ROI of the reference's shift-and-sum (base_dau_conv_layer.cpp:574-596): the overlap of the
 * image with itself shifted by (ax, ay) */
typedef struct { int w, h, sx, sy, dx, dy; } roi_t;
static roi_t make_roi(int W, int H, int ax, int ay) {
    roi_t r;
    r.w = (W + ax < W - ax) ? W + ax : W - ax;
    r.h = (H + ay < H - ay) ? H + ay : H - ay;
    r.sx = ax > 0 ? ax : 0; r.sy = ay > 0 ? ay : 0;
    r.dx = ax < 0 ? -ax : 0; r.dy = ay < 0 ? -ay : 0;
    return r;
}

/* dst[n, cout] += sum over units of w * bilinear shift of src[n, cin]
 * sgf != 0: params indexed [cin=s][g][cout=f] (forward);  sgf == 0: [cout=... ] FGS, i.e. the
 * parameter tensor is still [S,G,F] but cin runs over F and cout over S, offsets negated by
 * the caller (base_dau_conv_layer.cpp:930-989). */
static void shift_and_sum(const float* src, float* dst, const float* w, const float* mu1, const float* mu2,
                          int N, int Cin, int Cout, int G, int g_active, int H, int W, int sgf,
                          float mu_sign, int interp) {
#pragma omp parallel for collapse(2) schedule(static)
    for (int n = 0; n < N; ++n)
        for (int co = 0; co < Cout; ++co) {
            float* d = dst + ((size_t)n * Cout + co) * H * W;
            memset(d, 0, sizeof(float) * (size_t)H * W);
            for (int ci = 0; ci < Cin; ++ci) {
                const float* s = src + ((size_t)n * Cin + ci) * H * W;
                for (int g = 0; g < g_active; ++g) {
                    const size_t pi = sgf ? ((size_t)ci * G + g) * Cout + co : ((size_t)co * G + g) * Cin + ci;
                    const float wv = w[pi];
                    const float ox = mu_sign * mu1[pi], oy = mu_sign * mu2[pi];
                    const int oxi = (int)floorf(ox), oyi = (int)floorf(oy);
                    const float fx = interp ? ox - oxi : 0.f, fy = interp ? oy - oyi : 0.f;
                    for (int dy = 0; dy < (interp ? 2 : 1); ++dy)
                        for (int dx = 0; dx < (interp ? 2 : 1); ++dx) {
                            const float a = wv * (dx == 0 ? 1.f - fx : fx) * (dy == 0 ? 1.f - fy : fy);
                            const roi_t r = make_roi(W, H, oxi + dx, oyi + dy);
                            if (r.w <= 0 || r.h <= 0) continue;
                            for (int y = 0; y < r.h; ++y) {
                                const float* sp = s + (size_t)(r.sy + y) * W + r.sx;
                                float* dp = d + (size_t)(r.dy + y) * W + r.dx;
                                for (int x = 0; x < r.w; ++x) dp[x] += a * sp[x];   /* fixed `+=` */
                            }
                        }
                }
            }
        }
}

int dau_cpu_forward(const float* x, const float* w, const float* mu1, const float* mu2, float sigma,
                    int N, int S, int F, int G, int H, int W, int units_ignore, int interp,
                    int unit_norm, int square_norm, int single_dim, int forbid_pos, float* y) {
    double filt[6 * MAXKB * MAXKB];
    int kb;
    if (dau_cpu_filters(sigma, unit_norm, square_norm, single_dim, forbid_pos, &kb, filt)) return 1;
    float* xb = (float*)malloc(sizeof(float) * (size_t)N * S * H * W);
    if (!xb) return 2;
#pragma omp parallel for schedule(static)
    for (long p = 0; p < (long)N * S; ++p) blur_plane(x + (size_t)p * H * W, xb + (size_t)p * H * W, H, W, filt, kb);
    shift_and_sum(xb, y, w, mu1, mu2, N, S, F, G, G - units_ignore, H, W, 1, 1.f, interp);
    free(xb);
    return 0;
}

/* grads[k][s,g,f] = sum_{n,y,x} err * bilinear shift of (x blurred with filter 1+k); then x w */
int dau_cpu_backward(const float* x, const float* err, const float* w, const float* mu1, const float* mu2,
                     float sigma, int N, int S, int F, int G, int H, int W, int units_ignore, int interp,
                     int unit_norm, int square_norm, int single_dim, int forbid_pos,
                     int mask_last_row, int mask_last_col,
                     float* dx, float* dw, float* dmu1, float* dmu2, float* dsigma) {
    double filt[6 * MAXKB * MAXKB];
    int kb;
    if (dau_cpu_filters(sigma, unit_norm, square_norm, single_dim, forbid_pos, &kb, filt)) return 1;
    const size_t plane = (size_t)H * W, nin = (size_t)N * S * plane, nout = (size_t)N * F * plane;
    const int g_active = G - units_ignore;
    /* (1) dL/dx: blur the error with the mirrored filter, gather with -mu, params read FGS.
     * All units take part, like the reference model (dau_conv_test.py:232-238). */
    if (dx) {
        float* eb = (float*)malloc(sizeof(float) * nout);
        if (!eb) return 2;
#pragma omp parallel for schedule(static)
        for (long p = 0; p < (long)N * F; ++p)
            blur_plane(err + (size_t)p * plane, eb + (size_t)p * plane, H, W, filt + 5 * kb * kb, kb);
        shift_and_sum(eb, dx, w, mu1, mu2, N, F, S, G, G, H, W, 0, -1.f, interp);
        free(eb);
    }
    /* (2) parameter gradients */
    float* xk = (float*)malloc(sizeof(float) * nin);
    float* em = (float*)malloc(sizeof(float) * nout);
    if (!xk || !em) { free(xk); free(em); return 2; }
    memcpy(em, err, sizeof(float) * nout);
    if (mask_last_col || mask_last_row)                    /* base_dau_conv_layer.cpp:754-791 */
        for (size_t p = 0; p < (size_t)N * F; ++p) {
            float* e = em + p * plane;
            if (mask_last_col) for (int yy = 0; yy < H; ++yy) e[(size_t)yy * W + W - 1] = 0.f;
            if (mask_last_row) for (int xx = 0; xx < W; ++xx) e[(size_t)(H - 1) * W + xx] = 0.f;
        }
    float* outs[4] = {dw, dmu1, dmu2, dsigma};
    for (int k = 0; k < 4; ++k) {
        if (!outs[k]) continue;
#pragma omp parallel for schedule(static)
        for (long p = 0; p < (long)N * S; ++p)
            blur_plane(x + (size_t)p * plane, xk + (size_t)p * plane, H, W, filt + (1 + k) * kb * kb, kb);
#pragma omp parallel for collapse(2) schedule(dynamic, 4)
        for (int s = 0; s < S; ++s)
            for (int f = 0; f < F; ++f)
                for (int g = 0; g < G; ++g) {
                    const size_t pi = ((size_t)s * G + g) * F + f;
                    if (g >= g_active) { outs[k][pi] = 0.f; continue; }
                    const float ox = mu1[pi], oy = mu2[pi];
                    const int oxi = (int)floorf(ox), oyi = (int)floorf(oy);
                    const float fx = interp ? ox - oxi : 0.f, fy = interp ? oy - oyi : 0.f;
                    double tot = 0.0;
                    for (int dy = 0; dy < (interp ? 2 : 1); ++dy)
                        for (int dxx = 0; dxx < (interp ? 2 : 1); ++dxx) {
                            const double a = (double)(dxx == 0 ? 1.f - fx : fx) * (dy == 0 ? 1.f - fy : fy);
                            const roi_t r = make_roi(W, H, oxi + dxx, oyi + dy);
                            if (r.w <= 0 || r.h <= 0) continue;
                            double dot = 0.0;
                            for (int n = 0; n < N; ++n) {
                                const float* sp = xk + ((size_t)n * S + s) * plane;
                                const float* ep = em + ((size_t)n * F + f) * plane;
                                for (int yy = 0; yy < r.h; ++yy) {
                                    const float* a1 = sp + (size_t)(r.sy + yy) * W + r.sx;
                                    const float* b1 = ep + (size_t)(r.dy + yy) * W + r.dx;
                                    float row = 0.f;
                                    for (int xx = 0; xx < r.w; ++xx) row += a1[xx] * b1[xx];
                                    dot += row;
                                }
                            }
                            tot += a * dot;
                        }
                    /* x w on mu1, mu2, sigma (base_dau_conv_layer.cpp:1082-1090) */
                    outs[k][pi] = (float)(k == 0 ? tot : tot * (double)w[pi]);
                }
    }
    free(xk); free(em);
    return 0;
}
