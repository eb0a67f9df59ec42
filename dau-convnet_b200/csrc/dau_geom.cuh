// Compile-time tile geometry and cp.async helpers shared by the gather and gradient kernels.
#pragma once
#include "dau_kernels.cuh"

namespace dau {
namespace geom {

constexpr int ilog2(int v) { return v <= 1 ? 0 : 1 + ilog2(v >> 1); }

// geometry of one instantiation, all compile-time
template <int R_, int C_, int LX_, int LY_, int LN_, int NW_, int HALO_>
struct GatherGeom {
    static constexpr int R = R_, C = C_, LX = LX_, LY = LY_, LN = LN_, NW = NW_, HALO = HALO_;
    static_assert(LX * LY * LN == 32, "lanes");
    static_assert((C & (C - 1)) == 0, "C must be a power of two");
    static constexpr int TW = LX * C, TH = LY * R;
    static constexpr int HX = (HALO + C - 1) / C * C;   // x halo rounded to the interleave factor
    static constexpr int HY = HALO;
    static constexpr int TWH = TW + 2 * HX, THH = TH + 2 * HY;
    static constexpr int QS = TWH / C;                  // words per de-interleaved column plane

    static constexpr bool conflict_free(int rs, int is) {
        unsigned used = 0;
        for (int lane = 0; lane < 32; ++lane) {
            const int lx = lane % LX, ly = (lane / LX) % LY, ln = lane / (LX * LY);
            const int bank = (ln * is + ly * R * rs + lx) & 31;
            if (used & (1u << bank)) return false;
            used |= 1u << bank;
        }
        return true;
    }
    static constexpr int pick_rs() {
        for (int pad = 0; pad < 32; ++pad)
            for (int ipad = 0; ipad < 32; ++ipad)
                if (conflict_free(C * QS + pad, THH * (C * QS + pad) + ipad)) return C * QS + pad;
        return C * QS;
    }
    static constexpr int RS = pick_rs();
    static constexpr int pick_is() {
        for (int ipad = 0; ipad < 32; ++ipad)
            if (conflict_free(RS, THH * RS + ipad)) return THH * RS + ipad;
        return THH * RS;
    }
    static constexpr int IS = pick_is();
    static constexpr int TILE_WORDS = (LN * IS + 3) / 4 * 4;   // keep the records 16-byte aligned
};

__device__ __forceinline__ void cp_async4_zfill(float* dst, const float* src, bool ok) {
    const unsigned d = (unsigned)__cvta_generic_to_shared(dst);
    const int sz = ok ? 4 : 0;
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;\n" ::"r"(d), "l"(src), "r"(sz) : "memory");
}
__device__ __forceinline__ void cp_async16(void* dst, const void* src) {
    const unsigned d = (unsigned)__cvta_generic_to_shared(dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(d), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;\n" ::: "memory"); }


}  // namespace geom
}  // namespace dau
