// Register-resident radix-2 DFT building blocks for the folded negacyclic transform.
// All loops are compile-time unrolled: every twiddle is a literal (constant-bank operand of
// DMUL/DFMA) and trivial twiddles (1, +-i, (1+-i)/sqrt2) are special-cased, so a DFT-16 costs
// 128 DADD + 40 DMUL/DFMA.  Outputs are left in bit-reversed register order; callers index
// them through brev<R>() which is resolved at compile time.
#pragma once
#include <cuda_runtime.h>
#include "twiddles.h"

namespace hb {

__host__ __device__ constexpr int ilog2c(int x) { return x <= 1 ? 0 : 1 + ilog2c(x >> 1); }

template <int R> __host__ __device__ constexpr int brev(int v) {
    int r = 0;
    for (int b = 0; b < ilog2c(R); ++b) r |= ((v >> b) & 1) << (ilog2c(R) - 1 - b);
    return r;
}

// x * e^{SIGN * 2 pi i * NUM / 256} with NUM a compile-time value after unrolling
template <int SIGN> __device__ __forceinline__ double2 cmul_tw(double2 x, int num) {
    num &= 255;
    if (num == 0) return x;
    if (num == 128) return make_double2(-x.x, -x.y);
    if (num == 64) return SIGN > 0 ? make_double2(-x.y, x.x) : make_double2(x.y, -x.x);
    if (num == 192) return SIGN > 0 ? make_double2(x.y, -x.x) : make_double2(-x.y, x.x);
    if ((num & 31) == 0) {   // odd multiple of 32: (+-1 +- i)/sqrt2
        const double h = 0.70710678118654752440;
        const double c = (num == 32 || num == 224) ? h : -h;
        const double s0 = (num == 32 || num == 96) ? h : -h;
        const double s = SIGN > 0 ? s0 : -s0;
        // (x.x + i x.y)(c + i s), |c| = |s| = h
        return make_double2(x.x * c - x.y * s, x.x * s + x.y * c);
    }
    const double c = HB_TW_COS[num];
    const double s = SIGN > 0 ? HB_TW_SIN[num] : -HB_TW_SIN[num];
    return make_double2(fma(x.x, c, -(x.y * s)), fma(x.x, s, x.y * c));
}

__device__ __forceinline__ double2 cmul(double2 a, double2 b) {
    return make_double2(fma(a.x, b.x, -(a.y * b.y)), fma(a.x, b.y, a.y * b.x));
}
__device__ __forceinline__ double2 cmul_conj(double2 a, double2 b) {   // a * conj(b)
    return make_double2(fma(a.x, b.x, a.y * b.y), fma(a.y, b.x, -(a.x * b.y)));
}
__device__ __forceinline__ void cfma(double2& acc, double2 a, double2 b) {
    acc.x = fma(a.x, b.x, acc.x); acc.x = fma(-a.y, b.y, acc.x);
    acc.y = fma(a.x, b.y, acc.y); acc.y = fma(a.y, b.x, acc.y);
}

// In-register DFT of length R, kernel e^{SIGN 2 pi i jk/R}, DIF radix-2.
// Input natural order, output k is found in x[brev<R>(k)].
template <int R, int SIGN> __device__ __forceinline__ void dft_regs(double2 (&x)[R]) {
#pragma unroll
    for (int half = R / 2; half >= 1; half >>= 1) {
#pragma unroll
        for (int base = 0; base < R; base += 2 * half) {
#pragma unroll
            for (int j = 0; j < half; ++j) {
                const double2 a = x[base + j], b = x[base + j + half];
                x[base + j] = make_double2(a.x + b.x, a.y + b.y);
                const double2 d = make_double2(a.x - b.x, a.y - b.y);
                x[base + j + half] = cmul_tw<SIGN>(d, j * (HB_TW_DEN / (2 * half)));
            }
        }
    }
}

}  // namespace hb
