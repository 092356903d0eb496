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
    const double c = HB_COS(num);
    const double s = SIGN > 0 ? HB_SIN(num) : -HB_SIN(num);
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

// Same transform as dft_regs (same input / output register order), decimation in time with every twiddled butterfly
// written as 6 fused multiply-adds (Linzer-Feig / Goedecker form: w B = c (B + i t B) with t = tan, so A +- w B is
// two FMAs for the rotated B and four for the sums) instead of 4 adds + a 4-instruction complex multiply:
// a DFT-16 is 88 DADD + 60 DFMA = 148 FP64 instructions against 168.  Rounding differs from dft_regs in the last bits.
template <int SIGN> __device__ __forceinline__ void bfly_dit(double2& A, double2& B, int num) {   // num/256 of a turn, [0,128)
    const double2 a = A, b = B;
    if (num == 0) { A = make_double2(a.x + b.x, a.y + b.y); B = make_double2(a.x - b.x, a.y - b.y); return; }
    if (num == 64) {
        if (SIGN > 0) { A = make_double2(a.x - b.y, a.y + b.x); B = make_double2(a.x + b.y, a.y - b.x); }
        else          { A = make_double2(a.x + b.y, a.y - b.x); B = make_double2(a.x - b.y, a.y + b.x); }
        return;
    }
    const double t = SIGN > 0 ? HB_TAN(num) : -HB_TAN(num);      // tan or cot of the angle, whichever is <= 1 in magnitude
    if (num <= 32 || num >= 96) {          // |c| >= |s|:  w b = c (b + i t b)
        const double c = HB_COS(num);
        const double tr = fma(-t, b.y, b.x), ti = fma(t, b.x, b.y);
        A = make_double2(fma(c, tr, a.x), fma(c, ti, a.y));
        B = make_double2(fma(-c, tr, a.x), fma(-c, ti, a.y));
    } else {                               // w b = s (t b + i b)
        const double s = SIGN > 0 ? HB_SIN(num) : -HB_SIN(num);
        const double tr = fma(t, b.x, -b.y), ti = fma(t, b.y, b.x);
        A = make_double2(fma(s, tr, a.x), fma(s, ti, a.y));
        B = make_double2(fma(-s, tr, a.x), fma(-s, ti, a.y));
    }
}
template <int R, int SIGN> __device__ __forceinline__ void dft_regs_dit(double2 (&x)[R]) {
#pragma unroll
    for (int st = 0; st < ilog2c(R); ++st) {
#pragma unroll
        for (int blk = 0; blk < (R >> (st + 1)); ++blk) {
#pragma unroll
            for (int j = 0; j < (1 << st); ++j) {
                const int p = (blk << (st + 1)) + j;
                bfly_dit<SIGN>(x[brev<R>(p)], x[brev<R>(p + (1 << st))], j * (HB_TW_DEN >> (st + 1)));
            }
        }
    }
}

// dft_regs_dit of x[j] * e^{2 pi i j TWA / 256} (the negacyclic twist, angles below a quarter turn) with the twist's
// scale factor carried into the first butterfly stage: x[j] w_j = s_j (x[j] + i t_j x[j]) costs two FMAs for the
// rotated point and the scales ride on the stage-0 sums (2 DMUL + 4 DFMA per butterfly), 10 FP64 instructions per
// butterfly instead of 2 x 4 (twists) + 4 (adds).
template <int SIGN> __device__ __forceinline__ double2 rot_unscaled(double2 v, int num, double& scale) {   // num in [0, 64)
    if (num == 0) { scale = 1.0; return v; }
    const double t = SIGN > 0 ? HB_TAN(num) : -HB_TAN(num);
    if (num <= 32) { scale = HB_COS(num); return make_double2(fma(-t, v.y, v.x), fma(t, v.x, v.y)); }
    scale = SIGN > 0 ? HB_SIN(num) : -HB_SIN(num);
    return make_double2(fma(t, v.x, -v.y), fma(t, v.y, v.x));
}
template <int R, int SIGN, int TWA> __device__ __forceinline__ void dft_regs_dit_twisted(double2 (&x)[R]) {
    static_assert((R - 1) * TWA < 64, "twist angles stay in the first quadrant");
#pragma unroll
    for (int blk = 0; blk < R / 2; ++blk) {
        const int ja = brev<R>(2 * blk), jb = brev<R>(2 * blk + 1);
        double sa, sb;
        const double2 a = rot_unscaled<SIGN>(x[ja], ja * TWA, sa), b = rot_unscaled<SIGN>(x[jb], jb * TWA, sb);
        const double2 m = make_double2(sb * b.x, sb * b.y);
        if (ja * TWA == 0) {
            x[ja] = make_double2(a.x + m.x, a.y + m.y);
            x[jb] = make_double2(a.x - m.x, a.y - m.y);
        } else {
            x[ja] = make_double2(fma(sa, a.x, m.x), fma(sa, a.y, m.y));
            x[jb] = make_double2(fma(sa, a.x, -m.x), fma(sa, a.y, -m.y));
        }
    }
#pragma unroll
    for (int st = 1; st < ilog2c(R); ++st) {
#pragma unroll
        for (int blk = 0; blk < (R >> (st + 1)); ++blk) {
#pragma unroll
            for (int j = 0; j < (1 << st); ++j) {
                const int p = (blk << (st + 1)) + j;
                bfly_dit<SIGN>(x[brev<R>(p)], x[brev<R>(p + (1 << st))], j * (HB_TW_DEN >> (st + 1)));
            }
        }
    }
}

}  // namespace hb
