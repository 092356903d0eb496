// Signed gadget decomposition of u32 torus values for pbs16_kernel: the scalar closed form and the four-lane form.
// Self-contained (only __byte_perm and 32-bit integer arithmetic) so that tests/test_digit_lanes_cpu.py can compile it
// for the host with stub qualifiers and compare both forms exhaustively with the tfhe-rs recurrence.
#pragma once
#include <cstdint>

namespace hb {

#ifdef HB_PBS16_MAGIC_CVT
#define HB_DIG_BIAS 64u
#else
#define HB_DIG_BIAS 0u
#endif

// Signed decomposition of a u32 torus value into L biased bytes (digit + 64), closed form of tfhe-rs
// SignedDecomposer (App. A.3): with b = bit (BL-1) of the next residue, carry = (res + 2^(BL-1) - 1 + b) >> BL, hence
// u = state + 2^(BL-1) - 1 + b gives digit = (u mod 2^BL) - (2^(BL-1) - 1) - b and next state = u >> BL.
// xr = value + 2^(31 - BL*L) (the closest-representable rounding add; the caller folds it into the rotate-subtract).
// Checked exhaustively against digits_all (tests/test_oracle_kat.py restates it on the CPU side).
template <int BL, int L> __device__ __forceinline__ void digit_bytes_u32(uint32_t xr, uint32_t (&byte)[L]) {
    static_assert(BL * L < 32 && BL <= 7, "byte-staged digits");
    constexpr uint32_t HALF1 = (1u << (BL - 1)) - 1, MASK = (1u << BL) - 1;
    uint32_t s = xr >> (32 - BL * L);
#pragma unroll
    for (int l = L - 1; l >= 0; --l) {
        const uint32_t b = (s >> (2 * BL - 1)) & 1u;
        const uint32_t u = s + HALF1 + b;
        byte[l] = (u & MASK) + (HB_DIG_BIAS - HALF1) - b;   // biased byte, or the signed digit (low byte = two's complement)
        s = u >> BL;
    }
}

// The same decomposition for four torus values at once, one value per byte lane: the raw base-2^BL digits of the four
// words are transposed into lane words (7 PRMT + shifts/masks), then the carry rule above runs on all four lanes with
// plain 32-bit adds -- every lane value stays below 2^7, so nothing crosses a lane.  pk[lev] receives the level-lev digits
// of x[0..3] as two's-complement bytes (the operand format of I2F.F64.S8).  About 42 ALU instructions per four values
// against 85 for four scalar chains + packing; tests/test_digit_lanes_cpu.py checks both against the recurrence exhaustively.
// SIGNED = false: bytes biased by 64 instead (the DigPack<true> staging format of pbs_kernel), which also drops the final XOR.
template <int BL, int L, bool SIGNED = true> __device__ __forceinline__ void digit_lanes_u32(const uint32_t (&x)[4], uint32_t (&pk)[L]) {
    static_assert(BL * L <= 24 && BL >= 2 && BL <= 7 && L >= 2, "fields must lie in bytes 1..3; lane sums must stay below 2^8");
    constexpr uint32_t M1 = 0x01010101u, MASK = ((1u << BL) - 1) * M1, HALF1 = ((1u << (BL - 1)) - 1) * M1;
    const uint32_t p01 = __byte_perm(x[0], x[1], 0x7632), p23 = __byte_perm(x[2], x[3], 0x7632);
    uint32_t G[4];
    G[0] = 0;
    G[1] = __byte_perm(__byte_perm(x[0], x[1], 0x0051), __byte_perm(x[2], x[3], 0x0051), 0x5410);
    G[2] = __byte_perm(p01, p23, 0x6420);
    G[3] = __byte_perm(p01, p23, 0x7531);
    uint32_t W[L];                                             // raw digits, W[lev], lev 0 = most significant
#pragma unroll
    for (int lev = 0; lev < L; ++lev) {
        const int lo = 32 - BL * (lev + 1), k = lo >> 3, sh = lo & 7;
        if (sh + BL <= 8) W[lev] = (G[k] >> sh) & MASK;
        else {
            const uint32_t mlo = ((1u << (8 - sh)) - 1) * M1;
            W[lev] = (((G[k] >> sh) & mlo) | ((G[k + 1] << (8 - sh)) & ~mlo)) & MASK;
        }
    }
    uint32_t E = W[L - 1];
#pragma unroll
    for (int lev = L - 1; lev >= 0; --lev) {
        uint32_t Nx = 0, B = 0;
        if (lev > 0) {
            Nx = lev == L - 1 ? W[lev - 1] : W[lev - 1] + ((E >> BL) & M1);   // next residue with this one's overflow
            B = (Nx >> (BL - 1)) & M1;
        }
        const uint32_t U = (E & MASK) + HALF1 + B;
        if (SIGNED) pk[lev] = ((U & MASK) + (0x80u * M1 - HALF1) - B) ^ 0x80808080u;
        else pk[lev] = (U & MASK) + (0x40u * M1 - HALF1) - B;
        if (lev > 0) E = Nx + ((U >> BL) & M1);
    }
}

}  // namespace hb
