// Host check of helm_b200/csrc/digits_u32.cuh (compiled by tests/test_digit_lanes_cpu.py with stub CUDA qualifiers):
// the scalar closed form and the four-lane form against the tfhe-rs SignedDecomposer recurrence, every BL*L-bit state
// (with random discarded low bits) in every lane position.  Prints "ok <count>" or the first mismatch.
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#define __device__
#define __forceinline__ inline
static inline uint32_t __byte_perm(uint32_t a, uint32_t b, uint32_t s) {
    const uint64_t v = ((uint64_t)b << 32) | a;
    uint32_t r = 0;
    for (int i = 0; i < 4; ++i) {
        const uint32_t sel = (s >> (4 * i)) & 0xf;
        uint32_t byte = (uint32_t)(v >> (8 * (sel & 7))) & 0xff;
        if (sel & 8) byte = (byte & 0x80) ? 0xff : 0x00;
        r |= byte << (8 * i);
    }
    return r;
}
#include "digits_u32.cuh"

// tfhe-rs SignedDecomposer (closest representable, then the carry recurrence), digits[lev], lev 0 = most significant
template <int BL, int L> static void reference_digits(uint32_t x, int32_t (&d)[L]) {
    const int nr = 32 - BL * L;
    uint32_t state = ((x >> (nr - 1)) + 1) >> 1;
    state &= (1u << (BL * L)) - 1;
    const uint32_t mask = (1u << BL) - 1;
    for (int lev = L - 1; lev >= 0; --lev) {
        uint32_t res = state & mask;
        state >>= BL;
        uint32_t carry = ((res - 1u) | state) & res;
        carry >>= BL - 1;
        state += carry;
        d[lev] = (int32_t)(res - (carry << BL));
    }
}

template <int BL, int L> static long check() {
    const uint32_t RND = 1u << (31 - BL * L);
    long n = 0;
    uint32_t lcg = 12345u;
    for (uint32_t s = 0; s < (1u << (BL * L)); ++s) {
        uint32_t x[4], xr[4], pk[L];
        for (int u = 0; u < 4; ++u) {
            lcg = lcg * 1664525u + 1013904223u;
            const uint32_t other = (u == (int)(s & 3)) ? s : (lcg >> 7) & ((1u << (BL * L)) - 1);
            x[u] = (other << (32 - BL * L)) | ((lcg >> 3) & ((1u << (32 - BL * L)) - 1));
            xr[u] = x[u] + RND;
        }
        uint32_t pb[L];
        hb::digit_lanes_u32<BL, L>(xr, pk);
        hb::digit_lanes_u32<BL, L, false>(xr, pb);      // bytes biased by 64 (pbs_kernel's staging format)
        for (int u = 0; u < 4; ++u) {
            int32_t want[L];
            uint32_t by[L];
            reference_digits<BL, L>(x[u], want);
            hb::digit_bytes_u32<BL, L>(xr[u], by);
            for (int lev = 0; lev < L; ++lev) {
                const int32_t lane = (int8_t)(pk[lev] >> (8 * u));
                const int32_t scal = (int8_t)(by[lev] & 0xff);
                const int32_t biased = (int32_t)((pb[lev] >> (8 * u)) & 0xff) - 64;
                if (lane != want[lev] || scal != want[lev] || biased != want[lev]) {
                    printf("mismatch BL=%d L=%d x=%08x lane %d lev %d: lanes %d scalar %d reference %d\n", BL, L, x[u], u, lev, lane, scal, want[lev]);
                    exit(1);
                }
                ++n;
            }
        }
    }
    return n;
}

int main() {
    long n = check<6, 3>() + check<7, 3>() + check<4, 4>() + check<8 - 1, 2>() + check<2, 5>() + check<3, 8>();
    printf("ok %ld\n", n);
    return 0;
}
