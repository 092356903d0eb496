// Wide-batch LWE keyswitch on the tensor cores (u32 torus, ks_level = 4), same contract as ks_kernel.cuh:
//   out = (0,...,0,b_in) - sum_i sum_l d_{i,l} * KSK[i][l]     (wrapping mod 2^32, digits in [-B/2, B/2])
// For a level of W gates this is a dense integer contraction [W x kN*l] . [kN*l x (n+1)] (W = 65 536: 194 G multiply-adds),
// 3.2 % of the level in the IMAD form.  Here the 32-bit key words are split once, at hb_load_ksk, into four unsigned byte
// planes, the signed digits are int8, and every plane is accumulated exactly in int32 by mma.sync m16n8k32 (s8 x u8):
// |sum| <= kN*l * 128 * 255 < 2^31.  The planes are recombined with shifts, which wrap exactly like the 32-bit product, so
// the result is bit-identical to ks_kernel's.  (mma.sync, not tcgen05: the keyswitch was 3.2 % of the step and the legacy
// tensor path already brings it to 0.6 %: 18.7 -> 3.2 ms per 64k-gate level.)
//
// ks_digits_kernel writes the digits once to a scratch array [gates][K] (int8; the job's linear part fused, as in
// ks_kernel).  ks_mma2_kernel: CTA = 64 gates x 64 output columns, 8 warps as 4 (gates) x 2 (columns), warp tile
// 16 x 32 x 4 planes = 64 accumulator registers; K = kN*l is walked in chunks of 128 through a two-stage cp.async ring
// (A tile [64][128 B] + key tile [4 planes][64 cols][128 B] per stage, 2 CTAs per SM).  Rows are padded to 144 B so that
// the fragment loads are bank-conflict free.
#pragma once
#include "ks_kernel.cuh"

namespace hb {

// Generalised (round 2) to both torus widths and any ks_level: the key is split into sizeof(T) byte planes (4 for u32, 8 for
// u64), the CTA tile is 64 gates x (256 / planes) columns so that the accumulators stay at 64 registers per thread, and the
// planes are recombined in T.  u64 parameter sets (LUT / arithmetic mode) spent 27 % (M1C1) / 13 % (M2C2) of a wide level in the
// IMAD keyswitch; exactness bound per plane: K * 2^(base_log-1) * 255 < 2^31 (checked on the host).
constexpr int KM_TM = 64, KM_KC = 128, KM_NT = 256, KM_PITCH = 144;
template <class T> struct KmCfg {
    static constexpr int P = (int)sizeof(T);            // byte planes
    static constexpr int TN = 256 / P;                  // output columns per CTA: 64 (u32), 32 (u64)
    static constexpr int NB = TN / 16;                  // 8-column mma blocks per warp (2 warps across the columns)
    static constexpr int STAGE = (KM_TM + P * TN) * KM_PITCH;      // one ring stage: A tile + P plane tiles
};

struct KsMmaArgs {
    KsArgs a;
    const uint8_t* planes;     // [P][npad][K] bytes, K = kN * level, k = i * level + (level index)
    int32_t npad;              // columns padded to a multiple of the CTA tile (zero filled)
};

// KSK [kN][l][n+1] -> byte planes [P][npad][K]
template <class T> __global__ void ksk_planes_kernel(const T* ksk, uint8_t* planes, int n, int K, int npad) {
    const size_t total = (size_t)npad * K;
    for (size_t x = (size_t)blockIdx.x * blockDim.x + threadIdx.x; x < total; x += (size_t)gridDim.x * blockDim.x) {
        const int k = (int)(x % K), col = (int)(x / K);
        const T w = col <= n ? ksk[(size_t)k * (n + 1) + col] : (T)0;
#pragma unroll
        for (int p = 0; p < (int)sizeof(T); ++p) planes[((size_t)p * npad + col) * K + k] = (uint8_t)(w >> (8 * p));
    }
}

__device__ __forceinline__ void mma_s8u8(int (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.s8.u8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}

// digits of every mask word of every job, tfhe-rs SignedDecomposer (App. A.3), int8 at [g][i * level + l]; rows beyond n_jobs are zero
template <class T> __global__ void __launch_bounds__(256) ks_digits_kernel(const KsArgs A, uint8_t* dig, uint32_t rows_pad) {
    constexpr int BITS = (int)sizeof(T) * 8;
    const int kN = A.kN, BL = A.base_log, L = A.level, tot = BL * L;
    const T mask = (((T)1) << BL) - 1;
    const T* in = reinterpret_cast<const T*>(A.in);
    for (uint32_t g = blockIdx.x; g < rows_pad; g += gridDim.x) {
        KsJob j = {{0, 0, 0, 0}, {0, 0, 0, 0}, 0, 0, 0};
        if (g < A.n_jobs) j = A.jobs[g];
        uint8_t* row = dig + (size_t)g * kN * L;
        for (int i = threadIdx.x; i < kN; i += blockDim.x) {
            T v = 0;
#pragma unroll
            for (int s = 0; s < 4; ++s)
                if (j.c[s]) v += (T)(int64_t)j.c[s] * in[(size_t)j.src[s] * A.in_stride + i];
            T state = v;
            if (tot < BITS) {
                state = v >> (BITS - tot - 1);
                state = (state + 1) >> 1;
                state &= (((T)1) << tot) - 1;
            }
            if (L == 4) {                                                   // the common shape: one 32-bit store
                uint32_t packed = 0;
#pragma unroll
                for (int l = 3; l >= 0; --l) {
                    T res = state & mask;
                    state >>= BL;
                    T carry = ((res - 1) | state) & res;
                    carry >>= (BL - 1);
                    state += carry;
                    packed |= (uint32_t)(uint8_t)(int8_t)((int32_t)res - (int32_t)(carry << BL)) << (8 * l);
                }
                reinterpret_cast<uint32_t*>(row)[i] = packed;
            } else {
                for (int l = L - 1; l >= 0; --l) {
                    T res = state & mask;
                    state >>= BL;
                    T carry = ((res - 1) | state) & res;
                    carry >>= (BL - 1);
                    state += carry;
                    row[(size_t)i * L + l] = (uint8_t)(int8_t)((int32_t)res - (int32_t)(carry << BL));
                }
            }
        }
    }
}

template <class T> __global__ void __launch_bounds__(KM_NT, 2) ks_mma2_kernel(const KsMmaArgs P, const uint8_t* dig) {
    using C = KmCfg<T>;
    constexpr int NP = C::P, TN = C::TN, NB = C::NB;
    const KsArgs& A = P.a;
    extern __shared__ __align__(16) unsigned char km_smem[];
    __shared__ KsJob jobs[KM_TM];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int n = A.n, kN = A.kN, K = kN * A.level;
    const int job0 = blockIdx.x * KM_TM, col0 = blockIdx.y * TN;
    const int njob = min(KM_TM, (int)A.n_jobs - job0);
    if (tid < KM_TM) {
        KsJob j = {{0, 0, 0, 0}, {0, 0, 0, 0}, 0, 0, 0};
        if (tid < njob) j = A.jobs[job0 + tid];
        jobs[tid] = j;
    }
    const int wm = warp & 3, wn = warp >> 2;
    int acc[NB][NP][4];
#pragma unroll
    for (int nb = 0; nb < NB; ++nb)
#pragma unroll
        for (int p = 0; p < NP; ++p)
#pragma unroll
            for (int e = 0; e < 4; ++e) acc[nb][p][e] = 0;
    const uint32_t sm_addr = (uint32_t)__cvta_generic_to_shared(km_smem);
    auto issue = [&](int k0, int stage) {
        const uint32_t base = sm_addr + (uint32_t)stage * C::STAGE;
#pragma unroll
        for (int r = 0; r < 2; ++r) {                                   // A: 64 rows x 128 bytes
            const int q = tid + r * KM_NT, row = q >> 3, part = q & 7;
            cp_async16(base + (uint32_t)(row * KM_PITCH + part * 16), dig + (size_t)(job0 + row) * K + k0 + part * 16);
        }
#pragma unroll
        for (int r = 0; r < 8; ++r) {                                   // B: P planes x TN columns x 128 bytes = 2048 16-byte parts
            const int q = tid + r * KM_NT, pc = q >> 3, part = q & 7, p = pc / TN, col = pc % TN;
            cp_async16(base + (uint32_t)((KM_TM + pc) * KM_PITCH + part * 16),
                       P.planes + ((size_t)p * P.npad + col0 + col) * K + k0 + part * 16);
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    issue(0, 0);
    const int nchunk = K / KM_KC;
    for (int ch = 0; ch < nchunk; ++ch) {
        if (ch + 1 < nchunk) { issue((ch + 1) * KM_KC, (ch + 1) & 1); asm volatile("cp.async.wait_group 1;" ::: "memory"); }
        else asm volatile("cp.async.wait_group 0;" ::: "memory");
        __syncthreads();
        const unsigned char* As = km_smem + (ch & 1) * C::STAGE;
        const unsigned char* Bs = As + KM_TM * KM_PITCH;
#pragma unroll
        for (int ks = 0; ks < KM_KC / 32; ++ks) {
            uint32_t a[4];
            const unsigned char* ap = As + (wm * 16 + (lane >> 2)) * KM_PITCH + ks * 32 + (lane & 3) * 4;
            a[0] = *reinterpret_cast<const uint32_t*>(ap);
            a[1] = *reinterpret_cast<const uint32_t*>(ap + 8 * KM_PITCH);
            a[2] = *reinterpret_cast<const uint32_t*>(ap + 16);
            a[3] = *reinterpret_cast<const uint32_t*>(ap + 8 * KM_PITCH + 16);
#pragma unroll
            for (int p = 0; p < NP; ++p)
#pragma unroll
                for (int nb = 0; nb < NB; ++nb) {
                    const unsigned char* bp = Bs + (p * TN + wn * (TN / 2) + nb * 8 + (lane >> 2)) * KM_PITCH + ks * 32 + (lane & 3) * 4;
                    mma_s8u8(acc[nb][p], a, *reinterpret_cast<const uint32_t*>(bp), *reinterpret_cast<const uint32_t*>(bp + 16));
                }
        }
        __syncthreads();
    }
    const T* in = reinterpret_cast<const T*>(A.in);
    T* out = reinterpret_cast<T*>(A.out);
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const int g = wm * 16 + (lane >> 2) + 8 * h;
        if (g >= njob) continue;
        const KsJob& j = jobs[g];
        T* o = out + (size_t)j.dst * A.out_stride;
#pragma unroll
        for (int nb = 0; nb < NB; ++nb)
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                const int c = col0 + wn * (TN / 2) + nb * 8 + (lane & 3) * 2 + e;
                if (c > n) continue;
                T sum = 0;
#pragma unroll
                for (int p = 0; p < NP; ++p) sum += (T)(int64_t)acc[nb][p][2 * h + e] << (8 * p);    // wraps like the T-wide product
                T v = 0;
                if (c == n) {
#pragma unroll
                    for (int s = 0; s < 4; ++s)
                        if (j.c[s]) v += (T)(int64_t)j.c[s] * in[(size_t)j.src[s] * A.in_stride + kN];
                    v += (T)j.cst;
                }
                o[c] = v - sum;
            }
    }
}

}  // namespace hb
