// Batched LWE keyswitch (big key kN -> small key n) for sm_100a.
//
// Replaces tfhe-rs keyswitch_lwe_ciphertext (reached from reference src/gates.rs:255-274 after the
// bootstrap for boolean gates, before it for shortint LUTs):
//   out = (0,...,0,b_in) - sum_i sum_l d_{i,l} * KSK[i][l]     (wrapping, digits in [-B/2,B/2])
// Batched over TB gates per CTA it is an integer GEMM [TB x kN*l] . [kN*l x (n+1)]: every KSK word
// fetched from L2 is used TB times from registers; the TB digits of one (i,l) are one broadcast
// 128-bit shared-memory load.  The gate's linear pre/post-processing (MUX: sum of two bootstraps
// + 1/8; shortint bivariate: l*factor + r) is fused into the digit prologue.
#pragma once
#include <stdint.h>
#include <type_traits>

namespace hb {

struct KsJob {             // in = sum_s c[s]*src[s] + (cst on the body); c[s]==0 disables operand s
    uint32_t src[4];
    int32_t c[4];
    uint64_t cst;
    uint32_t dst;
    uint32_t pad;
};

struct KsArgs {
    const void* in;  uint32_t in_stride;     // big LWE rows (kN+1 words used)
    void* out;       uint32_t out_stride;    // small LWE rows (n+1 words used)
    const KsJob* jobs; uint32_t n_jobs;
    const void* ksk;                         // [kN][l][n+1]
    int32_t n, kN, base_log, level;
    int32_t split;                           // > 1: gridDim.y CTAs share one tile, each takes kN/split mask words and adds
                                             // its partial sum atomically into rows prepared by ks_init_rows (narrow batches)
};

constexpr int KS_TB = 16;      // gates per CTA
constexpr int KS_NT = 256;     // threads per CTA
constexpr int KS_QC = 3;       // output columns per thread: (n+1) <= KS_NT*KS_QC
constexpr int KS_IC = 128;     // input mask words per digit chunk
constexpr int KS_MAXL = 8;

template <class T, bool SPLIT> __global__ void __launch_bounds__(KS_NT) ks_kernel(const KsArgs A) {
    constexpr int BITS = sizeof(T) * 8;
    __shared__ __align__(16) int8_t dig[KS_IC * KS_MAXL * KS_TB];      // [ii*L + lev][g]
    __shared__ KsJob jobs[KS_TB];
    const int tid = threadIdx.x, n = A.n, kN = A.kN, L = A.level, BL = A.base_log;
    const int job0 = blockIdx.x * KS_TB;
    const int njob = min(KS_TB, (int)A.n_jobs - job0);
    if (tid < KS_TB) {
        KsJob j = {{0, 0, 0, 0}, {0, 0, 0, 0}, 0, 0, 0};
        if (tid < njob) j = A.jobs[job0 + tid];
        jobs[tid] = j;
    }
    __syncthreads();
    const T* in = reinterpret_cast<const T*>(A.in);
    const T* ksk = reinterpret_cast<const T*>(A.ksk);
    T acc[KS_TB][KS_QC];
#pragma unroll
    for (int g = 0; g < KS_TB; ++g)
#pragma unroll
        for (int q = 0; q < KS_QC; ++q) acc[g][q] = 0;

    const int tot = BL * L;
    const T mask = (((T)1) << BL) - 1;
    const int split = SPLIT ? A.split : 1;
    const int per = ((kN + split - 1) / split + 3) & ~3;
    const int ibeg = blockIdx.y * per, ilim = min(kN, ibeg + per);
    for (int i0 = ibeg; i0 < ilim; i0 += KS_IC) {
        // digits of this chunk for all TB gates (tfhe-rs SignedDecomposer, App. A.3), stored [ii*L + l][g]
        for (int t = tid; t < KS_IC * KS_TB; t += KS_NT) {
            const int ii = t % KS_IC, g = t / KS_IC, i = i0 + ii;
            T v = 0;
            if (i < ilim && g < njob) {
                const KsJob& j = jobs[g];
#pragma unroll
                for (int s = 0; s < 4; ++s)
                    if (j.c[s]) v += (T)(int64_t)j.c[s] * in[(size_t)j.src[s] * A.in_stride + i];
            }
            T state = v;
            if (tot < BITS) {
                state = v >> (BITS - tot - 1);
                state = (state + 1) >> 1;
                state &= (((T)1) << tot) - 1;
            }
            for (int l = L - 1; l >= 0; --l) {
                T res = state & mask;
                state >>= BL;
                T carry = ((res - 1) | state) & res;
                carry >>= (BL - 1);
                state += carry;
                dig[(ii * L + l) * KS_TB + g] = (int8_t)((int32_t)res - (int32_t)(carry << BL));
            }
        }
        __syncthreads();
        // rows (i, l) of the key are contiguous: one flat loop, unrolled so several rows' loads are in flight
        const int rend = min(KS_IC, ilim - i0) * L;
        const T* rowp = ksk + (size_t)i0 * L * (n + 1);
#pragma unroll(SPLIT ? 4 : 1)
        for (int rr = 0; rr < rend; ++rr) {
            const T* row = rowp + (size_t)rr * (n + 1);
            T kv[KS_QC];
#pragma unroll
            for (int q = 0; q < KS_QC; ++q) {
                const int c = tid + q * KS_NT;
                kv[q] = c <= n ? __ldg(row + c) : (T)0;
            }
            const int4 dv = *reinterpret_cast<const int4*>(&dig[rr * KS_TB]);
            const int words[4] = {dv.x, dv.y, dv.z, dv.w};
#pragma unroll
            for (int g = 0; g < KS_TB; ++g) {
                const int d = (int)(int8_t)((words[g >> 2] >> (8 * (g & 3))) & 0xff);
                const T dT = (T)(int64_t)d;
#pragma unroll
                for (int q = 0; q < KS_QC; ++q) acc[g][q] += dT * kv[q];
            }
        }
        __syncthreads();
    }
    T* out = reinterpret_cast<T*>(A.out);
#pragma unroll
    for (int g = 0; g < KS_TB; ++g) {
        if (g < njob) {
            const KsJob& j = jobs[g];
#pragma unroll
            for (int q = 0; q < KS_QC; ++q) {
                const int c = tid + q * KS_NT;
                if (c <= n) {
                    T* o = out + (size_t)j.dst * A.out_stride + c;
                    if (SPLIT) {                          // wrapping integer adds commute: the result is bit-identical
                        using AT = typename std::conditional<sizeof(T) == 4, unsigned int, unsigned long long>::type;
                        atomicAdd(reinterpret_cast<AT*>(o), (AT)(T)(0 - acc[g][q]));
                    } else {
                        T v = 0;
                        if (c == n) {
#pragma unroll
                            for (int s = 0; s < 4; ++s)
                                if (j.c[s]) v += (T)(int64_t)j.c[s] * in[(size_t)j.src[s] * A.in_stride + kN];
                            v += (T)j.cst;
                        }
                        *o = v - acc[g][q];
                    }
                }
            }
        }
    }
}

// split mode, first pass: out row = (0, ..., 0, body of the fused linear combination)
template <class T> __global__ void ks_init_rows(const KsArgs A) {
    const T* in = reinterpret_cast<const T*>(A.in);
    T* out = reinterpret_cast<T*>(A.out);
    for (uint32_t r = blockIdx.x; r < A.n_jobs; r += gridDim.x) {
        const KsJob j = A.jobs[r];
        for (int c = threadIdx.x; c <= A.n; c += blockDim.x) {
            T v = 0;
            if (c == A.n) {
#pragma unroll
                for (int s = 0; s < 4; ++s)
                    if (j.c[s]) v += (T)(int64_t)j.c[s] * in[(size_t)j.src[s] * A.in_stride + A.kN];
                v += (T)j.cst;
            }
            out[(size_t)j.dst * A.out_stride + c] = v;
        }
    }
}

}  // namespace hb
