// Batched programmable bootstrap (blind rotation + sample extract) for sm_100a.
//
// Replaces, for a whole netlist level at once, what HELM reaches one gate at a time through
// tfhe-rs (reference src/gates.rs:255-274 -> boolean::ServerKey::{and,...}; src/gates.rs:754-785
// -> shortint apply_lookup_table): programmable_bootstrap_lwe_ciphertext = modulus switch,
// n sequential CMUX steps (rotate-subtract, signed gadget decomposition, (k+1)l forward folded
// negacyclic FFTs, pointwise MAC with the Fourier BSK, (k+1) inverse FFTs, round, wrapping add)
// and extraction of coefficient 0.
//
// Mapping: one CTA evaluates G gates step-synchronously so one read of BSK_i from L2 feeds G
// external products.  The accumulator ((k+1) x N torus words per gate) and the (k+1)l spectra
// stay in shared memory for all n steps.  The M = N/2 point complex FFT is split M = RA x RB:
// each thread runs a whole radix-RA (then radix-RB) DFT in registers (fft_regs.cuh), with a
// single shared-memory transpose in between (padded rows, conflict-free 128-bit accesses).
// The twist omega^j is factored as omega^{j0} * omega^{RB j1}: the second factor is a literal,
// the first is merged with the inter-stage twiddle into one table W(j0,k0).
#pragma once
#include <stdint.h>
#include "fft_regs.cuh"
#include "digits_u32.cuh"

// PBS transforms: the FMA-form decimation-in-time butterflies (fft_regs.cuh) unless -DHB_PBS_DIF; the one-off key
// transform (bsk_fft_kernel) keeps dft_regs.
#ifdef HB_PBS_DIF
#define HB_DFT dft_regs
#else
#define HB_DFT dft_regs_dit
#endif

namespace hb {

struct PbsJob {            // one bootstrap: in = c0*src0 + c1*src1 + (cst on the body)
    uint32_t src0, src1;   // row indices into the input array
    int32_t c0, c1;        // small integer coefficients (0 disables the operand)
    uint64_t cst;          // torus constant added to the body
    uint32_t dst;          // row index into the output array
    uint32_t tv_kind;      // 0: constant body tv_cst (boolean 1/8); 1: boxed LUT (shortint); 2: explicit body at tv_ptr
    uint64_t tv_cst;       // kind 0: the constant; kind 1: delta
    uint64_t lut;          // kind 1: 4 bits per table entry, entry i at bits [4i,4i+4)
    uint32_t lut_boxes;    // kind 1: number of boxes (msg*carry)
    uint32_t pad;
    const void* tv_ptr;    // kind 2: N torus words (body polynomial, zero mask)
};

struct PbsArgs {
    const void* in;   uint32_t in_stride;    // words per input row (small LWE, n+1 used)
    void* out;        uint32_t out_stride;   // words per output row (big LWE, kN+1 used)
    const PbsJob* jobs; uint32_t n_jobs;
    const double2* bsk;                      // [n][ROWS][K+1][M]
    const double2* wtab;                     // [2][M]: WA[k0*RB+j0], WB[j0*RA+k0]
    int32_t n;  int32_t base_log;
    uint16_t* atil; uint32_t atil_stride;    // pbs16_kernel: switched masks, [jobs padded to the CTA tile][atil_stride]
    const double2* roots;                    // multi-bit instances: e^{i pi t / N}, t < 2N (spectra of the monomials X^e)
};

template <class T> struct TorusOps;
template <> struct TorusOps<uint32_t> {
    static constexpr int BITS = 32;
    __device__ static __forceinline__ uint32_t round_add(double z) {   // z in LSB units, |z| < 2^51
#ifndef HB_ROUND_F2I
        return (uint32_t)__double2loint(z + 6755399441055744.0);       // DADD on the FP64 pipe
#else
        return (uint32_t)__double2ll_rn(z);      // F2I.S64.F64: measured slower (519.4 vs 515.2 ms per 64k-gate level)
#endif
    }
};
template <> struct TorusOps<uint64_t> {
    static constexpr int BITS = 64;
    // z in torus units (the Fourier BSK carries the 2^-64 scale); |z| reaches ~2^27 before reduction, so reduce
    // mod 1 in floating point first (exact), then scale to 64-bit fixed point -- tfhe-rs's from_torus (App. A.4)
    __device__ static __forceinline__ uint64_t round_add(double z) {
        const double f = z - rint(z);
        return (uint64_t)__double2ll_rn(f * 18446744073709551616.0);
    }
};

__device__ __forceinline__ double int2double_exact(int32_t d) {
    return __hiloint2double(0x43300000, (int)((uint32_t)d ^ 0x80000000u)) - 4503601774854144.0;  // 2^52+2^31
}

// signed digit of level `lev` (0-based: level lev+1), tfhe-rs SignedDecomposer order (App. A.3)
template <class T, int L> __device__ __forceinline__ int32_t digit_of(T x, int base_log, int lev) {
    constexpr int BITS = TorusOps<T>::BITS;
    const int tot = base_log * L;
    T state = x;
    if (tot < BITS) {
        state = x >> (BITS - tot - 1);
        state = (state + 1) >> 1;
        state &= (((T)1) << tot) - 1;
    }
    const T mask = (((T)1) << base_log) - 1;
    int32_t dig = 0;
#pragma unroll
    for (int l = L - 1; l >= 0; --l) {
        if (l >= lev) {
            T res = state & mask;
            state >>= base_log;
            T carry = ((res - 1) | state) & res;
            carry >>= (base_log - 1);
            state += carry;
            dig = (int32_t)res - (int32_t)(carry << base_log);
        }
    }
    return dig;
}

// all L signed digits of x at once (dig[lev] = digit of level lev+1); same recurrence as digit_of
template <class T, int L> __device__ __forceinline__ void digits_all(T x, int base_log, int32_t (&dig)[L]) {
    constexpr int BITS = TorusOps<T>::BITS;
    const int tot = base_log * L;
    T state = x;
    if (tot < BITS) {
        state = x >> (BITS - tot - 1);
        state = (state + 1) >> 1;
        state &= (((T)1) << tot) - 1;
    }
    const T mask = (((T)1) << base_log) - 1;
#pragma unroll
    for (int l = L - 1; l >= 0; --l) {
        T res = state & mask;
        state >>= base_log;
        T carry = ((res - 1) | state) & res;
        carry >>= (base_log - 1);
        state += carry;
        dig[l] = (int32_t)res - (int32_t)(carry << base_log);
    }
}

// Digit staging in shared memory between the decomposition pass and the forward-FFT pass.  One pack = the digits of
// coefficients (c, c+M) for two consecutive j1 (four values): biased bytes when base_log <= 7 (boolean), int32 otherwise.
template <bool SMALL> struct DigPack;
template <> struct DigPack<true> {
    using type = uint32_t;
    static constexpr int BIAS = 64;
    __device__ static __forceinline__ type make(int32_t a, int32_t b, int32_t c, int32_t d) {
        return (uint32_t)(a + BIAS) | ((uint32_t)(b + BIAS) << 8) | ((uint32_t)(c + BIAS) << 16) | ((uint32_t)(d + BIAS) << 24);
    }
    template <int K4> __device__ static __forceinline__ double get(type p) {      // exact: 2^52 + byte - (2^52 + 64)
        return __hiloint2double(0x43300000, (int)__byte_perm(p, 0, 0x4440 | K4)) - 4503599627370560.0;
    }
};
template <> struct DigPack<false> {
    using type = int4;
    __device__ static __forceinline__ type make(int32_t a, int32_t b, int32_t c, int32_t d) { return make_int4(a, b, c, d); }
    template <int K4> __device__ static __forceinline__ double get(type p) {
        return int2double_exact(K4 == 0 ? p.x : K4 == 1 ? p.y : K4 == 2 ? p.z : p.w);
    }
};

template <class T> __device__ __forceinline__ T rot_read(const T* p, int N, int c, int a) {
    int t = c - a; if (t < 0) t += 2 * N;           // a in [0,2N)
    return t < N ? p[t] : (T)(0 - p[t - N]);
}

template <class T> __device__ __forceinline__ T tv_body(const PbsJob& jb, int N, int idx) {
    if (jb.tv_kind == 0) return (T)jb.tv_cst;
    if (jb.tv_kind == 2) return reinterpret_cast<const T*>(jb.tv_ptr)[idx];
    const int box = N / (int)jb.lut_boxes, half = box >> 1;
    int c = idx + half; bool neg = false;
    if (c >= N) { c -= N; neg = true; }
    const T v = (T)((jb.lut >> (4 * (c / box))) & 15) * (T)jb.tv_cst;
    return neg ? (T)(0 - v) : v;
}

// ---- mbarrier / bulk-copy helpers (shared-memory addresses as 32-bit shared-window offsets)
__device__ __forceinline__ void mbar_init(uint32_t addr, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(addr), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t addr) {
    asm volatile("{ .reg .b64 st; mbarrier.arrive.shared::cta.b64 st, [%0]; }" ::"r"(addr) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t addr, uint32_t parity) {
    uint32_t ok;
    do {
        asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                     : "=r"(ok) : "r"(addr), "r"(parity) : "memory");
    } while (!ok);
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t addr, uint32_t bytes) {
    asm volatile("{ .reg .b64 st; mbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1; }" ::"r"(addr), "r"(bytes) : "memory");
}
// one contiguous global -> shared copy by the TMA unit (no tensor map), completion counted in bytes on an mbarrier
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t mbar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(mbar) : "memory");
}

template <class T, int N, int K, int L, int RA, int RB, int G>
struct PbsCfg {
    static constexpr bool SMALLDIG = sizeof(T) == 4;      // u32 parameter sets have base_log <= 7 (checked on the host)
    using Dig = DigPack<SMALLDIG>;
    static constexpr int PARTS = RA / 2;
    static constexpr int M = N / 2;
    static constexpr int ROWS = (K + 1) * L;
    static constexpr int XROW = RA * (RB + 1);
    static constexpr int TWA = HB_TW_DEN / (4 * RA);     // omega^{RB j1} = e^{2 pi i j1 TWA/256}
    static_assert(RA * RB == M, "FFT split");
    static_assert(HB_TW_DEN % (4 * RA) == 0 && HB_TW_DEN % RB == 0, "twiddle table too small");
    __host__ __device__ static constexpr size_t smem_bytes(int n) {
        return (size_t)G * ROWS * XROW * sizeof(double2) + 2 * (size_t)M * sizeof(double2) +
               (size_t)G * (K + 1) * N * sizeof(T) + (size_t)G * ROWS * PARTS * RB * sizeof(typename Dig::type) +
               (((size_t)G * (n + 1) * 2 + 15) & ~(size_t)15);
    }
    // STAGE instances: one step's Fourier GGSW (ROWS x (K+1) spectra, contiguous in the key) + the mbarrier
    static constexpr size_t KEY_BYTES = (size_t)ROWS * (K + 1) * M * sizeof(double2);
    __host__ __device__ static constexpr size_t smem_bytes_staged(int n) { return smem_bytes(n) + KEY_BYTES + 16; }
};

// ------------------------------------------------------------------ the kernel
// -DHB_PHASE_CLOCKS: thread 0 of CTA 0 accumulates clock64() deltas at the phase boundaries of every CMUX step into
// g_phase_clk (read back with hb_debug_phase_clocks): the critical-path breakdown of the narrow-batch instances, which
// PC sampling cannot give (most warps idle at barriers most of the time).  Not compiled into the product build.
#ifdef HB_PHASE_CLOCKS
__device__ unsigned long long g_phase_clk[16];
#define HB_PC_DECL unsigned long long pc_t = clock64(), pc_acc[16] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0}
#define HB_PC_MARK(slot) do { const unsigned long long pc_n = clock64(); pc_acc[slot] += pc_n - pc_t; pc_t = pc_n; } while (0)
#define HB_PC_FLUSH() do { if (threadIdx.x == 0 && blockIdx.x == 0) for (int q = 0; q < 16; ++q) g_phase_clk[q] = pc_acc[q]; } while (0)
#else
#define HB_PC_DECL
#define HB_PC_MARK(slot)
#define HB_PC_FLUSH()
#endif
// STAGE (narrow-batch instances): BSK_{i+1} is copied into shared memory by one cp.async.bulk while step i runs its inverse
// transforms and step i+1 its rotation / forward transforms, so the MAC reads the key at shared-memory speed instead of
// waiting on L2 (with one gate per CTA nothing else amortises that latency; it was 37 % of the step, profiles/r01_v6_*).
// MB > 1: multi-bit blind rotation with grouping factor MB (tfhe-rs multi_bit PBS, the reference's arithmetic-mode parameter set,
// src/bin/helm.rs:81-86).  The key holds 2^MB GGSWs per group of MB mask elements; per group, pattern q carries the monomial
// X^{e_q}, e_q = sum of the switched mask elements selected by the bits of q, and ONE external product is made with
// sum_q GGSW_q X^{e_q} (combined per gate in the Fourier domain, where X^e is the spectrum roots[e (4f+1) mod 2N] at frequency
// f): acc <- ExtProd(combined, acc).  n / MB sequential steps instead of n; the digits are those of the accumulator itself (no
// rotate-subtract) and the result replaces the accumulator.  The key slices (2^MB x 64 KB) are read through L2, once per CTA.
template <class T, int N, int K, int L, int RA, int RB, int G, int NT, int MINB, bool STAGE = false, int MB = 0>
__global__ void __launch_bounds__(NT, MINB) pbs_kernel(const PbsArgs A) {
    static_assert(MB == 0 || (MB >= 2 && MB <= 4 && !STAGE), "multi-bit instances read the key through L2");
    using C = PbsCfg<T, N, K, L, RA, RB, G>;
    constexpr int M = C::M, ROWS = C::ROWS, XROW = C::XROW;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    double2* X = reinterpret_cast<double2*>(smem_raw);                  // [G][ROWS][XROW]
    double2* WA = X + (size_t)G * ROWS * XROW;                          // [RA][RB]  W(j0,k0) at k0*RB+j0
    double2* WB = WA + M;                                               // [RB][RA]  W(j0,k0) at j0*RA+k0
    T* acc = reinterpret_cast<T*>(WB + M);                              // [G][K+1][N]
    using Dig = typename C::Dig;
    using DigT = typename Dig::type;
    constexpr int PARTS = C::PARTS;
    constexpr bool CHAIN_F = RA == RB && RA <= 32 && G * ROWS * RB <= NT;
    constexpr bool CHAIN_I = RA == RB && RA <= 32 && G * (K + 1) * RA <= NT;
#ifndef HB_TWCH
#define HB_TWCH 8
#endif
#ifndef HB_SPLIT5
#define HB_SPLIT5 1
#endif
    // inverse-transform tail on all threads: pays for the 32-point split (M2C2 step 14.0 -> 13.0 k cycles), not for 16 x 16
    constexpr bool SPLIT5 = HB_SPLIT5 && RA >= 32 && G * (K + 1) * RB * 4 <= NT;
    __shared__ double2 UT[RA];                                          // conj untwist omega^{-RB j1}
    if (SPLIT5 && threadIdx.x == 0) {
#pragma unroll
        for (int j1 = 0; j1 < RA; ++j1) UT[j1] = make_double2(HB_TW_COS[(j1 * C::TWA) & 255], -HB_TW_SIN[(j1 * C::TWA) & 255]);
    }
    constexpr int TWCH = HB_TWCH < RA && HB_TWCH < RB ? HB_TWCH : (RA < RB ? RA : RB);   // inter-stage twiddles fetched per batch
#ifndef HB_SPREAD_I
#define HB_SPREAD_I 1
#endif
    constexpr bool SPREAD_I = HB_SPREAD_I && CHAIN_I && RA == 16 && G * (K + 1) * 32 <= NT;
    DigT* digs = reinterpret_cast<DigT*>(acc + (size_t)G * (K + 1) * N);           // [G][ROWS][PARTS][RB]
    uint16_t* atil = reinterpret_cast<uint16_t*>(digs + (size_t)G * ROWS * PARTS * RB);   // [G][n+1]

    const int tid = threadIdx.x;
    const int n = A.n;
    // staged key block [ROWS][K+1][M] and its mbarrier, after the (16-byte padded) switched masks
    double2* KB = reinterpret_cast<double2*>(smem_raw + C::smem_bytes(n));
    const uint32_t kb_addr = (uint32_t)__cvta_generic_to_shared(KB);
    const uint32_t kbar = kb_addr + (uint32_t)C::KEY_BYTES;
    uint32_t kphase = 0;
    if (STAGE && tid == 0) {
        mbar_init(kbar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        mbar_expect_tx(kbar, (uint32_t)C::KEY_BYTES);
        bulk_g2s(kb_addr, A.bsk, (uint32_t)C::KEY_BYTES, kbar);
    }
    const int job0 = blockIdx.x * G;
    const T* in = reinterpret_cast<const T*>(A.in);

    for (int x = tid; x < 2 * M; x += NT) WA[x] = A.wtab[x];
    // ---- prologue: linear combination + modulus switch (App. A.5)
    for (int g = 0; g < G; ++g) {
        const int jid = job0 + g;
        if (jid >= (int)A.n_jobs) { for (int x = tid; x <= n; x += NT) atil[g * (n + 1) + x] = 0; continue; }
        const PbsJob jb = A.jobs[jid];
        const T* s0 = in + (size_t)jb.src0 * A.in_stride;
        const T* s1 = in + (size_t)jb.src1 * A.in_stride;
        for (int x = tid; x <= n; x += NT) {
            T v = 0;
            if (jb.c0) v += (T)(int64_t)jb.c0 * s0[x];
            if (jb.c1) v += (T)(int64_t)jb.c1 * s1[x];
            if (x == n) v += (T)jb.cst;
            const T sh = v >> (TorusOps<T>::BITS - ilog2c(N) - 2);
            atil[g * (n + 1) + x] = (uint16_t)(((sh + 1) >> 1) & (2 * N - 1));
        }
    }
    __syncthreads();
    // ---- accumulator = TV * X^{-b~}
    for (int g = 0; g < G; ++g) {
        const int jid = job0 + g;
        T* ag = acc + (size_t)g * (K + 1) * N;
        if (jid >= (int)A.n_jobs) { for (int x = tid; x < (K + 1) * N; x += NT) ag[x] = 0; continue; }
        const PbsJob jb = A.jobs[jid];
        const int bt = atil[g * (n + 1) + n];
        const int a0 = (2 * N - bt) & (2 * N - 1);
        for (int x = tid; x < K * N; x += NT) ag[x] = 0;
        for (int c = tid; c < N; c += NT) {
            int t = c - a0; if (t < 0) t += 2 * N;
            const T v = t < N ? tv_body<T>(jb, N, t) : (T)(0 - tv_body<T>(jb, N, t - N));
            ag[K * N + c] = v;
        }
    }
    __syncthreads();

    // ---- n CMUX steps (multi-bit: n / MB external products)
    HB_PC_DECL;
    const int n_steps = MB > 1 ? n / MB : n;
    for (int i = 0; i < n_steps; ++i) {
        HB_PC_MARK(7);
        int any = MB > 1 ? 1 : 0;                     // a multi-bit step is an external product even when every exponent is 0
#pragma unroll
        for (int g = 0; g < G; ++g) any |= atil[g * (n + 1) + (MB > 1 ? 0 : i)];
        if (any == 0) {                               // uniform across the CTA
            if (STAGE) {                              // consume the block in flight, start the next one
                mbar_wait(kbar, kphase); kphase ^= 1u;
                __syncthreads();
                if (tid == 0 && i + 1 < n) {
                    mbar_expect_tx(kbar, (uint32_t)C::KEY_BYTES);
                    bulk_g2s(kb_addr, A.bsk + (size_t)(i + 1) * ROWS * (K + 1) * M, (uint32_t)C::KEY_BYTES, kbar);
                }
            }
            continue;
        }

        // P1a: rotate-subtract and signed decomposition, once per coefficient for all L levels; digits staged in
        // shared memory in the order the FFT pass reads them ([row][part][j0], part = a pair of j1)
        for (int t = tid; t < G * (K + 1) * PARTS * RB; t += NT) {
            const int j0 = t % RB, part = (t / RB) % PARTS, r = (t / (RB * PARTS)) % (K + 1), g = t / (RB * PARTS * (K + 1));
            const int a = MB > 1 ? 0 : atil[g * (n + 1) + i];
            const T* ap = acc + ((size_t)g * (K + 1) + r) * N;
            if constexpr (MB > 1) {                   // digits of the accumulator itself
                int32_t dgm[4][L];
#pragma unroll
                for (int u = 0; u < 2; ++u) {
                    const int c = j0 + RB * (2 * part + u);
                    digits_all<T, L>(ap[c], A.base_log, dgm[2 * u]);
                    digits_all<T, L>(ap[c + M], A.base_log, dgm[2 * u + 1]);
                }
#pragma unroll
                for (int lev = 0; lev < L; ++lev)
                    digs[(((size_t)g * ROWS + lev * (K + 1) + r) * PARTS + part) * RB + j0] = Dig::make(dgm[0][lev], dgm[1][lev], dgm[2][lev], dgm[3][lev]);
                continue;
            }
#ifndef HB_PBS_GENERIC_DIGITS
            if constexpr (sizeof(T) == 4 && 7 * L <= 24) {
                // u32 sets with the usual bases: all four values decomposed at once in byte lanes (digits_u32.cuh);
                // c7552 gates mode 79.8 -> 76.3 ms
                if (A.base_log == 6 || A.base_log == 7) {
                    uint32_t xv[4], pl[L];
                    const uint32_t rnd = 1u << (31 - A.base_log * L);
#pragma unroll
                    for (int u = 0; u < 2; ++u) {
                        const int c = j0 + RB * (2 * part + u);
                        xv[2 * u] = (uint32_t)(rot_read<T>(ap, N, c, a) - ap[c]) + rnd;
                        xv[2 * u + 1] = (uint32_t)(rot_read<T>(ap, N, c + M, a) - ap[c + M]) + rnd;
                    }
                    if (A.base_log == 6) digit_lanes_u32<6, L, false>(xv, pl); else digit_lanes_u32<7, L, false>(xv, pl);
#pragma unroll
                    for (int lev = 0; lev < L; ++lev)
                        digs[(((size_t)g * ROWS + lev * (K + 1) + r) * PARTS + part) * RB + j0] = (DigT)pl[lev];
                    continue;
                }
            }
#endif
            int32_t dg[4][L];
#pragma unroll
            for (int u = 0; u < 2; ++u) {
                const int c = j0 + RB * (2 * part + u);
                digits_all<T, L>(rot_read<T>(ap, N, c, a) - ap[c], A.base_log, dg[2 * u]);
                digits_all<T, L>(rot_read<T>(ap, N, c + M, a) - ap[c + M], A.base_log, dg[2 * u + 1]);
            }
#pragma unroll
            for (int lev = 0; lev < L; ++lev)
                digs[(((size_t)g * ROWS + lev * (K + 1) + r) * PARTS + part) * RB + j0] = Dig::make(dg[0][lev], dg[1][lev], dg[2][lev], dg[3][lev]);
        }
        __syncthreads();
        HB_PC_MARK(0);
        // P1b: twist, radix-RA DFT over j1, inter-stage twiddle
        for (int t = tid; t < G * ROWS * RB; t += NT) {
            const int j0 = t % RB, grow = t / RB;
            const DigT* dp = digs + (size_t)grow * PARTS * RB + j0;
            double2 x[RA];
#pragma unroll
            for (int part = 0; part < PARTS; ++part) {
                const DigT pk = dp[part * RB];
#ifndef HB_PBS_DIF
                x[2 * part] = make_double2(Dig::template get<0>(pk), Dig::template get<1>(pk));      // twisted in the DFT's first stage
                x[2 * part + 1] = make_double2(Dig::template get<2>(pk), Dig::template get<3>(pk));
#else
                x[2 * part] = cmul_tw<1>(make_double2(Dig::template get<0>(pk), Dig::template get<1>(pk)), (2 * part) * C::TWA);
                x[2 * part + 1] = cmul_tw<1>(make_double2(Dig::template get<2>(pk), Dig::template get<3>(pk)), (2 * part + 1) * C::TWA);
#endif
            }
            HB_PC_MARK(8);
#ifndef HB_PBS_DIF
            dft_regs_dit_twisted<RA, 1, C::TWA>(x);
#else
            HB_DFT<RA, 1>(x);
#endif
            HB_PC_MARK(9);
            double2* xr = X + (size_t)grow * XROW;
            // twiddle loads are batched ahead of the stores: X and WA share the dynamic shared array, so the compiler must
            // assume a store to X may change WA and otherwise serialises load -> multiply -> store 16 times (1 180 of the
            // 7 800 cycles of a boolean step, measured with HB_PHASE_CLOCKS)
#pragma unroll
            for (int c0 = 0; c0 < RA; c0 += TWCH) {
                double2 w[TWCH], y[TWCH];
#pragma unroll
                for (int u = 0; u < TWCH; ++u) w[u] = WA[(c0 + u) * RB + j0];
#pragma unroll
                for (int u = 0; u < TWCH; ++u) y[u] = cmul(x[brev<RA>(c0 + u)], w[u]);
#pragma unroll
                for (int u = 0; u < TWCH; ++u) xr[(c0 + u) * (RB + 1) + j0] = y[u];
            }
            HB_PC_MARK(10);
        }
        // the two stages of one row run on the same RA == RB consecutive threads when a single round covers every row:
        // then the transpose only needs a warp-level sync (RA, RB <= 32 divide the warp)
        if (CHAIN_F) __syncwarp(); else __syncthreads();
        HB_PC_MARK(1);
        // P2: radix-RB DFT over j0 (column k0), in place
        for (int t = tid; t < G * ROWS * RA; t += NT) {
            const int k0 = t % RA, grow = t / RA;
            double2* xc = X + (size_t)grow * XROW + k0 * (RB + 1);
            double2 x[RB];
#pragma unroll
            for (int j0 = 0; j0 < RB; ++j0) x[j0] = xc[j0];
            HB_DFT<RB, 1>(x);
#pragma unroll
            for (int k1 = 0; k1 < RB; ++k1) xc[k1] = x[brev<RB>(k1)];
        }
        // P3: pointwise MAC with BSK_i; one read of BSK feeds all G gates.  When every position has its own thread the
        // first rows' key loads are issued before the barrier (they do not depend on the spectra)
        constexpr bool PRE = !STAGE && M <= NT;
        constexpr int NPRE = ROWS < 3 ? ROWS : 3;
        double2 bpre[NPRE][K + 1];
        if (PRE && tid < M) {
            const double2* bp0 = A.bsk + (size_t)i * ROWS * (K + 1) * M + tid;
#pragma unroll
            for (int row = 0; row < NPRE; ++row)
#pragma unroll
                for (int j = 0; j <= K; ++j) bpre[row][j] = __ldg(bp0 + (size_t)(row * (K + 1) + j) * M);
        }
        __syncthreads();
        HB_PC_MARK(2);
        if (STAGE) { mbar_wait(kbar, kphase); kphase ^= 1u; }
        HB_PC_MARK(3);
        if constexpr (MB > 1) {
            // exponents of this group's patterns per gate (pattern 0 is the constant monomial)
            int ex[G][1 << MB];
#pragma unroll
            for (int g = 0; g < G; ++g)
#pragma unroll
                for (int q = 0; q < (1 << MB); ++q) {
                    int e = 0;
#pragma unroll
                    for (int j = 0; j < MB; ++j) if ((q >> j) & 1) e += atil[g * (n + 1) + i * MB + j];
                    ex[g][q] = e & (2 * N - 1);
                }
            for (int pos = tid; pos < M; pos += NT) {
                const int s = (pos / RB) * (RB + 1) + (pos % RB);
                const int f4 = 4 * ((pos / RB) + RA * (pos % RB)) + 1;        // position k0*RB + k1 holds frequency k0 + RA*k1
                double2 cmb[G][ROWS][K + 1];                                  // combined key entry per gate
#pragma unroll
                for (int g = 0; g < G; ++g)
#pragma unroll
                    for (int row = 0; row < ROWS; ++row)
#pragma unroll
                        for (int j = 0; j <= K; ++j) cmb[g][row][j] = make_double2(0.0, 0.0);
                const double2* bp = A.bsk + (size_t)i * (1 << MB) * ROWS * (K + 1) * M + pos;
#pragma unroll 2
                for (int q = 0; q < (1 << MB); ++q) {
                    double2 b[ROWS][K + 1];
#pragma unroll
                    for (int row = 0; row < ROWS; ++row)
#pragma unroll
                        for (int j = 0; j <= K; ++j) b[row][j] = __ldg(bp + ((size_t)q * ROWS * (K + 1) + row * (K + 1) + j) * M);
#pragma unroll
                    for (int g = 0; g < G; ++g) {
                        const double2 m = __ldg(A.roots + ((ex[g][q] * f4) & (2 * N - 1)));
#pragma unroll
                        for (int row = 0; row < ROWS; ++row)
#pragma unroll
                            for (int j = 0; j <= K; ++j) cfma(cmb[g][row][j], b[row][j], m);
                    }
                }
#pragma unroll
                for (int g = 0; g < G; ++g) {
                    double2 o[K + 1];
#pragma unroll
                    for (int j = 0; j <= K; ++j) o[j] = make_double2(0.0, 0.0);
#pragma unroll
                    for (int row = 0; row < ROWS; ++row) {
                        const double2 f = X[((size_t)g * ROWS + row) * XROW + s];
#pragma unroll
                        for (int j = 0; j <= K; ++j) cfma(o[j], f, cmb[g][row][j]);
                    }
#pragma unroll
                    for (int j = 0; j <= K; ++j) X[((size_t)g * ROWS + j) * XROW + s] = o[j];
                }
            }
        } else
        for (int pos = tid; pos < M; pos += NT) {
            const int s = (pos / RB) * (RB + 1) + (pos % RB);
            double2 o[G][K + 1];
#pragma unroll
            for (int g = 0; g < G; ++g)
#pragma unroll
                for (int j = 0; j <= K; ++j) o[g][j] = make_double2(0.0, 0.0);
            const double2* bp = A.bsk + (size_t)i * ROWS * (K + 1) * M + pos;
#pragma unroll
            for (int row = 0; row < ROWS; ++row) {
                double2 b[K + 1];
#pragma unroll
                for (int j = 0; j <= K; ++j)
                    b[j] = STAGE ? KB[(size_t)(row * (K + 1) + j) * M + pos]
                                 : (PRE && row < NPRE) ? bpre[row][j] : __ldg(bp + (size_t)(row * (K + 1) + j) * M);
#pragma unroll
                for (int g = 0; g < G; ++g) {
                    const double2 f = X[((size_t)g * ROWS + row) * XROW + s];
#pragma unroll
                    for (int j = 0; j <= K; ++j) cfma(o[g][j], f, b[j]);
                }
            }
#pragma unroll
            for (int g = 0; g < G; ++g)
#pragma unroll
                for (int j = 0; j <= K; ++j) X[((size_t)g * ROWS + j) * XROW + s] = o[g][j];
        }
        __syncthreads();
        HB_PC_MARK(4);
        if (STAGE && tid == 0 && i + 1 < n) {         // every read of the staged block is behind the barrier above
            mbar_expect_tx(kbar, (uint32_t)C::KEY_BYTES);
            bulk_g2s(kb_addr, A.bsk + (size_t)(i + 1) * ROWS * (K + 1) * M, (uint32_t)C::KEY_BYTES, kbar);
        }
        // P4: inverse radix-RB DFT over k1 (column k0), conj inter-stage twiddle
        // SPREAD_I: the G(K+1) half-warps of the inverse transform sit in different warps (lanes 0..15), i.e. on different
        // schedulers, instead of back to back in the first warps (c7552: 97.4 -> 95.4 ms).  The same spreading of the
        // (K+1)L forward half-warps was measured slower (109.8 ms): half-filled warps cost full FP64 issue slots
        for (int t = SPREAD_I ? ((tid & 31) < RA ? (tid >> 5) * RA + (tid & 31) : G * (K + 1) * RA) : tid; t < G * (K + 1) * RA;
             t += SPREAD_I ? G * (K + 1) * RA : NT) {
            const int k0 = t % RA, j = (t / RA) % (K + 1), g = t / (RA * (K + 1));
            double2* xc = X + ((size_t)g * ROWS + j) * XROW + k0 * (RB + 1);
            double2 x[RB];
#pragma unroll
            for (int k1 = 0; k1 < RB; ++k1) x[k1] = xc[k1];
            HB_PC_MARK(11);
            HB_DFT<RB, -1>(x);
            HB_PC_MARK(12);
#pragma unroll
            for (int c0 = 0; c0 < RB; c0 += TWCH) {              // as in P1b: loads batched ahead of the stores
                double2 w[TWCH], y[TWCH];
#pragma unroll
                for (int u = 0; u < TWCH; ++u) w[u] = WB[(c0 + u) * RA + k0];
#pragma unroll
                for (int u = 0; u < TWCH; ++u) y[u] = cmul_conj(x[brev<RB>(c0 + u)], w[u]);
#pragma unroll
                for (int u = 0; u < TWCH; ++u) xc[c0 + u] = y[u];
            }
            HB_PC_MARK(13);
        }
        if (CHAIN_I) __syncwarp(); else __syncthreads();
        HB_PC_MARK(5);
        // P5: inverse radix-RA DFT over k0, untwist, round, wrapping add into the accumulator
        for (int t = SPREAD_I ? ((tid & 31) < RB ? (tid >> 5) * RB + (tid & 31) : G * (K + 1) * RB) : tid; t < G * (K + 1) * RB;
             t += SPREAD_I ? G * (K + 1) * RB : NT) {
            const int j0 = t % RB, j = (t / RB) % (K + 1), g = t / (RB * (K + 1));
            double2* xr = X + ((size_t)g * ROWS + j) * XROW;
            double2 x[RA];
#pragma unroll
            for (int k0 = 0; k0 < RA; ++k0) x[k0] = xr[k0 * (RB + 1) + j0];
            HB_PC_MARK(14);
            HB_DFT<RA, -1>(x);
            HB_PC_MARK(15);
            if (SPLIT5) {                                  // the tail runs on every thread below: leave z in place
#pragma unroll
                for (int j1 = 0; j1 < RA; ++j1) xr[j1 * (RB + 1) + j0] = x[brev<RA>(j1)];
                continue;
            }
            T* ap = acc + ((size_t)g * (K + 1) + j) * N;
            // accumulator words are fetched a batch ahead of the stores (same reason as the twiddles in P1b / P4)
#pragma unroll
            for (int c0 = 0; c0 < RA; c0 += TWCH) {
                T lo[TWCH], hi[TWCH];
#pragma unroll
                for (int u = 0; u < TWCH; ++u) { lo[u] = ap[j0 + RB * (c0 + u)]; hi[u] = ap[j0 + RB * (c0 + u) + M]; }
#pragma unroll
                for (int u = 0; u < TWCH; ++u) {
                    const double2 z = cmul_tw<-1>(x[brev<RA>(c0 + u)], (c0 + u) * C::TWA);
                    if (MB > 1) { lo[u] = TorusOps<T>::round_add(z.x); hi[u] = TorusOps<T>::round_add(z.y); }   // acc <- product
                    else { lo[u] += TorusOps<T>::round_add(z.x); hi[u] += TorusOps<T>::round_add(z.y); }
                }
#pragma unroll
                for (int u = 0; u < TWCH; ++u) { ap[j0 + RB * (c0 + u)] = lo[u]; ap[j0 + RB * (c0 + u) + M] = hi[u]; }
            }
        }
        if (SPLIT5) {
            // P5b: untwist, round and the wrapping add, spread over all NT threads (the transform above keeps only
            // G(K+1)RB of them busy; this tail was 2 250 of 14 000 cycles of an M2C2 step on 64 threads)
            __syncthreads();
            constexpr int PER = (G * (K + 1) * M + NT - 1) / NT;
            T lo[PER], hi[PER];
            double2 z[PER];
#pragma unroll
            for (int q = 0; q < PER; ++q) {
                const int e = tid + q * NT;
                if (e < G * (K + 1) * M) {
                    const int pos = e % M, row = e / M, j1 = pos / RB, j0 = pos % RB;   // row = g * (K+1) + j
                    const double2 v = X[((size_t)(row / (K + 1)) * ROWS + row % (K + 1)) * XROW + j1 * (RB + 1) + j0];
                    const double2 w = UT[j1];
                    z[q] = make_double2(fma(v.x, w.x, -(v.y * w.y)), fma(v.x, w.y, v.y * w.x));
                    const T* ap = acc + (size_t)row * N;
                    lo[q] = ap[pos]; hi[q] = ap[pos + M];
                }
            }
#pragma unroll
            for (int q = 0; q < PER; ++q) {
                const int e = tid + q * NT;
                if (e < G * (K + 1) * M) {
                    T* ap = acc + (size_t)(e / M) * N;
                    ap[e % M] = (MB > 1 ? (T)0 : lo[q]) + TorusOps<T>::round_add(z[q].x);
                    ap[e % M + M] = (MB > 1 ? (T)0 : hi[q]) + TorusOps<T>::round_add(z[q].y);
                }
            }
        }
        __syncthreads();
        HB_PC_MARK(6);
    }
    HB_PC_FLUSH();

    // ---- epilogue: sample extract coefficient 0 (App. A.5)
    T* out = reinterpret_cast<T*>(A.out);
    for (int g = 0; g < G; ++g) {
        const int jid = job0 + g;
        if (jid >= (int)A.n_jobs) continue;
        const uint32_t dst = A.jobs[jid].dst;
        const T* ag = acc + (size_t)g * (K + 1) * N;
        T* o = out + (size_t)dst * A.out_stride;
        for (int x = tid; x < K * N; x += NT) {
            const int j = x / N, c = x % N;
            o[x] = c == 0 ? ag[j * N] : (T)(0 - ag[j * N + N - c]);
        }
        if (tid == 0) o[K * N] = ag[K * N];
    }
}

// ------------------------------------------------------------------ BSK: standard -> Fourier (once per key)
// One CTA per polynomial; same two-stage transform and the same output position order as the PBS kernel.
// scale = 2^(32-BITS) / M folds the inverse-FFT normalisation and the torus scaling into the key.
template <class T, int N, int RA, int RB>
__global__ void bsk_fft_kernel(const T* __restrict__ std_bsk, double2* __restrict__ out, const double2* __restrict__ wtab, double scale) {
    constexpr int M = N / 2, TWA = HB_TW_DEN / (4 * RA);
    __shared__ double2 X[RA * (RB + 1)];
    const T* p = std_bsk + (size_t)blockIdx.x * N;
    using ST = typename std::conditional<sizeof(T) == 4, int32_t, int64_t>::type;
    for (int j0 = threadIdx.x; j0 < RB; j0 += blockDim.x) {
        double2 x[RA];
#pragma unroll
        for (int j1 = 0; j1 < RA; ++j1) {
            const int c = j0 + RB * j1;
            const double2 v = make_double2((double)(ST)p[c] * scale, (double)(ST)p[c + M] * scale);
            x[j1] = cmul_tw<1>(v, j1 * TWA);
        }
        dft_regs<RA, 1>(x);
#pragma unroll
        for (int k0 = 0; k0 < RA; ++k0) X[k0 * (RB + 1) + j0] = cmul(x[brev<RA>(k0)], wtab[k0 * RB + j0]);
    }
    __syncthreads();
    for (int k0 = threadIdx.x; k0 < RA; k0 += blockDim.x) {
        double2 x[RB];
#pragma unroll
        for (int j0 = 0; j0 < RB; ++j0) x[j0] = X[k0 * (RB + 1) + j0];
        dft_regs<RB, 1>(x);
#pragma unroll
        for (int k1 = 0; k1 < RB; ++k1) out[(size_t)blockIdx.x * M + k0 * RB + k1] = x[brev<RB>(k1)];
    }
}

// roots[t] = e^{i pi t / N}, t < 2N: the spectrum of the monomial X^e at frequency f is roots[e (4f+1) mod 2N] (multi-bit instances)
__global__ void roots_kernel(double2* roots, int N) {
    for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < 2 * N; t += gridDim.x * blockDim.x) {
        double s, c;
        sincospi((double)t / (double)N, &s, &c);
        roots[t] = make_double2(c, s);
    }
}

// W(j0,k0) = e^{2 pi i j0 (4 k0 + 1) / (4M)} in both access orders
__global__ void wtab_kernel(double2* wtab, int RA, int RB) {
    const int M = RA * RB;
    for (int x = blockIdx.x * blockDim.x + threadIdx.x; x < M; x += gridDim.x * blockDim.x) {
        const int k0 = x / RB, j0 = x % RB;
        const long num = ((long)j0 * (4 * k0 + 1)) % (4L * M);
        double s, c;
        sincospi((double)num / (double)(2 * M), &s, &c);
        wtab[k0 * RB + j0] = make_double2(c, s);
        wtab[M + j0 * RA + k0] = make_double2(c, s);
    }
}

}  // namespace hb
