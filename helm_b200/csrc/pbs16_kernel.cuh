// Batched programmable bootstrap for N = 512 (M = 256 = 16 x 16), second-generation mapping (boolean default and
// PARAM_MESSAGE_1_CARRY_1): same arithmetic as pbs_kernel.cuh (bit-identical results), different work distribution.
//
// Why (profiles/r01_v2_*): in pbs_kernel every FFT stage ends in a CTA-wide barrier, the inverse transforms run on a
// third of the threads and only 4 gates fit one SM; barrier stalls were the top stall reason at 42 % FP64-pipe use.
//
// Mapping here: thread = (gate g, polynomial r, idx in [0,16)); the 16 threads of one (g, r) own that accumulator
// polynomial for the whole blind rotation.  They rotate/decompose it, run its forward FFTs, and later its inverse FFT
// and the rounding add -- the 16x16 transposes between FFT stages stay inside a half-warp, so those stages are chained
// with __syncwarp() only.  The single CTA-wide exchange is the pointwise MAC with BSK_i, which mixes polynomials:
// it is done level by level ((K+1) spectra per gate in shared memory instead of (K+1)L, so 8 boolean gates fit one
// SM), accumulating the (K+1) output spectra in registers across levels; BSK_i is read once per CTA and feeds G gates,
// and its loads are issued before the barrier they would otherwise wait behind.  The "spectra consumed" side of that
// exchange is a split arrive / wait (mbarrier), so the next level's register-resident DFT overlaps other warps' MAC.
//
// u32 torus: the accumulator is kept as its negacyclic extension E[t], t in [0, 2N), E[t+N] = -E[t], so the rotated
// read X^a * acc is E[(c - a) mod 2N] with no sign logic, and the signed decomposition uses a closed form of the
// tfhe-rs carry rule with compile-time base_log (profiles/r01_v3_*: the integer ALU pipe bounded that phase).
#pragma once
#include "pbs_kernel.cuh"
#include "digits_u32.cuh"

// Left-over MAC items (position, output polynomial): position-minor, so a warp's key loads are one contiguous run.
// -DHB_PBS16_ITEM_BY_POSITION gives consecutive threads the K+1 outputs of one position instead (the spectrum load becomes
// a shared-memory broadcast, but the key loads split into K+1 runs): measured slower, 530.1 vs 516.4 ms per 64k-gate level.
#ifndef HB_PBS16_ITEM_BY_POSITION
#define HB_ITEM_POS(item) ((item) % (REM > 0 ? REM : 1))
#define HB_ITEM_OUT(item) ((item) / (REM > 0 ? REM : 1))
#else
#define HB_ITEM_POS(item) ((item) / (K + 1))
#define HB_ITEM_OUT(item) ((item) % (K + 1))
#endif
#ifdef HB_PBS16_DIF
#define HB_DFT16 dft_regs
#else
#define HB_DFT16 dft_regs_dit
#endif

namespace hb {

__device__ __forceinline__ double2 ldg_nc(const double2* p) {
    double2 v;
    asm volatile("ld.global.nc.v2.f64 {%0, %1}, [%2];" : "=d"(v.x), "=d"(v.y) : "l"(p));
    return v;
}
// ---- tensor memory as per-thread scratch.  The kernel issues no MMA; TMEM (256 KB per SM, its own datapath) holds the
// 2 x 16 inter-stage twiddles every thread needs in each forward / inverse transform, which otherwise are 15 % of the
// shared-memory wavefronts (profiles/r01_v4_*).  Shape 32x32b: lane i of a warp reads/writes TMEM lane 32*(warp%4)+i,
// consecutive 32-bit columns -- i.e. thread-private storage.
__device__ __forceinline__ void tmem_alloc(uint32_t smem_dst, uint32_t ncols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_dst), "r"(ncols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tmem_st4(uint32_t taddr, double2 v) {      // one double2 = 4 columns
    asm volatile("tcgen05.st.sync.aligned.32x32b.x4.b32 [%0], {%1, %2, %3, %4};" ::"r"(taddr),
                 "r"(__double2loint(v.x)), "r"(__double2hiint(v.x)), "r"(__double2loint(v.y)), "r"(__double2hiint(v.y)) : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, double2 (&w)[4]) {  // four double2 = 16 columns
    uint32_t r[16];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                   "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
                 : "r"(taddr) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int q = 0; q < 4; ++q)
        w[q] = make_double2(__hiloint2double((int)r[4 * q + 1], (int)r[4 * q]), __hiloint2double((int)r[4 * q + 3], (int)r[4 * q + 2]));
}

// The same load split into issue and wait, so the next 16 columns are in flight while the current ones are used
// (-DHB_PBS16_TMEM_PIPE).  The wait names the registers as in/out operands: nothing may read them before it.
__device__ __forceinline__ void tmem_ld16_issue(uint32_t taddr, uint32_t (&r)[16]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                   "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
                 : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_ld16_wait(uint32_t (&r)[16]) {
    asm volatile("tcgen05.wait::ld.sync.aligned;"
                 : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]),
                   "+r"(r[8]), "+r"(r[9]), "+r"(r[10]), "+r"(r[11]), "+r"(r[12]), "+r"(r[13]), "+r"(r[14]), "+r"(r[15])
                 :: "memory");
}
__device__ __forceinline__ double2 tmem_pair(const uint32_t (&r)[16], int u) {
    return make_double2(__hiloint2double((int)r[4 * u + 1], (int)r[4 * u]), __hiloint2double((int)r[4 * u + 3], (int)r[4 * u + 2]));
}
// Byte B of a packed digit word (two's-complement digits) as a double: one I2F.F64.S8 with a byte selector (conversion
// unit, not the FP64 pipe) instead of PRMT + MOV + DADD for the 2^52 magic-number form of DigPack<true>::get.
template <int B> __device__ __forceinline__ double cvt_s8(uint32_t p) {
    if constexpr (B == 0) {      // the cast form compiles byte 0 to two PRMTs + I2F.F64.S16
        double d;
        asm("cvt.rn.f64.s8 %0, %1;" : "=d"(d) : "r"(p));
        return d;
    } else return (double)(int8_t)(p >> (8 * B));
}
template <class T, int K, int L, int G, bool TMEMW>
struct Pbs16Cfg {
    static constexpr int N = 512, M = 256, R = 16;
    static constexpr int NT = G * (K + 1) * R;
    static constexpr int XROW = R * (R + 1);               // one spectrum: [k0][j0], rows padded to 17 double2
    static constexpr bool EXT = sizeof(T) == 4;            // negacyclic-extended accumulator (2N words per polynomial)
    static constexpr int ASTRIDE = (EXT ? 2 * N : N) + 16;  // +16 words: odd polynomials land on the other 16 banks
    // MAC work split: PP whole positions per thread (all gates, all outputs) + the REM left-over positions as
    // (position, output) pairs, I2 per thread (all gates) -- every thread does the same number of complex FMAs
    static constexpr int PP = NT >= M ? 1 : M / NT;
    static constexpr int REM = NT >= M ? 0 : M - PP * NT;
    static constexpr int I2 = REM * (K + 1) / NT;
    static constexpr int NGRP = PP > I2 ? PP : I2;
    static_assert(NT >= M ? NT == M : (REM * (K + 1)) % NT == 0, "MAC split must divide evenly");
    static_assert(NT % 32 == 0, "whole warps");
    __host__ __device__ static constexpr size_t smem_bytes(int) {
        return (size_t)G * (K + 1) * XROW * sizeof(double2) + (TMEMW ? 0 : (size_t)XROW * sizeof(double2)) +
               (size_t)G * (K + 1) * ASTRIDE * sizeof(T) + 16;
    }
};

template <class T, int K, int L, int BL, int G, int MINB, bool TMEMW>
__global__ void __launch_bounds__(G * (K + 1) * 16, MINB) pbs16_kernel(const PbsArgs A) {
    using C = Pbs16Cfg<T, K, L, G, TMEMW>;
    constexpr int N = C::N, M = C::M, R = 16, NT = C::NT, XROW = C::XROW, AS = C::ASTRIDE, ROWS = (K + 1) * L;
    constexpr int TWA = HB_TW_DEN / (4 * R);
    constexpr bool EXT = C::EXT;
    constexpr int PP = C::PP, I2 = C::I2, NGRP = C::NGRP;
    [[maybe_unused]] constexpr int REM = C::REM;
    using Dig = DigPack<sizeof(T) == 4>;
    using DigT = typename Dig::type;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    double2* X = reinterpret_cast<double2*>(smem_raw);                  // [G][K+1][XROW]
    double2* W = X + (size_t)G * (K + 1) * XROW;                        // W(j0,k0) at k0*17+j0 (only without TMEM)
    T* acc = reinterpret_cast<T*>(W + (TMEMW ? 0 : XROW));              // [G][K+1][AS]

    const int tid = threadIdx.x;
    const int n = A.n;
    const int job0 = blockIdx.x * G;
    const T* in = reinterpret_cast<const T*>(A.in);
    const uint32_t mbar = (uint32_t)__cvta_generic_to_shared(acc + (size_t)G * (K + 1) * AS);
    if (tid == 0) mbar_init(mbar, NT);
    constexpr uint32_t TM_COLS = ((NT / 32 + 3) / 4) * 128;           // 2 x 64 columns per thread, warps sharing a lane quadrant side by side
    static_assert(!TMEMW || (TM_COLS == 128 || TM_COLS == 256), "TMEM allocation must be a power of two");
    if (TMEMW && tid < 32) tmem_alloc(mbar + 8, TM_COLS);

    if (!TMEMW) for (int x = tid; x < M; x += NT) W[(x >> 4) * (R + 1) + (x & 15)] = A.wtab[x];
    // ---- prologue: linear combination + modulus switch (App. A.5); the switched mask a~ goes to a global scratch row per
    // job (read back one value per step through L1), the switched body b~ is used right away
    for (int g = 0; g < G; ++g) {
        const int jid = job0 + g;
        uint16_t* at = A.atil + (size_t)jid * A.atil_stride;            // rows exist for the padded job count
        if (jid >= (int)A.n_jobs) { for (int x = tid; x <= n; x += NT) at[x] = 0; continue; }
        const PbsJob jb = A.jobs[jid];
        const T* s0 = in + (size_t)jb.src0 * A.in_stride;
        const T* s1 = in + (size_t)jb.src1 * A.in_stride;
        for (int x = tid; x <= n; x += NT) {
            T v = 0;
            if (jb.c0) v += (T)(int64_t)jb.c0 * s0[x];
            if (jb.c1) v += (T)(int64_t)jb.c1 * s1[x];
            if (x == n) v += (T)jb.cst;
            const T sh = v >> (TorusOps<T>::BITS - ilog2c(N) - 2);
            at[x] = (uint16_t)(((sh + 1) >> 1) & (2 * N - 1));
        }
    }
    __syncthreads();                                                    // global writes of this CTA visible to this CTA
    // ---- accumulator = TV * X^{-b~}
    for (int g = 0; g < G; ++g) {
        const int jid = job0 + g;
        T* ag = acc + (size_t)g * (K + 1) * AS;
        if (jid >= (int)A.n_jobs) { for (int x = tid; x < (K + 1) * AS; x += NT) ag[x] = 0; continue; }
        const PbsJob jb = A.jobs[jid];
        const int bt = A.atil[(size_t)jid * A.atil_stride + n];
        const int a0 = (2 * N - bt) & (2 * N - 1);
        for (int x = tid; x < K * AS; x += NT) ag[x] = 0;
        for (int c = tid; c < N; c += NT) {
            int t = c - a0; if (t < 0) t += 2 * N;
            const T v = t < N ? tv_body<T>(jb, N, t) : (T)(0 - tv_body<T>(jb, N, t - N));
            ag[K * AS + c] = v;
            if (EXT) ag[K * AS + N + c] = (T)(0 - v);
        }
    }
    __syncthreads();

    // ---- roles
    const int idx = tid & 15;                 // j0 in the first FFT stage, k0 in the second
    const int gr = tid >> 4;                  // g*(K+1) + r
    const int g_own = gr / (K + 1);
    T* ap = acc + (size_t)gr * AS;            // this half-warp's accumulator polynomial
    double2* xr = X + (size_t)gr * XROW;      // and its spectrum buffer
    const uint16_t* at_own = A.atil + (size_t)(job0 + g_own) * A.atil_stride;
    // MAC roles: group q = whole position tid + q*NT (q < PP) and left-over item tid + q*NT (q < I2), an item being
    // (position PP*NT + item % REM, output polynomial item / REM)
    auto spos = [](int pos) { return (pos >> 4) * (R + 1) + (pos & 15); };
    uint32_t parity = 0;
    uint32_t tw_b = 0, tw_e = 0;                                      // TMEM addresses of this thread's W(idx, .) and W(., idx)
    if constexpr (TMEMW) {
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncthreads();
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        uint32_t tbase;
        asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tbase) : "r"(mbar + 8) : "memory");
        const int warp = tid >> 5;
        tw_b = tbase + ((uint32_t)((warp & 3) * 32) << 16) + (uint32_t)(warp >> 2) * 128u;
        tw_e = tw_b + 64u;
#pragma unroll
        for (int q = 0; q < R; ++q) {                                   // wtab[k0*16 + j0] = W(j0, k0)
            tmem_st4(tw_b + 4 * q, A.wtab[q * R + idx]);
            tmem_st4(tw_e + 4 * q, A.wtab[idx * R + q]);
        }
        asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    }

    int a_next = at_own[0];
    HB_PC_DECL;
    for (int i = 0; i < n; ++i) {
        HB_PC_MARK(10);
        const int a = a_next;
        a_next = at_own[i + 1];                                         // index n is the body: harmless, never used as a mask
        // A: rotate-subtract and signed decomposition of this thread's 32 coefficients c = idx + 16 m, all L levels.
        // pk[lev][part] packs the digits of coefficients (c, c+M) for j1 = 2 part and 2 part + 1 (DigPack order).
        DigT pk[L][R / 2];
        {
            if constexpr (EXT) {
                const uint32_t b4 = (uint32_t)((idx - a) & (2 * N - 1)) << 2;     // byte offset of E[(idx - a) mod 2N]
                const unsigned char* eb = reinterpret_cast<const unsigned char*>(ap);
                constexpr uint32_t RND = 1u << (31 - BL * L);
#pragma unroll
                for (int part = 0; part < R / 2; ++part) {
#if !defined(HB_PBS16_MAGIC_CVT) && !defined(HB_PBS16_SCALAR_DIGITS)
                    uint32_t xv[4], pl[L];
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        const int m = 2 * part + (u >> 1) + 16 * (u & 1);        // u: (c), (c+M), (c+16), (c+16+M)
                        const uint32_t rot = *reinterpret_cast<const uint32_t*>(eb + ((b4 + 64u * m) & (8u * N - 1)));
                        xv[u] = rot - (uint32_t)ap[idx + R * m] + RND;
                    }
                    digit_lanes_u32<BL, L>(xv, pl);
#pragma unroll
                    for (int lev = 0; lev < L; ++lev) pk[lev][part] = pl[lev];
#else
                    uint32_t by[4][L];
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        const int m = 2 * part + (u >> 1) + 16 * (u & 1);        // u: (c), (c+M), (c+16), (c+16+M)
                        const uint32_t rot = *reinterpret_cast<const uint32_t*>(eb + ((b4 + 64u * m) & (8u * N - 1)));
                        digit_bytes_u32<BL, L>(rot - (uint32_t)ap[idx + R * m] + RND, by[u]);
                    }
#pragma unroll
                    for (int lev = 0; lev < L; ++lev)
#ifdef HB_PBS16_MAGIC_CVT
                        pk[lev][part] = by[0][lev] + (by[1][lev] << 8) + (by[2][lev] << 16) + (by[3][lev] << 24);
#else
                        pk[lev][part] = __byte_perm(__byte_perm(by[0][lev], by[1][lev], 0x0040), __byte_perm(by[2][lev], by[3][lev], 0x0040), 0x5410);
#endif
#endif
                }
            } else {
#pragma unroll
                for (int part = 0; part < R / 2; ++part) {
                    int32_t dg[4][L];
#pragma unroll
                    for (int u = 0; u < 2; ++u) {
                        const int c = idx + R * (2 * part + u);
                        digits_all<T, L>(rot_read<T>(ap, N, c, a) - ap[c], BL, dg[2 * u]);
                        digits_all<T, L>(rot_read<T>(ap, N, c + M, a) - ap[c + M], BL, dg[2 * u + 1]);
                    }
#pragma unroll
                    for (int lev = 0; lev < L; ++lev) pk[lev][part] = Dig::make(dg[0][lev], dg[1][lev], dg[2][lev], dg[3][lev]);
                }
            }
        }
        HB_PC_MARK(0);
        double2 o1[PP][G][K + 1];
        double2 o2[I2 > 0 ? I2 : 1][G];
#pragma unroll
        for (int g = 0; g < G; ++g) {
#pragma unroll
            for (int q = 0; q < (I2 > 0 ? I2 : 1); ++q) o2[q][g] = make_double2(0.0, 0.0);
#pragma unroll
            for (int q = 0; q < PP; ++q)
#pragma unroll
                for (int j = 0; j <= K; ++j) o1[q][g][j] = make_double2(0.0, 0.0);
        }

#ifdef HB_PBS16_UNROLL_LEV
#pragma unroll
#else
#pragma unroll 1
#endif
        for (int lev = 0; lev < L; ++lev) {
            // B: twist, radix-16 DFT over j1, inter-stage twiddle (all in registers) ...
            {
                double2 x[R];
#pragma unroll
                for (int part = 0; part < R / 2; ++part) {
                    const DigT p = pk[0][part];
#ifndef HB_PBS16_MAGIC_CVT
                    if constexpr (EXT) {
#if !defined(HB_PBS16_SEPARATE_TWIST) && !defined(HB_PBS16_DIF)
                        x[2 * part] = make_double2(cvt_s8<0>(p), cvt_s8<1>(p));          // twisted inside the DFT's first stage
                        x[2 * part + 1] = make_double2(cvt_s8<2>(p), cvt_s8<3>(p));
#else
                        x[2 * part] = cmul_tw<1>(make_double2(cvt_s8<0>(p), cvt_s8<1>(p)), (2 * part) * TWA);
                        x[2 * part + 1] = cmul_tw<1>(make_double2(cvt_s8<2>(p), cvt_s8<3>(p)), (2 * part + 1) * TWA);
#endif
                    } else
#endif
                    {
                        x[2 * part] = cmul_tw<1>(make_double2(Dig::template get<0>(p), Dig::template get<1>(p)), (2 * part) * TWA);
                        x[2 * part + 1] = cmul_tw<1>(make_double2(Dig::template get<2>(p), Dig::template get<3>(p)), (2 * part + 1) * TWA);
                    }
                }
#if !defined(HB_PBS16_SEPARATE_TWIST) && !defined(HB_PBS16_MAGIC_CVT) && !defined(HB_PBS16_DIF)
                if constexpr (EXT) dft_regs_dit_twisted<R, 1, TWA>(x);
                else
#endif
                HB_DFT16<R, 1>(x);
                double2 y[R];
                if constexpr (TMEMW) {
#ifdef HB_PBS16_TMEM_PIPE
                    uint32_t rw[2][16];
                    tmem_ld16_issue(tw_b, rw[0]);
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        tmem_ld16_wait(rw[q & 1]);
                        if (q < 3) tmem_ld16_issue(tw_b + 16 * (q + 1), rw[(q + 1) & 1]);
#pragma unroll
                        for (int u = 0; u < 4; ++u) y[4 * q + u] = cmul(x[brev<R>(4 * q + u)], tmem_pair(rw[q & 1], u));
                    }
#else
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        double2 w4[4];
                        tmem_ld16(tw_b + 16 * q, w4);
#pragma unroll
                        for (int u = 0; u < 4; ++u) y[4 * q + u] = cmul(x[brev<R>(4 * q + u)], w4[u]);
                    }
#endif
                } else {
#pragma unroll
                    for (int k0 = 0; k0 < R; ++k0) y[k0] = cmul(x[brev<R>(k0)], W[k0 * (R + 1) + idx]);
                }
                HB_PC_MARK(1);
                // ... then wait until every thread has consumed the previous level's spectra, and publish the row
                if (lev > 0) { mbar_wait(mbar, parity); parity ^= 1u; }
#pragma unroll
                for (int k0 = 0; k0 < R; ++k0) xr[k0 * (R + 1) + idx] = y[k0];
            }
            __syncwarp();
            HB_PC_MARK(2);
            // C: radix-16 DFT over j0 of column k0 = idx, in place
            {
                double2* xc = xr + idx * (R + 1);
                double2 x[R];
#pragma unroll
                for (int j0 = 0; j0 < R; ++j0) x[j0] = xc[j0];
                HB_DFT16<R, 1>(x);
#pragma unroll
                for (int k1 = 0; k1 < R; ++k1) xc[k1] = x[brev<R>(k1)];
            }
            HB_PC_MARK(3);
            // D: pointwise MAC of this level's (K+1) spectra per gate with rows lev*(K+1)+r of BSK_i, one group of
            // (whole position, left-over item) at a time.  The key loads do not depend on the spectra: the first group's
            // are issued before the barrier.
            const double2* bp = A.bsk + ((size_t)i * ROWS + lev * (K + 1)) * (K + 1) * M;
            double2 b1[K + 1][K + 1];
            double2 b2[K + 1];
            auto load_group = [&](int q) {
                if (q < PP) {
                    const int pos = NT >= M ? tid : tid + q * NT;
#pragma unroll
                    for (int r = 0; r <= K; ++r)
#pragma unroll
                        for (int j = 0; j <= K; ++j) b1[r][j] = ldg_nc(bp + (size_t)(r * (K + 1) + j) * M + pos);
                }
                if (q < I2) {
                    const int item = tid + q * NT, pos = PP * NT + HB_ITEM_POS(item), j2 = HB_ITEM_OUT(item);
#pragma unroll
                    for (int r = 0; r <= K; ++r) b2[r] = ldg_nc(bp + (size_t)(r * (K + 1) + j2) * M + pos);
                }
            };
            load_group(0);
            __syncthreads();
            HB_PC_MARK(4);
#pragma unroll
            for (int q = 0; q < NGRP; ++q) {
                if (q > 0) load_group(q);
                if (q < PP) {
                    const int s1 = spos(NT >= M ? tid : tid + q * NT);
#pragma unroll
                    for (int r = 0; r <= K; ++r) {
#pragma unroll
                        for (int g = 0; g < G; ++g) {
                            const double2 f = X[((size_t)g * (K + 1) + r) * XROW + s1];
#pragma unroll
                            for (int j = 0; j <= K; ++j) cfma(o1[q][g][j], f, b1[r][j]);
                        }
                    }
                }
                if (q < I2) {
                    const int item = tid + q * NT, s2 = spos(PP * NT + HB_ITEM_POS(item));
#pragma unroll
                    for (int r = 0; r <= K; ++r) {
#pragma unroll
                        for (int g = 0; g < G; ++g) cfma(o2[q][g], X[((size_t)g * (K + 1) + r) * XROW + s2], b2[r]);
                    }
                }
            }
            HB_PC_MARK(5);
            mbar_arrive(mbar);                       // this thread no longer reads this level's spectra
#pragma unroll
            for (int l = 0; l + 1 < L; ++l)
#pragma unroll
                for (int part = 0; part < R / 2; ++part) pk[l][part] = pk[l + 1][part];
        }
        // output spectra -> the buffers of the polynomials they belong to (once everyone is done reading them)
        HB_PC_MARK(11);
        mbar_wait(mbar, parity); parity ^= 1u;
        HB_PC_MARK(6);
#pragma unroll
        for (int q = 0; q < PP; ++q) {
            const int s1 = spos(NT >= M ? tid : tid + q * NT);
#pragma unroll
            for (int g = 0; g < G; ++g)
#pragma unroll
                for (int j = 0; j <= K; ++j) X[((size_t)g * (K + 1) + j) * XROW + s1] = o1[q][g][j];
        }
#pragma unroll
        for (int q = 0; q < I2; ++q) {
            const int item = tid + q * NT, s2 = spos(PP * NT + HB_ITEM_POS(item)), j2 = HB_ITEM_OUT(item);
#pragma unroll
            for (int g = 0; g < G; ++g) X[((size_t)g * (K + 1) + j2) * XROW + s2] = o2[q][g];
        }
        __syncthreads();
        HB_PC_MARK(7);
        // E: inverse radix-16 DFT over k1 of column k0 = idx, conj inter-stage twiddle
        {
            double2* xc = xr + idx * (R + 1);
            double2 x[R];
#pragma unroll
            for (int k1 = 0; k1 < R; ++k1) x[k1] = xc[k1];
            HB_DFT16<R, -1>(x);
            if constexpr (TMEMW) {
#ifdef HB_PBS16_TMEM_PIPE
                uint32_t rw[2][16];
                tmem_ld16_issue(tw_e, rw[0]);
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    tmem_ld16_wait(rw[q & 1]);
                    if (q < 3) tmem_ld16_issue(tw_e + 16 * (q + 1), rw[(q + 1) & 1]);
#pragma unroll
                    for (int u = 0; u < 4; ++u) xc[4 * q + u] = cmul_conj(x[brev<R>(4 * q + u)], tmem_pair(rw[q & 1], u));
                }
#else
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    double2 w4[4];
                    tmem_ld16(tw_e + 16 * q, w4);
#pragma unroll
                    for (int u = 0; u < 4; ++u) xc[4 * q + u] = cmul_conj(x[brev<R>(4 * q + u)], w4[u]);
                }
#endif
            } else {
#pragma unroll
                for (int j0 = 0; j0 < R; ++j0) xc[j0] = cmul_conj(x[brev<R>(j0)], W[idx * (R + 1) + j0]);
            }
        }
        __syncwarp();
        HB_PC_MARK(8);
        // F: inverse radix-16 DFT over k0, untwist, round, wrapping add into the accumulator
        {
            double2 x[R];
#pragma unroll
            for (int k0 = 0; k0 < R; ++k0) x[k0] = xr[k0 * (R + 1) + idx];
            HB_DFT16<R, -1>(x);
#pragma unroll
            for (int j1 = 0; j1 < R; ++j1) {
                const double2 z = cmul_tw<-1>(x[brev<R>(j1)], j1 * TWA);
                const int c = idx + R * j1;
                const T v0 = ap[c] + TorusOps<T>::round_add(z.x);
                const T v1 = ap[c + M] + TorusOps<T>::round_add(z.y);
                ap[c] = v0; ap[c + M] = v1;
                if constexpr (EXT) { ap[c + N] = (T)(0 - v0); ap[c + M + N] = (T)(0 - v1); }
            }
        }
        __syncwarp();
        HB_PC_MARK(9);
    }
    HB_PC_FLUSH();
    __syncthreads();

    // ---- epilogue: sample extract coefficient 0 (App. A.5)
    T* out = reinterpret_cast<T*>(A.out);
    for (int g = 0; g < G; ++g) {
        const int jid = job0 + g;
        if (jid >= (int)A.n_jobs) continue;
        const uint32_t dst = A.jobs[jid].dst;
        const T* ag = acc + (size_t)g * (K + 1) * AS;
        T* o = out + (size_t)dst * A.out_stride;
        for (int x = tid; x < K * N; x += NT) {
            const int j = x / N, c = x % N;
            o[x] = c == 0 ? ag[j * AS] : (T)(0 - ag[j * AS + N - c]);
        }
        if (tid == 0) o[K * N] = ag[K * AS];
    }
    if constexpr (TMEMW) {
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncthreads();
        if (tid < 32) {
            uint32_t tbase;
            asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tbase) : "r"(mbar + 8) : "memory");
            tmem_dealloc(tbase, TM_COLS);
        }
    }
}

}  // namespace hb
