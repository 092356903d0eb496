// Narrow-batch programmable bootstrap (one gate per CTA), third generation: three-stage transforms, 8 points per thread.
//
// Why: a netlist level narrower than the SM count (every level of c880 / c7552, every batch of arithmetic mode) is latency
// bound -- one bootstrap is n sequential CMUX steps and what matters is the time of ONE step of ONE CTA.  In pbs_kernel.cuh a
// transform of M = N/2 points is two stages of radix 16 (N = 512) or 32 (N = 2048) on 16 / 32 threads: a serial chain of ~420 /
// ~930 dependent-ish FP64 instructions per thread and stage pair, with one warp per scheduler (profiles/r01_v6_narrow_phase_clocks.txt:
// 2 790 + 1 895 of 6 860 cycles per boolean step are the forward / inverse transforms).  Here M = R1 x R2 x R3 (8 x 8 x 4 for
// N = 512, 8 x 8 x 16 for N = 2048) and a transform runs on M / 8 threads (one warp for N = 512, four for N = 2048): the chain per
// thread is 3 short DFTs (52 + 52 + 16 instructions for 8 x 8 x 4) and 2.25 warps per scheduler cover each other's latencies.
//
// Index algebra (j = input index of the folded polynomial z_j = (p_j + i p_{j+M}) omega^j, omega = e^{i pi / N}; f = frequency):
//   j = j3 + R3 (j2 + R2 j1),  f = k1 + R1 (k2 + R2 k3)
//   stage 1: DFT over j1 (twist omega^{R2 R3 j1} is a literal per register), then x T2 = e^{2 pi i R3 j2 (4 k1 + 1) / 4M}
//   stage 2: DFT over j2, then x T3 = e^{2 pi i j3 (4 (k1 + R1 k2) + 1) / 4M}
//   stage 3: DFT over j3.
// All three stages work in place on one padded array addressed by the digit triple, addr(d1, d2, d3) = d1 A1 + d2 A2 + d3; the
// twiddle tables use the same addressing, so a thread reads a twiddle with the bank pattern of the datum it multiplies.
// A2 = R3 + 1 and A1 = 4 (mod 8) when R3 < 8 make every stage's 128-bit accesses (almost) conflict free, including the MAC's,
// which walks the Fourier key in ITS order (position k0 * KB + k1 holds frequency k0 + KA k1: the key layout of bsk_fft_kernel is
// shared with the throughput kernels) and finds the spectrum value of that frequency at addr(f % R1, f / R1 % R2, f / (R1 R2)).
// Same arithmetic as pbs_kernel / pbs16_kernel: u32 results are bit-identical (the f64 transforms round to the exact integers).
#pragma once
#include "pbs_kernel.cuh"

namespace hb {

template <class T, int N, int K, int L, int R1, int R2, int R3, int KA, int NT, int G = 1>
struct Pbs3Cfg {
    static constexpr bool SMALLDIG = sizeof(T) == 4;
    using Dig = DigPack<SMALLDIG>;
    static constexpr int M = N / 2, ROWS = (K + 1) * L, RB = R2 * R3, PARTS = R1 / 2, KB2 = M / KA;
    static constexpr int TPT = M / 8;                                   // threads per transform
    static constexpr int A2 = R3 + 1;
    static constexpr int A1 = R3 >= 8 ? R2 * A2 : ((R2 * A2 + 3) / 8) * 8 + 4;      // = 4 (mod 8), >= R2 * A2
    static constexpr int XSZ = R1 * A1;                                 // double2 per transform buffer
    static constexpr int TWA = HB_TW_DEN / (4 * R1);                    // omega^{RB j1} = e^{2 pi i j1 TWA / 256}
    static_assert(R1 * R2 * R3 == M && R1 == 8 && R2 == 8, "stages 1 and 2 are radix 8: M / 8 threads per transform");
    static constexpr int NTG = NT / G;                                  // threads of one gate
    static_assert(A1 >= R2 * A2 && ROWS * TPT <= NTG && NT % (32 * G) == 0, "thread / buffer geometry");
    static_assert(TPT % 32 == 0, "whole warps per transform");
    static constexpr size_t KEY_BYTES = (size_t)ROWS * (K + 1) * M * sizeof(double2);
    __host__ __device__ static constexpr size_t smem_core(int n) {
        return 2 * (size_t)XSZ * sizeof(double2) +
               G * ((size_t)ROWS * XSZ * sizeof(double2) + (size_t)(K + 1) * N * sizeof(T) +
                    (size_t)ROWS * PARTS * RB * sizeof(typename Dig::type) + (((size_t)(n + 1) * 2 + 15) & ~(size_t)15));
    }
    __host__ __device__ static constexpr size_t smem_bytes(int n, bool stage) { return smem_core(n) + (stage ? KEY_BYTES + 16 : 0); }
    __host__ __device__ static constexpr int addr(int d1, int d2, int d3) { return d1 * A1 + d2 * A2 + d3; }
};

// STAGE: BSK_{i+1} staged in shared memory by one cp.async.bulk while step i finishes (as in pbs_kernel).  MB > 1: multi-bit blind
// rotation with grouping factor MB (see pbs_kernel.cuh): key read through L2, combined per step with the monomial spectra.
// G > 1 (wide batches of the N = 2048 shapes): G gates per CTA, NT / G threads each, in lock step; the MAC reads every key word once
// for all G gates (no staging: the key streams through L2 once per CTA and step).
template <class T, int N, int K, int L, int R1, int R2, int R3, int KA, int NT, bool STAGE, int MB = 0, int G = 1>
__global__ void __launch_bounds__(NT, 1) pbs3_kernel(const PbsArgs A) {
    using C = Pbs3Cfg<T, N, K, L, R1, R2, R3, KA, NT, G>;
    static_assert(G == 1 || (!STAGE && MB == 0), "several gates per CTA: key through L2, classic blind rotation");
    constexpr int NTG = C::NTG;
    using Dig = typename C::Dig;
    using DigT = typename Dig::type;
    constexpr int M = C::M, ROWS = C::ROWS, RB = C::RB, PARTS = C::PARTS, TPT = C::TPT, XSZ = C::XSZ, A1 = C::A1, A2 = C::A2, KB2 = C::KB2;
    static_assert(MB == 0 || (MB >= 2 && MB <= 4 && !STAGE), "multi-bit instances read the key through L2");
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int tid = threadIdx.x, n = A.n;
    const int gg = tid / NTG, gt = tid % NTG;                           // gate of this thread, thread within the gate
    const size_t atil_pad = (((size_t)(n + 1) * 2 + 15) & ~(size_t)15) / 2;          // uint16 per gate
    double2* TW2 = reinterpret_cast<double2*>(smem_raw);                // [XSZ] at addr(k1, j2, j3)
    double2* TW3 = TW2 + XSZ;                                           // [XSZ] at addr(k1, k2, j3)
    double2* Xall = TW3 + XSZ;                                          // [G][ROWS][XSZ]
    T* accall = reinterpret_cast<T*>(Xall + (size_t)G * ROWS * XSZ);    // [G][K+1][N]
    DigT* digsall = reinterpret_cast<DigT*>(accall + (size_t)G * (K + 1) * N);       // [G][ROWS][PARTS][RB]
    uint16_t* atilall = reinterpret_cast<uint16_t*>(digsall + (size_t)G * ROWS * PARTS * RB);   // [G][atil_pad]
    double2* X = Xall + (size_t)gg * ROWS * XSZ;
    T* acc = accall + (size_t)gg * (K + 1) * N;
    DigT* digs = digsall + (size_t)gg * ROWS * PARTS * RB;
    uint16_t* atil = atilall + (size_t)gg * atil_pad;
    double2* KBUF = reinterpret_cast<double2*>(smem_raw + C::smem_core(n));
    const uint32_t kb_addr = (uint32_t)__cvta_generic_to_shared(KBUF);
    const uint32_t kbar = kb_addr + (uint32_t)C::KEY_BYTES;
    uint32_t kphase = 0;
    if (STAGE && tid == 0) {
        mbar_init(kbar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        mbar_expect_tx(kbar, (uint32_t)C::KEY_BYTES);
        bulk_g2s(kb_addr, A.bsk, (uint32_t)C::KEY_BYTES, kbar);
    }
    const int jid = blockIdx.x * G + gg;
    const bool live = jid < (int)A.n_jobs;                              // the last CTA of a G > 1 launch may hold padding gates
    const T* in = reinterpret_cast<const T*>(A.in);
    PbsJob jb = {};
    if (live) jb = A.jobs[jid];

    // inter-stage twiddles from the table of 4M-th roots of unity (roots[t] = e^{2 pi i t / 4M}, 4M = 2N)
    for (int e = tid; e < M; e += NT) {
        const int d3 = e % R3, d2 = (e / R3) % R2, d1 = e / (R2 * R3);
        TW2[C::addr(d1, d2, d3)] = A.roots[(R3 * d2 * (4 * d1 + 1)) & (2 * N - 1)];
        TW3[C::addr(d1, d2, d3)] = A.roots[(d3 * (4 * (d1 + R1 * d2) + 1)) & (2 * N - 1)];
    }
    // ---- prologue: linear combination + modulus switch (App. A.5)
    {
        const T* s0 = in + (size_t)jb.src0 * A.in_stride;
        const T* s1 = in + (size_t)jb.src1 * A.in_stride;
        for (int x = gt; x <= n; x += NTG) {
            T v = 0;
            if (jb.c0) v += (T)(int64_t)jb.c0 * s0[x];
            if (jb.c1) v += (T)(int64_t)jb.c1 * s1[x];
            if (x == n) v += (T)jb.cst;
            const T sh = v >> (TorusOps<T>::BITS - ilog2c(N) - 2);
            atil[x] = (uint16_t)(((sh + 1) >> 1) & (2 * N - 1));
        }
    }
    __syncthreads();
    // ---- accumulator = TV * X^{-b~}
    {
        const int a0 = (2 * N - atil[n]) & (2 * N - 1);
        for (int x = gt; x < K * N; x += NTG) acc[x] = 0;
        for (int c = gt; c < N; c += NTG) {
            int t = c - a0; if (t < 0) t += 2 * N;
            acc[K * N + c] = !live ? (T)0 : t < N ? tv_body<T>(jb, N, t) : (T)(0 - tv_body<T>(jb, N, t - N));
        }
    }
    __syncthreads();

    // roles in the transforms: transform `tr` (forward: row of the decomposition; inverse: output polynomial), lane `lt` in it
    const int tr = gt / TPT, lt = gt % TPT;
    auto tsync = [&]() { if (TPT <= 32) __syncwarp(); else __syncthreads(); };

    const int n_steps = MB > 1 ? n / MB : n;
    HB_PC_DECL;
    for (int i = 0; i < n_steps; ++i) {
        HB_PC_MARK(9);
        const int a = MB > 1 ? 0 : atil[i];
        int any = a;
#pragma unroll
        for (int g2 = 0; g2 < G; ++g2) any |= atilall[(size_t)g2 * atil_pad + i];
        if (MB == 0 && any == 0) {                    // uniform across the CTA: acc X^0 - acc = 0 for every gate
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
        // ---- P1: rotate-subtract (multi-bit: the accumulator itself) and signed decomposition, once per coefficient for all L
        // levels; digits staged in the order stage 1 reads them: [row][part][j0], coefficient c = j0 + RB (2 part + u) (and c + M)
        for (int t = gt; t < (K + 1) * PARTS * RB; t += NTG) {
            const int j0 = t % RB, part = (t / RB) % PARTS, r = t / (RB * PARTS);
            const T* ap = acc + (size_t)r * N;
            T v[4];
#pragma unroll
            for (int u = 0; u < 2; ++u) {
                const int c = j0 + RB * (2 * part + u);
                if (MB > 1) { v[2 * u] = ap[c]; v[2 * u + 1] = ap[c + M]; }
                else { v[2 * u] = rot_read<T>(ap, N, c, a) - ap[c]; v[2 * u + 1] = rot_read<T>(ap, N, c + M, a) - ap[c + M]; }
            }
            bool done = false;
            if constexpr (sizeof(T) == 4 && 7 * L <= 24) {
                if (A.base_log == 6 || A.base_log == 7) {             // four values at once in byte lanes (digits_u32.cuh)
                    uint32_t xv[4], pl[L];
                    const uint32_t rnd = 1u << (31 - A.base_log * L);
#pragma unroll
                    for (int u = 0; u < 4; ++u) xv[u] = (uint32_t)v[u] + rnd;
                    if (A.base_log == 6) digit_lanes_u32<6, L, false>(xv, pl); else digit_lanes_u32<7, L, false>(xv, pl);
#pragma unroll
                    for (int lev = 0; lev < L; ++lev) digs[((size_t)(lev * (K + 1) + r) * PARTS + part) * RB + j0] = (DigT)pl[lev];
                    done = true;
                }
            }
            if (!done) {
                int32_t dg[4][L];
#pragma unroll
                for (int u = 0; u < 4; ++u) digits_all<T, L>(v[u], A.base_log, dg[u]);
#pragma unroll
                for (int lev = 0; lev < L; ++lev)
                    digs[((size_t)(lev * (K + 1) + r) * PARTS + part) * RB + j0] = Dig::make(dg[0][lev], dg[1][lev], dg[2][lev], dg[3][lev]);
            }
        }
        __syncthreads();
        HB_PC_MARK(0);
        // ---- forward transforms of the ROWS digit polynomials
        if (tr < ROWS) {                              // stage 1: thread j0 = lt, DFT over j1 with the twist folded in, x T2
            const int j3 = lt % R3, j2 = lt / R3;
            const DigT* dp = digs + (size_t)tr * PARTS * RB + lt;
            double2 x[R1], w[R1];
#pragma unroll
            for (int part = 0; part < PARTS; ++part) {
                const DigT pk = dp[part * RB];
                x[2 * part] = make_double2(Dig::template get<0>(pk), Dig::template get<1>(pk));
                x[2 * part + 1] = make_double2(Dig::template get<2>(pk), Dig::template get<3>(pk));
            }
#pragma unroll
            for (int k1 = 0; k1 < R1; ++k1) w[k1] = TW2[C::addr(k1, j2, j3)];
            dft_regs_dit_twisted<R1, 1, C::TWA>(x);
            double2* xr = X + (size_t)tr * XSZ;
#pragma unroll
            for (int k1 = 0; k1 < R1; ++k1) xr[C::addr(k1, j2, j3)] = cmul(x[brev<R1>(k1)], w[k1]);
        }
        tsync();
        HB_PC_MARK(1);
        if (tr < ROWS) {                              // stage 2: (k1, j3), DFT over j2, x T3, in place
            double2* xr = X + (size_t)tr * XSZ;
            for (int u = lt; u < M / R2; u += TPT) {
                const int j3 = u % R3, k1 = u / R3;
                double2 x[R2], w[R2];
#pragma unroll
                for (int j2 = 0; j2 < R2; ++j2) x[j2] = xr[C::addr(k1, j2, j3)];
#pragma unroll
                for (int k2 = 0; k2 < R2; ++k2) w[k2] = TW3[C::addr(k1, k2, j3)];
                HB_DFT<R2, 1>(x);
#pragma unroll
                for (int k2 = 0; k2 < R2; ++k2) xr[C::addr(k1, k2, j3)] = cmul(x[brev<R2>(k2)], w[k2]);
            }
        }
        tsync();
        HB_PC_MARK(2);
        if (tr < ROWS) {                              // stage 3: (k1, k2), DFT over j3, in place; addr(k1, k2, k3) holds f = k1 + R1 (k2 + R2 k3)
            double2* xr = X + (size_t)tr * XSZ;
            for (int v = lt; v < M / R3; v += TPT) {
                const int k2 = v % R2, k1 = v / R2;
                double2 x[R3];
#pragma unroll
                for (int j3 = 0; j3 < R3; ++j3) x[j3] = xr[C::addr(k1, k2, j3)];
                HB_DFT<R3, 1>(x);
#pragma unroll
                for (int k3 = 0; k3 < R3; ++k3) xr[C::addr(k1, k2, k3)] = x[brev<R3>(k3)];
            }
        }
        __syncthreads();
        HB_PC_MARK(3);
        // ---- pointwise MAC with BSK_i: key position pos = k0 * KB2 + k1 holds frequency f = k0 + KA k1
        if (STAGE) { mbar_wait(kbar, kphase); kphase ^= 1u; }
        HB_PC_MARK(4);
        if constexpr (MB > 1) {
            int ex[1 << MB];
#pragma unroll
            for (int q = 0; q < (1 << MB); ++q) {
                int e = 0;
#pragma unroll
                for (int j = 0; j < MB; ++j) if ((q >> j) & 1) e += atil[i * MB + j];
                ex[q] = e & (2 * N - 1);
            }
            for (int pos = tid; pos < M; pos += NT) {
                const int f = (pos / KB2) + KA * (pos % KB2);
                const int xa = C::addr(f % R1, (f / R1) % R2, f / (R1 * R2));
                double2 cmb[ROWS][K + 1];
#pragma unroll
                for (int row = 0; row < ROWS; ++row)
#pragma unroll
                    for (int j = 0; j <= K; ++j) cmb[row][j] = make_double2(0.0, 0.0);
                const double2* bp = A.bsk + (size_t)i * (1 << MB) * ROWS * (K + 1) * M + pos;
#pragma unroll 4
                for (int q = 0; q < (1 << MB); ++q) {
                    const double2 m = __ldg(A.roots + ((ex[q] * (4 * f + 1)) & (2 * N - 1)));
#pragma unroll
                    for (int row = 0; row < ROWS; ++row)
#pragma unroll
                        for (int j = 0; j <= K; ++j) cfma(cmb[row][j], __ldg(bp + ((size_t)q * ROWS * (K + 1) + row * (K + 1) + j) * M), m);
                }
                double2 o[K + 1];
#pragma unroll
                for (int j = 0; j <= K; ++j) o[j] = make_double2(0.0, 0.0);
#pragma unroll
                for (int row = 0; row < ROWS; ++row) {
                    const double2 fx = X[(size_t)row * XSZ + xa];
#pragma unroll
                    for (int j = 0; j <= K; ++j) cfma(o[j], fx, cmb[row][j]);
                }
#pragma unroll
                for (int j = 0; j <= K; ++j) X[(size_t)j * XSZ + xa] = o[j];
            }
        } else if constexpr (G == 1) {
            for (int pos = tid; pos < M; pos += NT) {
                const int f = (pos / KB2) + KA * (pos % KB2);
                const int xa = C::addr(f % R1, (f / R1) % R2, f / (R1 * R2));
                double2 o[K + 1];
#pragma unroll
                for (int j = 0; j <= K; ++j) o[j] = make_double2(0.0, 0.0);
                const double2* bp = A.bsk + (size_t)i * ROWS * (K + 1) * M + pos;
#pragma unroll
                for (int row = 0; row < ROWS; ++row) {
                    const double2 fx = X[(size_t)row * XSZ + xa];
#pragma unroll
                    for (int j = 0; j <= K; ++j)
                        cfma(o[j], fx, STAGE ? KBUF[(size_t)(row * (K + 1) + j) * M + pos] : __ldg(bp + (size_t)(row * (K + 1) + j) * M));
                }
#pragma unroll
                for (int j = 0; j <= K; ++j) X[(size_t)j * XSZ + xa] = o[j];
            }
        } else {
            // every thread of the CTA takes positions; one read of BSK_i (registers) feeds all G gates
            for (int pos = tid; pos < M; pos += NT) {
                const int f = (pos / KB2) + KA * (pos % KB2);
                const int xa = C::addr(f % R1, (f / R1) % R2, f / (R1 * R2));
                const double2* bp = A.bsk + (size_t)i * ROWS * (K + 1) * M + pos;
                double2 b[ROWS][K + 1];
#pragma unroll
                for (int row = 0; row < ROWS; ++row)
#pragma unroll
                    for (int j = 0; j <= K; ++j) b[row][j] = __ldg(bp + (size_t)(row * (K + 1) + j) * M);
#pragma unroll
                for (int g2 = 0; g2 < G; ++g2) {
                    double2* xg = Xall + (size_t)g2 * ROWS * XSZ;
                    double2 o[K + 1];
#pragma unroll
                    for (int j = 0; j <= K; ++j) o[j] = make_double2(0.0, 0.0);
#pragma unroll
                    for (int row = 0; row < ROWS; ++row) {
                        const double2 fx = xg[(size_t)row * XSZ + xa];
#pragma unroll
                        for (int j = 0; j <= K; ++j) cfma(o[j], fx, b[row][j]);
                    }
#pragma unroll
                    for (int j = 0; j <= K; ++j) xg[(size_t)j * XSZ + xa] = o[j];
                }
            }
        }
        __syncthreads();
        HB_PC_MARK(5);
        if (STAGE && tid == 0 && i + 1 < n) {         // every read of the staged block is behind the barrier above
            mbar_expect_tx(kbar, (uint32_t)C::KEY_BYTES);
            bulk_g2s(kb_addr, A.bsk + (size_t)(i + 1) * ROWS * (K + 1) * M, (uint32_t)C::KEY_BYTES, kbar);
        }
        // ---- inverse transforms of the K+1 output spectra: the forward stages backwards with conjugated twiddles
        if (tr <= K) {                                // stage 3': (k1, k2), inverse DFT over k3, x conj T3
            double2* xr = X + (size_t)tr * XSZ;
            for (int v = lt; v < M / R3; v += TPT) {
                const int k2 = v % R2, k1 = v / R2;
                double2 x[R3], w[R3];
#pragma unroll
                for (int k3 = 0; k3 < R3; ++k3) x[k3] = xr[C::addr(k1, k2, k3)];
#pragma unroll
                for (int j3 = 0; j3 < R3; ++j3) w[j3] = TW3[C::addr(k1, k2, j3)];
                HB_DFT<R3, -1>(x);
#pragma unroll
                for (int j3 = 0; j3 < R3; ++j3) xr[C::addr(k1, k2, j3)] = cmul_conj(x[brev<R3>(j3)], w[j3]);
            }
        }
        tsync();
        HB_PC_MARK(6);
        if (tr <= K) {                                // stage 2': (k1, j3), inverse DFT over k2, x conj T2
            double2* xr = X + (size_t)tr * XSZ;
            for (int u = lt; u < M / R2; u += TPT) {
                const int j3 = u % R3, k1 = u / R3;
                double2 x[R2], w[R2];
#pragma unroll
                for (int k2 = 0; k2 < R2; ++k2) x[k2] = xr[C::addr(k1, k2, j3)];
#pragma unroll
                for (int j2 = 0; j2 < R2; ++j2) w[j2] = TW2[C::addr(k1, j2, j3)];
                HB_DFT<R2, -1>(x);
#pragma unroll
                for (int j2 = 0; j2 < R2; ++j2) xr[C::addr(k1, j2, j3)] = cmul_conj(x[brev<R2>(j2)], w[j2]);
            }
        }
        tsync();
        HB_PC_MARK(7);
        if (tr <= K) {                                // stage 1': thread j0, inverse DFT over k1, untwist, round, wrapping add
            const int j3 = lt % R3, j2 = lt / R3;
            const double2* xr = X + (size_t)tr * XSZ;
            T* ap = acc + (size_t)tr * N;
            double2 x[R1];
            T lo[R1], hi[R1];
#pragma unroll
            for (int k1 = 0; k1 < R1; ++k1) x[k1] = xr[C::addr(k1, j2, j3)];
#pragma unroll
            for (int j1 = 0; j1 < R1; ++j1) { lo[j1] = MB > 1 ? (T)0 : ap[lt + RB * j1]; hi[j1] = MB > 1 ? (T)0 : ap[lt + RB * j1 + M]; }
            HB_DFT<R1, -1>(x);
#pragma unroll
            for (int j1 = 0; j1 < R1; ++j1) {
                const double2 z = cmul_tw<-1>(x[brev<R1>(j1)], j1 * C::TWA);
                ap[lt + RB * j1] = lo[j1] + TorusOps<T>::round_add(z.x);          // multi-bit: acc <- product
                ap[lt + RB * j1 + M] = hi[j1] + TorusOps<T>::round_add(z.y);
            }
        }
        __syncthreads();
        HB_PC_MARK(8);
    }
    HB_PC_FLUSH();

    // ---- epilogue: sample extract coefficient 0 (App. A.5)
    if (live) {
        T* o = reinterpret_cast<T*>(A.out) + (size_t)jb.dst * A.out_stride;
        for (int x = gt; x < K * N; x += NTG) {
            const int j = x / N, c = x % N;
            o[x] = c == 0 ? acc[j * N] : (T)(0 - acc[j * N + N - c]);
        }
        if (gt == 0) o[K * N] = acc[K * N];
    }
}

}  // namespace hb
