// One bootstrap split over a thread-block cluster of K+1 CTAs (one per accumulator polynomial), for the narrowest levels.
//
// Why: on ONE SM a CMUX step of the boolean parameter set is bound by that SM's shared-memory bandwidth and FP64 pipe (~510 KB of
// shared-memory traffic and ~1 850 FP64-pipe cycles per step against 7 000 measured, profiles/r02_v6_narrow_phase_clocks.txt), so a
// level with fewer gates than SMs / (K+1) is faster when each gate uses K+1 SMs.  (SURVEY section 7 "hard parts"; VERDICT r1 #5.)
// CTA r of the cluster owns accumulator polynomial r for the whole blind rotation:
//   decomposes it (L levels) -> L forward transforms (pbs3_kernel's three-stage form) -> multiplies them with ITS rows of BSK_i
//   (rows (lev, r), staged by L bulk copies: a third of the key per CTA) into K+1 partial output spectra, keeps partial r and
//   SENDS partial j to CTA j (st.async into CTA j's shared memory, completion counted in bytes on CTA j's mbarrier: distributed
//   shared memory, no cluster-wide barrier in the loop) -> sums the K+1 partials of its own polynomial while loading them into the
//   inverse transform -> rounding add into its polynomial.
// The receive buffers are double buffered by executed-step parity: a producer can only be two steps ahead of a consumer after
// it has received that consumer's next partial, which the consumer sends after it finished reading the older buffer.
// u32 results stay bit-identical to the other kernels (every f64 sum rounds to the same exact integer).
#pragma once
#include "pbs3_kernel.cuh"

namespace hb {

__device__ __forceinline__ uint32_t cluster_ctarank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ uint32_t mapa_shared(uint32_t local_addr, uint32_t rank) {
    uint32_t a; asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(a) : "r"(local_addr), "r"(rank)); return a;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// 16 bytes into another CTA's shared memory; the destination CTA's mbarrier counts the bytes
__device__ __forceinline__ void st_async_f64x2(uint32_t remote_addr, double2 v, uint32_t remote_mbar) {
    asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v2.f64 [%0], {%1, %2}, [%3];"
                 ::"r"(remote_addr), "d"(v.x), "d"(v.y), "r"(remote_mbar) : "memory");
}

template <class T, int N, int K, int L, int R3, int KA, int NT>
struct Pbs3cCfg {
    using B = Pbs3Cfg<T, N, K, L, 8, 8, R3, KA, (K + 1) * L * (N / 16) < 1024 ? 1024 : 2048>;   // geometry only (addr, XSZ, TPT ...)
    using Dig = typename B::Dig;
    static constexpr int M = B::M, TPT = B::TPT, XSZ = B::XSZ, RB = B::RB, PARTS = B::PARTS;
    static_assert(L * TPT <= NT && NT % 32 == 0, "L forward transforms per CTA");
    static constexpr size_t KEY_BYTES = (size_t)L * (K + 1) * M * sizeof(double2);     // rows (lev, r) of one step
    __host__ __device__ static constexpr size_t smem_bytes(int n) {
        return (size_t)L * XSZ * sizeof(double2) + 2 * (size_t)XSZ * sizeof(double2) +            // X, TW2, TW3
               (size_t)XSZ * sizeof(double2) + 2 * (size_t)K * XSZ * sizeof(double2) +            // OUT, recv[2][K]
               (size_t)N * sizeof(T) + (size_t)L * PARTS * RB * sizeof(typename Dig::type) +      // acc, digs
               (((size_t)(n + 1) * 2 + 15) & ~(size_t)15) + KEY_BYTES + 32;                       // atil, key block, 3 mbarriers
    }
};

template <class T, int N, int K, int L, int R3, int KA, int NT>
__global__ void __cluster_dims__(K + 1, 1, 1) __launch_bounds__(NT, 1) pbs3c_kernel(const PbsArgs A) {
    using CC = Pbs3cCfg<T, N, K, L, R3, KA, NT>;
    using C = typename CC::B;
    using Dig = typename C::Dig;
    using DigT = typename Dig::type;
    constexpr int R1 = 8, R2 = 8;
    constexpr int M = C::M, ROWS = C::ROWS, RB = C::RB, PARTS = C::PARTS, TPT = C::TPT, XSZ = C::XSZ, KB2 = C::KB2;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    double2* X = reinterpret_cast<double2*>(smem_raw);                  // [L][XSZ] spectra of this polynomial's digit levels
    double2* TW2 = X + (size_t)L * XSZ;
    double2* TW3 = TW2 + XSZ;
    double2* OUT = TW3 + XSZ;                                           // [XSZ] own partial of the own output polynomial
    double2* RECV = OUT + XSZ;                                          // [2][K][XSZ] partials from the other CTAs
    T* acc = reinterpret_cast<T*>(RECV + 2 * (size_t)K * XSZ);          // [N] accumulator polynomial r
    DigT* digs = reinterpret_cast<DigT*>(acc + N);                      // [L][PARTS][RB]
    uint16_t* atil = reinterpret_cast<uint16_t*>(digs + (size_t)L * PARTS * RB);   // [n+1]
    const int tid = threadIdx.x, n = A.n;
    double2* KBUF = reinterpret_cast<double2*>(reinterpret_cast<unsigned char*>(atil) + (((size_t)(n + 1) * 2 + 15) & ~(size_t)15));   // [L][K+1][M]
    const uint32_t kb_addr = (uint32_t)__cvta_generic_to_shared(KBUF);
    const uint32_t kbar = kb_addr + (uint32_t)CC::KEY_BYTES;            // key block arrived
    const uint32_t rbar = kbar + 8;                                     // rbar + 8 b: partials of buffer b arrived
    const uint32_t recv_addr = (uint32_t)__cvta_generic_to_shared(RECV);
    const int r = (int)cluster_ctarank();                               // polynomial of this CTA
    const int jid = blockIdx.x / (K + 1);
    const PbsJob jb = A.jobs[jid];
    const T* in = reinterpret_cast<const T*>(A.in);
    auto stage_key = [&](int step) {                                    // rows (lev, r) of BSK_step: L contiguous blocks
        mbar_expect_tx(kbar, (uint32_t)CC::KEY_BYTES);
#pragma unroll
        for (int lev = 0; lev < L; ++lev)
            bulk_g2s(kb_addr + (uint32_t)(lev * (K + 1) * M * sizeof(double2)),
                     A.bsk + ((size_t)step * ROWS + lev * (K + 1) + r) * (K + 1) * M, (uint32_t)((K + 1) * M * sizeof(double2)), kbar);
    };
    if (tid == 0) {
        mbar_init(kbar, 1); mbar_init(rbar, 1); mbar_init(rbar + 8, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        stage_key(0);
    }
    for (int e = tid; e < M; e += NT) {
        const int d3 = e % R3, d2 = (e / R3) % R2, d1 = e / (R2 * R3);
        TW2[C::addr(d1, d2, d3)] = A.roots[(R3 * d2 * (4 * d1 + 1)) & (2 * N - 1)];
        TW3[C::addr(d1, d2, d3)] = A.roots[(d3 * (4 * (d1 + R1 * d2) + 1)) & (2 * N - 1)];
    }
    {   // linear combination + modulus switch: every CTA of the cluster computes the whole switched mask
        const T* s0 = in + (size_t)jb.src0 * A.in_stride;
        const T* s1 = in + (size_t)jb.src1 * A.in_stride;
        for (int x = tid; x <= n; x += NT) {
            T v = 0;
            if (jb.c0) v += (T)(int64_t)jb.c0 * s0[x];
            if (jb.c1) v += (T)(int64_t)jb.c1 * s1[x];
            if (x == n) v += (T)jb.cst;
            const T sh = v >> (TorusOps<T>::BITS - ilog2c(N) - 2);
            atil[x] = (uint16_t)(((sh + 1) >> 1) & (2 * N - 1));
        }
    }
    __syncthreads();
    {   // accumulator polynomial r of TV * X^{-b~}: zero mask polynomials, rotated body
        const int a0 = (2 * N - atil[n]) & (2 * N - 1);
        for (int c = tid; c < N; c += NT) {
            int t = c - a0; if (t < 0) t += 2 * N;
            acc[c] = r < K ? (T)0 : (t < N ? tv_body<T>(jb, N, t) : (T)(0 - tv_body<T>(jb, N, t - N)));
        }
    }
    cluster_sync_all();                               // every CTA's mbarriers exist before anyone sends; also a CTA barrier
    const int tr = tid / TPT, lt = tid % TPT;
    auto tsync = [&]() { if (TPT <= 32) __syncwarp(); else __syncthreads(); };
    uint32_t kphase = 0, ex_steps = 0;                // executed (non-skipped) steps so far

    for (int i = 0; i < n; ++i) {
        const int a = atil[i];
        if (a == 0) {                                 // uniform across the cluster: nothing to add; keep the key pipeline moving
            mbar_wait(kbar, kphase); kphase ^= 1u;
            __syncthreads();
            if (tid == 0 && i + 1 < n) stage_key(i + 1);
            continue;
        }
        const uint32_t buf = ex_steps & 1u, rparity = (ex_steps >> 1) & 1u;
        ++ex_steps;
        if (tid == 0) mbar_expect_tx(rbar + 8 * buf, (uint32_t)(K * M * sizeof(double2)));   // this step's K incoming partials
        // ---- P1: rotate-subtract and signed decomposition of polynomial r (all L levels)
        for (int t = tid; t < PARTS * RB; t += NT) {
            const int j0 = t % RB, part = t / RB;
            T v[4];
#pragma unroll
            for (int u = 0; u < 2; ++u) {
                const int c = j0 + RB * (2 * part + u);
                v[2 * u] = rot_read<T>(acc, N, c, a) - acc[c];
                v[2 * u + 1] = rot_read<T>(acc, N, c + M, a) - acc[c + M];
            }
            bool done = false;
            if constexpr (sizeof(T) == 4 && 7 * L <= 24) {
                if (A.base_log == 6 || A.base_log == 7) {
                    uint32_t xv[4], pl[L];
                    const uint32_t rnd = 1u << (31 - A.base_log * L);
#pragma unroll
                    for (int u = 0; u < 4; ++u) xv[u] = (uint32_t)v[u] + rnd;
                    if (A.base_log == 6) digit_lanes_u32<6, L, false>(xv, pl); else digit_lanes_u32<7, L, false>(xv, pl);
#pragma unroll
                    for (int lev = 0; lev < L; ++lev) digs[((size_t)lev * PARTS + part) * RB + j0] = (DigT)pl[lev];
                    done = true;
                }
            }
            if (!done) {
                int32_t dg[4][L];
#pragma unroll
                for (int u = 0; u < 4; ++u) digits_all<T, L>(v[u], A.base_log, dg[u]);
#pragma unroll
                for (int lev = 0; lev < L; ++lev)
                    digs[((size_t)lev * PARTS + part) * RB + j0] = Dig::make(dg[0][lev], dg[1][lev], dg[2][lev], dg[3][lev]);
            }
        }
        __syncthreads();
        // ---- forward transforms of the L digit polynomials (as pbs3_kernel)
        if (tr < L) {
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
        if (tr < L) {
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
        if (tr < L) {
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
        // ---- MAC with rows (lev, r) of BSK_i: K+1 partial spectra; partial r stays, partial j goes to CTA j
        mbar_wait(kbar, kphase); kphase ^= 1u;
        for (int pos = tid; pos < M; pos += NT) {
            const int f = (pos / KB2) + KA * (pos % KB2);
            const int xa = C::addr(f % R1, (f / R1) % R2, f / (R1 * R2));
            double2 o[K + 1];
#pragma unroll
            for (int j = 0; j <= K; ++j) o[j] = make_double2(0.0, 0.0);
#pragma unroll
            for (int lev = 0; lev < L; ++lev) {
                const double2 fx = X[(size_t)lev * XSZ + xa];
#pragma unroll
                for (int j = 0; j <= K; ++j) cfma(o[j], fx, KBUF[(size_t)(lev * (K + 1) + j) * M + pos]);
            }
#pragma unroll
            for (int j = 0; j <= K; ++j) {
                if (j == r) OUT[xa] = o[j];
                else {
                    const int slot = r < j ? r : r - 1;
                    const uint32_t dst = recv_addr + (uint32_t)((((size_t)buf * K + slot) * XSZ + xa) * sizeof(double2));
                    st_async_f64x2(mapa_shared(dst, (uint32_t)j), o[j], mapa_shared(rbar + 8 * buf, (uint32_t)j));
                }
            }
        }
        __syncthreads();                              // OUT complete; every read of the staged key block done
        if (tid == 0 && i + 1 < n) stage_key(i + 1);
        // ---- inverse transform of the own polynomial: sum of the K+1 partials on load
        if (tr == 0) {
            mbar_wait(rbar + 8 * buf, rparity);       // the K partials of the other CTAs have landed
            const double2* rv = RECV + (size_t)buf * K * XSZ;
            for (int v = lt; v < M / R3; v += TPT) {
                const int k2 = v % R2, k1 = v / R2;
                double2 x[R3], w[R3];
#pragma unroll
                for (int k3 = 0; k3 < R3; ++k3) {
                    const int xa = C::addr(k1, k2, k3);
                    double2 s = OUT[xa];
#pragma unroll
                    for (int q = 0; q < K; ++q) { const double2 p = rv[(size_t)q * XSZ + xa]; s.x += p.x; s.y += p.y; }
                    x[k3] = s;
                }
#pragma unroll
                for (int j3 = 0; j3 < R3; ++j3) w[j3] = TW3[C::addr(k1, k2, j3)];
                HB_DFT<R3, -1>(x);
#pragma unroll
                for (int j3 = 0; j3 < R3; ++j3) OUT[C::addr(k1, k2, j3)] = cmul_conj(x[brev<R3>(j3)], w[j3]);
            }
        }
        tsync();
        if (tr == 0) {
            for (int u = lt; u < M / R2; u += TPT) {
                const int j3 = u % R3, k1 = u / R3;
                double2 x[R2], w[R2];
#pragma unroll
                for (int k2 = 0; k2 < R2; ++k2) x[k2] = OUT[C::addr(k1, k2, j3)];
#pragma unroll
                for (int j2 = 0; j2 < R2; ++j2) w[j2] = TW2[C::addr(k1, j2, j3)];
                HB_DFT<R2, -1>(x);
#pragma unroll
                for (int j2 = 0; j2 < R2; ++j2) OUT[C::addr(k1, j2, j3)] = cmul_conj(x[brev<R2>(j2)], w[j2]);
            }
        }
        tsync();
        if (tr == 0) {
            const int j3 = lt % R3, j2 = lt / R3;
            double2 x[R1];
            T lo[R1], hi[R1];
#pragma unroll
            for (int k1 = 0; k1 < R1; ++k1) x[k1] = OUT[C::addr(k1, j2, j3)];
#pragma unroll
            for (int j1 = 0; j1 < R1; ++j1) { lo[j1] = acc[lt + RB * j1]; hi[j1] = acc[lt + RB * j1 + M]; }
            HB_DFT<R1, -1>(x);
#pragma unroll
            for (int j1 = 0; j1 < R1; ++j1) {
                const double2 z = cmul_tw<-1>(x[brev<R1>(j1)], j1 * C::TWA);
                acc[lt + RB * j1] = lo[j1] + TorusOps<T>::round_add(z.x);
                acc[lt + RB * j1 + M] = hi[j1] + TorusOps<T>::round_add(z.y);
            }
        }
        __syncthreads();
    }
    // ---- epilogue: sample extract (App. A.5): CTA r < K writes mask polynomial r, CTA K the body
    T* o = reinterpret_cast<T*>(A.out) + (size_t)jb.dst * A.out_stride;
    if (r < K) { for (int c = tid; c < N; c += NT) o[r * N + c] = c == 0 ? acc[0] : (T)(0 - acc[N - c]); }
    else if (tid == 0) o[K * N] = acc[0];
    cluster_sync_all();                               // no CTA leaves while a peer could still write into its shared memory
}

}  // namespace hb
