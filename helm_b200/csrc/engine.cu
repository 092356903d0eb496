// helm_b200 engine: context, resident keys / wires, per-level batch scheduler, C ABI (include/helm_b200.h).
// No CPU fallback anywhere in this file: without a CUDA device every compute entry point fails loudly.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <algorithm>
#include <atomic>
#include <string>
#include <type_traits>
#include <vector>
#include "../../include/helm_b200.h"
#include "pbs_kernel.cuh"
#include "pbs16_kernel.cuh"
#include "pbs3_kernel.cuh"
#include "pbs3c_kernel.cuh"
#ifndef HB_TMEMW
#define HB_TMEMW true
#endif
#ifndef HB_PBS16_G
#define HB_PBS16_G 4
#endif
#include "ks_kernel.cuh"
#include "ks_mma_kernel.cuh"

using namespace hb;

namespace {
thread_local std::string g_create_error;

// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) is a per-DEVICE property of a kernel: one flag per (kernel, device), so a
// second context on another GPU of the same process opts in too (the current device is the launching context's).
struct SmemOptIn {
    std::atomic<uint64_t> done{0};
    template <class K> cudaError_t ensure(K kern, size_t bytes) {
        int dev = 0;
        cudaError_t e = cudaGetDevice(&dev);
        if (e != cudaSuccess) return e;
        if ((done.load(std::memory_order_acquire) >> dev) & 1ull) return cudaSuccess;
        e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
        if (e == cudaSuccess) done.fetch_or(1ull << dev, std::memory_order_release);
        return e;
    }
};

struct DevBuf {
    void* p = nullptr; size_t cap = 0;
    cudaError_t ensure(size_t bytes) {
        if (bytes <= cap) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr; cap = 0;
        size_t want = bytes + bytes / 4 + 256;
        cudaError_t e = cudaMalloc(&p, want);
        if (e == cudaSuccess) cap = want;
        return e;
    }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
};
struct PinBuf {
    void* p = nullptr; size_t cap = 0;
    cudaError_t ensure(size_t bytes) {
        if (bytes <= cap) return cudaSuccess;
        if (p) cudaFreeHost(p);
        p = nullptr; cap = 0;
        size_t want = bytes + bytes / 4 + 256;
        cudaError_t e = cudaMallocHost(&p, want);
        if (e == cudaSuccess) cap = want;
        return e;
    }
    void release() { if (p) cudaFreeHost(p); p = nullptr; cap = 0; }
};
}  // namespace

constexpr int NSLOT = 4;
struct JobSlot {
    PinBuf pin; DevBuf dev; cudaEvent_t done = nullptr;
    cudaEvent_t t[3] = {nullptr, nullptr, nullptr};   // brackets of the two phases of one level
    bool pending = false; int first_is_pbs = 1;
};
struct CtxExtra { JobSlot slots[NSLOT]; uint64_t next = 0; };

struct hb_ctx {
    int device = 0;
    CtxExtra extra;                        // job-descriptor ring (owned by the context: contexts share no mutable state)
    cudaStream_t stream = nullptr;
    cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
    hb_params p{};
    bool have_params = false, have_bsk = false, have_ksk = false;
    int shape = -1;                        // index into the kernel shape table
    DevBuf bsk_f, ksk, wtab;               // resident keys
    DevBuf ksk_planes, ks_dig; int ksk_npad = 0;   // byte planes of the KSK for the tensor-core keyswitch
    DevBuf scratch_a, scratch_b;           // PBS->KS (order 0) or KS->PBS (order 1) intermediates
    DevBuf atil;                           // switched masks of the running PBS batch (pbs16_kernel)
    DevBuf lin_stage;                      // staging rows of a linear batch with read-after-write between its jobs
    DevBuf jobs_dev, stage_dev, idx_dev;
    PinBuf jobs_pin, stage_pin;
    hb_stats_t stats{};
    std::string err;
    int sm_count = 0;
    // asynchronous row IO (hb_wires_*_rows_async): copy stream, its staging buffers, the two ordering events
    cudaStream_t io_stream = nullptr, io_down = nullptr;   // uploads / downloads: separate streams, so that the upload of batch s+1 does not queue behind the download of batch s
    cudaEvent_t io_written = nullptr, io_computed = nullptr;
    DevBuf io_in, io_out;
    // CUDA-graph capture of a sequence of batches (hb_graph_begin / hb_graph_end): job descriptors go to an arena owned by the graph
    bool capturing = false;
    PinBuf cap_pin; DevBuf cap_dev; size_t cap_used = 0;
    JobSlot cap_slot;                      // stand-in for the ring slot while capturing (its timing events are recorded but never read)
    uint64_t scratch_gen = 0;              // bumped whenever a scratch buffer is reallocated: a graph captured before that is stale
    hb_stats_t cap_stats0{};
    int cluster_cap = -1;                  // resident clusters of the cluster-split kernel on this device (-1: not asked yet)
    // multi-device group (hb_ctx_create_multi): the group context owns one ordinary context per device and fans every call
    // out; `level_ev` of a member = "my share of the current level is evaluated and pushed to every peer"
    std::vector<hb_ctx*> subs;
    cudaEvent_t level_ev = nullptr;
    uint32_t shard_min = 0;                // gates per device below which a level is not split further
    bool is_group() const { return !subs.empty(); }
};

constexpr int HB_MAX_DEV = 8;
struct PushArgs { void* dst[HB_MAX_DEV]; int n_dst; };

struct hb_wires {
    hb_ctx* ctx = nullptr;
    void* dev = nullptr;
    uint32_t n_wires = 0;
    bool owned = false;
    std::vector<hb_wires*> subs;           // group wires: one replica per device, kept coherent level by level
    PushArgs peers{};                      // member of a group: the other replicas (peer-mapped over NVLink)
};

namespace {

hb_status fail(hb_ctx* c, hb_status s, const std::string& m) { if (c) c->err = m; else g_create_error = m; return s; }
#define CU(c, call) do { cudaError_t e__ = (call); if (e__ != cudaSuccess) return fail((c), HB_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e__)); } while (0)

inline size_t tsize(const hb_ctx* c) { return c->p.torus_bits / 8; }
inline uint32_t small_words(const hb_ctx* c) { return c->p.n + 1; }
inline uint32_t big_words(const hb_ctx* c) { return c->p.k * c->p.N + 1; }
inline uint32_t pad4(uint32_t w) { return (w + 3u) & ~3u; }
inline uint32_t ct_words(const hb_ctx* c) { return c->p.order == 0 ? small_words(c) : big_words(c); }

// ---- elementwise helpers
template <class T> __global__ void scatter_rows(T* dst, uint32_t stride, const T* src, const uint32_t* idx, uint32_t cnt, uint32_t words) {
    for (uint32_t r = blockIdx.y; r < cnt; r += gridDim.y) {
        T* d = dst + (size_t)idx[r] * stride; const T* s = src + (size_t)r * words;
        for (uint32_t x = blockIdx.x * blockDim.x + threadIdx.x; x < words; x += gridDim.x * blockDim.x) d[x] = s[x];
    }
}
template <class T> __global__ void gather_rows(const T* src, uint32_t stride, T* dst, const uint32_t* idx, uint32_t cnt, uint32_t words) {
    for (uint32_t r = blockIdx.y; r < cnt; r += gridDim.y) {
        const T* s = src + (size_t)idx[r] * stride; T* d = dst + (size_t)r * words;
        for (uint32_t x = blockIdx.x * blockDim.x + threadIdx.x; x < words; x += gridDim.x * blockDim.x) d[x] = s[x];
    }
}
// packed rows (words each) <-> strided rows, contiguous range: the device side of a large host copy, so that the PCIe transfer
// itself is one 1-D copy (a 2-D copy of 2.9 KB rows reaches about half the link rate)
template <class T> __global__ void repack_rows(T* dst, uint32_t dstride, const T* src, uint32_t sstride, uint32_t cnt, uint32_t words) {
    for (uint32_t r = blockIdx.y; r < cnt; r += gridDim.y) {
        T* d = dst + (size_t)r * dstride; const T* q = src + (size_t)r * sstride;
        for (uint32_t x = blockIdx.x * blockDim.x + threadIdx.x; x < words; x += gridDim.x * blockDim.x) d[x] = q[x];
    }
}
struct LinJob { uint32_t src, dst; int32_t c; uint32_t pad; uint64_t cst; };   // dst = c*src + cst(body); c==0: trivial
// MODE 0: in place on the wire array.  MODE 1 / 2: the two halves of a staged batch (results to scratch row r, then scratch row r
// to its wire) -- used when a job reads a row another job of the batch writes (chained DFFs q2 <= q1 <= d in the DFF level), so that
// every job sees the pre-batch values like the reference's clone-then-write loop (src/circuit.rs:532-540).
template <class T, int MODE> __global__ void linear_rows(T* w, uint32_t stride, T* scratch, const LinJob* jobs, uint32_t cnt, uint32_t words) {
    for (uint32_t r = blockIdx.y; r < cnt; r += gridDim.y) {
        const LinJob j = jobs[r];
        const T* s = MODE == 2 ? scratch + (size_t)r * stride : w + (size_t)j.src * stride;
        T* d = MODE == 1 ? scratch + (size_t)r * stride : w + (size_t)j.dst * stride;
        for (uint32_t x = blockIdx.x * blockDim.x + threadIdx.x; x < words; x += gridDim.x * blockDim.x) {
            if (MODE == 2) { d[x] = s[x]; continue; }
            T v = j.c ? (T)(int64_t)j.c * s[x] : (T)0;
            if (x == words - 1) v += (T)j.cst;
            d[x] = v;
        }
    }
}

// dst = sum_s c[s]*src[s] + cst(body), up to 4 sources (job layout shared with the keyswitch prologue)
// MODE as in linear_rows (0 in place, 1 results to scratch row r, 2 scratch row r to its wire)
template <class T, int MODE> __global__ void lincomb_rows(T* w, uint32_t stride, T* scratch, const KsJob* jobs, uint32_t cnt, uint32_t words) {
    for (uint32_t r = blockIdx.y; r < cnt; r += gridDim.y) {
        const KsJob j = jobs[r];
        T* d = MODE == 1 ? scratch + (size_t)r * stride : w + (size_t)j.dst * stride;
        for (uint32_t x = blockIdx.x * blockDim.x + threadIdx.x; x < words; x += gridDim.x * blockDim.x) {
            if (MODE == 2) { d[x] = scratch[(size_t)r * stride + x]; continue; }
            T v = 0;
#pragma unroll
            for (int s = 0; s < 4; ++s) if (j.c[s]) v += (T)(int64_t)j.c[s] * w[(size_t)j.src[s] * stride + x];
            if (x == words - 1) v += (T)j.cst;
            d[x] = v;
        }
    }
}

// Level exchange of a multi-device group: every output row of this device's share of the level is stored into the
// replicas of all peers over NVLink (peer-mapped memory, 16-byte stores, a row per blockIdx.y), right behind the kernels that
// produced it on the same stream; a per-level event per device then orders the peers' next level after it.  The rows are
// named by the `dst` field of the job descriptors that are already on the device.
template <class J> __global__ void push_rows(const void* src, PushArgs P, uint32_t row_bytes, const J* jobs, uint32_t cnt) {
    const uint32_t q = row_bytes >> 4;
    for (uint32_t r = blockIdx.y; r < cnt; r += gridDim.y) {
        const size_t off = (size_t)jobs[r].dst * row_bytes;
        const uint4* s = reinterpret_cast<const uint4*>(reinterpret_cast<const char*>(src) + off);
        for (uint32_t x = blockIdx.x * blockDim.x + threadIdx.x; x < q; x += gridDim.x * blockDim.x) {
            const uint4 v = s[x];
            for (int p = 0; p < P.n_dst; ++p) reinterpret_cast<uint4*>(reinterpret_cast<char*>(P.dst[p]) + off)[x] = v;
        }
    }
}

__global__ void dfma_peak_kernel(double* out, int iters) {
    double a0 = threadIdx.x * 1e-9, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6, a7 = a0 + 7;
    const double b = 1.0000001, c = 1e-7;
    for (int i = 0; i < iters; ++i) {
        a0 = fma(a0, b, c); a1 = fma(a1, b, c); a2 = fma(a2, b, c); a3 = fma(a3, b, c);
        a4 = fma(a4, b, c); a5 = fma(a5, b, c); a6 = fma(a6, b, c); a7 = fma(a7, b, c);
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
}

// ---- kernel shape table: (torus, N, k, l) -> template instance
struct Shape {
    int bits, N, K, L, RA, RB, G, NT;
    size_t (*smem)(int n);
    cudaError_t (*launch)(const PbsArgs&, cudaStream_t, size_t smem);
    cudaError_t (*bsk_fft)(const void* std_bsk, double2* out, const double2* wtab, size_t polys, cudaStream_t);
    int BL = 0;                            // compile-time pbs_base_log of the instance (0: run-time)
    int MB = 0;                            // multi-bit grouping factor of the instance (0: classic blind rotation)
    // narrow batches (<= 3 x SMs bootstraps): one gate per CTA, every transform on its own group of threads
    size_t (*small_smem)(int n) = nullptr;
    cudaError_t (*small_launch)(const PbsArgs&, cudaStream_t, size_t smem) = nullptr;
    // between one and small2_waves x SMs x 2 bootstraps: two gates per CTA (one wave of those beats two waves of the above)
    size_t (*small2_smem)(int n) = nullptr;
    cudaError_t (*small2_launch)(const PbsArgs&, cudaStream_t, size_t smem) = nullptr;
    // third-generation one-gate instance (pbs3_kernel.cuh); preferred over small / small2 unless HB_PBS3=0
    size_t (*small3_smem)(int n) = nullptr;
    cudaError_t (*small3_launch)(const PbsArgs&, cudaStream_t, size_t smem) = nullptr;
    // one bootstrap per cluster of K+1 CTAs (pbs3c_kernel.cuh): batches of at most cluster_max() bootstraps
    size_t (*cluster_smem)(int n) = nullptr;
    cudaError_t (*cluster_launch)(const PbsArgs&, cudaStream_t, size_t smem) = nullptr;
    int (*cluster_max)(size_t smem) = nullptr;
    size_t (*wide3_smem)(int n) = nullptr;
    cudaError_t (*wide3_launch)(const PbsArgs&, cudaStream_t, size_t smem) = nullptr;
};
template <class Small> constexpr Shape with_small(Shape s) { s.small_smem = &Small::smem; s.small_launch = &Small::launch; return s; }
template <class Small2> constexpr Shape with_small2(Shape s) { s.small2_smem = &Small2::smem; s.small2_launch = &Small2::launch; return s; }

// pbs_kernel keeps a small static table (untwist factors) beside its dynamic shared memory: 227 KB per CTA minus 1 KB
constexpr size_t PBS_DYN_SMEM_MAX = 226 * 1024;
#ifndef HB_SMALL_NT
#define HB_SMALL_NT 384        // threads of the one-gate boolean instance (288: 100.2 ms, 384: 95.4 ms, 512: 97.9 ms on c7552)
#endif
#ifndef HB_STAGE_KEY
#define HB_STAGE_KEY true      // narrow-batch instances stage BSK_i in shared memory (cp.async.bulk), see pbs_kernel.cuh
#endif
template <class T, int N, int K, int L, int RA, int RB, int G, int NT, int MINB, bool STAGE = false, int MB = 0>
struct Inst {
    using C = PbsCfg<T, N, K, L, RA, RB, G>;
    static size_t smem(int n) { return STAGE ? C::smem_bytes_staged(n) : C::smem_bytes(n); }
    static cudaError_t launch(const PbsArgs& a, cudaStream_t s, size_t smem) {
        auto kern = pbs_kernel<T, N, K, L, RA, RB, G, NT, MINB, STAGE, MB>;
        static SmemOptIn opt;
        if (cudaError_t e = opt.ensure(kern, PBS_DYN_SMEM_MAX)) return e;
        const unsigned grid = (a.n_jobs + G - 1) / G;
        kern<<<grid, NT, smem, s>>>(a);
        return cudaGetLastError();
    }
    static cudaError_t bsk_fft(const void* std_bsk, double2* out, const double2* wtab, size_t polys, cudaStream_t s) {
        const double scale = (sizeof(T) == 4 ? 1.0 : 1.0 / 18446744073709551616.0) / (double)(N / 2);
        bsk_fft_kernel<T, N, RA, RB><<<(unsigned)polys, RA > RB ? RA : RB, 0, s>>>((const T*)std_bsk, out, wtab, scale);
        return cudaGetLastError();
    }
    static constexpr Shape shape() { Shape s{(int)sizeof(T) * 8, N, K, L, RA, RB, G, NT, &smem, &launch, &bsk_fft}; s.MB = MB; return s; }
};

// narrow batches, third generation (pbs3_kernel.cuh): one gate per CTA, three-stage transforms on M / 8 threads each
template <class T, int N, int K, int L, int R3, int KA, int NT, bool STAGE, int MB = 0, int G = 1>
struct Inst3 {
    using C = Pbs3Cfg<T, N, K, L, 8, 8, R3, KA, NT, G>;
    static size_t smem(int n) { return C::smem_bytes(n, STAGE); }
    static cudaError_t launch(const PbsArgs& a, cudaStream_t s, size_t smem) {
        auto kern = pbs3_kernel<T, N, K, L, 8, 8, R3, KA, NT, STAGE, MB, G>;
        static SmemOptIn opt;
        if (cudaError_t e = opt.ensure(kern, PBS_DYN_SMEM_MAX)) return e;
        kern<<<(a.n_jobs + G - 1) / G, NT, smem, s>>>(a);
        return cudaGetLastError();
    }
};
// wide batches on the three-stage kernel with several gates per CTA (replaces the shape's first-generation wide instance unless HB_PBS3_WIDE=0)
template <class Wide> constexpr Shape with_wide3(Shape s) { s.wide3_smem = &Wide::smem; s.wide3_launch = &Wide::launch; return s; }
// narrowest batches: one bootstrap per cluster of K+1 CTAs (pbs3c_kernel.cuh).  The request is padded beyond half an SM's shared
// memory so that the CTAs of a cluster sit on different SMs (the point of the split).
template <class T, int N, int K, int L, int R3, int KA, int NT>
struct Inst3c {
    using C = Pbs3cCfg<T, N, K, L, R3, KA, NT>;
    static size_t smem(int n) { const size_t b = C::smem_bytes(n); return b < 116 * 1024 ? 116 * 1024 : b; }
    static cudaError_t launch(const PbsArgs& a, cudaStream_t s, size_t smem) {
        auto kern = pbs3c_kernel<T, N, K, L, R3, KA, NT>;
        static SmemOptIn opt;
        if (cudaError_t e = opt.ensure(kern, PBS_DYN_SMEM_MAX)) return e;
        kern<<<a.n_jobs * (K + 1), NT, smem, s>>>(a);
        return cudaGetLastError();
    }
    // clusters of this kernel that can be resident at once on the current device (0 on error)
    static int max_clusters(size_t smem) {
        auto kern = pbs3c_kernel<T, N, K, L, R3, KA, NT>;
        static SmemOptIn opt;
        if (opt.ensure(kern, PBS_DYN_SMEM_MAX) != cudaSuccess) return 0;
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3((K + 1) * 64); cfg.blockDim = dim3(NT); cfg.dynamicSmemBytes = smem;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = K + 1; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
        cfg.attrs = at; cfg.numAttrs = 1;
        int nc = 0;
        if (cudaOccupancyMaxActiveClusters(&nc, kern, &cfg) != cudaSuccess) { (void)cudaGetLastError(); return 0; }
        return nc;
    }
};
template <class Cl> constexpr Shape with_cluster(Shape s) { s.cluster_smem = &Cl::smem; s.cluster_launch = &Cl::launch; s.cluster_max = &Cl::max_clusters; return s; }
template <class Small> constexpr Shape with_small3(Shape s) { s.small3_smem = &Small::smem; s.small3_launch = &Small::launch; return s; }

// N = 512 shapes run the half-warp-per-polynomial kernel (pbs16_kernel.cuh); the key transform and the Fourier key layout are shared
template <class T, int K, int L, int BL, int G, int MINB, bool TMEMW>
struct Inst16 {
    using C = Pbs16Cfg<T, K, L, G, TMEMW>;
    static size_t smem(int n) { return C::smem_bytes(n); }
    static cudaError_t launch(const PbsArgs& a, cudaStream_t s, size_t smem) {
        auto kern = pbs16_kernel<T, K, L, BL, G, MINB, TMEMW>;
        static SmemOptIn opt;
        if (cudaError_t e = opt.ensure(kern, 227 * 1024)) return e;
        const unsigned grid = (a.n_jobs + G - 1) / G;
        kern<<<grid, C::NT, smem, s>>>(a);
        return cudaGetLastError();
    }
    static constexpr Shape shape() { return Shape{(int)sizeof(T) * 8, 512, K, L, 16, 16, G, C::NT, &smem, &launch, &Inst<T, 512, K, L, 16, 16, 1, 32, 1>::bsk_fft, BL}; }
};

#ifndef HB_PBS3C_NT
#define HB_PBS3C_NT 128        // threads per CTA of the cluster kernel (A/B switch)
#endif
const Shape g_shapes[] = {
    // tfhe boolean DEFAULT_PARAMETERS
    // pbs3 measured per shape (tests/tools/narrow_sweep.py, 100 bootstraps): boolean 2.63 vs 2.50 ms and M1C1 2.02 vs 1.97 ms (slower: on one
    // SM the N = 512 step is bound by shared-memory bandwidth and the FP64 pipe, not by the chain length), M2C2 3.34 vs 5.00 ms, multi-bit
    // 3.05 vs 4.27 ms (faster) => used for the N = 2048 shapes; -DHB_PBS3_ALL builds it for every shape (A/B)
    with_cluster<Inst3c<uint32_t, 512, 2, 3, 4, 16, HB_PBS3C_NT>>(
#ifdef HB_PBS3_ALL
    with_small3<Inst3<uint32_t, 512, 2, 3, 4, 16, 384, true>>(
#else
    (
#endif
        with_small2<Inst<uint32_t, 512, 2, 3, 16, 16, 2, HB_SMALL_NT, 1, HB_STAGE_KEY>>(
        with_small<Inst<uint32_t, 512, 2, 3, 16, 16, 1, HB_SMALL_NT, 1, HB_STAGE_KEY>>(Inst16<uint32_t, 2, 3, 6, HB_PBS16_G, 8 / HB_PBS16_G, HB_TMEMW>::shape())))),
    Inst<uint32_t, 1024, 1, 3, 16, 32, 1, 192, 2>::shape(),    // concrete-core GPU path of the reference
    // PARAM_MESSAGE_1_CARRY_1_KS_PBS / PARAM_MESSAGE_2_CARRY_1_KS_PBS
#ifndef HB_PBS16_U64_G
#define HB_PBS16_U64_G 2
#endif
#if HB_PBS16_U64_G > 0
#ifdef HB_PBS3_ALL
    with_small3<Inst3<uint64_t, 512, 3, 1, 4, 16, 256, true>>(
#else
    (
#endif
        with_small<Inst<uint64_t, 512, 3, 1, 16, 16, 1, 256, 1, HB_STAGE_KEY>>(Inst16<uint64_t, 3, 1, 18, HB_PBS16_U64_G, 6 / HB_PBS16_U64_G, HB_TMEMW>::shape())),
#else
    with_small<Inst<uint64_t, 512, 3, 1, 16, 16, 1, 256, 1, HB_STAGE_KEY>>(Inst<uint64_t, 512, 3, 1, 16, 16, 2, 128, 2>::shape()),
#endif
    // PARAM_MESSAGE_2_CARRY_2_KS_PBS
    with_wide3<Inst3<uint64_t, 2048, 1, 1, 16, 32, 512, false, 0, 2>>(with_small3<Inst3<uint64_t, 2048, 1, 1, 16, 32, 512, true>>(
        with_small<Inst<uint64_t, 2048, 1, 1, 32, 32, 1, 256, 1, HB_STAGE_KEY>>(Inst<uint64_t, 2048, 1, 1, 32, 32, 2, 128, 1>::shape()))),
    // PARAM_MULTI_BIT_MESSAGE_2_CARRY_2_GROUP_3_KS_PBS (the reference's arithmetic mode, src/bin/helm.rs:81-86): multi-bit, grouping 3
    with_small3<Inst3<uint64_t, 2048, 1, 1, 16, 32, 512, false, 3>>(
        with_small<Inst<uint64_t, 2048, 1, 1, 32, 32, 1, 256, 1, false, 3>>(Inst<uint64_t, 2048, 1, 1, 32, 32, 2, 128, 1, false, 3>::shape())),
};

int find_shape(const hb_params& p) {
    for (size_t i = 0; i < sizeof(g_shapes) / sizeof(g_shapes[0]); ++i) {
        const Shape& s = g_shapes[i];
        if (s.bits == p.torus_bits && s.N == p.N && s.K == p.k && s.L == p.pbs_level && (s.BL == 0 || s.BL == p.pbs_base_log) &&
            s.MB == (p.grouping > 1 ? p.grouping : 0)) return (int)i;
    }
    return -1;
}

}  // namespace

// ================================================================ internal launch helpers
namespace {

CtxExtra* extra_of(hb_ctx* c) { return &c->extra; }

// folds a finished slot's phase timings into the context totals (slot->done must have completed)
void fold_timing(hb_ctx* c, JobSlot& s, JobSlot* next = nullptr) {
    if (!s.pending) return;
    s.pending = false;
    float a = 0, b = 0, g = 0;
    // stream time between the end of this batch's kernels and the start of the next batch's (job upload, launch gaps)
    if (next && next != &s && next->pending && cudaEventQuery(next->t[0]) == cudaSuccess &&
        cudaEventElapsedTime(&g, s.t[2], next->t[0]) == cudaSuccess && g > 0) c->stats.gap_ms_total += g;
    else (void)cudaGetLastError();
    if (cudaEventElapsedTime(&a, s.t[0], s.t[1]) != cudaSuccess || cudaEventElapsedTime(&b, s.t[1], s.t[2]) != cudaSuccess) return;
    const double pbs = s.first_is_pbs ? a : b, ks = s.first_is_pbs ? b : a;
    c->stats.pbs_ms = pbs; c->stats.ks_ms = ks;
    c->stats.pbs_ms_total += pbs; c->stats.ks_ms_total += ks;
}

hb_status check_ready(hb_ctx* c) {
    if (!c) return HB_ERR_ARG;
    if (!c->have_params) return fail(c, HB_ERR_STATE, "hb_set_params not called");
    if (!c->have_bsk || !c->have_ksk) return fail(c, HB_ERR_STATE, "keys not loaded (hb_load_bsk / hb_load_ksk)");
    return HB_OK;
}

hb_status ensure_scratch(hb_ctx* c, DevBuf& b, size_t bytes) {
    if (bytes <= b.cap) return HB_OK;
    if (c->capturing) return fail(c, HB_ERR_STATE, "graph capture: scratch buffers must be sized by an eager evaluation of the same batches first");
    CU(c, b.ensure(bytes));
    c->scratch_gen++;
    return HB_OK;
}

// uploads the three job arrays of one level through a pinned ring slot; returns device pointers
hb_status upload_jobs(hb_ctx* c, const std::vector<PbsJob>& pj, const std::vector<KsJob>& kj, const std::vector<LinJob>& lj,
                      const PbsJob** dp, const KsJob** dk, const LinJob** dl, JobSlot** slot_out) {
    const size_t a = pj.size() * sizeof(PbsJob), b = kj.size() * sizeof(KsJob), d = lj.size() * sizeof(LinJob);
    const size_t oa = 0, ob = (a + 15) & ~(size_t)15, od = ob + ((b + 15) & ~(size_t)15), tot = od + d + 16;
    if (c->capturing) {
        // recorded, not executed: the descriptors live in the graph's arena (the copy node re-reads the pinned half on every replay)
        const size_t off = (c->cap_used + 255) & ~(size_t)255;
        if (off + tot > c->cap_pin.cap || off + tot > c->cap_dev.cap)
            return fail(c, HB_ERR_ARG, "graph capture: job arena too small (hb_graph_begin job_bytes)");
        char* h = (char*)c->cap_pin.p + off; char* dv = (char*)c->cap_dev.p + off;
        if (a) memcpy(h + oa, pj.data(), a);
        if (b) memcpy(h + ob, kj.data(), b);
        if (d) memcpy(h + od, lj.data(), d);
        CU(c, cudaMemcpyAsync(dv, h, tot, cudaMemcpyHostToDevice, c->stream));
        c->cap_used = off + tot;
        *dp = (const PbsJob*)(dv + oa); *dk = (const KsJob*)(dv + ob); *dl = (const LinJob*)(dv + od);
        *slot_out = &c->cap_slot;
        return HB_OK;
    }
    CtxExtra* ex = extra_of(c);
    JobSlot& s = ex->slots[ex->next++ % NSLOT];
    CU(c, cudaEventSynchronize(s.done));
    fold_timing(c, s, &ex->slots[ex->next % NSLOT]);
    if (s.pin.cap < tot || s.dev.cap < tot) {
        // grow every slot of the ring at once: otherwise the pinned / device allocations of slots 1 .. NSLOT-1 (tens of ms
        // each on a busy host) land in the second .. fourth batch, i.e. inside a caller's timed region after a short warm-up
        for (JobSlot& o : ex->slots) {
            if (o.pin.cap >= tot && o.dev.cap >= tot) continue;
            if (&o != &s && o.done) CU(c, cudaEventSynchronize(o.done));
            CU(c, o.pin.ensure(tot));
            CU(c, o.dev.ensure(tot));
        }
    }
    char* h = (char*)s.pin.p;
    if (a) memcpy(h + oa, pj.data(), a);
    if (b) memcpy(h + ob, kj.data(), b);
    if (d) memcpy(h + od, lj.data(), d);
    CU(c, cudaMemcpyAsync(s.dev.p, h, tot, cudaMemcpyHostToDevice, c->stream));
    *dp = (const PbsJob*)((char*)s.dev.p + oa);
    *dk = (const KsJob*)((char*)s.dev.p + ob);
    *dl = (const LinJob*)((char*)s.dev.p + od);
    *slot_out = &s;
    return HB_OK;
}

hb_status launch_pbs(hb_ctx* c, const void* in, uint32_t in_stride, void* out, uint32_t out_stride, const PbsJob* jobs, uint32_t n_jobs) {
    if (!n_jobs) return HB_OK;
    const Shape& sh = g_shapes[c->shape];
    PbsArgs a{};
    a.in = in; a.in_stride = in_stride; a.out = out; a.out_stride = out_stride; a.jobs = jobs; a.n_jobs = n_jobs;
    a.bsk = (const double2*)c->bsk_f.p; a.wtab = (const double2*)c->wtab.p; a.n = c->p.n; a.base_log = c->p.pbs_base_log;
    a.roots = (const double2*)c->wtab.p + 2 * (size_t)(sh.RA * sh.RB);
    // Narrow batches are latency-bound (one bootstrap is n sequential CMUX steps, so what matters is the time of ONE step of
    // ONE CTA).
    // Measured on boolean default (tests/tools/narrow_sweep.py, ms per launch): one gate per CTA 2.87 up to 1 x SMs bootstraps;
    // two gates per CTA 4.2 up to 2 x SMs; the throughput kernel 6.97 up to 4 x SMs and 10.6 up to 8 x SMs.  So: G = 1 for one
    // wave, G = 2 for at most one wave of pairs, the throughput kernel beyond.  The shortint shapes have no pair instance: G = 1 for one
    // wave (M2C2 6.3 ms, M1C1 2.3 ms), then their wide instances (9.4 / 3.9 ms up to 2 x SMs).  HB_PBS_SMALL=0 disables both, HB_PBS_SMALL_MAX / HB_PBS_SMALL2_MAX override the wave counts.
    static const int small_on = getenv("HB_PBS_SMALL") ? atoi(getenv("HB_PBS_SMALL")) : 1;
    static const int small_max_env = getenv("HB_PBS_SMALL_MAX") ? atoi(getenv("HB_PBS_SMALL_MAX")) : 0;
    static const int small2_max = getenv("HB_PBS_SMALL2_MAX") ? atoi(getenv("HB_PBS_SMALL2_MAX")) : 2;
    // the narrowest batches: one bootstrap per cluster of K+1 SMs (HB_PBS3C=0 disables)
    static const int cluster_on = getenv("HB_PBS3C") ? atoi(getenv("HB_PBS3C")) : 1;
    if (small_on && cluster_on && sh.cluster_launch) {
        const size_t sm = sh.cluster_smem(c->p.n);
        if (c->cluster_cap < 0) c->cluster_cap = sm <= PBS_DYN_SMEM_MAX ? sh.cluster_max(sm) : 0;
        if (n_jobs <= (uint32_t)c->cluster_cap) {
            CU(c, sh.cluster_launch(a, c->stream, sm));
            c->stats.kernel_launches++; c->stats.pbs_count += n_jobs;
            return HB_OK;
        }
    }
    // third-generation one-gate kernel: up to small3_max waves of one CTA per SM (HB_PBS3=0 falls back to the instances below)
    static const int small3_on = getenv("HB_PBS3") ? atoi(getenv("HB_PBS3")) : 1;
    // waves of it that still beat the shape's wide kernel (narrow_sweep, M2C2: a wave of 148 takes 3.34 ms, the two-gate wide instance
    // 6.57 ms per 296 => one-gate waves win up to the point where the wide kernel's better tail packing matters; multi-bit: always)
    static const int small3_env = getenv("HB_PBS3_MAX") ? atoi(getenv("HB_PBS3_MAX")) : 0;
    const int small3_max = small3_env ? small3_env : sh.MB ? (1 << 20) : sh.wide3_launch ? 8 : 1;
    if (small_on && small3_on && sh.small3_launch && n_jobs <= (uint32_t)small3_max * (uint32_t)c->sm_count) {
        const size_t sm = sh.small3_smem(c->p.n);
        if (sm <= PBS_DYN_SMEM_MAX) {
            CU(c, sh.small3_launch(a, c->stream, sm));
            c->stats.kernel_launches++; c->stats.pbs_count += n_jobs;
            return HB_OK;
        }
    }
    const bool pair_ok = small_on && sh.small2_launch && small2_max > 0;
    const int small_max = small_max_env ? small_max_env : 1;
    if (pair_ok && n_jobs > (uint32_t)(small_max * c->sm_count) && n_jobs <= (uint32_t)(small2_max * c->sm_count)) {
        const size_t sm = sh.small2_smem(c->p.n);
        if (sm <= PBS_DYN_SMEM_MAX) {
            CU(c, sh.small2_launch(a, c->stream, sm));
            c->stats.kernel_launches++; c->stats.pbs_count += n_jobs;
            return HB_OK;
        }
    }
    if (small_on && sh.small_launch && n_jobs <= (uint32_t)(small_max * c->sm_count)) {
        const size_t sm = sh.small_smem(c->p.n);
        if (sm <= PBS_DYN_SMEM_MAX) {
            CU(c, sh.small_launch(a, c->stream, sm));
            c->stats.kernel_launches++; c->stats.pbs_count += n_jobs;
            return HB_OK;
        }
    }
    static const int wide3_on = getenv("HB_PBS3_WIDE") ? atoi(getenv("HB_PBS3_WIDE")) : 1;
    if (wide3_on && sh.wide3_launch) {
        const size_t sm = sh.wide3_smem(c->p.n);
        if (sm <= PBS_DYN_SMEM_MAX) {
            CU(c, sh.wide3_launch(a, c->stream, sm));
            c->stats.kernel_launches++; c->stats.pbs_count += n_jobs;
            return HB_OK;
        }
    }
    const size_t smem = sh.smem(c->p.n);
    if (smem > 227 * 1024) return fail(c, HB_ERR_UNSUPPORTED, "PBS shared memory footprint exceeds 227 KB");
    a.atil_stride = (uint32_t)((c->p.n + 2 + 7) & ~7);
    if (hb_status s = ensure_scratch(c, c->atil, ((size_t)n_jobs + 16) * a.atil_stride * sizeof(uint16_t))) return s;
    a.atil = (uint16_t*)c->atil.p;
    CU(c, sh.launch(a, c->stream, smem));
    c->stats.kernel_launches++; c->stats.pbs_count += n_jobs;
    return HB_OK;
}

hb_status launch_ks(hb_ctx* c, const void* in, uint32_t in_stride, void* out, uint32_t out_stride, const KsJob* jobs, uint32_t n_jobs) {
    if (!n_jobs) return HB_OK;
    KsArgs a;
    a.in = in; a.in_stride = in_stride; a.out = out; a.out_stride = out_stride; a.jobs = jobs; a.n_jobs = n_jobs;
    a.ksk = c->ksk.p; a.n = c->p.n; a.kN = c->p.k * c->p.N; a.base_log = c->p.ks_base_log; a.level = c->p.ks_level;
    // batches from ~100 jobs on the tensor cores (bit-identical; HB_KS_MMA=0 keeps the IMAD kernel)
    static const int mma_on = getenv("HB_KS_MMA") ? atoi(getenv("HB_KS_MMA")) : 1;
    static const int mma_min = getenv("HB_KS_MMA_MIN") ? atoi(getenv("HB_KS_MMA_MIN")) : 96;   // tests/tools/narrow_sweep.py: 0.050 ms up to 592 jobs, IMAD 0.044 (64) .. 0.195 (592)
    const bool too_wide = c->p.n + 1 > KS_NT * KS_QC;           // the IMAD tile covers 768 output columns: wider keys always take the tensor path
    if (too_wide && !c->ksk_npad) return fail(c, HB_ERR_UNSUPPORTED, "keyswitch: n+1 > 768 needs the tensor-core path");
    if (c->ksk_npad && (too_wide || (mma_on && n_jobs >= (uint32_t)mma_min))) {
        KsMmaArgs m;
        m.a = a; m.a.split = 1; m.planes = (const uint8_t*)c->ksk_planes.p; m.npad = c->ksk_npad;
        const int tn = c->p.torus_bits == 32 ? KmCfg<uint32_t>::TN : KmCfg<uint64_t>::TN;
        const dim3 grid((n_jobs + KM_TM - 1) / KM_TM, c->ksk_npad / tn);
        const uint32_t rows_pad = grid.x * KM_TM;
        if (hb_status s = ensure_scratch(c, c->ks_dig, (size_t)rows_pad * a.kN * a.level)) return s;
        const unsigned dg = rows_pad < 65535u ? rows_pad : 65535u;
        if (c->p.torus_bits == 32) {
            static SmemOptIn opt2;
            const size_t smem2 = 2 * (size_t)KmCfg<uint32_t>::STAGE;
            CU(c, opt2.ensure(ks_mma2_kernel<uint32_t>, smem2));
            ks_digits_kernel<uint32_t><<<dg, 256, 0, c->stream>>>(m.a, (uint8_t*)c->ks_dig.p, rows_pad);
            ks_mma2_kernel<uint32_t><<<grid, KM_NT, smem2, c->stream>>>(m, (const uint8_t*)c->ks_dig.p);
        } else {
            static SmemOptIn opt3;
            const size_t smem2 = 2 * (size_t)KmCfg<uint64_t>::STAGE;
            CU(c, opt3.ensure(ks_mma2_kernel<uint64_t>, smem2));
            ks_digits_kernel<uint64_t><<<dg, 256, 0, c->stream>>>(m.a, (uint8_t*)c->ks_dig.p, rows_pad);
            ks_mma2_kernel<uint64_t><<<grid, KM_NT, smem2, c->stream>>>(m, (const uint8_t*)c->ks_dig.p);
        }
        CU(c, cudaGetLastError());
        c->stats.kernel_launches++;
        c->stats.kernel_launches++; c->stats.ks_count += n_jobs;
        return HB_OK;
    }
    const unsigned tiles = (n_jobs + KS_TB - 1) / KS_TB;
    // narrow batches: fewer tiles than SMs would leave each CTA streaming the whole key alone -- split the mask
    // dimension over gridDim.y CTAs (partial sums added atomically; wrapping integer adds, so still bit-exact)
    int split = 1;
    if (tiles * 2 <= (unsigned)c->sm_count) {
        split = (2 * c->sm_count) / (int)tiles;
        const int max_split = a.kN / 16;
        if (split > max_split) split = max_split;
        if (split < 1) split = 1;
    }
    a.split = split;
    const dim3 grid(tiles, split);
    if (c->p.torus_bits == 32) {
        if (split > 1) ks_init_rows<uint32_t><<<n_jobs < 1024 ? n_jobs : 1024, 256, 0, c->stream>>>(a);
        if (split > 1) ks_kernel<uint32_t, true><<<grid, KS_NT, 0, c->stream>>>(a);
        else ks_kernel<uint32_t, false><<<grid, KS_NT, 0, c->stream>>>(a);
    } else {
        if (split > 1) ks_init_rows<uint64_t><<<n_jobs < 1024 ? n_jobs : 1024, 256, 0, c->stream>>>(a);
        if (split > 1) ks_kernel<uint64_t, true><<<grid, KS_NT, 0, c->stream>>>(a);
        else ks_kernel<uint64_t, false><<<grid, KS_NT, 0, c->stream>>>(a);
    }
    CU(c, cudaGetLastError());
    if (split > 1) c->stats.kernel_launches++;
    c->stats.kernel_launches++; c->stats.ks_count += n_jobs;
    return HB_OK;
}

// true iff some job of the batch reads a row that another job of the batch writes
bool lin_hazard(const std::vector<LinJob>& lj) {
    if (lj.size() < 2) return false;
    std::vector<uint32_t> dst(lj.size());
    for (size_t i = 0; i < lj.size(); ++i) dst[i] = lj[i].dst;
    std::sort(dst.begin(), dst.end());
    for (const LinJob& j : lj)
        if (j.c && j.src != j.dst && std::binary_search(dst.begin(), dst.end(), j.src)) return true;
    return false;
}

template <class T> hb_status launch_lin_t(hb_ctx* c, void* w, uint32_t stride, const LinJob* jobs, uint32_t cnt, uint32_t words, bool staged) {
    dim3 grid((words + 255) / 256, cnt < 32768 ? cnt : 32768);
    if (!staged) {
        linear_rows<T, 0><<<grid, 256, 0, c->stream>>>((T*)w, stride, nullptr, jobs, cnt, words);
        c->stats.kernel_launches++;
    } else {
        if (hb_status s = ensure_scratch(c, c->lin_stage, (size_t)cnt * stride * sizeof(T))) return s;
        linear_rows<T, 1><<<grid, 256, 0, c->stream>>>((T*)w, stride, (T*)c->lin_stage.p, jobs, cnt, words);
        linear_rows<T, 2><<<grid, 256, 0, c->stream>>>((T*)w, stride, (T*)c->lin_stage.p, jobs, cnt, words);
        c->stats.kernel_launches += 2;
    }
    CU(c, cudaGetLastError());
    return HB_OK;
}
hb_status launch_lin(hb_ctx* c, void* w, uint32_t stride, const LinJob* jobs, uint32_t cnt, uint32_t words, bool staged = false) {
    if (!cnt) return HB_OK;
    return c->p.torus_bits == 32 ? launch_lin_t<uint32_t>(c, w, stride, jobs, cnt, words, staged)
                                 : launch_lin_t<uint64_t>(c, w, stride, jobs, cnt, words, staged);
}

bool lincomb_hazard(const std::vector<KsJob>& kj) {
    if (kj.size() < 2) return false;
    std::vector<uint32_t> dst(kj.size());
    for (size_t i = 0; i < kj.size(); ++i) dst[i] = kj[i].dst;
    std::sort(dst.begin(), dst.end());
    for (const KsJob& j : kj)
        for (int s = 0; s < 4; ++s)
            if (j.c[s] && j.src[s] != j.dst && std::binary_search(dst.begin(), dst.end(), j.src[s])) return true;
    return false;
}
template <class T> hb_status launch_lincomb_t(hb_ctx* c, void* w, uint32_t stride, const KsJob* jobs, uint32_t cnt, uint32_t words, bool staged) {
    dim3 grid((words + 255) / 256, cnt < 32768 ? cnt : 32768);
    if (!staged) {
        lincomb_rows<T, 0><<<grid, 256, 0, c->stream>>>((T*)w, stride, nullptr, jobs, cnt, words);
        c->stats.kernel_launches++;
    } else {
        if (hb_status s = ensure_scratch(c, c->lin_stage, (size_t)cnt * stride * sizeof(T))) return s;
        lincomb_rows<T, 1><<<grid, 256, 0, c->stream>>>((T*)w, stride, (T*)c->lin_stage.p, jobs, cnt, words);
        lincomb_rows<T, 2><<<grid, 256, 0, c->stream>>>((T*)w, stride, (T*)c->lin_stage.p, jobs, cnt, words);
        c->stats.kernel_launches += 2;
    }
    CU(c, cudaGetLastError());
    return HB_OK;
}
hb_status launch_lincomb(hb_ctx* c, void* w, uint32_t stride, const KsJob* jobs, uint32_t cnt, uint32_t words, bool staged = false) {
    if (!cnt) return HB_OK;
    return c->p.torus_bits == 32 ? launch_lincomb_t<uint32_t>(c, w, stride, jobs, cnt, words, staged)
                                 : launch_lincomb_t<uint64_t>(c, w, stride, jobs, cnt, words, staged);
}

template <class J> hb_status launch_push(hb_ctx* c, hb_wires* w, const J* jobs, uint32_t cnt) {
    if (!cnt || !w->peers.n_dst) return HB_OK;
    const uint32_t row_bytes = (uint32_t)(pad4(ct_words(c)) * tsize(c));
    const dim3 grid((row_bytes / 16 + 255) / 256, cnt < 32768 ? cnt : 32768);
    push_rows<J><<<grid, 256, 0, c->stream>>>(w->dev, w->peers, row_bytes, jobs, cnt);
    CU(c, cudaGetLastError());
    c->stats.kernel_launches++;
    c->stats.push_bytes += (uint64_t)cnt * row_bytes * (uint64_t)w->peers.n_dst;
    return HB_OK;
}

// error text of a group member -> the group context
hb_status lift(hb_ctx* g, hb_ctx* sub, hb_status s) { if (s != HB_OK) g->err = "device " + std::to_string(sub->device) + ": " + sub->err; return s; }

}  // namespace

// ================================================================ multi-device group (implemented at the end of this file)
namespace {
hb_status group_destroy(hb_ctx* g);
hb_status group_each(hb_ctx* g, hb_status (*f)(hb_ctx*, const void*, size_t), const void* a, size_t b);
hb_status group_wires_create(hb_ctx* g, uint32_t n_wires, hb_wires** out);
hb_status group_eval_level(hb_ctx* g, hb_wires* w, const hb_gate* gates, uint32_t n_gates);
hb_status group_eval_sops(hb_ctx* g, hb_wires* w, const hb_sop* ops, uint32_t n_ops);
hb_status group_stats(hb_ctx* g, hb_stats_t* out);
}  // namespace

// ================================================================ C ABI
extern "C" {

const char* hb_last_error(const hb_ctx* c) { return c ? c->err.c_str() : g_create_error.c_str(); }

hb_status hb_ctx_create(int32_t device, hb_ctx** out) {
    if (!out) return HB_ERR_ARG;
    *out = nullptr;
    int cnt = 0;
    cudaError_t e = cudaGetDeviceCount(&cnt);
    if (e != cudaSuccess || cnt == 0)
        return fail(nullptr, HB_ERR_CUDA, std::string("no CUDA device (helm_b200 has no CPU fallback): ") + cudaGetErrorString(e));
    if (device < 0 || device >= cnt) return fail(nullptr, HB_ERR_ARG, "device index out of range");
    CU(nullptr, cudaSetDevice(device));
    hb_ctx* c = new hb_ctx();
    c->device = device;
    CU(nullptr, cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
    for (auto& ev : c->ev) CU(nullptr, cudaEventCreate(&ev));
    CU(nullptr, cudaDeviceGetAttribute(&c->sm_count, cudaDevAttrMultiProcessorCount, device));
    for (auto& s : c->extra.slots) {
        CU(nullptr, cudaEventCreateWithFlags(&s.done, cudaEventDisableTiming));
        for (auto& t : s.t) CU(nullptr, cudaEventCreate(&t));
    }
    CU(nullptr, cudaEventCreateWithFlags(&c->cap_slot.done, cudaEventDisableTiming));
    for (auto& t : c->cap_slot.t) CU(nullptr, cudaEventCreateWithFlags(&t, cudaEventDisableTiming));
    *out = c;
    return HB_OK;
}

hb_status hb_ctx_destroy(hb_ctx* c) {
    if (!c) return HB_ERR_ARG;
    if (c->is_group()) return group_destroy(c);
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    for (DevBuf* b : {&c->bsk_f, &c->ksk, &c->ksk_planes, &c->ks_dig, &c->wtab, &c->scratch_a, &c->scratch_b, &c->atil, &c->lin_stage, &c->jobs_dev, &c->stage_dev, &c->idx_dev}) b->release();
    c->jobs_pin.release(); c->stage_pin.release();
    for (auto& s : c->extra.slots) {
        s.pin.release(); s.dev.release(); cudaEventDestroy(s.done);
        for (auto& t : s.t) cudaEventDestroy(t);
    }
    for (auto& ev : c->ev) cudaEventDestroy(ev);
    cudaEventDestroy(c->cap_slot.done);
    for (auto& t : c->cap_slot.t) cudaEventDestroy(t);
    c->cap_pin.release(); c->cap_dev.release();
    if (c->level_ev) cudaEventDestroy(c->level_ev);
    if (c->io_stream) {
        cudaStreamSynchronize(c->io_stream); cudaStreamSynchronize(c->io_down);
        cudaStreamDestroy(c->io_stream); cudaStreamDestroy(c->io_down); cudaEventDestroy(c->io_written); cudaEventDestroy(c->io_computed);
    }
    c->io_in.release(); c->io_out.release();
    cudaStreamDestroy(c->stream);
    delete c;
    return HB_OK;
}

hb_status hb_set_params(hb_ctx* c, const hb_params* p) {
    if (!c || !p) return HB_ERR_ARG;
    if (c->is_group()) {
        for (hb_ctx* m : c->subs) if (hb_status s = lift(c, m, hb_set_params(m, p))) return s;
        c->p = *p; c->have_params = true; c->have_bsk = c->have_ksk = false;
        return HB_OK;
    }
    if (p->torus_bits != 32 && p->torus_bits != 64) return fail(c, HB_ERR_ARG, "torus_bits must be 32 or 64");
    const int sh = find_shape(*p);
    if (sh < 0) return fail(c, HB_ERR_UNSUPPORTED, "no sm_100a kernel instance for this (torus_bits, N, k, pbs_level)");
    const int64_t ksK = (int64_t)p->k * p->N * p->ks_level;
    const bool ks_mma_ok = ksK % KM_KC == 0 && p->ks_base_log <= 7 && ksK * (1ll << (p->ks_base_log - 1)) * 255 < (1ll << 31);
    if (p->n + 1 > KS_NT * KS_QC && !ks_mma_ok) return fail(c, HB_ERR_UNSUPPORTED, "n+1 exceeds the keyswitch tile (768) and the key shape rules out the tensor-core keyswitch");
    if (p->grouping > 1 && p->n % p->grouping) return fail(c, HB_ERR_ARG, "multi-bit: n must be a multiple of the grouping factor");
    if (p->ks_level > KS_MAXL || p->ks_base_log > 7) return fail(c, HB_ERR_UNSUPPORTED, "keyswitch decomposition too wide (level<=8, base_log<=7)");
    if (p->pbs_base_log * p->pbs_level > p->torus_bits || p->pbs_base_log > 31) return fail(c, HB_ERR_ARG, "bad PBS decomposition");
    if (p->torus_bits == 32 && p->pbs_base_log > 7) return fail(c, HB_ERR_UNSUPPORTED, "u32 torus: pbs_base_log must be <= 7 (digits are staged as bytes)");
    if (p->order == 1 && (p->msg_mod * p->carry_mod < 2 || p->msg_mod * p->carry_mod > 16)) return fail(c, HB_ERR_ARG, "msg_mod*carry_mod must be in [2,16]");
    CU(c, cudaSetDevice(c->device));
    c->p = *p; c->shape = sh; c->have_params = true; c->have_bsk = c->have_ksk = false; c->cluster_cap = -1;
    const Shape& s = g_shapes[sh];
    const int M = s.RA * s.RB;
    CU(c, c->wtab.ensure((2 * (size_t)M + 2 * (size_t)p->N) * sizeof(double2)));
    wtab_kernel<<<(M + 255) / 256, 256, 0, c->stream>>>((double2*)c->wtab.p, s.RA, s.RB);
    roots_kernel<<<(2 * p->N + 255) / 256, 256, 0, c->stream>>>((double2*)c->wtab.p + 2 * M, p->N);      // multi-bit monomial spectra
    CU(c, cudaGetLastError());
    c->stats.kernel_launches += 2;
    return HB_OK;
}

hb_status hb_load_bsk(hb_ctx* c, const void* bsk, size_t bytes) {
    if (!c || !bsk) return HB_ERR_ARG;
    if (c->is_group()) { if (hb_status s = group_each(c, &hb_load_bsk, bsk, bytes)) return s; c->have_bsk = true; return HB_OK; }
    if (!c->have_params) return fail(c, HB_ERR_STATE, "hb_set_params not called");
    CU(c, cudaSetDevice(c->device));
    const hb_params& p = c->p;
    const size_t n_ggsw = p.grouping > 1 ? (size_t)(p.n / p.grouping) << p.grouping : (size_t)p.n;
    const size_t polys = n_ggsw * p.pbs_level * (p.k + 1) * (p.k + 1);
    const size_t want = polys * p.N * tsize(c);
    if (bytes != want) return fail(c, HB_ERR_ARG, "hb_load_bsk: byte size does not match [n][l][k+1][k+1][N] (multi-bit: [n/g][2^g][l][k+1][k+1][N])");
    const Shape& s = g_shapes[c->shape];
    CU(c, c->bsk_f.ensure(polys * (p.N / 2) * sizeof(double2)));
    // stage the standard-domain key in chunks, transform on device
    const size_t chunk_polys = 4096;
    CU(c, c->stage_dev.ensure(chunk_polys * p.N * tsize(c)));
    for (size_t q = 0; q < polys; q += chunk_polys) {
        const size_t m = polys - q < chunk_polys ? polys - q : chunk_polys;
        CU(c, cudaMemcpyAsync(c->stage_dev.p, (const char*)bsk + q * p.N * tsize(c), m * p.N * tsize(c), cudaMemcpyHostToDevice, c->stream));
        CU(c, s.bsk_fft(c->stage_dev.p, (double2*)c->bsk_f.p + q * (p.N / 2), (const double2*)c->wtab.p, m, c->stream));
        c->stats.kernel_launches++;
        CU(c, cudaStreamSynchronize(c->stream));
    }
    c->have_bsk = true;
    return HB_OK;
}

hb_status hb_load_ksk(hb_ctx* c, const void* ksk, size_t bytes) {
    if (!c || !ksk) return HB_ERR_ARG;
    if (c->is_group()) { if (hb_status s = group_each(c, &hb_load_ksk, ksk, bytes)) return s; c->have_ksk = true; return HB_OK; }
    if (!c->have_params) return fail(c, HB_ERR_STATE, "hb_set_params not called");
    CU(c, cudaSetDevice(c->device));
    const hb_params& p = c->p;
    const size_t want = (size_t)p.k * p.N * p.ks_level * (p.n + 1) * tsize(c);
    if (bytes != want) return fail(c, HB_ERR_ARG, "hb_load_ksk: byte size does not match [kN][l_ks][n+1]");
    CU(c, c->ksk.ensure(want));
    CU(c, cudaMemcpyAsync(c->ksk.p, ksk, want, cudaMemcpyHostToDevice, c->stream));
    c->ksk_npad = 0;
    // tensor-core keyswitch (ks_mma_kernel.cuh): byte planes of the key; every int32 plane sum must be exact
    const int64_t K = (int64_t)p.k * p.N * p.ks_level;
    if (K % KM_KC == 0 && p.ks_base_log <= 7 && K * (1ll << (p.ks_base_log - 1)) * 255 < (1ll << 31)) {
        const int planes = p.torus_bits / 8, tn = 256 / planes, npad = ((p.n + 1 + tn - 1) / tn) * tn;
        CU(c, c->ksk_planes.ensure((size_t)planes * npad * K));
        if (p.torus_bits == 32) ksk_planes_kernel<uint32_t><<<4 * c->sm_count, 256, 0, c->stream>>>((const uint32_t*)c->ksk.p, (uint8_t*)c->ksk_planes.p, p.n, (int)K, npad);
        else ksk_planes_kernel<uint64_t><<<4 * c->sm_count, 256, 0, c->stream>>>((const uint64_t*)c->ksk.p, (uint8_t*)c->ksk_planes.p, p.n, (int)K, npad);
        CU(c, cudaGetLastError());
        c->ksk_npad = npad;
    }
    CU(c, cudaStreamSynchronize(c->stream));
    c->have_ksk = true;
    return HB_OK;
}

uint32_t hb_ct_words(const hb_ctx* c) { return c && c->have_params ? ct_words(c) : 0; }
uint32_t hb_wire_stride(const hb_ctx* c) { return c && c->have_params ? pad4(ct_words(c)) : 0; }
void* hb_stream(hb_ctx* c) { return !c ? nullptr : c->is_group() ? (void*)c->subs[0]->stream : (void*)c->stream; }

hb_status hb_wires_create(hb_ctx* c, uint32_t n_wires, hb_wires** out) {
    if (!c || !out || !n_wires) return HB_ERR_ARG;
    if (!c->have_params) return fail(c, HB_ERR_STATE, "hb_set_params not called");
    if (c->is_group()) return group_wires_create(c, n_wires, out);
    CU(c, cudaSetDevice(c->device));
    hb_wires* w = new hb_wires();
    w->ctx = c; w->n_wires = n_wires; w->owned = true;
    const size_t bytes = (size_t)n_wires * hb_wire_stride(c) * tsize(c);
    cudaError_t e = cudaMalloc(&w->dev, bytes);
    if (e != cudaSuccess) { delete w; return fail(c, HB_ERR_CUDA, std::string("cudaMalloc wires: ") + cudaGetErrorString(e)); }
    CU(c, cudaMemsetAsync(w->dev, 0, bytes, c->stream));
    *out = w;
    return HB_OK;
}
hb_status hb_wires_attach(hb_ctx* c, void* dev, uint32_t n_wires, hb_wires** out) {
    if (!c || !out || !dev || !n_wires) return HB_ERR_ARG;
    if (c->is_group()) return fail(c, HB_ERR_UNSUPPORTED, "hb_wires_attach: a multi-device context owns its wire replicas");
    if (!c->have_params) return fail(c, HB_ERR_STATE, "hb_set_params not called");
    hb_wires* w = new hb_wires();
    w->ctx = c; w->n_wires = n_wires; w->owned = false; w->dev = dev;
    *out = w;
    return HB_OK;
}
hb_status hb_wires_destroy(hb_wires* w) {
    if (!w) return HB_ERR_ARG;
    for (hb_wires* m : w->subs) hb_wires_destroy(m);
    if (!w->subs.empty()) { delete w; return HB_OK; }
    if (w->owned && w->dev) { cudaSetDevice(w->ctx->device); cudaStreamSynchronize(w->ctx->stream); cudaFree(w->dev); }
    delete w;
    return HB_OK;
}
void* hb_wires_device_ptr(hb_wires* w) { return !w ? nullptr : w->subs.empty() ? w->dev : w->subs[0]->dev; }

static hb_status check_idx(hb_wires* w, const uint32_t* idx, uint32_t cnt) {
    if (idx) { for (uint32_t i = 0; i < cnt; ++i) if (idx[i] >= w->n_wires) return fail(w->ctx, HB_ERR_ARG, "wire index out of range"); }
    else if (cnt > w->n_wires) return fail(w->ctx, HB_ERR_ARG, "wire count out of range");
    return HB_OK;
}

hb_status hb_wires_write(hb_wires* w, const uint32_t* idx, uint32_t cnt, const void* host) {
    if (!w || (!host && cnt)) return HB_ERR_ARG;
    if (!cnt) return HB_OK;
    if (!w->subs.empty()) {                                   // every replica
        for (hb_wires* m : w->subs) if (hb_status s = lift(w->ctx, m->ctx, hb_wires_write(m, idx, cnt, host))) return s;
        return HB_OK;
    }
    hb_ctx* c = w->ctx;
    if (hb_status s = check_idx(w, idx, cnt)) return s;
    CU(c, cudaSetDevice(c->device));
    const uint32_t words = ct_words(c), stride = hb_wire_stride(c);
    const size_t ts = tsize(c);
    if (!idx) {
        CU(c, cudaMemcpy2DAsync(w->dev, stride * ts, host, words * ts, words * ts, cnt, cudaMemcpyHostToDevice, c->stream));
    } else {
        CU(c, c->stage_dev.ensure((size_t)cnt * words * ts));
        CU(c, c->idx_dev.ensure((size_t)cnt * 4));
        CU(c, cudaMemcpyAsync(c->stage_dev.p, host, (size_t)cnt * words * ts, cudaMemcpyHostToDevice, c->stream));
        CU(c, cudaMemcpyAsync(c->idx_dev.p, idx, (size_t)cnt * 4, cudaMemcpyHostToDevice, c->stream));
        dim3 grid((words + 255) / 256, cnt < 32768 ? cnt : 32768);
        if (ts == 4) scatter_rows<uint32_t><<<grid, 256, 0, c->stream>>>((uint32_t*)w->dev, stride, (const uint32_t*)c->stage_dev.p, (const uint32_t*)c->idx_dev.p, cnt, words);
        else scatter_rows<uint64_t><<<grid, 256, 0, c->stream>>>((uint64_t*)w->dev, stride, (const uint64_t*)c->stage_dev.p, (const uint32_t*)c->idx_dev.p, cnt, words);
        CU(c, cudaGetLastError());
        c->stats.kernel_launches++;
    }
    CU(c, cudaStreamSynchronize(c->stream));
    return HB_OK;
}

hb_status hb_wires_read(hb_wires* w, const uint32_t* idx, uint32_t cnt, void* host) {
    if (!w || (!host && cnt)) return HB_ERR_ARG;
    if (!cnt) return HB_OK;
    if (!w->subs.empty()) {                                   // replicas are coherent between levels: read the first
        if (hb_status s = hb_sync(w->ctx)) return s;
        return lift(w->ctx, w->subs[0]->ctx, hb_wires_read(w->subs[0], idx, cnt, host));
    }
    hb_ctx* c = w->ctx;
    if (hb_status s = check_idx(w, idx, cnt)) return s;
    CU(c, cudaSetDevice(c->device));
    const uint32_t words = ct_words(c), stride = hb_wire_stride(c);
    const size_t ts = tsize(c);
    if (!idx) {
        CU(c, cudaMemcpy2DAsync(host, words * ts, w->dev, stride * ts, words * ts, cnt, cudaMemcpyDeviceToHost, c->stream));
    } else {
        CU(c, c->stage_dev.ensure((size_t)cnt * words * ts));
        CU(c, c->idx_dev.ensure((size_t)cnt * 4));
        CU(c, cudaMemcpyAsync(c->idx_dev.p, idx, (size_t)cnt * 4, cudaMemcpyHostToDevice, c->stream));
        dim3 grid((words + 255) / 256, cnt < 32768 ? cnt : 32768);
        if (ts == 4) gather_rows<uint32_t><<<grid, 256, 0, c->stream>>>((const uint32_t*)w->dev, stride, (uint32_t*)c->stage_dev.p, (const uint32_t*)c->idx_dev.p, cnt, words);
        else gather_rows<uint64_t><<<grid, 256, 0, c->stream>>>((const uint64_t*)w->dev, stride, (uint64_t*)c->stage_dev.p, (const uint32_t*)c->idx_dev.p, cnt, words);
        CU(c, cudaGetLastError());
        c->stats.kernel_launches++;
        CU(c, cudaMemcpyAsync(host, c->stage_dev.p, (size_t)cnt * words * ts, cudaMemcpyDeviceToHost, c->stream));
    }
    CU(c, cudaStreamSynchronize(c->stream));
    return HB_OK;
}

static constexpr size_t BIG_COPY = 4u << 20;   // row-range copies from this size go through a 1-D transfer + device repack
hb_status hb_wires_write_rows(hb_wires* w, uint32_t first, uint32_t cnt, const void* host) {
    if (!w || (!host && cnt)) return HB_ERR_ARG;
    if (!cnt) return HB_OK;
    if (!w->subs.empty()) {
        // one PCIe copy into the first replica, then device-to-device copies over NVLink (not n_devices host copies)
        hb_wires* w0 = w->subs[0];
        if (hb_status s = lift(w->ctx, w0->ctx, hb_wires_write_rows(w0, first, cnt, host))) return s;
        const size_t ts = tsize(w0->ctx), stride = hb_wire_stride(w0->ctx), off = (size_t)first * stride * ts, len = (size_t)cnt * stride * ts;
        for (size_t d = 1; d < w->subs.size(); ++d) {
            hb_wires* m = w->subs[d];
            CU(w->ctx, cudaSetDevice(m->ctx->device));
            CU(w->ctx, cudaMemcpyPeerAsync((char*)m->dev + off, m->ctx->device, (const char*)w0->dev + off, w0->ctx->device, len, m->ctx->stream));
        }
        return hb_sync(w->ctx);
    }
    hb_ctx* c = w->ctx;
    if ((uint64_t)first + cnt > w->n_wires) return fail(c, HB_ERR_ARG, "wire row range out of bounds");
    CU(c, cudaSetDevice(c->device));
    const size_t ts = tsize(c), words = ct_words(c), stride = hb_wire_stride(c);
    char* drow = (char*)w->dev + (size_t)first * stride * ts;
    if ((size_t)cnt * words * ts >= BIG_COPY) {
        CU(c, c->stage_dev.ensure((size_t)cnt * words * ts));
        CU(c, cudaMemcpyAsync(c->stage_dev.p, host, (size_t)cnt * words * ts, cudaMemcpyHostToDevice, c->stream));
        const dim3 grid((unsigned)((words + 255) / 256), cnt < 65535 ? cnt : 65535);
        if (ts == 4) repack_rows<uint32_t><<<grid, 256, 0, c->stream>>>((uint32_t*)drow, (uint32_t)stride, (const uint32_t*)c->stage_dev.p, (uint32_t)words, cnt, (uint32_t)words);
        else repack_rows<uint64_t><<<grid, 256, 0, c->stream>>>((uint64_t*)drow, (uint32_t)stride, (const uint64_t*)c->stage_dev.p, (uint32_t)words, cnt, (uint32_t)words);
        CU(c, cudaGetLastError());
        c->stats.kernel_launches++;
    } else {
        CU(c, cudaMemcpy2DAsync(drow, stride * ts, host, words * ts, words * ts, cnt, cudaMemcpyHostToDevice, c->stream));
    }
    CU(c, cudaStreamSynchronize(c->stream));
    return HB_OK;
}
hb_status hb_wires_read_rows(hb_wires* w, uint32_t first, uint32_t cnt, void* host) {
    if (!w || (!host && cnt)) return HB_ERR_ARG;
    if (!cnt) return HB_OK;
    if (!w->subs.empty()) {
        if (hb_status s = hb_sync(w->ctx)) return s;
        return lift(w->ctx, w->subs[0]->ctx, hb_wires_read_rows(w->subs[0], first, cnt, host));
    }
    hb_ctx* c = w->ctx;
    if ((uint64_t)first + cnt > w->n_wires) return fail(c, HB_ERR_ARG, "wire row range out of bounds");
    CU(c, cudaSetDevice(c->device));
    const size_t ts = tsize(c), words = ct_words(c), stride = hb_wire_stride(c);
    const char* drow = (const char*)w->dev + (size_t)first * stride * ts;
    if ((size_t)cnt * words * ts >= BIG_COPY) {
        CU(c, c->stage_dev.ensure((size_t)cnt * words * ts));
        const dim3 grid((unsigned)((words + 255) / 256), cnt < 65535 ? cnt : 65535);
        if (ts == 4) repack_rows<uint32_t><<<grid, 256, 0, c->stream>>>((uint32_t*)c->stage_dev.p, (uint32_t)words, (const uint32_t*)drow, (uint32_t)stride, cnt, (uint32_t)words);
        else repack_rows<uint64_t><<<grid, 256, 0, c->stream>>>((uint64_t*)c->stage_dev.p, (uint32_t)words, (const uint64_t*)drow, (uint32_t)stride, cnt, (uint32_t)words);
        CU(c, cudaGetLastError());
        c->stats.kernel_launches++;
        CU(c, cudaMemcpyAsync(host, c->stage_dev.p, (size_t)cnt * words * ts, cudaMemcpyDeviceToHost, c->stream));
    } else {
        CU(c, cudaMemcpy2DAsync(host, words * ts, drow, stride * ts, words * ts, cnt, cudaMemcpyDeviceToHost, c->stream));
    }
    CU(c, cudaStreamSynchronize(c->stream));
    return HB_OK;
}

hb_status hb_wires_set_trivial(hb_wires* w, const uint32_t* idx, uint32_t cnt, const uint64_t* pt) {
    if (!w || !idx || !pt) return HB_ERR_ARG;
    if (!w->subs.empty()) {
        for (hb_wires* m : w->subs) if (hb_status s = lift(w->ctx, m->ctx, hb_wires_set_trivial(m, idx, cnt, pt))) return s;
        return hb_sync(w->ctx);                               // a replica must not receive a peer's pushes before its own initialisation ran
    }
    hb_ctx* c = w->ctx;
    if (hb_status s = check_idx(w, idx, cnt)) return s;
    CU(c, cudaSetDevice(c->device));
    std::vector<LinJob> lj(cnt);
    for (uint32_t i = 0; i < cnt; ++i) lj[i] = LinJob{idx[i], idx[i], 0, 0, pt[i]};
    const PbsJob* dp; const KsJob* dk; const LinJob* dl; JobSlot* slot;
    if (hb_status s = upload_jobs(c, {}, {}, lj, &dp, &dk, &dl, &slot)) return s;
    if (hb_status s = launch_lin(c, w->dev, hb_wire_stride(c), dl, cnt, ct_words(c))) return s;
    CU(c, cudaEventRecord(slot->done, c->stream));
    return HB_OK;
}

hb_status hb_sync(hb_ctx* c) {
    if (!c) return HB_ERR_ARG;
    if (c->is_group()) {
        for (hb_ctx* m : c->subs) if (hb_status s = lift(c, m, hb_sync(m))) return s;
        return HB_OK;
    }
    if (c->capturing) return HB_OK;                           // between hb_graph_begin and hb_graph_end nothing runs: nothing to wait for
    CU(c, cudaSetDevice(c->device));
    CU(c, cudaStreamSynchronize(c->stream));
    return HB_OK;
}

#ifdef HB_PHASE_CLOCKS
// debug builds only (see pbs_kernel.cuh): cycles per phase of CTA 0 in the last pbs_kernel launch
hb_status hb_debug_phase_clocks(hb_ctx* c, double* out8) {
    unsigned long long v[16];
    CU(c, cudaDeviceSynchronize());
    CU(c, cudaMemcpyFromSymbol(v, g_phase_clk, sizeof v));
    for (int q = 0; q < 16; ++q) out8[q] = (double)v[q];
    return HB_OK;
}
#endif

hb_status hb_stats(hb_ctx* c, hb_stats_t* out) {
    if (!c || !out) return HB_ERR_ARG;
    if (c->is_group()) return group_stats(c, out);
    CU(c, cudaSetDevice(c->device));
    CtxExtra* ex = extra_of(c);
    // drain in submission order so pbs_ms / ks_ms end up describing the latest level
    for (uint64_t q = 0; q < NSLOT; ++q) {
        JobSlot& s = ex->slots[(ex->next + q) % NSLOT];
        if (!s.pending) continue;
        CU(c, cudaEventSynchronize(s.done));
        fold_timing(c, s, q + 1 < NSLOT ? &ex->slots[(ex->next + q + 1) % NSLOT] : nullptr);
    }
    c->stats.n_devices = 1;
    *out = c->stats;
    return HB_OK;
}

}  // extern "C"

// ================================================================ level evaluation
namespace {

// boolean linear parts: in = c*(a+b) + cst
struct BoolLin { int32_t c; int64_t cst_eighths; };
inline BoolLin bool_lin(int op) {
    switch (op) {
        case HB_AND: return {1, -1};
        case HB_NAND: return {-1, 1};
        case HB_OR: return {1, 1};
        case HB_NOR: return {-1, -1};
        case HB_XOR: return {2, 2};
        case HB_XNOR: return {-2, -2};
    }
    return {0, 0};
}
// boolean op -> 2-input truth table with index = a*2 + b (for the shortint order)
inline uint64_t bool_as_lut(int op) {
    switch (op) {
        case HB_AND: return 0x8; case HB_NAND: return 0x7; case HB_OR: return 0xE;
        case HB_NOR: return 0x1; case HB_XOR: return 0x6; case HB_XNOR: return 0x9;
    }
    return 0;
}

}  // namespace

extern "C" {

hb_status hb_eval_level(hb_ctx* c, hb_wires* w, const hb_gate* gates, uint32_t n_gates) {
    if (c && c->is_group()) return group_eval_level(c, w, gates, n_gates);
    if (hb_status s = check_ready(c)) return s;
    if (!w || w->ctx != c || (!gates && n_gates)) return fail(c, HB_ERR_ARG, "hb_eval_level: bad arguments");
    if (!n_gates) return HB_OK;
    CU(c, cudaSetDevice(c->device));
    const hb_params& p = c->p;
    const int bits = p.torus_bits;
    const uint64_t tmask = bits == 64 ? ~0ull : 0xffffffffull;
    const uint64_t eighth = 1ull << (bits - 3);
    std::vector<PbsJob> pj; std::vector<KsJob> kj; std::vector<LinJob> lj;
    pj.reserve(n_gates + 16); kj.reserve(n_gates);
    auto chk = [&](uint32_t wi) { return wi < w->n_wires; };

    if (p.order == 0) {
        for (uint32_t g = 0; g < n_gates; ++g) {
            const hb_gate& G = gates[g];
            if (!chk(G.out)) return fail(c, HB_ERR_ARG, "gate output wire out of range");
            switch (G.op) {
                case HB_NOT: case HB_COPY:
                    if (!chk(G.in[0])) return fail(c, HB_ERR_ARG, "gate input wire out of range");
                    lj.push_back(LinJob{G.in[0], G.out, G.op == HB_NOT ? -1 : 1, 0, 0});
                    break;
                case HB_AND: case HB_NAND: case HB_OR: case HB_NOR: case HB_XOR: case HB_XNOR: {
                    if (!chk(G.in[0]) || !chk(G.in[1])) return fail(c, HB_ERR_ARG, "gate input wire out of range");
                    const BoolLin bl = bool_lin(G.op);
                    PbsJob j{}; j.src0 = G.in[0]; j.src1 = G.in[1];
                    j.c0 = (G.flags & 1) ? -bl.c : bl.c; j.c1 = (G.flags & 2) ? -bl.c : bl.c;     // flags: input taken negated (folded NOT)
                    j.cst = (uint64_t)(bl.cst_eighths * (int64_t)eighth) & tmask;
                    j.dst = (uint32_t)pj.size(); j.tv_kind = 0; j.tv_cst = eighth;
                    KsJob k{}; k.src[0] = j.dst; k.c[0] = 1; k.dst = G.out;
                    pj.push_back(j); kj.push_back(k);
                    break;
                }
                case HB_MUX: {   // then=in[0], else=in[1], select=in[2]  (reference src/gates.rs:265)
                    if (!chk(G.in[0]) || !chk(G.in[1]) || !chk(G.in[2])) return fail(c, HB_ERR_ARG, "gate input wire out of range");
                    const int st = (G.flags & 1) ? -1 : 1, se = (G.flags & 2) ? -1 : 1, ss = (G.flags & 4) ? -1 : 1;   // then / else / select negated
                    PbsJob a{}; a.src0 = G.in[2]; a.src1 = G.in[0]; a.c0 = ss; a.c1 = st;
                    a.cst = (0 - eighth) & tmask; a.dst = (uint32_t)pj.size(); a.tv_kind = 0; a.tv_cst = eighth;
                    pj.push_back(a);
                    PbsJob b{}; b.src0 = G.in[1]; b.src1 = G.in[2]; b.c0 = se; b.c1 = -ss;
                    b.cst = (0 - eighth) & tmask; b.dst = (uint32_t)pj.size(); b.tv_kind = 0; b.tv_cst = eighth;
                    pj.push_back(b);
                    KsJob k{}; k.src[0] = a.dst; k.src[1] = b.dst; k.c[0] = 1; k.c[1] = 1; k.cst = eighth; k.dst = G.out;
                    kj.push_back(k);
                    break;
                }
                default:
                    return fail(c, HB_ERR_UNSUPPORTED, "gate op not supported with order 0 (boolean) parameters");
            }
        }
    } else {
        const int mc = p.msg_mod * p.carry_mod;
        int lg = 0; while ((1 << lg) < mc) ++lg;
        const uint64_t delta = 1ull << (bits - 1 - lg);
        for (uint32_t g = 0; g < n_gates; ++g) {
            hb_gate G = gates[g];
            if (!chk(G.out)) return fail(c, HB_ERR_ARG, "gate output wire out of range");
            if (G.flags) return fail(c, HB_ERR_UNSUPPORTED, "hb_gate.flags (negated inputs) need boolean (order 0) parameters");
            if (G.op == HB_COPY) {
                if (!chk(G.in[0])) return fail(c, HB_ERR_ARG, "gate input wire out of range");
                lj.push_back(LinJob{G.in[0], G.out, 1, 0, 0});
                continue;
            }
            if (G.op <= HB_XNOR) { G.lut = bool_as_lut(G.op); G.arity = 2; G.op = HB_LUT; }
            else if (G.op == HB_NOT) { G.lut = 0x1; G.arity = 1; G.op = HB_LUT; }
            else if (G.op == HB_MUX) {   // index = then*4 + else*2 + sel : sel ? then : else
                uint64_t t = 0; for (int x = 0; x < 8; ++x) if ((x & 1) ? (x >> 2) & 1 : (x >> 1) & 1) t |= 1ull << x;
                G.lut = t; G.arity = 3; G.op = HB_LUT;
            }
            if (G.op != HB_LUT) return fail(c, HB_ERR_UNSUPPORTED, "gate op not supported with order 1 (shortint) parameters");
            if (G.arity < 1 || G.arity > 4 || (1 << G.arity) > mc)
                return fail(c, HB_ERR_UNSUPPORTED, "LUT arity needs 2^arity <= msg_mod*carry_mod");
            KsJob k{}; 
            for (int s = 0; s < G.arity; ++s) {
                if (!chk(G.in[s])) return fail(c, HB_ERR_ARG, "gate input wire out of range");
                k.src[s] = G.in[s]; k.c[s] = 1 << (G.arity - 1 - s);
            }
            k.dst = (uint32_t)kj.size();
            PbsJob j{}; j.src0 = k.dst; j.c0 = 1; j.c1 = 0; j.cst = 0; j.dst = G.out;
            j.tv_kind = 1; j.tv_cst = delta; j.lut_boxes = mc;
            uint64_t packed = 0;
            for (int x = 0; x < mc && x < (1 << G.arity); ++x) packed |= ((G.lut >> x) & 1ull) << (4 * x);
            j.lut = packed;
            kj.push_back(k); pj.push_back(j);
        }
    }
    const PbsJob* dp; const KsJob* dk; const LinJob* dl; JobSlot* slot;
    if (hb_status s = upload_jobs(c, pj, kj, lj, &dp, &dk, &dl, &slot)) return s;
    const uint32_t wstride = hb_wire_stride(c);
    const size_t ts = tsize(c);
    CU(c, cudaEventRecord(slot->t[0], c->stream));
    if (p.order == 0) {
        const uint32_t bstride = pad4(big_words(c));
        if (hb_status s = ensure_scratch(c, c->scratch_a, (size_t)pj.size() * bstride * ts)) return s;
        if (hb_status s = launch_pbs(c, w->dev, wstride, c->scratch_a.p, bstride, dp, (uint32_t)pj.size())) return s;
        CU(c, cudaEventRecord(slot->t[1], c->stream));
        if (hb_status s = launch_ks(c, c->scratch_a.p, bstride, w->dev, wstride, dk, (uint32_t)kj.size())) return s;
    } else {
        const uint32_t sstride = pad4(small_words(c));
        if (hb_status s = ensure_scratch(c, c->scratch_b, (size_t)kj.size() * sstride * ts)) return s;
        if (hb_status s = launch_ks(c, w->dev, wstride, c->scratch_b.p, sstride, dk, (uint32_t)kj.size())) return s;
        CU(c, cudaEventRecord(slot->t[1], c->stream));
        if (hb_status s = launch_pbs(c, c->scratch_b.p, sstride, w->dev, wstride, dp, (uint32_t)pj.size())) return s;
    }
    CU(c, cudaEventRecord(slot->t[2], c->stream));
    slot->first_is_pbs = p.order == 0; slot->pending = true;
    if (hb_status s = launch_lin(c, w->dev, wstride, dl, (uint32_t)lj.size(), ct_words(c), lin_hazard(lj))) return s;
    if (w->peers.n_dst) {                                      // member of a multi-device group: exchange this share's output rows
        if (p.order == 0) { if (hb_status s = launch_push(c, w, dk, (uint32_t)kj.size())) return s; }
        else if (hb_status s = launch_push(c, w, dp, (uint32_t)pj.size())) return s;
        if (hb_status s = launch_push(c, w, dl, (uint32_t)lj.size())) return s;
    }
    CU(c, cudaEventRecord(slot->done, c->stream));
    c->stats.levels++;
    return HB_OK;
}

hb_status hb_eval_shortint_ops(hb_ctx* c, hb_wires* w, const hb_sop* ops, uint32_t n_ops) {
    if (c && c->is_group()) return group_eval_sops(c, w, ops, n_ops);
    if (hb_status s = check_ready(c)) return s;
    if (!w || w->ctx != c || (!ops && n_ops)) return fail(c, HB_ERR_ARG, "hb_eval_shortint_ops: bad arguments");
    if (c->p.order != 1) return fail(c, HB_ERR_UNSUPPORTED, "hb_eval_shortint_ops needs a shortint (order 1) parameter set");
    if (!n_ops) return HB_OK;
    CU(c, cudaSetDevice(c->device));
    const hb_params& p = c->p;
    const int mc = p.msg_mod * p.carry_mod;
    int lg = 0; while ((1 << lg) < mc) ++lg;
    const uint64_t delta = 1ull << (p.torus_bits - 1 - lg);
    std::vector<PbsJob> pj; std::vector<KsJob> kj, lin;
    for (uint32_t i = 0; i < n_ops; ++i) {
        const hb_sop& o = ops[i];
        if (o.n_in > 4 || o.out >= w->n_wires) return fail(c, HB_ERR_ARG, "shortint op: bad arity or output wire");
        KsJob k{};
        for (int s = 0; s < o.n_in; ++s) {
            if (o.in[s] >= w->n_wires) return fail(c, HB_ERR_ARG, "shortint op: input wire out of range");
            k.src[s] = o.in[s]; k.c[s] = o.coef[s];
        }
        k.cst = (uint64_t)((int64_t)o.cst) * delta;
        if (o.kind == HB_SOP_LINEAR) { k.dst = o.out; lin.push_back(k); }
        else if (o.kind == HB_SOP_LUT) {
            k.dst = (uint32_t)kj.size();
            PbsJob j{}; j.src0 = k.dst; j.c0 = 1; j.dst = o.out; j.tv_kind = 1; j.tv_cst = delta; j.lut_boxes = mc; j.lut = o.table;
            kj.push_back(k); pj.push_back(j);
        } else return fail(c, HB_ERR_ARG, "shortint op: unknown kind");
    }
    const uint32_t wstride = hb_wire_stride(c);
    const size_t ts = tsize(c);
    const PbsJob* dp; const KsJob* dk; const LinJob* dl; JobSlot* slot;
    if (!pj.empty()) {
        if (hb_status s = upload_jobs(c, pj, kj, {}, &dp, &dk, &dl, &slot)) return s;
        const uint32_t sstride = pad4(small_words(c));
        if (hb_status s = ensure_scratch(c, c->scratch_b, (size_t)kj.size() * sstride * ts)) return s;
        CU(c, cudaEventRecord(slot->t[0], c->stream));
        if (hb_status s = launch_ks(c, w->dev, wstride, c->scratch_b.p, sstride, dk, (uint32_t)kj.size())) return s;
        CU(c, cudaEventRecord(slot->t[1], c->stream));
        if (hb_status s = launch_pbs(c, c->scratch_b.p, sstride, w->dev, wstride, dp, (uint32_t)pj.size())) return s;
        CU(c, cudaEventRecord(slot->t[2], c->stream));
        slot->first_is_pbs = 0; slot->pending = true;
        if (hb_status s = launch_push(c, w, dp, (uint32_t)pj.size())) return s;
        CU(c, cudaEventRecord(slot->done, c->stream));
    }
    if (!lin.empty()) {
        if (hb_status s = upload_jobs(c, {}, lin, {}, &dp, &dk, &dl, &slot)) return s;
        if (hb_status s = launch_lincomb(c, w->dev, wstride, dk, (uint32_t)lin.size(), ct_words(c), lincomb_hazard(lin))) return s;
        if (hb_status s = launch_push(c, w, dk, (uint32_t)lin.size())) return s;
        CU(c, cudaEventRecord(slot->done, c->stream));
    }
    c->stats.levels++;
    return HB_OK;
}

hb_status hb_pbs_batch(hb_ctx* c, const void* lwe_in, const void* tv, int32_t shared_tv, uint32_t count, void* lwe_out) {
    if (c && c->is_group()) return lift(c, c->subs[0], hb_pbs_batch(c->subs[0], lwe_in, tv, shared_tv, count, lwe_out));
    if (hb_status s = check_ready(c)) return s;
    if (!lwe_in || !tv || !lwe_out) return fail(c, HB_ERR_ARG, "hb_pbs_batch: null buffer");
    if (!count) return HB_OK;
    CU(c, cudaSetDevice(c->device));
    const size_t ts = tsize(c);
    const uint32_t sw = small_words(c), bw = big_words(c), N = c->p.N;
    const size_t in_b = (size_t)count * sw * ts, tv_b = (size_t)(shared_tv ? 1 : count) * N * ts, out_b = (size_t)count * bw * ts;
    CU(c, c->scratch_a.ensure(in_b + 256 + tv_b));
    CU(c, c->scratch_b.ensure(out_b));
    char* tv_dev = (char*)c->scratch_a.p + ((in_b + 255) & ~(size_t)255);
    CU(c, cudaMemcpyAsync(c->scratch_a.p, lwe_in, in_b, cudaMemcpyHostToDevice, c->stream));
    CU(c, cudaMemcpyAsync(tv_dev, tv, tv_b, cudaMemcpyHostToDevice, c->stream));
    std::vector<PbsJob> pj(count);
    for (uint32_t i = 0; i < count; ++i) {
        PbsJob j{}; j.src0 = i; j.c0 = 1; j.dst = i; j.tv_kind = 2;
        j.tv_ptr = tv_dev + (shared_tv ? 0 : (size_t)i * N * ts);
        pj[i] = j;
    }
    const PbsJob* dp; const KsJob* dk; const LinJob* dl; JobSlot* slot;
    if (hb_status s = upload_jobs(c, pj, {}, {}, &dp, &dk, &dl, &slot)) return s;
    CU(c, cudaEventRecord(slot->t[0], c->stream));
    if (hb_status s = launch_pbs(c, c->scratch_a.p, sw, c->scratch_b.p, bw, dp, count)) return s;
    CU(c, cudaEventRecord(slot->t[1], c->stream));
    CU(c, cudaEventRecord(slot->t[2], c->stream));
    slot->first_is_pbs = 1; slot->pending = true;
    CU(c, cudaEventRecord(slot->done, c->stream));
    CU(c, cudaMemcpyAsync(lwe_out, c->scratch_b.p, out_b, cudaMemcpyDeviceToHost, c->stream));
    CU(c, cudaStreamSynchronize(c->stream));
    c->stats.levels++;
    return HB_OK;
}

hb_status hb_keyswitch_batch(hb_ctx* c, const void* lwe_in, uint32_t count, void* lwe_out) {
    if (c && c->is_group()) return lift(c, c->subs[0], hb_keyswitch_batch(c->subs[0], lwe_in, count, lwe_out));
    if (hb_status s = check_ready(c)) return s;
    if (!lwe_in || !lwe_out) return fail(c, HB_ERR_ARG, "hb_keyswitch_batch: null buffer");
    if (!count) return HB_OK;
    CU(c, cudaSetDevice(c->device));
    const size_t ts = tsize(c);
    const uint32_t sw = small_words(c), bw = big_words(c);
    CU(c, c->scratch_a.ensure((size_t)count * bw * ts));
    CU(c, c->scratch_b.ensure((size_t)count * sw * ts));
    CU(c, cudaMemcpyAsync(c->scratch_a.p, lwe_in, (size_t)count * bw * ts, cudaMemcpyHostToDevice, c->stream));
    std::vector<KsJob> kj(count);
    for (uint32_t i = 0; i < count; ++i) { KsJob k{}; k.src[0] = i; k.c[0] = 1; k.dst = i; kj[i] = k; }
    const PbsJob* dp; const KsJob* dk; const LinJob* dl; JobSlot* slot;
    if (hb_status s = upload_jobs(c, {}, kj, {}, &dp, &dk, &dl, &slot)) return s;
    if (hb_status s = launch_ks(c, c->scratch_a.p, bw, c->scratch_b.p, sw, dk, count)) return s;
    CU(c, cudaEventRecord(slot->done, c->stream));
    CU(c, cudaMemcpyAsync(lwe_out, c->scratch_b.p, (size_t)count * sw * ts, cudaMemcpyDeviceToHost, c->stream));
    CU(c, cudaStreamSynchronize(c->stream));
    return HB_OK;
}

hb_status hb_measure_fp64_peak(hb_ctx* c, double* tflops) {
    if (!c || !tflops) return HB_ERR_ARG;
    if (c->is_group()) return lift(c, c->subs[0], hb_measure_fp64_peak(c->subs[0], tflops));
    CU(c, cudaSetDevice(c->device));
    const int blocks = c->sm_count * 4, threads = 512, iters = 1 << 15;
    CU(c, c->stage_dev.ensure((size_t)blocks * threads * sizeof(double)));
    float best = 1e30f;
    for (int rep = 0; rep < 4; ++rep) {
        CU(c, cudaEventRecord(c->ev[3], c->stream));
        dfma_peak_kernel<<<blocks, threads, 0, c->stream>>>((double*)c->stage_dev.p, iters);
        CU(c, cudaGetLastError());
        cudaEvent_t e2 = c->ev[2];
        CU(c, cudaEventRecord(e2, c->stream));
        CU(c, cudaEventSynchronize(e2));
        float ms = 0; CU(c, cudaEventElapsedTime(&ms, c->ev[3], e2));
        if (rep > 0 && ms < best) best = ms;
        c->stats.kernel_launches++;
    }
    *tflops = 2.0 * 8.0 * (double)iters * blocks * threads / (best * 1e-3) / 1e12;
    return HB_OK;
}

}  // extern "C"

// ================================================================ multi-device group
// One host thread drives every device: all launches are asynchronous, so the per-device shares of a level run concurrently.
namespace {

hb_status group_destroy(hb_ctx* g) {
    for (hb_ctx* m : g->subs) hb_ctx_destroy(m);
    delete g;
    return HB_OK;
}

hb_status group_each(hb_ctx* g, hb_status (*f)(hb_ctx*, const void*, size_t), const void* a, size_t b) {
    for (hb_ctx* m : g->subs) if (hb_status s = lift(g, m, f(m, a, b))) return s;
    return HB_OK;
}

hb_status group_wires_create(hb_ctx* g, uint32_t n_wires, hb_wires** out) {
    hb_wires* w = new hb_wires();
    w->ctx = g; w->n_wires = n_wires;
    for (hb_ctx* m : g->subs) {
        hb_wires* r = nullptr;
        if (hb_status s = lift(g, m, hb_wires_create(m, n_wires, &r))) { hb_wires_destroy(w); return s; }
        w->subs.push_back(r);
    }
    for (size_t d = 0; d < w->subs.size(); ++d) {
        PushArgs& P = w->subs[d]->peers;
        P.n_dst = 0;
        for (size_t e = 0; e < w->subs.size(); ++e) if (e != d) P.dst[P.n_dst++] = w->subs[e]->dev;
    }
    if (hb_status s = hb_sync(g)) { hb_wires_destroy(w); return s; }      // every replica zeroed before any peer may push into it
    *out = w;
    return HB_OK;
}

// how many devices share a batch of `n` items: a slice narrower than shard_min cannot make the batch faster
uint32_t group_width(const hb_ctx* g, uint32_t n) {
    const uint32_t D = (uint32_t)g->subs.size(), per = g->shard_min ? g->shard_min : 1;
    uint32_t d = (n + per - 1) / per;
    return d < 1 ? 1 : d > D ? D : d;
}

// after every member enqueued its share (+ pushes): per-device "share done" events, and every stream waits for all of them,
// so no device starts the next level before all replicas hold this level's outputs
hb_status group_level_barrier(hb_ctx* g, uint32_t active) {
    for (uint32_t d = 0; d < active; ++d) {
        hb_ctx* m = g->subs[d];
        CU(g, cudaSetDevice(m->device));
        CU(g, cudaEventRecord(m->level_ev, m->stream));
    }
    for (hb_ctx* m : g->subs) {
        CU(g, cudaSetDevice(m->device));
        for (uint32_t e = 0; e < active; ++e)
            if (g->subs[e] != m) CU(g, cudaStreamWaitEvent(m->stream, g->subs[e]->level_ev, 0));
    }
    return HB_OK;
}

template <class Item, class F> hb_status group_run(hb_ctx* g, hb_wires* w, const Item* items, uint32_t n, F eval) {
    if (hb_status s = check_ready(g)) return s;
    if (!w || w->ctx != g || w->subs.size() != g->subs.size() || (!items && n)) return fail(g, HB_ERR_ARG, "multi-device batch: bad arguments");
    if (!n) return HB_OK;
    const uint32_t active = group_width(g, n), per = (n + active - 1) / active;
    for (uint32_t d = 0; d < active; ++d) {
        const uint32_t lo = d * per, hi = lo + per < n ? lo + per : n;
        if (lo >= hi) continue;
        if (hb_status s = lift(g, g->subs[d], eval(g->subs[d], w->subs[d], items + lo, hi - lo))) return s;
    }
    g->stats.levels++;
    if (active > 1) g->stats.sharded_levels++;
    return group_level_barrier(g, active);
}

hb_status group_eval_level(hb_ctx* g, hb_wires* w, const hb_gate* gates, uint32_t n_gates) {
    return group_run(g, w, gates, n_gates, [](hb_ctx* m, hb_wires* r, const hb_gate* p, uint32_t c) { return hb_eval_level(m, r, p, c); });
}
hb_status group_eval_sops(hb_ctx* g, hb_wires* w, const hb_sop* ops, uint32_t n_ops) {
    if (g->p.order != 1) return fail(g, HB_ERR_UNSUPPORTED, "hb_eval_shortint_ops needs a shortint (order 1) parameter set");
    return group_run(g, w, ops, n_ops, [](hb_ctx* m, hb_wires* r, const hb_sop* p, uint32_t c) { return hb_eval_shortint_ops(m, r, p, c); });
}

hb_status group_stats(hb_ctx* g, hb_stats_t* out) {
    hb_stats_t t{};
    for (hb_ctx* m : g->subs) {
        hb_stats_t s{};
        if (hb_status e = lift(g, m, hb_stats(m, &s))) return e;
        t.pbs_count += s.pbs_count; t.ks_count += s.ks_count; t.kernel_launches += s.kernel_launches; t.push_bytes += s.push_bytes;
        t.pbs_ms = s.pbs_ms > t.pbs_ms ? s.pbs_ms : t.pbs_ms; t.ks_ms = s.ks_ms > t.ks_ms ? s.ks_ms : t.ks_ms;
        t.pbs_ms_total = s.pbs_ms_total > t.pbs_ms_total ? s.pbs_ms_total : t.pbs_ms_total;
        t.ks_ms_total = s.ks_ms_total > t.ks_ms_total ? s.ks_ms_total : t.ks_ms_total;
        t.gap_ms_total = s.gap_ms_total > t.gap_ms_total ? s.gap_ms_total : t.gap_ms_total;
    }
    t.levels = g->stats.levels; t.sharded_levels = g->stats.sharded_levels; t.n_devices = (uint32_t)g->subs.size();
    *out = t;
    return HB_OK;
}

}  // namespace

extern "C" hb_status hb_ctx_create_multi(const int32_t* devices, int32_t n_devices, uint32_t min_gates_per_device, hb_ctx** out) {
    if (!out || !devices || n_devices < 1 || n_devices > HB_MAX_DEV) return fail(nullptr, HB_ERR_ARG, "hb_ctx_create_multi: 1..8 devices");
    *out = nullptr;
    for (int a = 0; a < n_devices; ++a)
        for (int b = a + 1; b < n_devices; ++b)
            if (devices[a] == devices[b]) return fail(nullptr, HB_ERR_ARG, "hb_ctx_create_multi: devices must be distinct");
    hb_ctx* g = new hb_ctx();
    for (int a = 0; a < n_devices; ++a) {
        hb_ctx* m = nullptr;
        if (hb_status s = hb_ctx_create(devices[a], &m)) { group_destroy(g); return s; }
        g->subs.push_back(m);
        if (cudaEventCreateWithFlags(&m->level_ev, cudaEventDisableTiming) != cudaSuccess) { group_destroy(g); return fail(nullptr, HB_ERR_CUDA, "cudaEventCreate"); }
    }
    // every pair must be peer-mapped: the level exchange is device-initiated stores into the peers' replicas
    for (int a = 0; a < n_devices; ++a)
        for (int b = 0; b < n_devices; ++b) {
            if (a == b) continue;
            int can = 0;
            cudaDeviceCanAccessPeer(&can, devices[a], devices[b]);
            if (!can) { group_destroy(g); return fail(nullptr, HB_ERR_UNSUPPORTED, "hb_ctx_create_multi: devices are not peer-accessible"); }
            cudaSetDevice(devices[a]);
            const cudaError_t e = cudaDeviceEnablePeerAccess(devices[b], 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) { group_destroy(g); return fail(nullptr, HB_ERR_CUDA, std::string("cudaDeviceEnablePeerAccess: ") + cudaGetErrorString(e)); }
            (void)cudaGetLastError();
        }
    g->device = devices[0];
    g->sm_count = g->subs[0]->sm_count;
    g->shard_min = min_gates_per_device ? min_gates_per_device : (uint32_t)g->sm_count;
    *out = g;
    return HB_OK;
}

// ================================================================ keyswitch without any allocation
// Added; it has no counterpart in the reference.  For callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it): the tensor-core keyswitch of ks_mma_kernel.cuh
// needs two scratch arrays -- the identity job descriptors and the int8 digit matrix [rows][kN * ks_level] -- and here both live in
// ONE file-scope device array of fixed size instead of the context's growable buffers.  Every GPU that loads the library holds
// HB_NOALLOC_SCRATCH_BYTES until exit; every call (of every context on that device) shares it, so calls that overlap on the GPU
// race; it keeps the previous call's contents, and each call rewrites every job and digit row it reads (rows up to the padded
// tile count are zeroed by the digit kernel), so nothing needs to start at zero.
namespace {
constexpr size_t NOALLOC_JOB_BYTES = 64 * 1024;                       // 1365 KsJob (48 B): bounds the row count together with the digits
static_assert(HB_NOALLOC_SCRATCH_BYTES > NOALLOC_JOB_BYTES, "scratch layout");
static __device__ __align__(256) unsigned char g_noalloc_scratch[HB_NOALLOC_SCRATCH_BYTES];

// identity jobs (row i -> row i) written straight into the static array: device code may name it directly
__global__ void noalloc_jobs_kernel(uint32_t rows) {
    KsJob* jobs = reinterpret_cast<KsJob*>(g_noalloc_scratch);
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < rows; i += gridDim.x * blockDim.x) {
        KsJob k = {{i, 0, 0, 0}, {1, 0, 0, 0}, 0, i, 0};
        jobs[i] = k;
    }
}

// In host code `g_noalloc_scratch` is only a stand-in for the symbol: its device address is asked for once per device, with that
// device current, and cached.
std::atomic<unsigned char*> g_noalloc_ptr[64];
cudaError_t noalloc_scratch_ptr(int device, unsigned char** out) {
    if (device < 0 || device >= 64) return cudaErrorInvalidDevice;
    unsigned char* p = g_noalloc_ptr[device].load(std::memory_order_acquire);
    if (!p) {
        void* q = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&q, g_noalloc_scratch);
        if (e != cudaSuccess) return e;
        p = static_cast<unsigned char*>(q);
        g_noalloc_ptr[device].store(p, std::memory_order_release);
    }
    *out = p;
    return cudaSuccess;
}

template <class T> hb_status ks_noalloc_t(hb_ctx* c, const void* in_dev, uint32_t count, void* out_dev, unsigned char* scratch) {
    using Cfg = KmCfg<T>;
    KsMmaArgs m;
    KsArgs& a = m.a;
    a.in = in_dev; a.in_stride = big_words(c); a.out = out_dev; a.out_stride = small_words(c);
    a.jobs = reinterpret_cast<const KsJob*>(scratch); a.n_jobs = count;
    a.ksk = c->ksk.p; a.n = c->p.n; a.kN = c->p.k * c->p.N; a.base_log = c->p.ks_base_log; a.level = c->p.ks_level; a.split = 1;
    m.planes = (const uint8_t*)c->ksk_planes.p; m.npad = c->ksk_npad;
    uint8_t* dig = scratch + NOALLOC_JOB_BYTES;
    const dim3 grid((count + KM_TM - 1) / KM_TM, c->ksk_npad / Cfg::TN);
    const uint32_t rows_pad = grid.x * KM_TM;
    static SmemOptIn opt;
    const size_t smem = 2 * (size_t)Cfg::STAGE;
    CU(c, opt.ensure(ks_mma2_kernel<T>, smem));
    noalloc_jobs_kernel<<<(count + 255) / 256, 256, 0, c->stream>>>(count);
    ks_digits_kernel<T><<<rows_pad, 256, 0, c->stream>>>(a, dig, rows_pad);
    ks_mma2_kernel<T><<<grid, KM_NT, smem, c->stream>>>(m, dig);
    CU(c, cudaGetLastError());
    c->stats.kernel_launches += 3; c->stats.ks_count += count;
    return HB_OK;
}
}  // namespace

extern "C" hb_status hb_keyswitch_batch_noalloc(hb_ctx* c, const void* lwe_in_dev, uint32_t count, void* lwe_out_dev) {
    if (c && c->is_group()) return lift(c, c->subs[0], hb_keyswitch_batch_noalloc(c->subs[0], lwe_in_dev, count, lwe_out_dev));
    if (hb_status s = check_ready(c)) return s;
    if (!lwe_in_dev || !lwe_out_dev) return fail(c, HB_ERR_ARG, "hb_keyswitch_batch_noalloc: null buffer");
    if (!count) return HB_OK;
    if (!c->ksk_npad) return fail(c, HB_ERR_UNSUPPORTED, "hb_keyswitch_batch_noalloc: this parameter set has no tensor-core keyswitch (the scratch-using kernel)");
    const size_t row_bytes = (size_t)c->p.k * c->p.N * c->p.ks_level;                     // int8 digits per row
    size_t max_rows = (HB_NOALLOC_SCRATCH_BYTES - NOALLOC_JOB_BYTES) / row_bytes / KM_TM * KM_TM;   // whole 64-row tiles
    if (max_rows > NOALLOC_JOB_BYTES / sizeof(KsJob)) max_rows = NOALLOC_JOB_BYTES / sizeof(KsJob) / KM_TM * KM_TM;
    if (count > max_rows)
        return fail(c, HB_ERR_ARG, "hb_keyswitch_batch_noalloc: count " + std::to_string(count) + " exceeds the static scratch (" +
                    std::to_string(HB_NOALLOC_SCRATCH_BYTES) + " bytes): at most " + std::to_string(max_rows) + " rows for this parameter set");
    CU(c, cudaSetDevice(c->device));
    unsigned char* scratch = nullptr;
    CU(c, noalloc_scratch_ptr(c->device, &scratch));
    return c->p.torus_bits == 32 ? ks_noalloc_t<uint32_t>(c, lwe_in_dev, count, lwe_out_dev, scratch)
                                 : ks_noalloc_t<uint64_t>(c, lwe_in_dev, count, lwe_out_dev, scratch);
}

// ================================================================ asynchronous row IO
// Pipelined serving loops: the host -> device copy of batch s+1 and the device -> host copy of batch s-1 run on the context's copy stream
// while batch s is evaluated on its compute stream (the reference stages every level synchronously, src/circuit.rs:745-761, 877-892).
namespace {
hb_status io_setup(hb_ctx* c) {
    if (c->io_stream) return HB_OK;
    // highest priority: the small repack kernels (and whatever the caller puts on the upload stream) must not queue behind the thousands of
    // pending CTAs of a running PBS kernel, or the copy of batch s+1 only completes after batch s and nothing overlaps (measured at N = 8:
    // 73.4 ms per step against 66.8 resident with default priority)
    int prio_lo = 0, prio_hi = 0;
    CU(c, cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
    CU(c, cudaStreamCreateWithPriority(&c->io_stream, cudaStreamNonBlocking, prio_hi));
    CU(c, cudaStreamCreateWithPriority(&c->io_down, cudaStreamNonBlocking, prio_hi));
    CU(c, cudaEventCreateWithFlags(&c->io_written, cudaEventDisableTiming));
    CU(c, cudaEventCreateWithFlags(&c->io_computed, cudaEventDisableTiming));
    return HB_OK;
}
template <class T> void io_repack(T* dst, uint32_t dstride, const T* src, uint32_t sstride, uint32_t cnt, uint32_t words, cudaStream_t st) {
    const dim3 grid((words + 255) / 256, cnt < 65535 ? cnt : 65535);
    repack_rows<T><<<grid, 256, 0, st>>>(dst, dstride, src, sstride, cnt, words);
}
}  // namespace

extern "C" hb_status hb_wires_write_rows_async(hb_wires* w, uint32_t first, uint32_t cnt, const void* host) {
    if (!w || (!host && cnt)) return HB_ERR_ARG;
    hb_ctx* c = w->ctx;
    if (c->is_group()) return fail(c, HB_ERR_UNSUPPORTED, "asynchronous row IO: single-device contexts only");
    if (!cnt) return HB_OK;
    if ((uint64_t)first + cnt > w->n_wires) return fail(c, HB_ERR_ARG, "wire row range out of bounds");
    CU(c, cudaSetDevice(c->device));
    if (hb_status s = io_setup(c)) return s;
    const size_t ts = tsize(c), words = ct_words(c), stride = hb_wire_stride(c), bytes = (size_t)cnt * words * ts;
    if (c->io_in.cap < bytes) { CU(c, cudaStreamSynchronize(c->io_stream)); CU(c, c->io_in.ensure(bytes)); }
    char* drow = (char*)w->dev + (size_t)first * stride * ts;
    CU(c, cudaMemcpyAsync(c->io_in.p, host, bytes, cudaMemcpyHostToDevice, c->io_stream));
    if (ts == 4) io_repack<uint32_t>((uint32_t*)drow, (uint32_t)stride, (const uint32_t*)c->io_in.p, (uint32_t)words, cnt, (uint32_t)words, c->io_stream);
    else io_repack<uint64_t>((uint64_t*)drow, (uint32_t)stride, (const uint64_t*)c->io_in.p, (uint32_t)words, cnt, (uint32_t)words, c->io_stream);
    CU(c, cudaGetLastError());
    c->stats.kernel_launches++;
    CU(c, cudaEventRecord(c->io_written, c->io_stream));
    CU(c, cudaStreamWaitEvent(c->stream, c->io_written, 0));           // everything enqueued on the compute stream from here on sees the rows
    return HB_OK;
}

extern "C" hb_status hb_wires_read_rows_async(hb_wires* w, uint32_t first, uint32_t cnt, void* host) {
    if (!w || (!host && cnt)) return HB_ERR_ARG;
    hb_ctx* c = w->ctx;
    if (c->is_group()) return fail(c, HB_ERR_UNSUPPORTED, "asynchronous row IO: single-device contexts only");
    if (!cnt) return HB_OK;
    if ((uint64_t)first + cnt > w->n_wires) return fail(c, HB_ERR_ARG, "wire row range out of bounds");
    CU(c, cudaSetDevice(c->device));
    if (hb_status s = io_setup(c)) return s;
    const size_t ts = tsize(c), words = ct_words(c), stride = hb_wire_stride(c), bytes = (size_t)cnt * words * ts;
    if (c->io_out.cap < bytes) { CU(c, cudaStreamSynchronize(c->io_down)); CU(c, c->io_out.ensure(bytes)); }
    const char* drow = (const char*)w->dev + (size_t)first * stride * ts;
    CU(c, cudaEventRecord(c->io_computed, c->stream));                   // the rows as of the work enqueued so far
    CU(c, cudaStreamWaitEvent(c->io_down, c->io_computed, 0));
    if (ts == 4) io_repack<uint32_t>((uint32_t*)c->io_out.p, (uint32_t)words, (const uint32_t*)drow, (uint32_t)stride, cnt, (uint32_t)words, c->io_down);
    else io_repack<uint64_t>((uint64_t*)c->io_out.p, (uint32_t)words, (const uint64_t*)drow, (uint32_t)stride, cnt, (uint32_t)words, c->io_down);
    CU(c, cudaGetLastError());
    c->stats.kernel_launches++;
    CU(c, cudaMemcpyAsync(host, c->io_out.p, bytes, cudaMemcpyDeviceToHost, c->io_down));
    return HB_OK;
}

// the upload stream (cudaStream_t as void*): a caller that has to post-process uploaded rows before they are evaluated (the all-gather of
// the ranks' input shares in a one-process-per-GPU deployment) enqueues that work here, then calls hb_io_fence
extern "C" void* hb_io_stream(hb_ctx* c) {
    if (!c || c->is_group()) return nullptr;
    if (cudaSetDevice(c->device) != cudaSuccess || io_setup(c) != HB_OK) return nullptr;
    return (void*)c->io_stream;
}
// everything enqueued on the upload stream so far happens before every hb_eval_* call issued after this
extern "C" hb_status hb_io_fence(hb_ctx* c) {
    if (!c) return HB_ERR_ARG;
    if (c->is_group()) return fail(c, HB_ERR_UNSUPPORTED, "asynchronous row IO: single-device contexts only");
    CU(c, cudaSetDevice(c->device));
    if (hb_status s = io_setup(c)) return s;
    CU(c, cudaEventRecord(c->io_written, c->io_stream));
    CU(c, cudaStreamWaitEvent(c->stream, c->io_written, 0));
    return HB_OK;
}

extern "C" hb_status hb_io_sync(hb_ctx* c) {
    if (!c) return HB_ERR_ARG;
    if (c->is_group() || !c->io_stream) return HB_OK;
    CU(c, cudaSetDevice(c->device));
    CU(c, cudaStreamSynchronize(c->io_stream));
    CU(c, cudaStreamSynchronize(c->io_down));
    return HB_OK;
}

// ================================================================ CUDA-graph capture of a batch sequence
// A cycle of a sequential circuit issues the same batches every time: hb_graph_begin .. hb_graph_end records the hb_eval_level /
// hb_eval_shortint_ops calls made in between (nothing runs) into one CUDA graph -- job-descriptor uploads included -- and
// hb_graph_launch replays it with one launch.  The reference has no counterpart (its cycle loop re-issues every gate,
// src/bin/helm.rs:256-270).
struct hb_graph {
    hb_ctx* ctx = nullptr;
    cudaGraph_t graph = nullptr;
    cudaGraphExec_t exec = nullptr;
    PinBuf pin; DevBuf dev;                // the job arena the graph's copy nodes read / write
    uint64_t scratch_gen = 0;
    hb_stats_t delta{};                    // what one replay adds to the context's counters
};

extern "C" hb_status hb_graph_begin(hb_ctx* c, size_t job_bytes) {
    if (hb_status s = check_ready(c)) return s;
    if (c->is_group()) return fail(c, HB_ERR_UNSUPPORTED, "graph capture: single-device contexts only");
    if (c->capturing) return fail(c, HB_ERR_STATE, "graph capture already in progress");
    CU(c, cudaSetDevice(c->device));
    CU(c, cudaStreamSynchronize(c->stream));
    const size_t want = job_bytes + 4096;
    CU(c, c->cap_pin.ensure(want));
    CU(c, c->cap_dev.ensure(want));
    c->cap_used = 0;
    c->cap_stats0 = c->stats;
    CU(c, cudaStreamBeginCapture(c->stream, cudaStreamCaptureModeRelaxed));
    c->capturing = true;
    return HB_OK;
}

extern "C" hb_status hb_graph_end(hb_ctx* c, hb_graph** out) {
    if (!c || !out) return HB_ERR_ARG;
    if (!c->capturing) return fail(c, HB_ERR_STATE, "hb_graph_end without hb_graph_begin");
    *out = nullptr;
    c->capturing = false;
    cudaGraph_t g = nullptr;
    CU(c, cudaStreamEndCapture(c->stream, &g));
    hb_graph* hg = new hb_graph();
    hg->ctx = c; hg->graph = g;
    cudaError_t e = cudaGraphInstantiate(&hg->exec, g, 0);
    if (e != cudaSuccess) { cudaGraphDestroy(g); delete hg; return fail(c, HB_ERR_CUDA, std::string("cudaGraphInstantiate: ") + cudaGetErrorString(e)); }
    hg->pin = c->cap_pin; hg->dev = c->cap_dev;                        // the arena now belongs to the graph
    c->cap_pin = PinBuf(); c->cap_dev = DevBuf();
    hg->scratch_gen = c->scratch_gen;
    hg->delta.pbs_count = c->stats.pbs_count - c->cap_stats0.pbs_count;
    hg->delta.ks_count = c->stats.ks_count - c->cap_stats0.ks_count;
    hg->delta.levels = c->stats.levels - c->cap_stats0.levels;
    hg->delta.kernel_launches = c->stats.kernel_launches - c->cap_stats0.kernel_launches;
    c->stats.pbs_count = c->cap_stats0.pbs_count; c->stats.ks_count = c->cap_stats0.ks_count;      // nothing ran yet
    c->stats.levels = c->cap_stats0.levels; c->stats.kernel_launches = c->cap_stats0.kernel_launches;
    *out = hg;
    return HB_OK;
}

extern "C" hb_status hb_graph_launch(hb_graph* g) {
    if (!g) return HB_ERR_ARG;
    hb_ctx* c = g->ctx;
    if (c->capturing) return fail(c, HB_ERR_STATE, "hb_graph_launch during a capture");
    if (g->scratch_gen != c->scratch_gen)
        return fail(c, HB_ERR_STATE, "graph is stale: a larger batch reallocated the scratch buffers it points into; capture it again");
    CU(c, cudaSetDevice(c->device));
    CU(c, cudaGraphLaunch(g->exec, c->stream));
    c->stats.pbs_count += g->delta.pbs_count; c->stats.ks_count += g->delta.ks_count;
    c->stats.levels += g->delta.levels; c->stats.kernel_launches += g->delta.kernel_launches;
    return HB_OK;
}

extern "C" hb_status hb_graph_destroy(hb_graph* g) {
    if (!g) return HB_ERR_ARG;
    cudaSetDevice(g->ctx->device);
    cudaStreamSynchronize(g->ctx->stream);
    if (g->exec) cudaGraphExecDestroy(g->exec);
    if (g->graph) cudaGraphDestroy(g->graph);
    g->pin.release(); g->dev.release();
    delete g;
    return HB_OK;
}
