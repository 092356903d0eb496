// Micro-benchmarks that decide the PBS kernel design on B200: FP64 FMA pipe, FP64 tensor (DMMA) pipe, whether they
// overlap, shared-memory bandwidth (LDS.128), SHFL rate.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3
#include <cstdio>
#include <cuda_runtime.h>
static __device__ int d_iters = 8192;
#define ITERS 8192
#define DITERS (*(volatile int*)&d_iters)
__global__ void k_dfma(double* out) {
    double a[8]; for (int i = 0; i < 8; ++i) a[i] = threadIdx.x * 1e-9 + i;
    const double b = 1.0000001, c = 1e-7;
    for (int it = 0, ni = DITERS; it < ni; ++it)
#pragma unroll
        for (int i = 0; i < 8; ++i) a[i] = fma(a[i], b, c);
    double s = 0; for (int i = 0; i < 8; ++i) s += a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__device__ __forceinline__ void dmma884(double& d0, double& d1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}
__global__ void k_dmma(double* out) {
    double d[8]; for (int i = 0; i < 8; ++i) d[i] = i;
    double a = threadIdx.x * 1e-9, b = 1.0000001;
    for (int it = 0, ni = DITERS; it < ni; ++it)
#pragma unroll
        for (int i = 0; i < 4; ++i) dmma884(d[2 * i], d[2 * i + 1], a, b);
    double s = 0; for (int i = 0; i < 8; ++i) s += d[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__device__ __forceinline__ void dmma16816(double (&d)[4], const double (&a)[8], const double (&b)[4]) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7,%8,%9,%10,%11}, {%12,%13,%14,%15}, {%0,%1,%2,%3};"
                 : "+d"(d[0]), "+d"(d[1]), "+d"(d[2]), "+d"(d[3])
                 : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(a[4]), "d"(a[5]), "d"(a[6]), "d"(a[7]), "d"(b[0]), "d"(b[1]), "d"(b[2]), "d"(b[3]));
}
__global__ void k_dmma16(double* out) {
    double d0[4] = {0, 1, 2, 3}, d1[4] = {1, 2, 3, 4}, a[8], b[4];
    for (int i = 0; i < 8; ++i) a[i] = threadIdx.x * 1e-9 + i; for (int i = 0; i < 4; ++i) b[i] = 1.0 + i * 1e-7;
    for (int it = 0, ni = DITERS; it < ni; ++it) { dmma16816(d0, a, b); dmma16816(d1, a, b); }
    out[blockIdx.x * blockDim.x + threadIdx.x] = d0[0] + d0[1] + d0[2] + d0[3] + d1[0] + d1[1] + d1[2] + d1[3];
}
__global__ void k_mixed(double* out) {   // 4 DMMA884 + 8 DFMA per iteration
    double d[8], a8[8]; for (int i = 0; i < 8; ++i) { d[i] = i; a8[i] = threadIdx.x * 1e-9 + i; }
    double a = threadIdx.x * 1e-9, b = 1.0000001, c = 1e-7;
    for (int it = 0, ni = DITERS; it < ni; ++it) {
#pragma unroll
        for (int i = 0; i < 4; ++i) { dmma884(d[2 * i], d[2 * i + 1], a, b); a8[2 * i] = fma(a8[2 * i], b, c); a8[2 * i + 1] = fma(a8[2 * i + 1], b, c); }
    }
    double s = 0; for (int i = 0; i < 8; ++i) s += d[i] + a8[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void k_lds128(double* out) {
    __shared__ double2 buf[1024];
    for (int i = threadIdx.x; i < 1024; i += blockDim.x) buf[i] = make_double2(i, -i);
    __syncthreads();
    double2 s = make_double2(0, 0); int idx = threadIdx.x;
    for (int it = 0, ni = DITERS; it < ni; ++it) {
#pragma unroll
        for (int u = 0; u < 8; ++u) { double2 v = buf[(idx + u * 32) & 1023]; s.x += v.x; s.y += v.y; }   // note: DADD too
        idx = (idx + 7) & 1023;
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s.x + s.y;
}
__global__ void k_lds128_i(unsigned* out) {   // integer consumer so the FP64 pipe is not the limiter
    __shared__ uint4 buf[1024];
    for (int i = threadIdx.x; i < 1024; i += blockDim.x) buf[i] = make_uint4(i, i + 1, i + 2, i + 3);
    __syncthreads();
    unsigned s = 0; int idx = threadIdx.x;
    for (int it = 0, ni = DITERS; it < ni; ++it) {
#pragma unroll
        for (int u = 0; u < 8; ++u) { uint4 v = buf[(idx + u * 32) & 1023]; s ^= v.x ^ v.y ^ v.z ^ v.w; }
        idx = (idx + 7) & 1023;
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void k_shfl(unsigned* out) {
    unsigned v[8]; for (int i = 0; i < 8; ++i) v[i] = threadIdx.x + i;
    for (int it = 0, ni = DITERS; it < ni; ++it)
#pragma unroll
        for (int i = 0; i < 8; ++i) v[i] = __shfl_xor_sync(0xffffffffu, v[i], 1 + (i & 3)) + 1;
    unsigned s = 0; for (int i = 0; i < 8; ++i) s += v[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void k_imad(unsigned* out) {
    unsigned v[8]; for (int i = 0; i < 8; ++i) v[i] = threadIdx.x + i;
    unsigned m = out[0] | 3;
    for (int it = 0, ni = DITERS; it < ni; ++it)
#pragma unroll
        for (int i = 0; i < 8; ++i) v[i] = v[i] * m + 12345u;
    unsigned s = 0; for (int i = 0; i < 8; ++i) s += v[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <class F> float timeit(F f) {
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    f(); cudaDeviceSynchronize();
    float best = 1e30f;
    for (int r = 0; r < 3; ++r) { cudaEventRecord(e0); f(); cudaEventRecord(e1); cudaEventSynchronize(e1); float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms; }
    return best;
}
int main() {
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    int clk; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    const int B = sms * 4, T = 512;
    double* o; cudaMalloc(&o, (size_t)B * T * 8);
    const double thr = (double)B * T, ghz = clk * 1e-6;
    printf("SMs %d, clock %.3f GHz (nominal max); per-SM-per-clk figures assume that clock\n", sms, ghz);
    float ms;
    ms = timeit([&] { k_dfma<<<B, T>>>(o); });
    printf("DFMA        : %.2f TFLOP/s  (%.1f DFMA/clk/SM)\n", 2.0 * 8 * ITERS * thr / ms / 1e9, 8.0 * ITERS * thr / (ms * 1e-3) / sms / (ghz * 1e9));
    ms = timeit([&] { k_dmma<<<B, T>>>(o); });
    printf("DMMA m8n8k4 : %.2f TFLOP/s  (%.2f warp-mma/clk/SM)\n", 2.0 * 256 * 4 * ITERS * thr / 32 / ms / 1e9, 4.0 * ITERS * thr / 32 / (ms * 1e-3) / sms / (ghz * 1e9));
    ms = timeit([&] { k_dmma16<<<B, T>>>(o); });
    printf("DMMA m16n8k16: %.2f TFLOP/s\n", 2.0 * 2048 * 2 * ITERS * thr / 32 / ms / 1e9);
    ms = timeit([&] { k_mixed<<<B, T>>>(o); });
    printf("mixed 4 DMMA884 + 8 DFMA / iter: %.3f ms -> DFMA part %.2f TF + DMMA part %.2f TF = %.2f TF total\n", ms,
           2.0 * 8 * ITERS * thr / ms / 1e9, 2.0 * 256 * 4 * ITERS * thr / 32 / ms / 1e9, (2.0 * 8 * ITERS * thr + 2.0 * 256 * 4 * ITERS * thr / 32) / ms / 1e9);
    ms = timeit([&] { k_lds128<<<B, T>>>(o); });
    printf("LDS.128 (+2 DADD each): %.1f B/clk/SM\n", 16.0 * 8 * ITERS * thr / (ms * 1e-3) / sms / (ghz * 1e9));
    ms = timeit([&] { k_lds128_i<<<B, T>>>((unsigned*)o); });
    printf("LDS.128 (int consumer): %.1f B/clk/SM\n", 16.0 * 8 * ITERS * thr / (ms * 1e-3) / sms / (ghz * 1e9));
    ms = timeit([&] { k_shfl<<<B, T>>>((unsigned*)o); });
    printf("SHFL.BFLY   : %.1f lanes/clk/SM\n", 8.0 * ITERS * thr / (ms * 1e-3) / sms / (ghz * 1e9));
    ms = timeit([&] { k_imad<<<B, T>>>((unsigned*)o); });
    printf("IMAD        : %.1f lanes/clk/SM\n", 8.0 * ITERS * thr / (ms * 1e-3) / sms / (ghz * 1e9));
    return 0;
}
