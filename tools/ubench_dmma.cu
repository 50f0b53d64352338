// ubench_dmma.cu -- is mma.sync.m8n8k4.f64 (and the wider sm_90+ f64 shapes) the k-ascending chain of fused multiply-adds
// the profile kernels are defined by, bit for bit?  And how fast is it next to the DFMA pipe?   (development aid)
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o ubench_dmma ubench_dmma.cu && ./ubench_dmma
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <cmath>
#include <cuda_runtime.h>

__device__ __forceinline__ void dmma884(double &d0, double &d1, double a, double b, double c0, double c1)
{
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%4,%5};"
                 : "=d"(d0), "=d"(d1) : "d"(a), "d"(b), "d"(c0), "d"(c1));
}
__device__ __forceinline__ void dmma1688(double (&d)[4], const double (&a)[4], const double (&b)[2], const double (&c)[4])
{
    asm volatile("mma.sync.aligned.m16n8k8.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%10,%11,%12,%13};"
                 : "=d"(d[0]), "=d"(d[1]), "=d"(d[2]), "=d"(d[3])
                 : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(b[0]), "d"(b[1]), "d"(c[0]), "d"(c[1]), "d"(c[2]), "d"(c[3]));
}
__device__ __forceinline__ void dmma1684(double (&d)[4], const double (&a)[2], double b, const double (&c)[4])
{
    asm volatile("mma.sync.aligned.m16n8k4.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%7,%8,%9,%10};"
                 : "=d"(d[0]), "=d"(d[1]), "=d"(d[2]), "=d"(d[3])
                 : "d"(a[0]), "d"(a[1]), "d"(b), "d"(c[0]), "d"(c[1]), "d"(c[2]), "d"(c[3]));
}

// A [M][K] row-major, B [K][8], C [M][8] in global memory; one warp.
__global__ void exact884(const double *A, const double *B, const double *C, double *D, int trials)
{
    const int lane = threadIdx.x, g = lane >> 2, t = lane & 3;
    for (int it = 0; it < trials; ++it) {
        const double *a = A + it * 32, *b = B + it * 32, *c = C + it * 64;
        double d0, d1;
        dmma884(d0, d1, a[g * 4 + t], b[t * 8 + g], c[g * 8 + 2 * t], c[g * 8 + 2 * t + 1]);
        D[it * 64 + g * 8 + 2 * t] = d0; D[it * 64 + g * 8 + 2 * t + 1] = d1;
    }
}
__global__ void exact1688(const double *A, const double *B, const double *C, double *D, int trials)
{
    // A 16x8: a0 (g, t) a1 (g+8, t) a2 (g, t+4) a3 (g+8, t+4); B 8x8: b0 (t, g) b1 (t+4, g); C: c0,c1 (g, 2t..) c2,c3 (g+8, 2t..)
    const int lane = threadIdx.x, g = lane >> 2, t = lane & 3;
    for (int it = 0; it < trials; ++it) {
        const double *a = A + it * 128, *b = B + it * 64, *c = C + it * 128;
        double av[4] = {a[g * 8 + t], a[(g + 8) * 8 + t], a[g * 8 + t + 4], a[(g + 8) * 8 + t + 4]};
        double bv[2] = {b[t * 8 + g], b[(t + 4) * 8 + g]};
        double cv[4] = {c[g * 8 + 2 * t], c[g * 8 + 2 * t + 1], c[(g + 8) * 8 + 2 * t], c[(g + 8) * 8 + 2 * t + 1]};
        double d[4];
        dmma1688(d, av, bv, cv);
        double *o = D + it * 128;
        o[g * 8 + 2 * t] = d[0]; o[g * 8 + 2 * t + 1] = d[1]; o[(g + 8) * 8 + 2 * t] = d[2]; o[(g + 8) * 8 + 2 * t + 1] = d[3];
    }
}

// throughput: NT independent 8x8 accumulator tiles per warp, operands in registers
template <int NT>
__global__ void __launch_bounds__(256) rate884(int iters, double seed, double *sink)
{
    double acc[NT][2];
#pragma unroll
    for (int i = 0; i < NT; ++i) { acc[i][0] = seed + threadIdx.x + i; acc[i][1] = seed - i; }
    double a = 1.0000001 + threadIdx.x * 1e-9, b = 0.9999999;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < 4; ++u)
#pragma unroll
            for (int i = 0; i < NT; ++i) dmma884(acc[i][0], acc[i][1], a, b, acc[i][0], acc[i][1]);
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < NT; ++i) s += acc[i][0] + acc[i][1];
    if (s == 12345.678) sink[0] = s;
}
template <int NT>
__global__ void __launch_bounds__(256) rate1688(int iters, double seed, double *sink)
{
    double acc[NT][4];
#pragma unroll
    for (int i = 0; i < NT; ++i) { acc[i][0] = seed + threadIdx.x + i; acc[i][1] = seed - i; acc[i][2] = seed; acc[i][3] = 1.0; }
    double a[4] = {1.0000001 + threadIdx.x * 1e-9, 1.0, 0.5, 0.25}, b[2] = {0.9999999, 1.0};
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < 4; ++u)
#pragma unroll
            for (int i = 0; i < NT; ++i) dmma1688(acc[i], a, b, acc[i]);
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < NT; ++i) s += acc[i][0] + acc[i][1] + acc[i][2] + acc[i][3];
    if (s == 12345.678) sink[0] = s;
}
__global__ void __launch_bounds__(256) rate_dfma(int iters, double seed, double *sink)
{
    double a0 = seed + threadIdx.x, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6, a7 = a0 + 7;
    const double m = 1.0000001, c = 1e-9;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int u = 0; u < 16; ++u) {
            a0 = __fma_rn(a0, m, c); a1 = __fma_rn(a1, m, c); a2 = __fma_rn(a2, m, c); a3 = __fma_rn(a3, m, c);
            a4 = __fma_rn(a4, m, c); a5 = __fma_rn(a5, m, c); a6 = __fma_rn(a6, m, c); a7 = __fma_rn(a7, m, c);
        }
    }
    const double s = ((a0 + a1) + (a2 + a3)) + ((a4 + a5) + (a6 + a7));
    if (s == 12345.678) sink[0] = s;
}

static double rnd(int mode)
{
    // mode 0: uniform(-1,1); mode 1: wide exponents; mode 2: float32-valued weights x small integers (the kernels' operands)
    double u = (double)rand() / RAND_MAX * 2 - 1;
    if (mode == 0) return u;
    if (mode == 1) return ldexp(u, rand() % 60 - 30);
    return (double)(float)u;
}

template <class F>
static float timeit(F f)
{
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    f(); cudaDeviceSynchronize();
    cudaEventRecord(e0); f(); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1); return ms;
}

int main()
{
    const int T = 20000;
    for (int mode = 0; mode < 3; ++mode) {
        // ---- m8n8k4
        double *A, *B, *C, *D;
        cudaMallocManaged(&A, T * 32 * 8); cudaMallocManaged(&B, T * 32 * 8); cudaMallocManaged(&C, T * 64 * 8); cudaMallocManaged(&D, T * 64 * 8);
        for (int i = 0; i < T * 32; ++i) { A[i] = mode == 2 ? (double)(rand() % 200) : rnd(mode); B[i] = rnd(mode); }
        for (int i = 0; i < T * 64; ++i) C[i] = rnd(mode);
        exact884<<<1, 32>>>(A, B, C, D, T); cudaDeviceSynchronize();
        long asc = 0, desc = 0, pair = 0, tot = 0;
        for (int it = 0; it < T; ++it)
            for (int i = 0; i < 8; ++i)
                for (int j = 0; j < 8; ++j) {
                    const double *a = A + it * 32 + i * 4, *b = B + it * 32 + j, c = C[it * 64 + i * 8 + j];
                    double r1 = c; for (int k = 0; k < 4; ++k) r1 = fma(a[k], b[k * 8], r1);
                    double r2 = c; for (int k = 3; k >= 0; --k) r2 = fma(a[k], b[k * 8], r2);
                    double r3 = fma(a[1], b[8], fma(a[0], b[0], c)) ; double r3b = fma(a[3], b[24], a[2] * b[16]); double r3s = r3 + r3b;
                    const double d = D[it * 64 + i * 8 + j];
                    tot++; asc += (d == r1 || (d != d && r1 != r1)); desc += (d == r2); pair += (d == r3s);
                }
        printf("m8n8k4  mode %d: == ascending chain %ld / %ld, == descending chain %ld, == pairwise %ld\n", mode, asc, tot, desc, pair);
        cudaFree(A); cudaFree(B); cudaFree(C); cudaFree(D);
        // ---- m16n8k8
        cudaMallocManaged(&A, T * 128 * 8); cudaMallocManaged(&B, T * 64 * 8); cudaMallocManaged(&C, T * 128 * 8); cudaMallocManaged(&D, T * 128 * 8);
        for (int i = 0; i < T * 128; ++i) { A[i] = mode == 2 ? (double)(rand() % 200) : rnd(mode); C[i] = rnd(mode); }
        for (int i = 0; i < T * 64; ++i) B[i] = rnd(mode);
        exact1688<<<1, 32>>>(A, B, C, D, T); cudaDeviceSynchronize();
        asc = 0; tot = 0; long asc44 = 0;
        for (int it = 0; it < T; ++it)
            for (int i = 0; i < 16; ++i)
                for (int j = 0; j < 8; ++j) {
                    const double *a = A + it * 128 + i * 8, *b = B + it * 64 + j, c = C[it * 128 + i * 8 + j];
                    double r1 = c; for (int k = 0; k < 8; ++k) r1 = fma(a[k], b[k * 8], r1);
                    const double d = D[it * 128 + i * 8 + j];
                    tot++; asc += (d == r1);
                }
        printf("m16n8k8 mode %d: == ascending chain %ld / %ld\n", mode, asc, tot);
        cudaFree(A); cudaFree(B); cudaFree(C); cudaFree(D);
    }
    // ---- rates
    int sms = 0; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    double *sink; cudaMalloc(&sink, 8);
    const int iters = 20000;
    {
        float ms = timeit([&] { rate_dfma<<<sms * 8, 256>>>(iters / 4, 1.0, sink); });
        printf("DFMA          : %.2f TFLOP/s\n", (double)sms * 8 * 256 * (iters / 4) * 16 * 8 * 2 / (ms * 1e-3) / 1e12);
    }
#define RATE(K, NT, FL)                                                                                          \
    for (int bps = 1; bps <= 8; bps *= 2) {                                                                      \
        float ms = timeit([&] { K<NT><<<sms * bps, 256>>>(iters, 1.0, sink); });                                  \
        printf(#K "<%d> %d x 8 warps/SM: %.2f TFLOP/s\n", NT, bps, (double)sms * bps * 8 * iters * 4 * NT * FL * 2 / (ms * 1e-3) / 1e12); \
    }
    RATE(rate884, 1, 256.0) RATE(rate884, 4, 256.0) RATE(rate884, 16, 256.0)
    RATE(rate1688, 1, 1024.0) RATE(rate1688, 4, 1024.0) RATE(rate1688, 8, 1024.0)
    return 0;
}
