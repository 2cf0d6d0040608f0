// FP64 issue-rate microbenchmark for B200: DFMA, DMMA m8n8k4, DMMA m16n8k16 (if it
// assembles), and DFMA+DMMA interleaved (do the two share one pipe?).  Prints TFLOP/s.
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o fp64_peak fp64_peak.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

#define ITERS 4096

__global__ void k_dfma(double *out, double a, double b) {
    double c[16];
#pragma unroll
    for (int i = 0; i < 16; i++) c[i] = threadIdx.x + i;
    for (int it = 0; it < ITERS; it++) {
#pragma unroll
        for (int i = 0; i < 16; i++) c[i] = fma(a, c[i], b);
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < 16; i++) s += c[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void k_dmma884(double *out, double a, double b) {
    double c[16][2];
#pragma unroll
    for (int i = 0; i < 16; i++) { c[i][0] = threadIdx.x; c[i][1] = i; }
    for (int it = 0; it < ITERS; it++) {
#pragma unroll
        for (int i = 0; i < 16; i++)
            asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                         : "+d"(c[i][0]), "+d"(c[i][1]) : "d"(a), "d"(b));
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < 16; i++) s += c[i][0] + c[i][1];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

#ifdef TRY_BIG
__global__ void k_dmma16816(double *out, double a, double b) {
    double c[8][4];
#pragma unroll
    for (int i = 0; i < 8; i++) { c[i][0] = threadIdx.x; c[i][1] = i; c[i][2] = 1; c[i][3] = 2; }
    for (int it = 0; it < ITERS; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++)
            asm volatile("mma.sync.aligned.m16n8k16.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, "
                         "{%4,%4,%4,%4,%4,%4,%4,%4}, {%5,%5,%5,%5}, {%0,%1,%2,%3};"
                         : "+d"(c[i][0]), "+d"(c[i][1]), "+d"(c[i][2]), "+d"(c[i][3]) : "d"(a), "d"(b));
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) s += c[i][0] + c[i][1] + c[i][2] + c[i][3];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
#endif

__global__ void k_mixed(double *out, double a, double b) {
    double c[8][2], f[8];
#pragma unroll
    for (int i = 0; i < 8; i++) { c[i][0] = threadIdx.x; c[i][1] = i; f[i] = i; }
    for (int it = 0; it < ITERS; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++) {
            asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                         : "+d"(c[i][0]), "+d"(c[i][1]) : "d"(a), "d"(b));
            f[i] = fma(a, f[i], b);
        }
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) s += c[i][0] + c[i][1] + f[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <typename K>
static double run(K kern, int blocks, int threads, double *out) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    kern<<<blocks, threads>>>(out, 1.0000001, 1e-9);
    cudaDeviceSynchronize();
    float best = 1e30f;
    for (int r = 0; r < 5; r++) {
        cudaEventRecord(e0);
        kern<<<blocks, threads>>>(out, 1.0000001, 1e-9);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    return best;
}

int main(int argc, char **argv) {
    int blocks = argc > 1 ? atoi(argv[1]) : 148 * 8, threads = argc > 2 ? atoi(argv[2]) : 256;
    printf("blocks=%d threads=%d\n", blocks, threads);
    double *out; cudaMalloc(&out, (size_t)blocks * threads * 8);
    double thr = (double)blocks * threads;
    double ms = run(k_dfma, blocks, threads, out);
    printf("DFMA           : %8.3f ms  %7.2f TFLOP/s\n", ms, thr * ITERS * 16 * 2 / ms / 1e9);
    ms = run(k_dmma884, blocks, threads, out);
    // one m8n8k4 per warp = 256 FMA = 8 per lane
    printf("DMMA m8n8k4    : %8.3f ms  %7.2f TFLOP/s\n", ms, thr * ITERS * 16 * 8 * 2 / ms / 1e9);
#ifdef TRY_BIG
    ms = run(k_dmma16816, blocks, threads, out);
    printf("DMMA m16n8k16  : %8.3f ms  %7.2f TFLOP/s\n", ms, thr * ITERS * 8 * 64 * 2 / ms / 1e9);
#endif
    ms = run(k_mixed, blocks, threads, out);
    printf("mixed 8 DMMA + 8 DFMA per iter: %8.3f ms  DMMA-part %7.2f TFLOP/s  DFMA-part %7.2f TFLOP/s\n", ms,
           thr * ITERS * 8 * 8 * 2 / ms / 1e9, thr * ITERS * 8 * 2 / ms / 1e9);
    cudaError_t e = cudaGetLastError();
    printf("status: %s\n", cudaGetErrorString(e));
    return e != cudaSuccess;
}
