// Issue-rate microbenchmark for the legacy (mma.sync) TF32 and BF16 tensor paths on B200.
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tf32_peak tf32_peak.cu
#include <cstdio>
#include <cuda_runtime.h>
#define ITERS 4096

__global__ void k_tf32(float *out, unsigned a, unsigned b) {
    float c[16][4];
#pragma unroll
    for (int i = 0; i < 16; i++) { c[i][0] = threadIdx.x; c[i][1] = i; c[i][2] = 1; c[i][3] = 2; }
    for (int it = 0; it < ITERS; it++) {
#pragma unroll
        for (int i = 0; i < 16; i++)
            asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%4,%4,%4}, {%5,%5}, {%0,%1,%2,%3};"
                         : "+f"(c[i][0]), "+f"(c[i][1]), "+f"(c[i][2]), "+f"(c[i][3]) : "r"(a), "r"(b));
    }
    float s = 0;
#pragma unroll
    for (int i = 0; i < 16; i++) s += c[i][0] + c[i][1] + c[i][2] + c[i][3];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void k_bf16(float *out, unsigned a, unsigned b) {
    float c[16][4];
#pragma unroll
    for (int i = 0; i < 16; i++) { c[i][0] = threadIdx.x; c[i][1] = i; c[i][2] = 1; c[i][3] = 2; }
    for (int it = 0; it < ITERS; it++) {
#pragma unroll
        for (int i = 0; i < 16; i++)
            asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%4,%4,%4}, {%5,%5}, {%0,%1,%2,%3};"
                         : "+f"(c[i][0]), "+f"(c[i][1]), "+f"(c[i][2]), "+f"(c[i][3]) : "r"(a), "r"(b));
    }
    float s = 0;
#pragma unroll
    for (int i = 0; i < 16; i++) s += c[i][0] + c[i][1] + c[i][2] + c[i][3];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <class K> void run(const char *name, K k, double flops_per_thread_iter, int threads) {
    float *out; cudaMalloc(&out, 148 * 8 * 1024 * sizeof(float));
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int blocks_per_sm = 1; blocks_per_sm <= 4; blocks_per_sm *= 2) {
        k<<<148 * blocks_per_sm, threads>>>(out, 0x3f800000u, 0x3f800000u);
        cudaDeviceSynchronize();
        cudaEventRecord(e0);
        k<<<148 * blocks_per_sm, threads>>>(out, 0x3f800000u, 0x3f800000u);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        double fl = (double)148 * blocks_per_sm * threads * ITERS * flops_per_thread_iter;
        printf("%s: %d blocks/SM x %d threads: %.1f TFLOP/s\n", name, blocks_per_sm, threads, fl / ms / 1e9);
    }
}
int main() {
    // per warp per mma: m16n8k8 = 2*16*8*8 = 2048 flop -> per thread 64; 16 mma per iter
    run("mma.sync m16n8k8 tf32", k_tf32, 16 * 64.0, 256);
    run("mma.sync m16n8k16 bf16", k_bf16, 16 * 128.0, 256);
    return 0;
}
