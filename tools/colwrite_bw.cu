// Write-bandwidth probe for the K3 output pattern: persistent CTAs, each writing whole dense columns
// (S doubles, contiguous) of a column-major S x ncols matrix, column = blockIdx.x + k * gridDim.x.
// nvcc -O3 -gencode arch=compute_100a,code=sm_100a tools/colwrite_bw.cu -o tools/colwrite_bw && ./tools/colwrite_bw
#include <cstdio>
#include <cuda_runtime.h>
__global__ void __launch_bounds__(512) k_cols(double *out, size_t S, size_t ncols, int chunk) {
    const double2 z = make_double2(1.5, 2.5);
    for (size_t col0 = (size_t)blockIdx.x * chunk; col0 < ncols; col0 += (size_t)gridDim.x * chunk)
        for (size_t col = col0; col < col0 + chunk && col < ncols; col++) {
            double2 *o = reinterpret_cast<double2 *>(out + col * S);
            for (size_t i = threadIdx.x; i < S / 2; i += blockDim.x) o[i] = z;
        }
}
// the same columns through the copy engine: cp.async.bulk.global.shared::cta of `pieces` pieces per column
__global__ void __launch_bounds__(512) k_cols_bulk(double *out, size_t S, size_t ncols, int pieces) {
    extern __shared__ __align__(16) double buf[];
    for (size_t i = threadIdx.x; i < S; i += blockDim.x) buf[i] = 1.5;
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    const unsigned b = (unsigned)__cvta_generic_to_shared(buf);
    const size_t P = S / pieces;
    if (threadIdx.x == 0)
        for (size_t col = blockIdx.x; col < ncols; col += gridDim.x)
            for (int q = 0; q < pieces; q++) {
                asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(out + col * S + q * P),
                             "r"(b + (unsigned)(q * P * 8)), "r"((unsigned)(P * 8)) : "memory");
                asm volatile("cp.async.bulk.commit_group;" ::: "memory");
                asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
            }
    if (threadIdx.x == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}
int main() {
    const size_t S = 20000, ncols = 100000;
    double *out; cudaMalloc(&out, S * ncols * 8);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int grids[] = {148, 296, 592, 1184, 148 * 16};
    for (int g : grids) for (int chunk : {1, 4}) {
        k_cols<<<g, 512>>>(out, S, ncols, chunk); cudaDeviceSynchronize();
        float best = 1e9;
        for (int r = 0; r < 3; r++) {
            cudaEventRecord(e0); k_cols<<<g, 512>>>(out, S, ncols, chunk); cudaEventRecord(e1); cudaEventSynchronize(e1);
            float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
        }
        printf("grid %5d x 512 threads, %d column(s) per turn: %.3f ms  %.0f GB/s\n", g, chunk, best, S * ncols * 8 / best / 1e6);
    }
    cudaFuncSetAttribute(k_cols_bulk, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(S * 8));
    for (int pieces : {1, 2, 4, 10}) {
        k_cols_bulk<<<148, 512, S * 8>>>(out, S, ncols, pieces); cudaDeviceSynchronize();
        float best = 1e9;
        for (int r = 0; r < 3; r++) {
            cudaEventRecord(e0); k_cols_bulk<<<148, 512, S * 8>>>(out, S, ncols, pieces); cudaEventRecord(e1); cudaEventSynchronize(e1);
            float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
        }
        printf("bulk copies from shared memory, 148 CTAs, %2d piece(s) per column: %.3f ms  %.0f GB/s  (%s)\n", pieces, best,
               S * ncols * 8 / best / 1e6, cudaGetErrorString(cudaGetLastError()));
    }
    return 0;
}
