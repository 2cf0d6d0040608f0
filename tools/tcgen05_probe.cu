// Probe: tcgen05.mma.kind::tf32 (M=128, N=128, K=8*KSTEPS) on operands written to shared
// memory by ordinary stores in the no-swizzle K-major core-matrix layout, accumulator read back
// from TMEM.  Checks descriptors / TMEM addressing before they go into the library.
// Build: nvcc -O2 -gencode arch=compute_100a,code=sm_100a -o tcgen05_probe tcgen05_probe.cu
// Result on B200: max error 0, 0 mismatches of 16384.
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cmath>
#include <cuda_runtime.h>

constexpr int M = 128, N = 128, KSTEPS = 4, K = 8 * KSTEPS;
// K-major, no swizzle: element (r, k) at (r/8)*SBO + (k/4)*LBO + (r%8)*16 + (k%4)*4 bytes
constexpr uint32_t LBO = 128, SBO = (K / 4) * 128;

__device__ __forceinline__ uint64_t make_desc(uint32_t saddr) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr >> 4) & 0x3fff);
    d |= (uint64_t)((LBO >> 4) & 0x3fff) << 16;
    d |= (uint64_t)((SBO >> 4) & 0x3fff) << 32;
    d |= (uint64_t)1 << 46;          // version = 1 (Blackwell)
    return d;                        // base_offset 0, lbo_mode 0, layout_type 0 (no swizzle)
}

__global__ void __launch_bounds__(128) k_probe(const float *A, const float *B, float *D) {
    extern __shared__ __align__(128) unsigned char smem[];
    float *As = reinterpret_cast<float *>(smem);                 // M*K floats
    float *Bs = As + M * K;
    __shared__ __align__(8) uint64_t mbar;
    __shared__ uint32_t tmem_base_s;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    for (int e = tid; e < M * K; e += 128) {
        const int r = e / K, k = e % K;
        const uint32_t off = (r / 8) * SBO + (k / 4) * LBO + (r % 8) * 16 + (k % 4) * 4;
        As[off / 4] = A[r * K + k];
        Bs[off / 4] = B[r * K + k];
    }
    const uint32_t mbar_a = (uint32_t)__cvta_generic_to_shared(&mbar);
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(mbar_a));
        asm volatile("fence.mbarrier_init.release.cluster;");
    }
    if (warp == 0) {
        const uint32_t dst = (uint32_t)__cvta_generic_to_shared(&tmem_base_s);
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 128;" ::"r"(dst));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    asm volatile("fence.proxy.async.shared::cta;");      // generic-proxy smem writes -> async proxy
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;");
    const uint32_t tmem = tmem_base_s;
    if (tid == 0) {
        const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
        const uint32_t a0 = (uint32_t)__cvta_generic_to_shared(As), b0 = (uint32_t)__cvta_generic_to_shared(Bs);
        for (int ks = 0; ks < KSTEPS; ks++) {
            const uint64_t ad = make_desc(a0 + ks * 2 * LBO), bd = make_desc(b0 + ks * 2 * LBO);
            const uint32_t accum = ks > 0;
            asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                         "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
                         ::"r"(tmem), "l"(ad), "l"(bd), "r"(idesc), "r"(accum));
        }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(mbar_a));
    }
    {
        uint32_t done = 0;
        while (!done)
            asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                         : "=r"(done) : "r"(mbar_a));
    }
    asm volatile("tcgen05.fence::after_thread_sync;");
    // warp w reads TMEM lanes [32w, 32w+32): lane = row m, column = n
    for (int c0 = 0; c0 < N; c0 += 8) {
        uint32_t r[8];
        const uint32_t taddr = tmem + ((uint32_t)(warp * 32) << 16) + (uint32_t)c0;
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                     : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                     : "r"(taddr));
        asm volatile("tcgen05.wait::ld.sync.aligned;");
        for (int e = 0; e < 8; e++) D[(warp * 32 + lane) * N + c0 + e] = __uint_as_float(r[e]);
    }
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 128;" ::"r"(tmem));
}

int main() {
    float *hA = (float *)malloc(M * K * 4), *hB = (float *)malloc(N * K * 4), *hD = (float *)malloc(M * N * 4);
    srand(1);
    for (int i = 0; i < M * K; i++) hA[i] = (float)((rand() % 17) - 8) / 8.0f;      // exact in tf32
    for (int i = 0; i < N * K; i++) hB[i] = (float)((rand() % 13) - 6) / 4.0f;
    float *A, *B, *D;
    cudaMalloc(&A, M * K * 4); cudaMalloc(&B, N * K * 4); cudaMalloc(&D, M * N * 4);
    cudaMemcpy(A, hA, M * K * 4, cudaMemcpyHostToDevice);
    cudaMemcpy(B, hB, N * K * 4, cudaMemcpyHostToDevice);
    cudaMemset(D, 0xff, M * N * 4);
    k_probe<<<1, 128, 2 * M * K * 4>>>(A, B, D);
    cudaError_t e = cudaDeviceSynchronize();
    printf("kernel: %s\n", cudaGetErrorString(e));
    cudaMemcpy(hD, D, M * N * 4, cudaMemcpyDeviceToHost);
    double maxerr = 0; int bad = 0;
    for (int m = 0; m < M; m++)
        for (int n = 0; n < N; n++) {
            double ref = 0;
            for (int k = 0; k < K; k++) ref += (double)hA[m * K + k] * hB[n * K + k];
            double err = fabs(ref - hD[m * N + n]);
            if (!(err < 1e-4)) { if (bad < 5) printf("  D[%d][%d] = %g, want %g\n", m, n, hD[m * N + n], ref); bad++; }
            if (err > maxerr) maxerr = err;
        }
    printf("max error %.3g, mismatches %d of %d\n", maxerr, bad, M * N);
    return 0;
}
