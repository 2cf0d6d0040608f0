// FJLT: SA = sqrt(N/S) * Sample_S( FUT( scale * D .* A ) )
//
// Reference pieces replaced:
//   FJLT_data_t::build / RFUT_data_t::build   sketch/FJLT_data.hpp:79-86, sketch/RFUT_data.hpp:27-50
//   RFUT_t::apply_impl_vdist (sign + scale)   sketch/RFUT_Elemental.hpp:165-184
//   FFTW REDFT10 (DCT-II), scale 1/sqrt(2N)   utility/fft/fftw_futs.h:44-46, 117-126
//   sampling and sqrt(N/S) scale              sketch/FJLT_Elemental.hpp:157-166
// plus a Walsh-Hadamard mode (scale 1/sqrt(N), natural order; sketch/FUT.hpp:25-113
// is what the reference would use) that the reference's FJLT does not have.
//
// The counter overlap of the reference is kept: D uses stream [c, c+N), the sample
// indices use [c, c+S) of the same stream.
#include "rng.cuh"
#include "skb_internal.h"

namespace skb {

__global__ void k_fjlt_tables(uint64_t key, uint64_t base, size_t N, size_t S, int order,
                              double *__restrict__ D, int32_t *__restrict__ samples) {
    size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t < N) {
        MicroUrng u(key, base + t);
        D[t] = boost_rademacher(u, order);
    }
    if (t < S) {
        MicroUrng u(key, base + t);
        samples[t] = (int32_t)boost_uniform_int(u, (uint64_t)N - 1);
    }
}

void fjlt_build_tables(int seed, uint64_t base, size_t N, size_t S, int order, double *D,
                       int32_t *samples, cudaStream_t st) {
    size_t n = N > S ? N : S;
    if (n == 0) return;
    LaunchScope ls(KC_TABLES, st);
    k_fjlt_tables<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(seed_to_key(seed), base, N, S,
                                                               order, D, samples);
    SKB_CUDA(cudaGetLastError());
}

// ---------------------------------------------------------------------------
// General-N path: the S sampled outputs evaluated directly from the definition,
//   DCT: Y[k] = 2 sum_n x_n cos(pi (2n+1) k / (2N)),   WHT: Y[k] = sum_n (-1)^popc(k&n) x_n
// O(S*N) per column; used for sizes the fast path does not cover.
// ---------------------------------------------------------------------------
__global__ void k_cos_table(double *T, size_t N) {
    size_t r = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r < 4 * N) T[r] = cospi((double)r / (double)(2 * N));
}

template <int FUT>
__global__ void __launch_bounds__(256)
k_fjlt_direct(const double *__restrict__ X, size_t sk, size_t sj, size_t N, size_t m,
              const double *__restrict__ D, const int32_t *__restrict__ samples, size_t S,
              const double *__restrict__ T, double inscale, double outscale,
              double *__restrict__ OUT, size_t oi, size_t oj) {
    const size_t i = blockIdx.x, j = blockIdx.y;
    const unsigned long long k = (unsigned long long)samples[i];
    const unsigned long long mod = 4ULL * N;
    double acc = 0.0;
    for (size_t n = threadIdx.x; n < N; n += blockDim.x) {
        double x = __dmul_rn(__dmul_rn(inscale, D[n]), X[n * sk + j * sj]);
        double w;
        if (FUT == FUT_DCT) w = T[((2ULL * n + 1ULL) * k) % mod];
        else w = (__popcll(k & (unsigned long long)n) & 1) ? -1.0 : 1.0;
        acc = fma(x, w, acc);
    }
    __shared__ double red[256];
    red[threadIdx.x] = acc;
    __syncthreads();
    for (int s = 128; s > 0; s >>= 1) {
        if ((int)threadIdx.x < s) red[threadIdx.x] += red[threadIdx.x + s];
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        double y = FUT == FUT_DCT ? 2.0 * red[0] : red[0];
        OUT[i * oi + j * oj] = outscale * y;
    }
}

size_t fjlt_workspace_bytes(size_t N, size_t m, int fut) {
    (void)m;
    return fut == FUT_DCT ? 4 * N * sizeof(double) + 256 : 256;
}

void fjlt_apply(int fut, bool layout_rows, const double *X, size_t ldx, size_t N, size_t m,
                const double *D, const int32_t *samples, size_t S, double *OUT, size_t ldo,
                void *workspace, size_t workspace_bytes, cudaStream_t st) {
    if (S == 0 || m == 0) return;
    if (fut == FUT_WHT && (N & (N - 1)) != 0)
        throw error(SKB_ERR_INVALID_PARAM, "FJLT (WHT mode): N must be a power of two");
    if (workspace_bytes < fjlt_workspace_bytes(N, m, fut))
        throw error(SKB_ERR_ALLOC, "fjlt apply: workspace too small");
    if (S > 0x7fffffffULL || m > 65535)
        throw error(SKB_ERR_INVALID_PARAM, "fjlt apply (general path): shape too large");
    const size_t sk = layout_rows ? ldx : 1, sj = layout_rows ? 1 : ldx;
    const size_t oi = layout_rows ? ldo : 1, oj = layout_rows ? 1 : ldo;
    const double outscale = sqrt((double)N / (double)S);
    double *T = reinterpret_cast<double *>(workspace);
    if (fut == FUT_DCT) {
        {
            LaunchScope ls(KC_FJLT, st);
            k_cos_table<<<(unsigned)((4 * N + 255) / 256), 256, 0, st>>>(T, N);
        }
        const double inscale = 1.0 / sqrt(2.0 * (double)N);
        LaunchScope ls(KC_FJLT, st);
        k_fjlt_direct<FUT_DCT><<<dim3((unsigned)S, (unsigned)m), 256, 0, st>>>(
            X, sk, sj, N, m, D, samples, S, T, inscale, outscale, OUT, oi, oj);
    } else {
        const double inscale = 1.0 / sqrt((double)N);
        LaunchScope ls(KC_FJLT, st);
        k_fjlt_direct<FUT_WHT><<<dim3((unsigned)S, (unsigned)m), 256, 0, st>>>(
            X, sk, sj, N, m, D, samples, S, T, inscale, outscale, OUT, oi, oj);
    }
    SKB_CUDA(cudaGetLastError());
}

}  // namespace skb
