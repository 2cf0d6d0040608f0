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
#include <algorithm>
#include <cstdlib>
#include <vector>

#include "rng.cuh"
#include "skb_internal.h"

namespace skb {

__global__ void k_fjlt_tables(uint64_t key, uint64_t base, size_t N, size_t S, int order,
                              double *__restrict__ D, int32_t *__restrict__ samples,
                              uint32_t *__restrict__ Dbits) {
    size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    bool neg = false;
    if (t < N) {
        MicroUrng u(key, base + t);
        double d = boost_rademacher(u, order);
        D[t] = d;
        neg = d < 0.0;
    }
    unsigned word = __ballot_sync(0xffffffffu, neg);     // bit l of word t/32 = sign of D[t]
    if ((threadIdx.x & 31) == 0 && t < ((N + 31) / 32) * 32) Dbits[t >> 5] = word;
    if (t < S) {
        MicroUrng u(key, base + t);
        samples[t] = (int32_t)boost_uniform_int(u, (uint64_t)N - 1);
    }
}

void fjlt_build_tables(int seed, uint64_t base, size_t N, size_t S, int order, double *D,
                       int32_t *samples, uint32_t *Dbits, cudaStream_t st) {
    size_t n = N > S ? N : S;
    if (n == 0) return;
    LaunchScope ls(KC_TABLES, st);
    k_fjlt_tables<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(seed_to_key(seed), base, N, S,
                                                               order, D, samples, Dbits);
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


// ---------------------------------------------------------------------------
// WHT mode, N = 2^p >= 8192, every column contiguous (K7).
// H_N = H_{N/SEG} (x) H_SEG in natural order, so a column is streamed ONCE in SEG = 8192-element
// segments (64 KB) and only the S sampled outputs are kept: sample r = (k1, k2) accumulates
// (-1)^popc(k1 & n1) * H_SEG(seg_n1)[k2] over the segments n1, in registers.  HBM traffic is the
// algorithmic n*d*8 + s*d*8 bytes; no scratch, no atomics.
//
// One CTA per SM, two roles (warp specialised, named barriers):
//  * 4 butterfly warps: thread t loads its 64 values of a segment straight from global memory
//    into registers (32 coalesced 16-byte loads: index bits {0, 8..12}), flips signs from a
//    per-thread 64-bit word of D, does 6 butterfly levels in registers and the level of index
//    bit 1 with its neighbour lane (shuffle), exchanges through shared memory ONCE, does the 6
//    levels of index bits 2..7 and leaves the segment in shared memory.  Two groups of 4 warps
//    take alternate segments, so one computes while the other waits for its loads.
//  * 8 gather warps hold the S sample accumulators (32 per thread).  Per segment each sample
//    reads its value, applies the cross-segment sign and adds.  They work on segment i while
//    the butterfly warps are already on segments i+1 and i+2 (ring of three shared buffers),
//    and one of them keeps an L2 prefetch ahead of the butterfly warps' loads.
//  Registers are split with setmaxnreg: 168 per butterfly thread, 88 per gather thread.
// ---------------------------------------------------------------------------
static constexpr int WSEG_LOG = 13, WSEG = 1 << WSEG_LOG;
static constexpr int WB_THREADS = 128;               // one butterfly group (there are two)
static constexpr int WG_THREADS = 256, WTHREADS = 2 * WB_THREADS + WG_THREADS;
static constexpr int WSPT = 32;      // sample slots per gather thread (WG_THREADS * WSPT = 8192 per pass)
static constexpr int WBUF_BYTES = WSEG * 8, WNBUF = 3;
static constexpr int WSMEM_BYTES = WNBUF * WBUF_BYTES + WG_THREADS * WSPT * 4;
// named barriers: FULL one id per ring buffer; EMPTY one id per (ring buffer, butterfly group
// that will refill it) -- with a shared id the other group, running ahead, could complete the
// barrier in place of the group it was meant for; INNER one per butterfly group
enum { WBAR_FULL = 1, WBAR_EMPTY = 4, WBAR_INNER = 10 };
static constexpr int WREG_B = 168, WREG_G = 88;      // 256*168 + 256*88 = the whole register file
#ifndef SKB_WHT_PF
#define SKB_WHT_PF 4                 // L2 prefetch distance in segments (0 = off)
#endif

// Segment element e lives at shared word swz(e): index bits 8..9 are XORed into bits 2..3.  The
// exchange store (16-byte pairs, consecutive lanes), the phase-B accesses (a half warp varies
// index bits {0, 1, 8, 9}) and the gather (slots arranged by the host) are all conflict-free.
__host__ __device__ __forceinline__ int swz(int e) { return e ^ (((e >> 8) & 3) << 2); }

__device__ __forceinline__ void bar_sync(int id, int n) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n) : "memory");
}
__device__ __forceinline__ void bar_arrive(int id, int n) {
    asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(n) : "memory");
}
__device__ __forceinline__ double2 ldg_stream2(const double *p) {
    double2 v;
    asm volatile("ld.global.nc.L1::no_allocate.v2.f64 {%0, %1}, [%2];" : "=d"(v.x), "=d"(v.y) : "l"(p));
    return v;
}
__device__ __forceinline__ double ldg_stream1(const double *p) {
    double v;
    asm volatile("ld.global.nc.L1::no_allocate.f64 %0, [%1];" : "=d"(v) : "l"(p));
    return v;
}
__device__ __forceinline__ double lds_f64(unsigned addr) {
    double v;
    asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(addr));
    return v;
}
__device__ __forceinline__ void sts_f64(unsigned addr, double v) {
    asm volatile("st.shared.f64 [%0], %1;" ::"r"(addr), "d"(v) : "memory");
}
__device__ __forceinline__ void sts_v2f64(unsigned addr, double a, double b) {
    asm volatile("st.shared.v2.f64 [%0], {%1, %2};" ::"r"(addr), "d"(a), "d"(b) : "memory");
}
__device__ __forceinline__ double xor_sign(double x, unsigned mask) {   // mask = 0 or 0x80000000
    return __hiloint2double(__double2hiint(x) ^ (int)mask, __double2loint(x));
}

__device__ __forceinline__ void wht64(double (&v)[64]) {      // 6 butterfly levels over 64 registers
#pragma unroll
    for (int h = 1; h < 64; h <<= 1)
#pragma unroll
        for (int i = 0; i < 64; i++)
            if (!(i & h)) { double a = v[i], b = v[i + h]; v[i] = a + b; v[i + h] = a - b; }
}

// Per-thread sign words of the butterfly warps: bit (2*j + b) of Dwht[seg*128 + t] is the sign
// bit of D at element seg*8192 + j*256 + t*2 + b.
__global__ void k_wht_sign_words(const uint32_t *__restrict__ Dbits, size_t nwords, uint64_t *__restrict__ Dwht) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nwords) return;
    const size_t seg = i >> 7; const int t = (int)(i & 127);
    uint64_t w = 0;
    for (int j = 0; j < 32; j++) {
        const size_t e = (seg << WSEG_LOG) + (size_t)j * 256 + (size_t)t * 2;
        const uint64_t two = (Dbits[e >> 5] >> (e & 31)) & 3u;
        w |= two << (2 * j);
    }
    Dwht[i] = w;
}

template <bool AL16>
__global__ void __launch_bounds__(WTHREADS, 1)
k_fjlt_wht_cols(const double *__restrict__ X, size_t ldx, size_t N, size_t m,
                const uint64_t *__restrict__ Dwht, const int32_t *__restrict__ slot_sample,
                const int32_t *__restrict__ slot_out, size_t seg0, double inscale, double outscale,
                double *__restrict__ OUT, size_t oi, size_t oj) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const unsigned sbase = (unsigned)__cvta_generic_to_shared(smem_raw);
    const int tid = threadIdx.x;
    const size_t nseg = N >> WSEG_LOG;      // segments present in X (a row shard starts at seg0)
    const size_t ncols = (m - blockIdx.x + gridDim.x - 1) / gridDim.x;
    const size_t total = ncols * nseg;      // (column, segment) items of this CTA, in order;
                                            // item it lives in ring buffer it % 3

    if (tid < 2 * WB_THREADS) {
        // ================= butterfly warps: group s takes the items with it % 2 == s =================
        asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(WREG_B));
        const int s = tid >> 7, t = tid & (WB_THREADS - 1);
        // exchange store: element pair e = j*256 + t*2 -> swz(e) = j*256 + ((t*2) ^ (c << 2)),
        // c = index bits 8..9 = j & 3
        unsigned offA[4];
#pragma unroll
        for (int c = 0; c < 4; c++) offA[c] = (unsigned)(((t * 2) ^ (c << 2)) * 8);
        // phase B: thread = (index bits 0..1, index bits 8..12), registers = index bits 2..7
        const int lo2 = t & 3, hi = t >> 2, h2 = hi & 3;
        unsigned offB[4];
#pragma unroll
        for (int rl = 0; rl < 4; rl++) offB[rl] = (unsigned)((hi * 256 + lo2 + ((rl ^ h2) << 2)) * 8);
        const unsigned odd = (unsigned)(t & 1) << 31;     // index bit 1 of this thread's phase-A elements
        size_t it = 0;
        int b = 0;                                  // it % 3
        for (size_t j = blockIdx.x; j < m; j += gridDim.x) {
            const double *col = X + j * ldx;
            for (size_t n1 = 0; n1 < nseg; n1++, it++, b = (b == WNBUF - 1) ? 0 : b + 1) {
                if ((int)(it & 1) != s) continue;
                const double *g = col + (n1 << WSEG_LOG) + t * 2;
                const uint64_t w = Dwht[((seg0 + n1) << 7) + t];
                double v[64];
#pragma unroll
                for (int jj = 0; jj < 32; jj++) {
                    if (AL16) {
                        double2 x = ldg_stream2(g + jj * 256);
                        v[2 * jj] = x.x; v[2 * jj + 1] = x.y;
                    } else {
                        v[2 * jj] = ldg_stream1(g + jj * 256);
                        v[2 * jj + 1] = ldg_stream1(g + jj * 256 + 1);
                    }
                }
                const unsigned wlo = (unsigned)w, whi = (unsigned)(w >> 32);
#pragma unroll
                for (int i = 0; i < 32; i++) {
                    v[i] = xor_sign(v[i], (wlo << (31 - i)) & 0x80000000u);
                    v[32 + i] = xor_sign(v[32 + i], (whi << (31 - i)) & 0x80000000u);
                }
                wht64(v);                                   // index bits {0, 8..12}
#pragma unroll
                for (int i = 0; i < 64; i++) {              // index bit 1: the partner is the next lane
                    const double o = __shfl_xor_sync(0xffffffffu, v[i], 1);
                    v[i] = o + xor_sign(v[i], odd);
                }
                const unsigned buf = sbase + (unsigned)b * WBUF_BYTES;
                if (it >= WNBUF) bar_sync(WBAR_EMPTY + 2 * b + s, WB_THREADS + WG_THREADS);   // gather of item it-3 is done
#pragma unroll
                for (int jj = 0; jj < 32; jj++)
                    sts_v2f64(buf + offA[jj & 3] + jj * 2048, v[2 * jj], v[2 * jj + 1]);
                bar_sync(WBAR_INNER + s, WB_THREADS);
#pragma unroll
                for (int r = 0; r < 64; r++) v[r] = lds_f64(buf + offB[r & 3] + (r >> 2) * 128);
                wht64(v);                                   // index bits 2..7
#pragma unroll
                for (int r = 0; r < 64; r++) sts_f64(buf + offB[r & 3] + (r >> 2) * 128, v[r]);
                __threadfence_block();
                bar_arrive(WBAR_FULL + b, WB_THREADS + WG_THREADS);
            }
        }
    } else {
        // ================= gather warps =================
        asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(WREG_G));
        const int g = tid - 2 * WB_THREADS;
        // slot table (fjlt_build_wht_slots): bits 0..15 byte offset of the sample's value, bits
        // 16..30 k1; empty slots read offset 0 and are never written out.  Kept in shared memory as [WSPT/4][WG_THREADS] int4 (the registers
        // belong to the accumulators).
        const unsigned ptab = sbase + WNBUF * WBUF_BYTES + (unsigned)g * 16;
#pragma unroll
        for (int q4 = 0; q4 < WSPT / 4; q4++) {
            int4 pp;
            pp.x = slot_sample[(q4 * 4 + 0) * WG_THREADS + g];
            pp.y = slot_sample[(q4 * 4 + 1) * WG_THREADS + g];
            pp.z = slot_sample[(q4 * 4 + 2) * WG_THREADS + g];
            pp.w = slot_sample[(q4 * 4 + 3) * WG_THREADS + g];
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(ptab + q4 * WG_THREADS * 16),
                         "r"(pp.x), "r"(pp.y), "r"(pp.z), "r"(pp.w) : "memory");
        }
        const double scale = inscale * outscale;   // both commute with the transform
        size_t it = 0;
        int b = 0;
        for (size_t j = blockIdx.x; j < m; j += gridDim.x) {
            double acc[WSPT];
#pragma unroll
            for (int q = 0; q < WSPT; q++) acc[q] = 0.0;
            for (size_t n1 = 0; n1 < nseg; n1++, it++, b = (b == WNBUF - 1) ? 0 : b + 1) {
                if (AL16 && SKB_WHT_PF > 0 && g == 0) {   // keep L2 ahead of the butterfly warps' loads
                    size_t jp = j, np = n1 + SKB_WHT_PF;
                    while (np >= nseg) { np -= nseg; jp += gridDim.x; }
                    if (jp < m) {
                        const double *nx = X + jp * ldx + (np << WSEG_LOG);
                        asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(nx), "r"(WBUF_BYTES));
                    }
                }
                const unsigned buf = sbase + (unsigned)b * WBUF_BYTES;
                const unsigned n1g = (unsigned)(seg0 + n1);
                bar_sync(WBAR_FULL + b, WB_THREADS + WG_THREADS);
#pragma unroll
                for (int q4 = 0; q4 < WSPT / 4; q4++) {
                    int pq[4];
                    asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];"
                                 : "=r"(pq[0]), "=r"(pq[1]), "=r"(pq[2]), "=r"(pq[3])
                                 : "r"(ptab + q4 * WG_THREADS * 16));
#pragma unroll
                    for (int k = 0; k < 4; k++) {
                        const double y = lds_f64(buf + ((unsigned)pq[k] & 0xFFFFu));
                        const unsigned par = (unsigned)__popc(((unsigned)pq[k] >> 16) & n1g) << 31;
                        acc[q4 * 4 + k] += xor_sign(y, par);
                    }
                }
                if (it + WNBUF < total)           // item it+3 (other parity) refills this buffer
                    bar_arrive(WBAR_EMPTY + 2 * b + (int)((it + 1) & 1), WB_THREADS + WG_THREADS);
            }
#pragma unroll
            for (int q = 0; q < WSPT; q++) {
                const int i = slot_out[q * WG_THREADS + g];
                if (i >= 0) OUT[(size_t)i * oi + j * oj] = scale * acc[q];
            }
        }
    }
}

// Host side of the gather layout: distributes the samples of one pass (<= 8192) over the
// [WSPT][WG_THREADS] slots so that each aligned group of 16 slots (= the lanes of a half warp in
// one gather step) holds samples from 16 different shared-memory bank pairs wherever the
// sample set allows it.  slot_sample / slot_out get npasses * WSPT * WG_THREADS entries.
void fjlt_build_wht_slots(const int32_t *samples, size_t S, std::vector<int32_t> &slot_sample,
                          std::vector<int32_t> &slot_out) {
    const size_t per_pass = (size_t)WG_THREADS * WSPT;
    const size_t npass = S == 0 ? 0 : (S + per_pass - 1) / per_pass;
    slot_sample.assign(npass * per_pass, 0);
    slot_out.assign(npass * per_pass, -1);
    auto first_word = [](int32_t r) { return swz(r & (WSEG - 1)); };
    for (size_t p = 0; p < npass; p++) {
        const size_t s0 = p * per_pass, s1 = std::min(S, s0 + per_pass);
        std::vector<std::vector<int32_t>> q(16);
        for (size_t i = s0; i < s1; i++) q[first_word(samples[i]) & 15].push_back((int32_t)i);
        std::vector<int32_t> leftovers;
        int32_t *ss = slot_sample.data() + p * per_pass, *so = slot_out.data() + p * per_pass;
        const size_t ngroups = per_pass / 16;
        std::vector<size_t> holes;
        for (size_t g = 0; g < ngroups; g++)
            for (int b = 0; b < 16; b++) {
                if (!q[b].empty()) { so[g * 16 + b] = q[b].back(); q[b].pop_back(); }
                else holes.push_back(g * 16 + b);
            }
        for (int b = 0; b < 16; b++) for (int32_t i : q[b]) leftovers.push_back(i);
        for (size_t h = 0; h < holes.size() && h < leftovers.size(); h++) so[holes[h]] = leftovers[h];
        for (size_t t = 0; t < per_pass; t++)
            if (so[t] >= 0) {
                const int32_t r = samples[so[t]];
                const uint32_t k1 = (uint32_t)(r >> WSEG_LOG);
                ss[t] = (int32_t)((k1 << 16) | (uint32_t)(first_word(r) * 8));
            }
    }
}

// Tiled transpose: T[k + j*N] = X[k*ldx + j]  (rows layout -> contiguous columns).  The long
// extent (N / 32 tiles) rides in grid.x; grid.y is capped and strides over the column tiles.
__global__ void k_transpose_to_cols(const double *__restrict__ X, size_t ldx, size_t N, size_t m,
                                    double *__restrict__ T) {
    __shared__ double tile[32][33];
    const size_t k0 = (size_t)blockIdx.x * 32;
    for (size_t j0 = (size_t)blockIdx.y * 32; j0 < m; j0 += (size_t)gridDim.y * 32) {
        for (int r = threadIdx.y; r < 32; r += blockDim.y) {
            size_t k = k0 + r, j = j0 + threadIdx.x;
            tile[r][threadIdx.x] = (k < N && j < m) ? X[k * ldx + j] : 0.0;
        }
        __syncthreads();
        for (int r = threadIdx.y; r < 32; r += blockDim.y) {
            size_t j = j0 + r, k = k0 + threadIdx.x;
            if (k < N && j < m) T[k + j * N] = tile[threadIdx.x][r];
        }
        __syncthreads();
    }
}
static void launch_transpose_to_cols(const double *X, size_t ldx, size_t N, size_t m, double *T,
                                     cudaStream_t st) {
    const size_t ty = (m + 31) / 32;
    dim3 grid((unsigned)((N + 31) / 32), (unsigned)(ty < 65535 ? ty : 65535));
    k_transpose_to_cols<<<grid, dim3(32, 8), 0, st>>>(X, ldx, N, m, T);
}


// ---------------------------------------------------------------------------
// DCT-II mode (what the reference computes), N = 2^p >= 1024, contiguous columns (K8).
//   v = Makhoul permutation of x' = scale * D .* x;   z[n] = v[2n] + i v[2n+1]  (N' = N/2)
//   Z = DFT_N'(z) by a two-pass Cooley-Tukey split N' = N1 * N2 (n = n1 + N1 n2,
//   k = N2 k1 + k2): pass 1 = length-N2 FFTs over the strided index + twiddle W_N'^(n1 k2),
//   pass 2 = length-N1 FFTs over contiguous rows; both are radix-4/2 Stockham FFTs in
//   shared memory.  Only the S sampled bins are finished:
//   V[k] = E[k] + W_N^k O[k] (real-input split of Z), Y[k] = 2 Re(W_4N^k V[k]).
// Columns are processed in batches small enough that the N*8-byte intermediate of every
// column in flight stays in the 126 MB L2, so HBM sees each input element once.
// FP64 FFT arithmetic (~40 flop per element) is comparable to the HBM time of the element
// on this part, so this mode is compute/shared-memory bound, not HBM bound.
// ---------------------------------------------------------------------------
struct __align__(16) cplx { double x, y; };
__device__ __forceinline__ cplx cmul(cplx a, cplx b) {
    cplx r; r.x = a.x * b.x - a.y * b.y; r.y = a.x * b.y + a.y * b.x; return r;
}
__device__ __forceinline__ cplx cadd(cplx a, cplx b) { cplx r; r.x = a.x + b.x; r.y = a.y + b.y; return r; }
__device__ __forceinline__ cplx csub(cplx a, cplx b) { cplx r; r.x = a.x - b.x; r.y = a.y - b.y; return r; }

struct DctPlan {
    size_t N, Np, N1, N2;     // Np = N/2 = N1*N2
    int lo_bits;              // twiddle exponent split: e = hi << lo_bits | lo
    cplx *W1, *W2;            // W_N2[m], W_N1[m]
    cplx *Tlo, *Thi;          // W_Np^(lo), W_Np^(hi << lo_bits)
    cplx *post;               // per sample: W_N^k, W_4N^k
    size_t G1, G2;            // sequences per CTA in pass 1 / pass 2
    // register-blocked engine only: pass 2 keeps just the bins the samples need.  need = bitmap
    // over positions of H (k2*N1 + k1), rank = number of needed positions before each word,
    // Zc = per column the needed bins in position order (stride zc_ld)
    uint32_t *need, *rank;
    cplx *Zc; size_t zc_ld;
};

__device__ __forceinline__ size_t dct_pos(const DctPlan &pl, size_t k) {     // bin k -> position in H
    return (k & (pl.N2 - 1)) * pl.N1 + k / pl.N2;
}
__global__ void k_dct_mark(DctPlan pl, const int32_t *__restrict__ samples, size_t S) {
    size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= S) return;
    const size_t k = (size_t)samples[t], kp = k & (pl.Np - 1), km = (pl.Np - kp) & (pl.Np - 1);
    const size_t pa = dct_pos(pl, kp), pb = dct_pos(pl, km);
    atomicOr(pl.need + (pa >> 5), 1u << (pa & 31));
    atomicOr(pl.need + (pb >> 5), 1u << (pb & 31));
}
// exclusive prefix of the popcounts of `need` (one 1024-thread block)
__global__ void __launch_bounds__(1024) k_dct_rank(DctPlan pl) {
    __shared__ uint32_t part[1024];
    const size_t nwords = pl.Np >> 5, per = (nwords + 1023) / 1024;
    const size_t w0 = (size_t)threadIdx.x * per, w1 = w0 + per < nwords ? w0 + per : nwords;
    uint32_t c = 0;
    for (size_t w = w0; w < w1; w++) c += __popc(pl.need[w]);
    part[threadIdx.x] = c;
    __syncthreads();
    for (int off = 1; off < 1024; off <<= 1) {
        uint32_t v = threadIdx.x >= (unsigned)off ? part[threadIdx.x - off] : 0;
        __syncthreads();
        part[threadIdx.x] += v;
        __syncthreads();
    }
    uint32_t run = part[threadIdx.x] - c;
    for (size_t w = w0; w < w1; w++) { pl.rank[w] = run; run += __popc(pl.need[w]); }
}
__device__ __forceinline__ uint32_t dct_rank_of(const DctPlan &pl, size_t p) {
    const uint32_t w = __ldg(pl.need + (p >> 5));
    return __ldg(pl.rank + (p >> 5)) + __popc(w & ((1u << (p & 31)) - 1u));
}

__global__ void k_dct_tables(DctPlan pl, const int32_t *__restrict__ samples, size_t S) {
    size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    double sn, cs;
    if (t < pl.N2) { sincospi(-2.0 * (double)t / (double)pl.N2, &sn, &cs); pl.W1[t].x = cs; pl.W1[t].y = sn; }
    if (t < pl.N1) { sincospi(-2.0 * (double)t / (double)pl.N1, &sn, &cs); pl.W2[t].x = cs; pl.W2[t].y = sn; }
    size_t nlo = (size_t)1 << pl.lo_bits;
    if (t < nlo) { sincospi(-2.0 * (double)t / (double)pl.Np, &sn, &cs); pl.Tlo[t].x = cs; pl.Tlo[t].y = sn; }
    size_t nhi = (pl.Np + nlo - 1) / nlo;
    if (t < nhi) {
        sincospi(-2.0 * (double)(t << pl.lo_bits) / (double)pl.Np, &sn, &cs);
        pl.Thi[t].x = cs; pl.Thi[t].y = sn;
    }
    if (t < S) {
        double k = (double)samples[t];
        sincospi(-2.0 * k / (double)pl.N, &sn, &cs); pl.post[2 * t].x = cs; pl.post[2 * t].y = sn;
        sincospi(-0.5 * k / (double)pl.N, &sn, &cs); pl.post[2 * t + 1].x = cs; pl.post[2 * t + 1].y = sn;
    }
}

// Shared-memory FFT buffers are addressed through pad(): one spare element after every 8,
// which spreads the power-of-two strides of the Stockham scatter over the banks.
__device__ __forceinline__ int pad(int i) { return i + (i >> 3); }
__host__ __device__ constexpr size_t padded(size_t n) { return n + (n >> 3) + 8; }

// One radix-R Stockham pass over G sequences of length L held in shared memory.
template <int R>
__device__ __forceinline__ void stockham_pass(const cplx *in, cplx *out, int L, int Ns, int G,
                                              const cplx *__restrict__ W) {
    const int per = L / R, units = G * per, step = L / (Ns * R), pshift = 31 - __clz(per);
    for (int u = threadIdx.x; u < units; u += blockDim.x) {
        const int g = u >> pshift, j = u & (per - 1), k = j & (Ns - 1);
        const int sbase = g * L + j;
        cplx a[R];
#pragma unroll
        for (int t = 0; t < R; t++) a[t] = in[pad(sbase + t * per)];
        if (Ns > 1) {
#pragma unroll
            for (int t = 1; t < R; t++) a[t] = cmul(a[t], W[t * k * step]);
        }
        const int dbase = g * L + (j - k) * R + k;
        if (R == 2) {
            out[pad(dbase)] = cadd(a[0], a[1]);
            out[pad(dbase + Ns)] = csub(a[0], a[1]);
        } else {
            cplx b0 = cadd(a[0], a[2]), b1 = csub(a[0], a[2]), b2 = cadd(a[1], a[3]);
            cplx d = csub(a[1], a[3]);
            cplx b3; b3.x = d.y; b3.y = -d.x;               // -i * (a1 - a3)
            out[pad(dbase)] = cadd(b0, b2);
            out[pad(dbase + Ns)] = cadd(b1, b3);
            out[pad(dbase + 2 * Ns)] = csub(b0, b2);
            out[pad(dbase + 3 * Ns)] = csub(b1, b3);
        }
    }
}

// Full FFT of G sequences of length L (power of two) in shared memory; returns the buffer
// that holds the result (ping-pong between a and b).
__device__ __forceinline__ cplx *smem_fft(cplx *a, cplx *b, int L, int G, const cplx *__restrict__ W) {
    int Ns = 1;
    cplx *in = a, *out = b;
    int logL = 31 - __clz(L);
    if (logL & 1) {
        stockham_pass<2>(in, out, L, Ns, G, W);
        __syncthreads();
        Ns = 2; cplx *t = in; in = out; out = t;
    }
    for (; Ns < L; Ns *= 4) {
        stockham_pass<4>(in, out, L, Ns, G, W);
        __syncthreads();
        cplx *t = in; in = out; out = t;
    }
    return in;
}

__device__ __forceinline__ double dct_xprime(const double *__restrict__ col,
                                             const uint32_t *__restrict__ Dbits, size_t q,
                                             double inscale) {
    uint32_t w = Dbits[q >> 5];
    return __dmul_rn((w >> (q & 31)) & 1 ? -inscale : inscale, col[q]);
}

// pass 1: grid (N1/G, columns in batch)
__global__ void __launch_bounds__(1024)
k_dct_pass1(const double *__restrict__ X, size_t ldx, size_t col0, DctPlan pl,
            const uint32_t *__restrict__ Dbits, double inscale, cplx *__restrict__ H) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int G = (int)pl.G1, N2 = (int)pl.N2;
    const size_t N = pl.N, N1 = pl.N1;
    cplx *a = reinterpret_cast<cplx *>(smem_raw), *b = a + padded((size_t)G * N2), *Ws = b + padded((size_t)G * N2);
    for (int i = threadIdx.x; i < N2; i += blockDim.x) Ws[i] = pl.W1[i];     // twiddles on chip
    const int gshift = 31 - __clz(G);                                         // G is a power of two
    const size_t n1_0 = (size_t)blockIdx.x * G;
    const double *col = X + (col0 + blockIdx.y) * ldx;
    cplx *Hc = H + (size_t)blockIdx.y * pl.Np;
    const int total = G * N2;
    // four elements per trip so that all their (independent) global loads are in flight together
    for (int base = threadIdx.x; base < total; base += 4 * blockDim.x) {
        size_t q0[4], q1[4]; uint32_t w0[4], w1[4]; double x0[4], x1[4];
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const int idx = base + u * blockDim.x;
            const int g = idx & (G - 1), n2 = idx >> gshift;
            const size_t n = n1_0 + g + N1 * (size_t)n2;
            const bool lowhalf = 4 * n < N;
            q0[u] = lowhalf ? 4 * n : 2 * N - 4 * n - 1;
            q1[u] = lowhalf ? 4 * n + 2 : 2 * N - 4 * n - 3;
            if (idx < total) {
                w0[u] = __ldg(Dbits + (q0[u] >> 5)); w1[u] = __ldg(Dbits + (q1[u] >> 5));
                x0[u] = __ldg(col + q0[u]); x1[u] = __ldg(col + q1[u]);
            }
        }
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const int idx = base + u * blockDim.x;
            if (idx < total) {
                cplx z;
                z.x = __dmul_rn((w0[u] >> (q0[u] & 31)) & 1 ? -inscale : inscale, x0[u]);
                z.y = __dmul_rn((w1[u] >> (q1[u] & 31)) & 1 ? -inscale : inscale, x1[u]);
                a[pad((idx & (G - 1)) * N2 + (idx >> gshift))] = z;
            }
        }
    }
    __syncthreads();
    cplx *res = smem_fft(a, b, N2, G, Ws);
    const size_t lo_mask = ((size_t)1 << pl.lo_bits) - 1;
#pragma unroll 4
    for (int idx = threadIdx.x; idx < total; idx += blockDim.x) {
        const int g = idx & (G - 1), k2 = idx >> gshift;
        const size_t n1 = n1_0 + g, e = n1 * (size_t)k2;       // < Np
        cplx tw = cmul(pl.Thi[e >> pl.lo_bits], pl.Tlo[e & lo_mask]);
        Hc[(size_t)k2 * N1 + n1] = cmul(res[pad(g * N2 + k2)], tw);
    }
}

// pass 2: grid (N2/G, columns in batch); rows of H are contiguous; in place
__global__ void __launch_bounds__(1024)
k_dct_pass2(DctPlan pl, cplx *__restrict__ H) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int G = (int)pl.G2, N1 = (int)pl.N1;
    cplx *a = reinterpret_cast<cplx *>(smem_raw), *b = a + padded((size_t)G * N1), *Ws = b + padded((size_t)G * N1);
    for (int i = threadIdx.x; i < N1; i += blockDim.x) Ws[i] = pl.W2[i];
    cplx *rows = H + (size_t)blockIdx.y * pl.Np + (size_t)blockIdx.x * G * N1;
#pragma unroll 4
    for (int idx = threadIdx.x; idx < G * N1; idx += blockDim.x) a[pad(idx)] = rows[idx];
    __syncthreads();
    cplx *res = smem_fft(a, b, N1, G, Ws);
#pragma unroll 4
    for (int idx = threadIdx.x; idx < G * N1; idx += blockDim.x) rows[idx] = res[pad(idx)];
}


// ---------------------------------------------------------------------------
// Register-blocked FFT engine used for N >= 8192: a CTA of 256 threads transforms 4096 complex
// points (G sequences of length L = 4096 / G) IN PLACE in shared memory.  Every thread holds 16
// points per pass and does a radix-16 butterfly in registers (radix 8 / 4 / 2 for the last
// pass), so a 1024-point transform is 3 passes (16 * 16 * 4) instead of 5 radix-4 ones, with
// 16 independent loads, twiddles and stores per thread in flight.
// ---------------------------------------------------------------------------
// one spare element after every 16, one more after every 256 (sequences of a CTA start on different banks)
__device__ __forceinline__ int pad16(int i) { return i + (i >> 4) + (i >> 8); }
#ifndef SKB_FFT_PPT
#define SKB_FFT_PPT 16          // points a thread holds per pass (16: radix-16 passes, 8: radix-8)
#endif
#ifndef SKB_FFT_MINB
#define SKB_FFT_MINB 2
#endif
#ifndef SKB_FFT_TWPOW
#define SKB_FFT_TWPOW 1         // twiddles of a butterfly as powers of one table entry
#endif
constexpr int FFT_PPT = SKB_FFT_PPT;
constexpr size_t FFT_POINTS = 4096, FFT_THREADS = FFT_POINTS / FFT_PPT;
constexpr size_t FFT_SMEM = (FFT_POINTS + (FFT_POINTS >> 4) + (FFT_POINTS >> 8) + 16) * sizeof(cplx);

__device__ __forceinline__ cplx ldg_cplx(const cplx *p) {
    double2 v = __ldg(reinterpret_cast<const double2 *>(p));
    cplx r; r.x = v.x; r.y = v.y; return r;
}
__device__ __forceinline__ cplx mul_mi(cplx a) { cplx r; r.x = a.y; r.y = -a.x; return r; }   // * (-i)

__device__ __forceinline__ void dft4(cplx &x0, cplx &x1, cplx &x2, cplx &x3) {
    cplx t0 = cadd(x0, x2), t1 = csub(x0, x2), t2 = cadd(x1, x3), t3 = mul_mi(csub(x1, x3));
    x0 = cadd(t0, t2); x1 = cadd(t1, t3); x2 = csub(t0, t2); x3 = csub(t1, t3);
}

template <int M> __device__ __forceinline__ cplx mul_w16(cplx a) {      // a * exp(-2 pi i M / 16)
    const double c = 0.92387953251128673848, s = 0.38268343236508978178, h = 0.70710678118654752440;
    cplx r;
    if (M == 0) return a;
    if (M == 1) { r.x = a.x * c + a.y * s; r.y = a.y * c - a.x * s; }
    if (M == 2) { r.x = h * (a.x + a.y); r.y = h * (a.y - a.x); }
    if (M == 3) { r.x = a.x * s + a.y * c; r.y = a.y * s - a.x * c; }
    if (M == 4) { r.x = a.y; r.y = -a.x; }
    if (M == 6) { r.x = h * (a.y - a.x); r.y = -h * (a.x + a.y); }
    if (M == 9) { r.x = -(a.x * c + a.y * s); r.y = -(a.y * c - a.x * s); }
    return r;
}

template <int R> __device__ __forceinline__ void butterfly(cplx *a);
template <> __device__ __forceinline__ void butterfly<2>(cplx *a) {
    cplx t = cadd(a[0], a[1]); a[1] = csub(a[0], a[1]); a[0] = t;
}
template <> __device__ __forceinline__ void butterfly<4>(cplx *a) { dft4(a[0], a[1], a[2], a[3]); }
template <> __device__ __forceinline__ void butterfly<8>(cplx *a) {
    cplx e0 = a[0], e1 = a[2], e2 = a[4], e3 = a[6], o0 = a[1], o1 = a[3], o2 = a[5], o3 = a[7];
    dft4(e0, e1, e2, e3);
    dft4(o0, o1, o2, o3);
    o1 = mul_w16<2>(o1); o2 = mul_w16<4>(o2); o3 = mul_w16<6>(o3);
    a[0] = cadd(e0, o0); a[4] = csub(e0, o0);
    a[1] = cadd(e1, o1); a[5] = csub(e1, o1);
    a[2] = cadd(e2, o2); a[6] = csub(e2, o2);
    a[3] = cadd(e3, o3); a[7] = csub(e3, o3);
}
template <> __device__ __forceinline__ void butterfly<16>(cplx *a) {
    // n = 4 n1 + n2, k = k1 + 4 k2:  B[n2][k1] = DFT4_n1 a[4 n1 + n2];  B *= W16^(n2 k1);
    // X[k1 + 4 k2] = DFT4_n2 B[n2][k1]
    dft4(a[0], a[4], a[8], a[12]);      // n2 = 0: a[4 k1 + 0] now holds B[0][k1]
    dft4(a[1], a[5], a[9], a[13]);
    dft4(a[2], a[6], a[10], a[14]);
    dft4(a[3], a[7], a[11], a[15]);
    a[5] = mul_w16<1>(a[5]);  a[9] = mul_w16<2>(a[9]);   a[13] = mul_w16<3>(a[13]);   // n2 = 1
    a[6] = mul_w16<2>(a[6]);  a[10] = mul_w16<4>(a[10]); a[14] = mul_w16<6>(a[14]);   // n2 = 2
    a[7] = mul_w16<3>(a[7]);  a[11] = mul_w16<6>(a[11]); a[15] = mul_w16<9>(a[15]);   // n2 = 3
    // now a[4 k1 + n2] = B'[n2][k1]; transform over n2 and place X[k1 + 4 k2]
    cplx x[16];
#pragma unroll
    for (int k1 = 0; k1 < 4; k1++) {
        cplx b0 = a[4 * k1], b1 = a[4 * k1 + 1], b2 = a[4 * k1 + 2], b3 = a[4 * k1 + 3];
        dft4(b0, b1, b2, b3);
        x[k1] = b0; x[k1 + 4] = b1; x[k1 + 8] = b2; x[k1 + 12] = b3;
    }
#pragma unroll
    for (int i = 0; i < 16; i++) a[i] = x[i];
}

template <int R>
__device__ __forceinline__ void inplace_pass(cplx *buf, int L, int Ns, const cplx *__restrict__ W) {
    constexpr int U = FFT_PPT / R;       // butterflies per thread
    const int per = L / R, pshift = 31 - __clz(per), step = L / (Ns * R);
    cplx a[U][R];
    int dbase[U];
#pragma unroll
    for (int uu = 0; uu < U; uu++) {
        const int u = (int)threadIdx.x + uu * (int)FFT_THREADS;
        const int g = u >> pshift, j = u & (per - 1), k = j & (Ns - 1);
        const int sbase = g * L + j;
#pragma unroll
        for (int t = 0; t < R; t++) a[uu][t] = buf[pad16(sbase + t * per)];
        if (Ns > 1) {
#if SKB_FFT_TWPOW
            // one table read per butterfly; the other R-2 twiddles are its powers (product tree of
            // depth <= 4, a few ulp): the FP64 pipe is idle here, the load latency is not
            cplx wp[R];
            wp[1] = ldg_cplx(W + k * step);
#pragma unroll
            for (int t = 2; t < R; t++) wp[t] = (t & 1) ? cmul(wp[t - 1], wp[1]) : cmul(wp[t / 2], wp[t / 2]);
#pragma unroll
            for (int t = 1; t < R; t++) a[uu][t] = cmul(a[uu][t], wp[t]);
#else
#pragma unroll
            for (int t = 1; t < R; t++) a[uu][t] = cmul(a[uu][t], ldg_cplx(W + t * k * step));
#endif
        }
        dbase[uu] = g * L + (j - k) * R + k;
    }
    __syncthreads();                     // every input is in registers: the buffer may be overwritten
#pragma unroll
    for (int uu = 0; uu < U; uu++) {
        butterfly<R>(a[uu]);
#pragma unroll
        for (int q = 0; q < R; q++) buf[pad16(dbase[uu] + q * Ns)] = a[uu][q];
    }
    __syncthreads();
}

// In-place FFT of 4096 / L sequences of length L (16 <= L <= 4096, power of two); W = W_L table.
__device__ __forceinline__ void smem_fft16(cplx *buf, int L, const cplx *__restrict__ W) {
    int Ns = 1;
    for (; L / Ns >= FFT_PPT; Ns *= FFT_PPT) inplace_pass<FFT_PPT>(buf, L, Ns, W);
    const int rem = L / Ns;
    if (FFT_PPT > 8 && rem == 8) inplace_pass<(FFT_PPT > 8 ? 8 : 2)>(buf, L, Ns, W);
    else if (rem == 4) inplace_pass<4>(buf, L, Ns, W);
    else if (rem == 2) inplace_pass<2>(buf, L, Ns, W);
}

// pass 1 (register-blocked engine): grid (N1 / G, columns in batch), G = 4096 / N2 sequences
// per CTA.  The Makhoul permutation puts z[n] = (x[4n], x[4n+2]) and its mirror
// z[N'-1-n] = (x[4n+3], x[4n+1]) in the same 32-byte sector, and the mirror of sequence n1,
// position n2 is sequence N1-1-n1, position N2-1-n2 -- so a CTA takes G/2 sequences and their
// G/2 mirrors and every sector it touches is used whole (one 4-double read per pair of points).
template <bool AL32>
__global__ void __launch_bounds__(FFT_THREADS, SKB_FFT_MINB)
k_dct_pass1_r16(const double *__restrict__ X, size_t ldx, size_t col0, DctPlan pl,
                const uint32_t *__restrict__ Dbits, double inscale, cplx *__restrict__ H) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    cplx *buf = reinterpret_cast<cplx *>(smem_raw);
    const int N2 = (int)pl.N2, N1 = (int)pl.N1, G = (int)(FFT_POINTS / pl.N2), gshift = 31 - __clz(G);
    const int Gh = G >> 1, a = (int)blockIdx.x * Gh;
    const double *col = X + (col0 + blockIdx.y) * ldx;
    cplx *Hc = H + (size_t)blockIdx.y * pl.Np;
    auto seq = [&](int g) { return g < Gh ? a + g : N1 - 1 - a - (g - Gh); };   // sequence id of slot g
    constexpr int PAIRS = (int)(FFT_POINTS / 2 / FFT_THREADS);      // 8 pairs of points per thread
#pragma unroll
    for (int b4 = 0; b4 < PAIRS / 4; b4++) {
        double x[4][4]; uint32_t w[4];
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const int p = (int)threadIdx.x + (b4 * 4 + u) * (int)FFT_THREADS;
            const int g = p & (G - 1), n2 = p >> gshift;                // n2 < N2/2: the lower half
            const unsigned q = 4u * (unsigned)(seq(g) + N1 * n2);       // < N <= 2^25
            w[u] = __ldg(Dbits + (q >> 5)) >> (q & 31);
            if (AL32) {
                double2 lo = ldg_stream2(col + q), hi = ldg_stream2(col + q + 2);
                x[u][0] = lo.x; x[u][1] = lo.y; x[u][2] = hi.x; x[u][3] = hi.y;
            } else {
#pragma unroll
                for (int e = 0; e < 4; e++) x[u][e] = ldg_stream1(col + q + e);
            }
        }
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const int p = (int)threadIdx.x + (b4 * 4 + u) * (int)FFT_THREADS;
            const int g = p & (G - 1), n2 = p >> gshift;
            double sx[4];
#pragma unroll
            for (int e = 0; e < 4; e++) sx[e] = __dmul_rn((w[u] >> e) & 1 ? -inscale : inscale, x[u][e]);
            cplx z, zm;
            z.x = sx[0]; z.y = sx[2];            // z[n]
            zm.x = sx[3]; zm.y = sx[1];          // z[N'-1-n]
            buf[pad16(g * N2 + n2)] = z;
            buf[pad16((g ^ Gh) * N2 + (N2 - 1 - n2))] = zm;
        }
    }
    __syncthreads();
    smem_fft16(buf, N2, pl.W1);
    // twiddle W_N'^(n1*k2): k2 advances by a fixed step per iteration, so the factor is refreshed
    // from the tables every fourth point and advanced by one multiplication in between
    const size_t lo_mask = ((size_t)1 << pl.lo_bits) - 1;
    const int g = (int)threadIdx.x & (G - 1), n1 = seq(g);
    const int k2_0 = (int)threadIdx.x >> gshift, dk = (int)FFT_THREADS >> gshift;
    auto table_tw = [&](size_t e) {
        e &= pl.Np - 1;
        return cmul(ldg_cplx(pl.Thi + (e >> pl.lo_bits)), ldg_cplx(pl.Tlo + (e & lo_mask)));
    };
    const cplx wstep = table_tw((size_t)n1 * (size_t)dk);
#pragma unroll
    for (int i4 = 0; i4 < FFT_PPT / 4; i4++) {
        cplx tw = table_tw((size_t)n1 * (size_t)(k2_0 + i4 * 4 * dk));
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const int k2 = k2_0 + (i4 * 4 + u) * dk;
            Hc[(size_t)k2 * N1 + n1] = cmul(buf[pad16(g * N2 + k2)], tw);
            tw = cmul(tw, wstep);
        }
    }
}

// pass 2 (register-blocked engine): grid (N2 / G, columns in batch), G = 4096 / N1 contiguous rows per CTA,
// keeping only the bins the samples need (see DctPlan::need)
__global__ void __launch_bounds__(FFT_THREADS, SKB_FFT_MINB)
k_dct_pass2_r16_sel(DctPlan pl, const cplx *__restrict__ H) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    cplx *buf = reinterpret_cast<cplx *>(smem_raw);
    const size_t base = (size_t)blockIdx.x * FFT_POINTS;
    const cplx *rows = H + (size_t)blockIdx.y * pl.Np + base;
#pragma unroll
    for (int it = 0; it < FFT_PPT; it++) {
        const int idx = (int)threadIdx.x + it * (int)FFT_THREADS;
        buf[pad16(idx)] = rows[idx];
    }
    __syncthreads();
    smem_fft16(buf, (int)pl.N1, pl.W2);
    cplx *zc = pl.Zc + (size_t)blockIdx.y * pl.zc_ld;
#pragma unroll
    for (int it = 0; it < FFT_PPT; it++) {
        const int idx = (int)threadIdx.x + it * (int)FFT_THREADS;
        const size_t p = base + idx;
        const uint32_t w = __ldg(pl.need + (p >> 5));
        if ((w >> (p & 31)) & 1u)
            zc[__ldg(pl.rank + (p >> 5)) + __popc(w & ((1u << (p & 31)) - 1u))] = buf[pad16(idx)];
    }
}

__global__ void k_dct_post_sel(DctPlan pl, const int32_t *__restrict__ samples,
                               size_t S, size_t col0, size_t ncols, double outscale,
                               double *__restrict__ OUT, size_t oi, size_t oj) {
    size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= S * ncols) return;
    const size_t i = t % S, c = t / S;
    const size_t k = (size_t)samples[i], kp = k & (pl.Np - 1), km = (pl.Np - kp) & (pl.Np - 1);
    const cplx *Z = pl.Zc + c * pl.zc_ld;
    const cplx za = Z[dct_rank_of(pl, dct_pos(pl, kp))];
    const cplx zb = Z[dct_rank_of(pl, dct_pos(pl, km))];
    cplx E, O;
    E.x = 0.5 * (za.x + zb.x); E.y = 0.5 * (za.y - zb.y);
    O.x = 0.5 * (za.y + zb.y); O.y = -0.5 * (za.x - zb.x);
    cplx V = cadd(E, cmul(pl.post[2 * i], O));
    cplx Yc = cmul(pl.post[2 * i + 1], V);
    OUT[i * oi + (col0 + c) * oj] = outscale * (2.0 * Yc.x);
}

// sampled bins -> DCT values -> OUT
__global__ void k_dct_post(DctPlan pl, const cplx *__restrict__ H, const int32_t *__restrict__ samples,
                           size_t S, size_t col0, size_t ncols, double outscale,
                           double *__restrict__ OUT, size_t oi, size_t oj) {
    size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= S * ncols) return;
    const size_t i = t % S, c = t / S;
    const size_t k = (size_t)samples[i], kp = k & (pl.Np - 1), km = (pl.Np - kp) & (pl.Np - 1);
    const cplx *Z = H + c * pl.Np;
    const cplx za = Z[(kp & (pl.N2 - 1)) * pl.N1 + kp / pl.N2];
    const cplx zb = Z[(km & (pl.N2 - 1)) * pl.N1 + km / pl.N2];
    cplx E, O;
    E.x = 0.5 * (za.x + zb.x); E.y = 0.5 * (za.y - zb.y);
    O.x = 0.5 * (za.y + zb.y); O.y = -0.5 * (za.x - zb.x);
    cplx V = cadd(E, cmul(pl.post[2 * i], O));
    cplx Yc = cmul(pl.post[2 * i + 1], V);
    OUT[i * oi + (col0 + c) * oj] = outscale * (2.0 * Yc.x);
}

static bool dct_fast_ok(size_t N) { return N >= 1024 && (N & (N - 1)) == 0 && N <= ((size_t)1 << 25); }

static size_t dct_batch_cols(size_t N, size_t m) {
    static size_t mb = 0;
    if (!mb) { const char *e = getenv("SKB_DCT_BATCH_MB"); mb = e && atol(e) > 0 ? (size_t)atol(e) : 512; }
    size_t b = (mb << 20) / (N * sizeof(double));
    if (b < 1) b = 1;
    if (b > m) b = m;
    if (b > 65535) b = 65535;      // a batch rides in grid.y
    return b;
}

static size_t dct_table_bytes(size_t N, size_t S) {
    return (N + 4096 + 2 * S + 1024) * sizeof(cplx);     // generous upper bound
}
static size_t dct_sel_bytes(size_t N, size_t S, size_t B) {     // need bitmap, ranks, kept bins
    return (N / 2 / 32 + 64) * 2 * sizeof(uint32_t) + (B * 2 * S + 64) * sizeof(cplx) + 512;
}

static bool wht_fast_ok(size_t N) { return N >= (size_t)WSEG && (N & (N - 1)) == 0 && N <= ((size_t)1 << 28); }

size_t fjlt_workspace_bytes(size_t N, size_t m, int fut, bool layout_rows, size_t S) {
    if (fut == FUT_DCT && dct_fast_ok(N))
        return (layout_rows ? N * m * sizeof(double) : 0) + dct_batch_cols(N, m) * N * sizeof(double) +
               dct_table_bytes(N, S) + dct_sel_bytes(N, S, dct_batch_cols(N, m)) + 1024;
    if (fut == FUT_DCT) return 4 * N * sizeof(double) + 256;
    if (fut == FUT_WHT && wht_fast_ok(N))         // sign words (+ transposed copy)
        return (layout_rows ? (N * m * sizeof(double) + 255) / 256 * 256 : 0) + N / 8 + 256;
    return 256;
}

bool fjlt_shard_ok(int fut, size_t N, size_t offset, size_t nloc) {
    return fut == FUT_WHT && N >= (size_t)WSEG && (N & (N - 1)) == 0 && N <= ((size_t)1 << 28) &&
           offset % WSEG == 0 && nloc % WSEG == 0 && nloc > 0;
}

void fjlt_apply(int fut, bool layout_rows, const double *X, size_t ldx, size_t N, size_t m,
                const double *D, const uint32_t *Dbits, const int32_t *samples, size_t S,
                const int32_t *wht_slot_sample, const int32_t *wht_slot_out,
                double *OUT, size_t ldo, void *workspace, size_t workspace_bytes,
                cudaStream_t st, size_t offset, size_t nloc) {
    if (S == 0 || m == 0) return;
    if (nloc != N && !fjlt_shard_ok(fut, N, offset, nloc))
        throw error(SKB_ERR_INVALID_USAGE,
                    "FJLT: a shard of the sketched dimension needs the Walsh-Hadamard mode, N a power "
                    "of two >= 8192, and shard offset and length that are multiples of 8192");
    if (fut == FUT_WHT && (N & (N - 1)) != 0)
        throw error(SKB_ERR_INVALID_PARAM, "FJLT (WHT mode): N must be a power of two");
    if (workspace_bytes < fjlt_workspace_bytes(N, m, fut, layout_rows, S))
        throw error(SKB_ERR_ALLOC, "fjlt apply: workspace too small");
    if (fut == FUT_DCT && dct_fast_ok(N)) {
        char *ws = reinterpret_cast<char *>(workspace);
        const double *Xc = X;
        size_t ldc = ldx;
        if (layout_rows) {      // the reference also transposes for this direction
            double *T = reinterpret_cast<double *>(ws);
            ws += ((N * m * sizeof(double) + 255) / 256) * 256;
            LaunchScope ls(KC_FJLT, st);
            launch_transpose_to_cols(X, ldx, N, m, T, st);
            SKB_CUDA(cudaGetLastError());
            Xc = T; ldc = N;
        }
        DctPlan pl;
        pl.N = N; pl.Np = N / 2;
        int lg = 0; while (((size_t)1 << lg) < pl.Np) lg++;
        pl.N1 = (size_t)1 << (lg / 2);
        // long contiguous rows for pass 2 and 256-point strided FFTs for pass 1 (16 sequences per
        // CTA -> 128-byte runs in its loads and stores) when the sizes allow: measured 18.1 ms
        // against 21.6 ms for the balanced split at N = 2^20
        if (lg >= 16 && lg - 8 <= 12) pl.N1 = (size_t)1 << (lg - 8);
        {   // debug knob: log2 of the contiguous (pass 2) FFT length
            static const char *e = getenv("SKB_DCT_N1LOG");
            if (e && atoi(e) > 0 && atoi(e) < lg) pl.N1 = (size_t)1 << atoi(e);
        }
        pl.N2 = pl.Np / pl.N1;
        pl.lo_bits = 10;
        const size_t B = dct_batch_cols(N, m);
        cplx *H = reinterpret_cast<cplx *>(ws);
        cplx *tab = H + B * pl.Np;
        pl.W1 = tab; tab += pl.N2;
        pl.W2 = tab; tab += pl.N1;
        pl.Tlo = tab; tab += (size_t)1 << pl.lo_bits;
        pl.Thi = tab; tab += (pl.Np >> pl.lo_bits) + 1;
        pl.post = tab; tab += 2 * S;
        pl.need = reinterpret_cast<uint32_t *>(tab);
        pl.rank = pl.need + (pl.Np / 32 + 64);
        pl.Zc = reinterpret_cast<cplx *>((reinterpret_cast<uintptr_t>(pl.rank + (pl.Np / 32 + 64)) + 255) / 256 * 256);
        pl.zc_ld = 2 * S;
        // sequences per CTA: G*L = 2048 points -> two 32 KB ping-pong buffers, 512 threads with
        // one radix-4 unit each per pass, three CTAs per SM
        auto seqs = [](size_t L, size_t other) {
            size_t g = 2048 / L; if (g < 1) g = 1; if (g > 16) g = 16;
            while (g > 1 && other % g) g /= 2;
            return g;
        };
        pl.G1 = seqs(pl.N2, pl.N1); pl.G2 = seqs(pl.N1, pl.N2);
        const size_t smem1 = (2 * padded(pl.G1 * pl.N2) + pl.N2) * sizeof(cplx), smem2 = (2 * padded(pl.G2 * pl.N1) + pl.N1) * sizeof(cplx);
        SKB_CUDA(cudaFuncSetAttribute(k_dct_pass1, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem1));
        SKB_CUDA(cudaFuncSetAttribute(k_dct_pass2, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem2));
        {
            size_t nt = pl.N2 > S ? pl.N2 : S;
            if (nt < pl.N1) nt = pl.N1;
            if (nt < 1024) nt = 1024;
            if (nt < (pl.Np >> pl.lo_bits) + 1) nt = (pl.Np >> pl.lo_bits) + 1;
            LaunchScope ls(KC_FJLT, st);
            k_dct_tables<<<(unsigned)((nt + 255) / 256), 256, 0, st>>>(pl, samples, S);
            SKB_CUDA(cudaGetLastError());
        }
        const size_t oi = layout_rows ? ldo : 1, oj = layout_rows ? 1 : ldo;
        const double inscale = 1.0 / sqrt(2.0 * (double)N), outscale = sqrt((double)N / (double)S);
        // one radix-4 unit per thread per pass where possible (latency hiding by occupancy)
        auto threads_for = [](size_t units) { size_t t = 128; while (t < units && t < 1024) t *= 2; return (unsigned)t; };
        const unsigned thr1 = threads_for(pl.G1 * pl.N2 / 4), thr2 = threads_for(pl.G2 * pl.N1 / 4);
        // register-blocked radix-16 engine when a CTA's 4096 points hold whole sequences
        const bool r16 = pl.N1 >= (size_t)FFT_PPT && pl.N2 >= (size_t)FFT_PPT && pl.N1 <= FFT_POINTS && pl.N2 <= FFT_POINTS / 2 &&
                         pl.N1 * pl.N2 >= FFT_POINTS;
        if (r16) {
            SKB_CUDA(cudaFuncSetAttribute(k_dct_pass1_r16<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)FFT_SMEM));
            SKB_CUDA(cudaFuncSetAttribute(k_dct_pass1_r16<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)FFT_SMEM));
            SKB_CUDA(cudaFuncSetAttribute(k_dct_pass2_r16_sel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)FFT_SMEM));
            SKB_CUDA(cudaMemsetAsync(pl.need, 0, (pl.Np / 32) * sizeof(uint32_t), st));
            {
                LaunchScope ls(KC_FJLT, st);
                k_dct_mark<<<(unsigned)((S + 255) / 256), 256, 0, st>>>(pl, samples, S);
            }
            {
                LaunchScope ls(KC_FJLT, st);
                k_dct_rank<<<1, 1024, 0, st>>>(pl);
            }
        }
        for (size_t c0 = 0; c0 < m; c0 += B) {
            const size_t nb = m - c0 < B ? m - c0 : B;
            if (r16) {
                {
                    LaunchScope ls(KC_FJLT, st);
                    const bool al32 = (ldc % 4 == 0) && ((uintptr_t)Xc % 32 == 0);
                    auto k1 = al32 ? k_dct_pass1_r16<true> : k_dct_pass1_r16<false>;
                    k1<<<dim3((unsigned)(pl.N1 * pl.N2 / FFT_POINTS), (unsigned)nb),
                         FFT_THREADS, FFT_SMEM, st>>>(Xc, ldc, c0, pl, Dbits, inscale, H);
                }
                {
                    LaunchScope ls(KC_FJLT, st);
                    k_dct_pass2_r16_sel<<<dim3((unsigned)(pl.N1 * pl.N2 / FFT_POINTS), (unsigned)nb),
                                          FFT_THREADS, FFT_SMEM, st>>>(pl, H);
                }
                {
                    LaunchScope ls(KC_FJLT, st);
                    k_dct_post_sel<<<(unsigned)((S * nb + 255) / 256), 256, 0, st>>>(
                        pl, samples, S, c0, nb, outscale, OUT, oi, oj);
                }
                SKB_CUDA(cudaGetLastError());
                continue;
            } else {
            {
                LaunchScope ls(KC_FJLT, st);
                k_dct_pass1<<<dim3((unsigned)(pl.N1 / pl.G1), (unsigned)nb), thr1, smem1, st>>>(
                    Xc, ldc, c0, pl, Dbits, inscale, H);
            }
            {
                LaunchScope ls(KC_FJLT, st);
                k_dct_pass2<<<dim3((unsigned)(pl.N2 / pl.G2), (unsigned)nb), thr2, smem2, st>>>(pl, H);
            }
            }
            {
                LaunchScope ls(KC_FJLT, st);
                k_dct_post<<<(unsigned)((S * nb + 255) / 256), 256, 0, st>>>(
                    pl, H, samples, S, c0, nb, outscale, OUT, oi, oj);
            }
            SKB_CUDA(cudaGetLastError());
        }
        return;
    }
    if (fut == FUT_WHT && wht_fast_ok(N)) {
        const double *Xc = X;
        size_t ldc = ldx;
        char *ws = reinterpret_cast<char *>(workspace);
        if (layout_rows) {      // the reference also transposes for this direction
            double *T = reinterpret_cast<double *>(ws);
            ws += (nloc * m * sizeof(double) + 255) / 256 * 256;
            LaunchScope ls(KC_FJLT, st);
            launch_transpose_to_cols(X, ldx, nloc, m, T, st);
            SKB_CUDA(cudaGetLastError());
            Xc = T; ldc = nloc;
        }
        uint64_t *Dwht = reinterpret_cast<uint64_t *>(ws);      // N/64 words, rebuilt per apply (~2 us)
        {
            LaunchScope ls(KC_FJLT, st);
            const size_t nwords = N >> 6;
            k_wht_sign_words<<<(unsigned)((nwords + 255) / 256), 256, 0, st>>>(Dbits, nwords, Dwht);
            SKB_CUDA(cudaGetLastError());
        }
        const size_t oi = layout_rows ? ldo : 1, oj = layout_rows ? 1 : ldo;
        const double inscale = 1.0 / sqrt((double)N), outscale = sqrt((double)N / (double)S);
        const size_t smem = (size_t)WSMEM_BYTES;
        const bool al16 = (ldc % 2 == 0) && ((uintptr_t)Xc % 16 == 0);
        auto kern = al16 ? k_fjlt_wht_cols<true> : k_fjlt_wht_cols<false>;
        SKB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        unsigned grid = (unsigned)(m < 148 ? m : 148);
        const size_t per_pass = (size_t)WG_THREADS * WSPT;               // 8192 samples per pass
        for (size_t s0 = 0, p = 0; s0 < S; s0 += per_pass, p++) {
            LaunchScope ls(KC_FJLT, st);
            kern<<<grid, WTHREADS, smem, st>>>(Xc, ldc, nloc, m, Dwht, wht_slot_sample + p * per_pass,
                                               wht_slot_out + p * per_pass, offset >> WSEG_LOG,
                                               inscale, outscale, OUT, oi, oj);
            SKB_CUDA(cudaGetLastError());
        }
        return;
    }
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
