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

// Fused reduce-scatter of a row-sharded sketch (multi-GPU): instead of its own OUT, the WHT kernel
// sends the finished column j of this rank's partial sketch to the GPU that owns column j, into slab
// `rank` of that GPU's staging buffer (plain stores to peer memory over NVLink).  The gather warps
// first assemble the column in a local scratch line (their per-sample stores are scattered), then
// copy it out in coalesced 16-byte pieces.  ptrs == nullptr: single-GPU behaviour.
struct WhtPeerOut {
    double *const *ptrs;     // device array: staging buffer of every rank, [world][cols_per_owner][S]
    double *scratch;         // (number of CTAs) * S doubles, local
    int cols_per_owner, rank;
    int S;
};
static thread_local WhtPeerOut g_wht_peer = {nullptr, nullptr, 0, 0, 0};     // set around one apply
void fjlt_set_peer_out(double *const *ptrs, double *scratch, int cols_per_owner, int rank, int S) {
    g_wht_peer.ptrs = ptrs; g_wht_peer.scratch = scratch; g_wht_peer.cols_per_owner = cols_per_owner;
    g_wht_peer.rank = rank; g_wht_peer.S = S;
}
static int g_sm_reserve = 0;
void fjlt_set_sm_reserve(int sms) { g_sm_reserve = sms; }
static unsigned fjlt_sms() { return (unsigned)(148 - (g_sm_reserve < 140 ? g_sm_reserve : 140)); }

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
enum { WBAR_FULL = 1, WBAR_EMPTY = 4, WBAR_INNER = 10, WBAR_OUT = 12 };
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
                double *__restrict__ OUT, size_t oi, size_t oj, WhtPeerOut peer) {
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
            if (peer.ptrs == nullptr) {
#pragma unroll
                for (int q = 0; q < WSPT; q++) {
                    const int i = slot_out[q * WG_THREADS + g];
                    if (i >= 0) OUT[(size_t)i * oi + j * oj] = scale * acc[q];
                }
            } else {
                double *loc = peer.scratch + (size_t)blockIdx.x * peer.S;
#pragma unroll
                for (int q = 0; q < WSPT; q++) {
                    const int i = slot_out[q * WG_THREADS + g];
                    if (i >= 0) loc[i] = scale * acc[q];
                }
                bar_sync(WBAR_OUT, WG_THREADS);                 // the column is complete in the scratch line
                double *dst = peer.ptrs[j / peer.cols_per_owner] +
                              ((size_t)peer.rank * peer.cols_per_owner + (j % peer.cols_per_owner)) * (size_t)peer.S;
                for (int r = g * 2; r + 1 < peer.S; r += WG_THREADS * 2) {
                    const double2 v = __ldcg(reinterpret_cast<const double2 *>(loc + r));
                    *reinterpret_cast<double2 *>(dst + r) = v;
                }
                if ((peer.S & 1) && g == 0) dst[peer.S - 1] = __ldcg(loc + peer.S - 1);
                bar_sync(WBAR_OUT, WG_THREADS);                 // before the next column reuses the line
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
// The same with 64 x 64 tiles: 512-byte runs on both sides and 16 loads in flight per thread
// (the 32 x 32 form moved 3.7 TB/s at C3; this one is what the rows layout uses when the shape
// allows it).  Shared rows are 65 doubles apart: column reads of a tile are conflict-free.
__global__ void __launch_bounds__(256)
k_transpose_to_cols64(const double *__restrict__ X, size_t ldx, size_t N, size_t m, double *__restrict__ T) {
    __shared__ double tile[64][65];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;       // 8 warps
    const size_t k0 = (size_t)blockIdx.x * 64;
    for (size_t j0 = (size_t)blockIdx.y * 64; j0 < m; j0 += (size_t)gridDim.y * 64) {
        double v[16];
#pragma unroll
        for (int i = 0; i < 16; i++) {                             // row w + 8 * (i / 2), columns lane + 32 * (i % 2)
            const size_t k = k0 + w + 8 * (i >> 1), j = j0 + lane + 32 * (i & 1);
            v[i] = (k < N && j < m) ? __ldcs(X + k * ldx + j) : 0.0;
        }
#pragma unroll
        for (int i = 0; i < 16; i++) tile[w + 8 * (i >> 1)][lane + 32 * (i & 1)] = v[i];
        __syncthreads();
#pragma unroll
        for (int i = 0; i < 16; i++) {                             // column w + 8 * (i / 2), rows lane + 32 * (i % 2)
            const size_t j = j0 + w + 8 * (i >> 1), k = k0 + lane + 32 * (i & 1);
            if (k < N && j < m) T[k + j * N] = tile[lane + 32 * (i & 1)][w + 8 * (i >> 1)];
        }
        __syncthreads();
    }
}
static void launch_transpose_to_cols(const double *X, size_t ldx, size_t N, size_t m, double *T,
                                     cudaStream_t st) {
    static const bool small_tiles = getenv("SKB_TRANSPOSE32") != nullptr;      // A/B: the 32 x 32 form
    if (!small_tiles) {
        const size_t ty = (m + 63) / 64;
        dim3 grid((unsigned)((N + 63) / 64), (unsigned)(ty < 65535 ? ty : 65535));
        k_transpose_to_cols64<<<grid, 256, 0, st>>>(X, ldx, N, m, T);
        return;
    }
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
    uint32_t *pmask;          // fused engine: per (pass-2 tile, thread) the needed last-pass outputs
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
#ifndef SKB_FFT_PPT
#define SKB_FFT_PPT 16          // points a thread holds per pass (16: radix-16 passes, 8: radix-8)
#endif
#ifndef SKB_FFT_MINB
#define SKB_FFT_MINB 2
#endif
constexpr int FFT_PPT = SKB_FFT_PPT;
#ifndef SKB_FFT_POINTS
#define SKB_FFT_POINTS 4096     // complex points a CTA transforms at a time (4096 or 2048)
#endif
constexpr size_t FFT_POINTS = SKB_FFT_POINTS, FFT_THREADS = FFT_POINTS / FFT_PPT;
constexpr int FFT_POINTS_LOG = FFT_POINTS == 4096 ? 12 : 11;
static_assert(FFT_POINTS == 4096 || FFT_POINTS == 2048, "tile of 4096 or 2048 points");

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


// ---------------------------------------------------------------------------
// K8, fused form: ONE persistent launch per apply does pass 1, pass 2 and the sampled post-pass
// of every column, with the intermediate H of a column handed from pass 1 to pass 2 through a
// ring of DCT_NSLOT column slots (32 MB at N = 2^20) that stays in the 126 MB L2 -- HBM sees each
// input element once and the s*d result once.
//
// Work items are numbered so that dependencies always point backwards: stage s holds the pass-1
// tiles of column s, the pass-2 tiles of column s - LAG (alternating, so HBM-bound and
// compute-bound tiles share the SMs) and the post item of column s - 2 LAG.  CTAs take items in
// order from a global counter; an item that needs another one to have finished (pass 2: all
// pass-1 tiles of its column; pass 1: the ring slot's previous user; post: all pass-2 tiles) spins
// on a per-column counter that only ever depends on lower-numbered items, which are already
// running or done -- no co-residency assumption, no deadlock.
//
// Inside a tile the first and last radix passes are fused with the global traffic: pass 2 loads
// its rows straight into butterfly registers and hands its last-pass outputs to the selection
// (only the bins the samples need are kept); pass 1 multiplies its last-pass outputs by the
// inter-pass twiddle in registers and stores them with the sequence index varying across lanes,
// i.e. in 256-byte runs of H.
// ---------------------------------------------------------------------------
// Pass 2 of a column runs `lag` stages behind its pass 1, the post-pass `lag` stages behind pass 2;
// the rings hold lag + 2 columns of H and lag + 14 columns of kept bins.  lag is chosen per launch so
// that a dependency is at least one full round of resident CTAs old when it is needed.
static constexpr int DCT_POST_CHUNK = 1024;                      // samples per post item
constexpr int FFT_PPT_LOG = FFT_PPT == 16 ? 4 : 3;
static_assert(FFT_PPT == 16 || FFT_PPT == 8, "fused DCT: 16 or 8 points per thread");
constexpr size_t FFT_SMEM2 = (FFT_POINTS + (FFT_POINTS >> FFT_PPT_LOG) + 64) * sizeof(cplx);
// Shared-memory twiddle tables of the fused kernel, filled once per (persistent) CTA -- every
// dependency wait ends in a fence that drops the L1, so tables read through L1 would miss on every
// tile.  Per pass kind a compact table of 16 + 256 butterfly twiddles; the inter-pass factors
// W_N'^e = Thi[e >> 10] * Tlo[e & 1023] when both tables fit (N <= 2^21).
constexpr int TW_COMPACT = FFT_PPT == 16 ? 16 + 256 : 8 + 64 + 512;
constexpr int TT_SMEM_MAX = (FFT_PPT == 16 && FFT_POINTS == 4096) ? 2048 : 0;     // otherwise the buffers take the room
constexpr size_t FFT_SMEM2_TW = 2 * TW_COMPACT * sizeof(cplx);
// W: table of W_L^m, m < L.  One region per twiddled pass, in pass order: a middle pass with
// sub-length Ns holds W_L^(k L / (Ns R)), k < Ns (R = FFT_PPT); the last pass W_L^j, j < L / R_last.
__device__ __forceinline__ void fused_twiddles(cplx *tw, const cplx *__restrict__ W, int L) {
    int Ns = FFT_PPT, off = 0;
    while (L / Ns > FFT_PPT) {
        const int step = L / (Ns * FFT_PPT);
        for (int k = threadIdx.x; k < Ns; k += blockDim.x) tw[off + k] = ldg_cplx(W + k * step);
        off += Ns; Ns *= FFT_PPT;
    }
    for (int j = threadIdx.x; j < Ns; j += blockDim.x) tw[off + j] = ldg_cplx(W + j);
}

struct DctFused {
    const double *X; size_t ldx; size_t m;
    DctPlan pl;
    const uint32_t *Dbits; const int32_t *samples; size_t S;
    double inscale, outscale;
    cplx *H;                 // ring: DCT_NSLOT * Np
    cplx *Zc;                // ring: DCT_NZ * zc_ld
    const int32_t *pidx;     // per sample: position of Z[k'] and of Z[N'-k'] in a column's kept bins
    const uint32_t *pmask;   // per (pass-2 tile, thread): which of its last-pass outputs are needed
    unsigned *ctr;           // [0] item counter | ready1[m] | done2[m] | postdone[m]
    double *OUT; size_t oi, oj;
    unsigned T;              // tiles per column per pass = Np / 4096
    unsigned PB;             // post items per column
    unsigned TPI;            // tiles per pass-1 / pass-2 item
    unsigned lag, nslot, nz; // see above
    int tt_smem;             // inter-pass twiddle tables (Tlo, Thi) fit in shared memory
    unsigned pf_dist;        // L2 prefetch distance in pass-1 tiles
    int dbg;                 // SKB_DCT_DBG bits (timing experiments): 1 skip pass-2 work, 2 skip pass-1 work, 4 skip post
};

__device__ __forceinline__ unsigned ld_acquire_u32(const unsigned *p) {
    unsigned v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned ld_relaxed_u32(const unsigned *p) {
    unsigned v;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
// One thread polls and fences; the barrier extends the ordering to the CTA.  What other SMs wrote is
// read through L2 only (ld.cg), so no L1 line can be stale.
__device__ __forceinline__ void fence_acq_rel_gpu() { asm volatile("fence.acq_rel.gpu;" ::: "memory"); }
__device__ __forceinline__ void red_release_add(unsigned *p, unsigned v) {
    asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
// `seen` = a value of *p polled earlier by thread 0 (0 if none)
__device__ __forceinline__ void wait_counter(const unsigned *p, unsigned target, unsigned seen = 0) {
    if (threadIdx.x == 0) {
        while (seen < target) { seen = ld_relaxed_u32(p); if (seen < target) __nanosleep(40); }
        fence_acq_rel_gpu();
    }
    __syncthreads();
}
__device__ __forceinline__ void signal_counter(unsigned *p) {
    __syncthreads();
    if (threadIdx.x == 0) { __threadfence(); atomicAdd(p, 1u); }
}
__device__ __forceinline__ uint64_t l2_policy_evict_last() {
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
__device__ __forceinline__ uint64_t l2_policy_evict_first() {
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
__device__ __forceinline__ void stg_cplx_hint(cplx *p, cplx v, uint64_t pol) {
    asm volatile("st.global.L2::cache_hint.v2.f64 [%0], {%1, %2}, %3;" ::"l"(p), "d"(v.x), "d"(v.y), "l"(pol) : "memory");
}
__device__ __forceinline__ cplx ldcg_cplx_hint(const cplx *p, uint64_t pol) {      // L2 only: H is written by other SMs
    cplx v;
    asm volatile("ld.global.cg.L2::cache_hint.v2.f64 {%0, %1}, [%2], %3;" : "=d"(v.x), "=d"(v.y) : "l"(p), "l"(pol));
    return v;
}
__device__ __forceinline__ double2 ldg_stream2_hint(const double *p, uint64_t pol) {
    double2 v;
    asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v2.f64 {%0, %1}, [%2], %3;"
                 : "=d"(v.x), "=d"(v.y) : "l"(p), "l"(pol));
    return v;
}

// shared-memory index of point i: a spare element after every FFT_PPT; then either one per 256 (rows
// walked with the position varying across lanes) or one per sequence (pass 1, whose last pass
// walks the SEQUENCE index across lanes)
template <bool SEQPAD> __device__ __forceinline__ int padx(int i, int lshift) {
    return i + (i >> FFT_PPT_LOG) + (SEQPAD ? (i >> lshift) : (FFT_PPT == 16 ? (i >> 8) : 0));
}

// twiddles of one radix-R butterfly of a pass with sub-length Ns: powers of W_L^(k * step)
template <int R> __device__ __forceinline__ void twiddle_inputs(cplx *a, const cplx *W, int kstep) {
    cplx wp[R];
    wp[1] = W[kstep];
#pragma unroll
    for (int t = 2; t < R; t++) wp[t] = (t & 1) ? cmul(wp[t - 1], wp[1]) : cmul(wp[t / 2], wp[t / 2]);
#pragma unroll
    for (int t = 1; t < R; t++) a[t] = cmul(a[t], wp[t]);
}

// middle pass, in place in shared memory (all reads, barrier, all writes, barrier)
#ifdef SKB_K8_NOADD
constexpr bool K8_ADD = false;
#else
constexpr bool K8_ADD = true;
#endif
#ifdef SKB_K8_NOPMASK
constexpr bool K8_PMASK = false;
#else
constexpr bool K8_PMASK = true;
#endif
template <int R, bool SEQPAD>
__device__ __forceinline__ void mid_pass(cplx *buf, int L, int lshift, int Ns, const cplx *W) {
    constexpr int U = FFT_PPT / R;
    const int per = L / R, pshift = 31 - __clz(per);
    cplx a[U][R];
    int dbase[U];
#pragma unroll
    for (int uu = 0; uu < U; uu++) {
        const int u = (int)threadIdx.x + uu * (int)FFT_THREADS;
        const int g = u >> pshift, j = u & (per - 1), k = j & (Ns - 1);
#pragma unroll
        for (int t = 0; t < R; t++) a[uu][t] = buf[padx<SEQPAD>(g * L + j + t * per, lshift)];
        twiddle_inputs<R>(a[uu], W, k);           // compact table: entry k = W_L^(k * step)
        dbase[uu] = g * L + (j - k) * R + k;
    }
    __syncthreads();
#pragma unroll
    for (int uu = 0; uu < U; uu++) {
        butterfly<R>(a[uu]);
        if (K8_ADD && FFT_PPT == 16) {
            // the middle pass has Ns = 16: dbase + 16 q moves one more padding element per step, and
            // padx is additive there (no carry out of the low 8 bits, same sequence)
            cplx *dst = buf + padx<SEQPAD>(dbase[uu], lshift);
#pragma unroll
            for (int q = 0; q < R; q++) dst[q * 17] = a[uu][q];
        } else {
#pragma unroll
            for (int q = 0; q < R; q++) buf[padx<SEQPAD>(dbase[uu] + q * Ns, lshift)] = a[uu][q];
        }
    }
    __syncthreads();
}

// last pass (Ns * R == L): outputs go to sink(g, k, value), k = natural-order index inside
// sequence g.  GFAST: the sequence index varies fastest across lanes.
template <int R, bool SEQPAD, bool GFAST, class Sink>
__device__ __forceinline__ void last_pass(const cplx *buf, int L, int lshift, const cplx *W, Sink sink) {
    constexpr int U = FFT_PPT / R;
    const int Ns = L / R, nshift = 31 - __clz(Ns), gshift = FFT_POINTS_LOG - lshift;     // G = FFT_POINTS / L sequences
#pragma unroll
    for (int uu = 0; uu < U; uu++) {
        const int u = (int)threadIdx.x + uu * (int)FFT_THREADS;
        const int g = GFAST ? (u & ((1 << gshift) - 1)) : (u >> nshift);
        const int j = GFAST ? (u >> gshift) : (u & (Ns - 1));
        cplx a[R];
        if (K8_ADD && FFT_PPT == 16) {
            // Ns is 16 or 256 here, and padx is additive in t along that stride
            const cplx *src = buf + padx<SEQPAD>(g * L + j, lshift);
            const int c = Ns + (Ns >> 4) + ((!SEQPAD && Ns == 256) ? 1 : 0);
#pragma unroll
            for (int t = 0; t < R; t++) a[t] = src[t * c];
        } else {
#pragma unroll
            for (int t = 0; t < R; t++) a[t] = buf[padx<SEQPAD>(g * L + j + t * Ns, lshift)];
        }
        twiddle_inputs<R>(a, W, j);                 // step = 1
        butterfly<R>(a);
        sink(g, j, Ns, a, uu);
    }
}

template <bool SEQPAD, bool GFAST, class Sink>
__device__ __forceinline__ void fft_tail(cplx *buf, int L, int lshift, const cplx *W, Sink sink,
                                         const unsigned *slot_free, unsigned slot_free_target, unsigned slot_seen) {
    // after the first radix-16 pass (Ns = 16): at most one middle radix-16 pass (L >= 512), then the
    // last pass.  W is the compact shared-memory table of fused_twiddles(): 16 entries for the
    // middle pass, then the last pass's.
    int Ns = FFT_PPT;
    while (L / Ns > FFT_PPT) { mid_pass<FFT_PPT, SEQPAD>(buf, L, lshift, Ns, W); W += Ns; Ns *= FFT_PPT; }
    // the ring slot the last pass stores into must have been released by its previous user: checked
    // here, at the latest possible point, where it has long been the case
    if (slot_free) {         // write-after-read: ordered by the control dependency and the barrier, no fence
        if (threadIdx.x == 0)
            while (slot_seen < slot_free_target) { __nanosleep(20); slot_seen = ld_relaxed_u32(slot_free); }
        __syncthreads();
    }
    const int rem = L / Ns;
    if (FFT_PPT == 16 && rem == 16) last_pass<FFT_PPT, SEQPAD, GFAST>(buf, L, lshift, W, sink);
    else if (rem == 8) last_pass<8, SEQPAD, GFAST>(buf, L, lshift, W, sink);
    else if (rem == 4) last_pass<4, SEQPAD, GFAST>(buf, L, lshift, W, sink);
    else last_pass<2, SEQPAD, GFAST>(buf, L, lshift, W, sink);
}

template <bool AL32>
__device__ __forceinline__ void dct_tile_pass1(const DctFused &a, cplx *buf, size_t c, unsigned tile,
                                               uint64_t pol_first, uint64_t pol_last,
                                               const unsigned *slot_free, unsigned slot_free_target, unsigned slot_seen,
                                               const cplx *tw, const cplx *tt) {
    const DctPlan &pl = a.pl;
    const int N2 = (int)pl.N2, N1 = (int)pl.N1, lshift = 31 - __clz(N2);
    const int G = (int)(FFT_POINTS / pl.N2), gshift = FFT_POINTS_LOG - lshift, Gh = G >> 1, a0 = (int)tile * Gh;
    const double *col = a.X + c * a.ldx;
    cplx *Hc = a.H + (c % a.nslot) * pl.Np;
    auto seq = [&](int g) { return g < Gh ? a0 + g : N1 - 1 - a0 - (g - Gh); };
    constexpr int PAIRS = (int)(FFT_POINTS / 2 / FFT_THREADS);
#pragma unroll
    for (int b4 = 0; b4 < PAIRS / 4; b4++) {
        double x[4][4]; uint32_t w[4];
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const int p = (int)threadIdx.x + (b4 * 4 + u) * (int)FFT_THREADS;
            const int g = p & (G - 1), n2 = p >> gshift;                // n2 < N2/2: the lower half
            const unsigned q = 4u * (unsigned)(seq(g) + N1 * n2);
            w[u] = __ldg(a.Dbits + (q >> 5)) >> (q & 31);
            if (AL32) {
                double2 lo = ldg_stream2_hint(col + q, pol_first), hi = ldg_stream2_hint(col + q + 2, pol_first);
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
            for (int e = 0; e < 4; e++) sx[e] = __dmul_rn((w[u] >> e) & 1 ? -a.inscale : a.inscale, x[u][e]);
            cplx z, zm;
            z.x = sx[0]; z.y = sx[2];            // z[n]
            zm.x = sx[3]; zm.y = sx[1];          // z[N'-1-n]
            buf[padx<true>(g * N2 + n2, lshift)] = z;
            buf[padx<true>((g ^ Gh) * N2 + (N2 - 1 - n2), lshift)] = zm;
        }
    }
    __syncthreads();
    {   // first pass (radix FFT_PPT, Ns = 1), in place
        const int per = N2 >> FFT_PPT_LOG, pshift = 31 - __clz(per);
        const int u = (int)threadIdx.x, g = u >> pshift, j = u & (per - 1);
        cplx v[FFT_PPT];
#pragma unroll
        for (int t = 0; t < FFT_PPT; t++) v[t] = buf[padx<true>(g * N2 + j + t * per, lshift)];
        __syncthreads();
        butterfly<FFT_PPT>(v);
#pragma unroll
        {
            cplx *dst = buf + padx<true>(g * N2 + j * FFT_PPT, lshift);      // FFT_PPT consecutive slots
#pragma unroll
            for (int q = 0; q < FFT_PPT; q++) dst[q] = v[q];
        }
        __syncthreads();
    }
    const size_t lo_mask = ((size_t)1 << pl.lo_bits) - 1;
    auto table_tw = [&](size_t e) {
        e &= pl.Np - 1;
        if (a.tt_smem) return cmul(tt[((size_t)1 << pl.lo_bits) + (e >> pl.lo_bits)], tt[e & lo_mask]);
        return cmul(ldg_cplx(pl.Thi + (e >> pl.lo_bits)), ldg_cplx(pl.Tlo + (e & lo_mask)));
    };
    // last-pass outputs k2 = j + q*Ns of sequence g: times W_N'^(n1*k2), stored at H[k2*N1 + n1]
    auto sink = [&](int g, int j, int Ns, auto &v, int) {
        constexpr int R = (int)(sizeof(v) / sizeof(cplx));
        const size_t n1 = (size_t)seq(g);
        const cplx wstep = table_tw(n1 * (size_t)Ns);
        cplx tw = table_tw(n1 * (size_t)j);
        cplx *dst = Hc + (size_t)j * (size_t)N1 + n1;
        const size_t dstep = (size_t)Ns * (size_t)N1;
#pragma unroll
        for (int q = 0; q < R; q++) {
            if (q && (q & 7) == 0) tw = table_tw(n1 * (size_t)(j + q * Ns));     // bound the error of the running product
            stg_cplx_hint(dst, cmul(v[q], tw), pol_last);
            dst += dstep;
            tw = cmul(tw, wstep);
        }
    };
    fft_tail<true, true>(buf, N2, lshift, tw, sink, slot_free, slot_free_target, slot_seen);
}

__device__ __forceinline__ void dct_tile_pass2(const DctFused &a, cplx *buf, size_t c, unsigned tile, uint64_t pol_last,
                                               const unsigned *slot_free, unsigned slot_free_target, unsigned slot_seen,
                                               const cplx *tw, uint32_t *s_need, uint32_t *s_rank) {
    const DctPlan &pl = a.pl;
    const int L = (int)pl.N1, lshift = 31 - __clz(L);
    const size_t base = (size_t)tile * FFT_POINTS;
    const cplx *rows = a.H + (c % a.nslot) * pl.Np + base;
    for (int i = threadIdx.x; i < (int)(FFT_POINTS / 32); i += blockDim.x) {     // which of the tile's bins the samples need, and their ranks
        s_need[i] = __ldg(pl.need + (base >> 5) + i);
        s_rank[i] = __ldg(pl.rank + (base >> 5) + i);
    }
    {   // first pass (radix FFT_PPT) straight from global memory (rows are contiguous)
        const int per = L >> FFT_PPT_LOG, pshift = 31 - __clz(per);
        const int u = (int)threadIdx.x, g = u >> pshift, j = u & (per - 1);
        cplx v[FFT_PPT];
#pragma unroll
        for (int t = 0; t < FFT_PPT; t++) v[t] = ldcg_cplx_hint(rows + g * L + j + t * per, pol_last);
        butterfly<FFT_PPT>(v);
#pragma unroll
        {
            cplx *dst = buf + padx<false>(g * L + j * FFT_PPT, lshift);
#pragma unroll
            for (int q = 0; q < FFT_PPT; q++) dst[q] = v[q];
        }
        __syncthreads();
    }
    cplx *zc = a.Zc + (c % a.nz) * pl.zc_ld;
    // which of this thread's FFT_PPT last-pass outputs the samples need: one precomputed word per
    // (tile, thread) -- most threads hold none (the samples touch ~3 % of the bins) and skip the lot
    const uint32_t mine = __ldg(a.pmask + (size_t)tile * FFT_THREADS + threadIdx.x);
    auto sink = [&](int g, int j, int Ns, auto &v, int uu) {
        constexpr int R = (int)(sizeof(v) / sizeof(cplx));
        const uint32_t m = K8_PMASK ? ((mine >> (uu * R)) & ((1u << R) - 1u)) : 0xffffu;
        if (m == 0) return;
#pragma unroll
        for (int q = 0; q < R; q++) {
            if ((m >> q) & 1u) {
                const int p = g * L + j + q * Ns;                  // position inside the tile
                const uint32_t w = s_need[p >> 5];
                if (K8_PMASK || ((w >> (p & 31)) & 1u))
                    zc[s_rank[p >> 5] + __popc(w & ((1u << (p & 31)) - 1u))] = v[q];
            }
        }
    };
    fft_tail<false, false>(buf, L, lshift, tw, sink, slot_free, slot_free_target, slot_seen);
}

__device__ __forceinline__ void dct_post_chunk(const DctFused &a, size_t c, unsigned chunk) {
    const DctPlan &pl = a.pl;
    const cplx *Z = a.Zc + (c % a.nz) * pl.zc_ld;
    const size_t i0 = (size_t)chunk * DCT_POST_CHUNK, i1 = i0 + DCT_POST_CHUNK < a.S ? i0 + DCT_POST_CHUNK : a.S;
    for (size_t i = i0 + threadIdx.x; i < i1; i += blockDim.x) {
        const int2 pq = __ldg(reinterpret_cast<const int2 *>(a.pidx) + i);
        const double2 ra = __ldcg(reinterpret_cast<const double2 *>(Z + pq.x));
        const double2 rb = __ldcg(reinterpret_cast<const double2 *>(Z + pq.y));
        cplx E, O;
        E.x = 0.5 * (ra.x + rb.x); E.y = 0.5 * (ra.y - rb.y);
        O.x = 0.5 * (ra.y + rb.y); O.y = -0.5 * (ra.x - rb.x);
        cplx V = cadd(E, cmul(ldg_cplx(pl.post + 2 * i), O));
        cplx Yc = cmul(ldg_cplx(pl.post + 2 * i + 1), V);
        a.OUT[i * a.oi + c * a.oj] = a.outscale * (2.0 * Yc.x);
    }
}

template <bool AL32>
__global__ void __launch_bounds__(FFT_THREADS, SKB_FFT_MINB)
k_dct_fused(const DctFused a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    cplx *buf = reinterpret_cast<cplx *>(smem_raw);
    __shared__ unsigned s_item[2];
    __shared__ uint32_t s_need[FFT_POINTS / 32], s_rank[FFT_POINTS / 32];
    // an item is TPI consecutive tiles of one pass of one column (or one chunk of the post-pass);
    // the per-column counters count items
    const unsigned TPI = a.TPI, TI = a.T / TPI, PB = a.PB, per_stage = 2 * TI + PB;
    const unsigned long long total = (unsigned long long)(a.m + 2 * a.lag) * per_stage;
    unsigned *ready1 = a.ctr + 1, *done2 = ready1 + a.m, *postdone = done2 + a.m;
    const uint64_t pol_first = l2_policy_evict_first(), pol_last = l2_policy_evict_last();
    cplx *tw1 = reinterpret_cast<cplx *>(smem_raw + FFT_SMEM2), *tw2 = tw1 + TW_COMPACT, *tt = tw2 + TW_COMPACT;
    fused_twiddles(tw1, a.pl.W1, (int)a.pl.N2);
    fused_twiddles(tw2, a.pl.W2, (int)a.pl.N1);
    if (a.tt_smem) {
        const int nlo = 1 << a.pl.lo_bits, nhi = (int)(a.pl.Np >> a.pl.lo_bits);
        for (int i = threadIdx.x; i < nlo; i += blockDim.x) tt[i] = ldg_cplx(a.pl.Tlo + i);
        for (int i = threadIdx.x; i < nhi; i += blockDim.x) tt[nlo + i] = ldg_cplx(a.pl.Thi + i);
    }
    if (threadIdx.x == 0) s_item[0] = atomicAdd(a.ctr, 1u);
    __syncthreads();
    for (int par = 0;; par ^= 1) {
        const unsigned item = s_item[par];
        if (item >= total) break;
        // the next item is fetched while this one is worked on (items are still handed out in order)
        // duties are spread over three warps so that their memory round trips overlap: thread 0 polls
        // dependencies, thread 32 publishes completions, thread 64 fetches items
        unsigned nxt = 0;
        if (threadIdx.x == 64) nxt = atomicAdd(a.ctr, 1u);
        const size_t stage = item / per_stage;
        const unsigned r = item % per_stage;
        unsigned *done = nullptr;                  // counter this item bumps when its stores are out
        if (r >= 2 * TI) {
            if (stage >= 2 * a.lag && stage - (2 * a.lag) < a.m) {
                const size_t c = stage - (2 * a.lag);
                wait_counter(done2 + c, TI);
                if (!(a.dbg & 4)) dct_post_chunk(a, c, r - 2 * TI);
                done = postdone + c;
            }
        } else if ((r & 1) == 0) {
            if (stage < a.m) {
                const size_t c = stage;
                const unsigned *sf = c >= a.nslot ? done2 + (c - a.nslot) : nullptr;      // H ring slot free
                unsigned seen = (sf && threadIdx.x == 0) ? ld_relaxed_u32(sf) : 0u;            // polled early, needed late
                if (AL32 && !(a.dbg & 8)) {
                    // pull the input of the pass-1 tiles that run about one round of CTAs later into L2
                    const int N2 = (int)a.pl.N2, N1 = (int)a.pl.N1, Gh = (int)(FFT_POINTS / a.pl.N2) >> 1;
                    for (unsigned sub = 0; sub < TPI; sub++) {
                        const unsigned long long q = (unsigned long long)stage * a.T + (r >> 1) * TPI + sub + a.pf_dist;
                        const size_t cpf = (size_t)(q / a.T);
                        if (cpf >= a.m) break;
                        const int a0 = (int)(q % a.T) * Gh;
                        const double *colp = a.X + cpf * a.ldx;
                        for (int i = threadIdx.x; i < N2; i += blockDim.x) {        // N2/2 positions x 2 runs
                            const int n2 = i >> 1, first = (i & 1) ? N1 - a0 - Gh : a0;
                            const double *src = colp + 4 * ((size_t)first + (size_t)N1 * n2);
                            asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(src), "r"(Gh * 32));
                        }
                    }
                }
                for (unsigned sub = 0; sub < TPI; sub++) {
                    if (sub) __syncthreads();      // the previous tile's last pass has left the buffer
                    if (!(a.dbg & 2))
                        dct_tile_pass1<AL32>(a, buf, c, (r >> 1) * TPI + sub, pol_first, pol_last,
                                             sub == 0 ? sf : nullptr, TI, seen, tw1, tt);
                }
                done = ready1 + c;
            }
        } else {
            if (stage >= a.lag && stage - a.lag < a.m) {
                const size_t c = stage - a.lag;
                const unsigned *sf = c >= a.nz ? postdone + (c - a.nz) : nullptr;         // kept-bin ring slot free
                unsigned seen = (sf && threadIdx.x == 0) ? ld_relaxed_u32(sf) : 0u;
                wait_counter(ready1 + c, TI);
                for (unsigned sub = 0; sub < TPI; sub++) {
                    if (sub) __syncthreads();
                    if (!(a.dbg & 1))
                        dct_tile_pass2(a, buf, c, (r >> 1) * TPI + sub, pol_last, sub == 0 ? sf : nullptr, PB, seen,
                                       tw2, s_need, s_rank);
                }
                done = done2 + c;
            }
        }
        if (threadIdx.x == 64) s_item[par ^ 1] = nxt;
        __syncthreads();                           // stores are issued, the tile buffer is free, s_item[par ^ 1] is set
        if (done && threadIdx.x == 32) red_release_add(done, 1u);       // cumulative over the barrier; off the others' path
    }
}

// split N' = N1 * N2 of the fused form: 128-point strided FFTs (pass 1: 32 sequences per tile, 256-byte
// runs) and rows of up to 4096 points (pass 2: one tile per row)
static bool dct_fused_ok(size_t N) { return N >= 8192 && (N & (N - 1)) == 0 && N <= ((size_t)1 << 24); }
static void dct_fused_split(size_t N, size_t *N1, size_t *N2) {
    const size_t Np = N / 2;
    size_t n2 = 128;
    {   // debug knob: log2 of the strided (pass 1) FFT length
        static const char *e = getenv("SKB_DCT_N2LOG");
        if (e && atoi(e) >= 7 && atoi(e) <= 11) n2 = (size_t)1 << atoi(e);
    }
    while (Np / n2 > FFT_POINTS) n2 *= 2;
    while (Np / n2 < 32 && n2 > 128) n2 /= 2;
    *N2 = n2; *N1 = Np / n2;
}

static bool dct_legacy_forced() {
    static const bool v = getenv("SKB_DCT_LEGACY") != nullptr;      // A/B: the per-batch two-kernel form
    return v;
}
static size_t align256(size_t x) { return (x + 255) / 256 * 256; }

// Tables of the fused form: they depend on (N, S, samples) only, so a transform builds them once
// per device (fjlt_dct_plan_build) and every apply reuses them.
static DctPlan dct_plan_layout(size_t N, size_t S, void *buf) {
    DctPlan pl;
    pl.N = N; pl.Np = N / 2;
    dct_fused_split(N, &pl.N1, &pl.N2);
    pl.lo_bits = 10;
    pl.G1 = pl.G2 = 0;
    char *p = reinterpret_cast<char *>(buf);
    auto take = [&](size_t bytes) { char *r = p; p += align256(bytes); return r; };
    pl.W1 = reinterpret_cast<cplx *>(take(pl.N2 * sizeof(cplx)));
    pl.W2 = reinterpret_cast<cplx *>(take(pl.N1 * sizeof(cplx)));
    pl.Tlo = reinterpret_cast<cplx *>(take(((size_t)1 << pl.lo_bits) * sizeof(cplx)));
    pl.Thi = reinterpret_cast<cplx *>(take(((pl.Np >> pl.lo_bits) + 1) * sizeof(cplx)));
    pl.post = reinterpret_cast<cplx *>(take(2 * S * sizeof(cplx)));
    pl.need = reinterpret_cast<uint32_t *>(take((pl.Np / 32 + 64) * sizeof(uint32_t)));
    pl.rank = reinterpret_cast<uint32_t *>(take((pl.Np / 32 + 64) * sizeof(uint32_t)));
    pl.Zc = nullptr; pl.zc_ld = 2 * S;
    pl.G2 = (size_t)(take(2 * S * sizeof(int32_t)) - reinterpret_cast<char *>(buf));   // offset of pidx
    pl.pmask = reinterpret_cast<uint32_t *>(take((pl.Np / FFT_POINTS) * FFT_THREADS * sizeof(uint32_t)));
    pl.G1 = (size_t)(p - reinterpret_cast<char *>(buf));     // bytes used (host side only)
    return pl;
}
// bit (uu * R + q) of pmask[tile * FFT_THREADS + tid] = 1 iff the bin that last_pass<R, .., GFAST = false>
// hands to its sink as (unit uu, output q) of thread tid in pass-2 tile `tile` is needed by a sample
__global__ void k_dct_pmask(DctPlan pl, uint32_t *__restrict__ pmask) {
    const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t ntiles = pl.Np / FFT_POINTS;
    if (t >= ntiles * FFT_THREADS) return;
    const int tid = (int)(t % FFT_THREADS);
    const size_t base = (t / FFT_THREADS) * FFT_POINTS;
    const int L = (int)pl.N1;
    int Ns = FFT_PPT;
    while (L / Ns > FFT_PPT) Ns *= FFT_PPT;
    const int R = L / Ns, U = FFT_PPT / R, nshift = 31 - __clz(Ns);
    uint32_t m = 0;
    for (int uu = 0; uu < U; uu++) {
        const int u = tid + uu * (int)FFT_THREADS, g = u >> nshift, j = u & (Ns - 1);
        for (int q = 0; q < R; q++) {
            const size_t p = base + (size_t)(g * L + j + q * Ns);
            if ((pl.need[p >> 5] >> (p & 31)) & 1u) m |= 1u << (uu * R + q);
        }
    }
    pmask[t] = m;
}
__global__ void k_dct_pidx(DctPlan pl, const int32_t *__restrict__ samples, size_t S, int32_t *__restrict__ pidx) {
    size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= S) return;
    const size_t k = (size_t)samples[t], kp = k & (pl.Np - 1), km = (pl.Np - kp) & (pl.Np - 1);
    pidx[2 * t] = (int32_t)dct_rank_of(pl, dct_pos(pl, kp));
    pidx[2 * t + 1] = (int32_t)dct_rank_of(pl, dct_pos(pl, km));
}
size_t fjlt_dct_plan_bytes(size_t N, size_t S) {
    if (!dct_fused_ok(N) || dct_legacy_forced()) return 0;
    return dct_plan_layout(N, S, nullptr).G1;
}
void fjlt_dct_plan_build(size_t N, size_t S, const int32_t *samples, void *buf, cudaStream_t st) {
    DctPlan pl = dct_plan_layout(N, S, buf);
    size_t nt = pl.N2 > S ? pl.N2 : S;
    if (nt < pl.N1) nt = pl.N1;
    if (nt < 1024) nt = 1024;
    if (nt < (pl.Np >> pl.lo_bits) + 1) nt = (pl.Np >> pl.lo_bits) + 1;
    {
        LaunchScope ls(KC_TABLES, st);
        k_dct_tables<<<(unsigned)((nt + 255) / 256), 256, 0, st>>>(pl, samples, S);
        SKB_CUDA(cudaGetLastError());
    }
    SKB_CUDA(cudaMemsetAsync(pl.need, 0, (pl.Np / 32 + 64) * sizeof(uint32_t), st));
    {
        LaunchScope ls(KC_TABLES, st);
        k_dct_mark<<<(unsigned)((S + 255) / 256), 256, 0, st>>>(pl, samples, S);
        SKB_CUDA(cudaGetLastError());
    }
    {
        LaunchScope ls(KC_TABLES, st);
        k_dct_rank<<<1, 1024, 0, st>>>(pl);
        SKB_CUDA(cudaGetLastError());
    }
    {
        LaunchScope ls(KC_TABLES, st);
        const size_t nt2 = (pl.Np / FFT_POINTS) * FFT_THREADS;
        k_dct_pmask<<<(unsigned)((nt2 + 255) / 256), 256, 0, st>>>(pl, pl.pmask);
        SKB_CUDA(cudaGetLastError());
    }
    {
        LaunchScope ls(KC_TABLES, st);
        k_dct_pidx<<<(unsigned)((S + 255) / 256), 256, 0, st>>>(
            pl, samples, S, reinterpret_cast<int32_t *>(reinterpret_cast<char *>(buf) + pl.G2));
        SKB_CUDA(cudaGetLastError());
    }
}
static unsigned dct_fused_tpi(size_t N) {
    static const char *e = getenv("SKB_DCT_TPI");
    unsigned tpi = e ? (unsigned)atoi(e) : 1u;
    const unsigned T = (unsigned)(N / 2 / FFT_POINTS);
    if (tpi < 1) tpi = 1;
    while (tpi > 1 && T % tpi) tpi--;
    return tpi;
}
static unsigned dct_fused_lag(size_t N, size_t S) {
    static const char *e = getenv("SKB_DCT_LAG");
    if (e && atoi(e) > 0) return (unsigned)atoi(e);
    const unsigned long long per_stage = 2ULL * (N / 2 / FFT_POINTS / dct_fused_tpi(N)) + (S + DCT_POST_CHUNK - 1) / DCT_POST_CHUNK;
    const unsigned long long slots = 148ULL * SKB_FFT_MINB;
    unsigned long long lag = (slots + per_stage - 1) / per_stage + 1;
    if (lag < 2) lag = 2;
    // keep the H ring inside L2 where the sizes allow (a longer ring only costs residency, not correctness)
    const unsigned long long cap = ((unsigned long long)48 << 20) / ((N / 2) * sizeof(cplx));
    if (lag + 2 > cap && cap >= 4) lag = cap - 2;
    return (unsigned)lag;
}
static size_t dct_fused_ws_bytes(size_t N, size_t m, size_t S) {
    const size_t lag = dct_fused_lag(N, S);
    return align256((lag + 2) * (N / 2) * sizeof(cplx)) + align256((lag + 14) * 2 * S * sizeof(cplx)) +
           align256((1 + 3 * m) * sizeof(unsigned));
}
// columns [c0, c0 + mc) of Xc through the fused kernel
static void dct_fused_launch(const double *Xc, size_t ldc, size_t N, size_t mc, const uint32_t *Dbits,
                             const int32_t *samples, size_t S, const void *plan, char *ws, double *OUT,
                             size_t oi, size_t oj, cudaStream_t st) {
    DctFused a;
    a.X = Xc; a.ldx = ldc; a.m = mc;
    a.pl = dct_plan_layout(N, S, const_cast<void *>(plan));
    a.Dbits = Dbits; a.samples = samples; a.S = S;
    a.inscale = 1.0 / sqrt(2.0 * (double)N); a.outscale = sqrt((double)N / (double)S);
    a.lag = dct_fused_lag(N, S); a.nslot = a.lag + 2; a.nz = a.lag + 14;
    a.H = reinterpret_cast<cplx *>(ws); ws += align256(a.nslot * (N / 2) * sizeof(cplx));
    a.Zc = reinterpret_cast<cplx *>(ws); ws += align256(a.nz * 2 * S * sizeof(cplx));
    a.pidx = reinterpret_cast<const int32_t *>(reinterpret_cast<const char *>(plan) + a.pl.G2);
    a.pmask = a.pl.pmask;
    a.PB = (unsigned)((S + DCT_POST_CHUNK - 1) / DCT_POST_CHUNK);
    a.TPI = dct_fused_tpi(N);
    a.ctr = reinterpret_cast<unsigned *>(ws);
    a.OUT = OUT; a.oi = oi; a.oj = oj;
    a.T = (unsigned)(a.pl.Np / FFT_POINTS);
    { static const char *e = getenv("SKB_DCT_DBG"); a.dbg = e ? atoi(e) : 0; }
    SKB_CUDA(cudaMemsetAsync(a.ctr, 0, (1 + 3 * mc) * sizeof(unsigned), st));
    const bool al32 = (ldc % 4 == 0) && ((uintptr_t)Xc % 32 == 0);
    auto kern = al32 ? k_dct_fused<true> : k_dct_fused<false>;
    const size_t ntt = ((size_t)1 << a.pl.lo_bits) + (a.pl.Np >> a.pl.lo_bits);
    a.tt_smem = ntt <= (size_t)TT_SMEM_MAX ? 1 : 0;
    const size_t smem = FFT_SMEM2 + FFT_SMEM2_TW + (a.tt_smem ? ntt * sizeof(cplx) : 0);
    { static const char *e = getenv("SKB_DCT_PF"); a.pf_dist = e ? (unsigned)atoi(e) : 148u; }
    SKB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const unsigned long long items = (unsigned long long)(mc + 2 * a.lag) * (2ULL * (a.T / a.TPI) + a.PB);
    const unsigned long long slots = (unsigned long long)fjlt_sms() * SKB_FFT_MINB;
    LaunchScope ls(KC_FJLT, st);
    kern<<<(unsigned)(items < slots ? items : slots), FFT_THREADS, smem, st>>>(a);
    SKB_CUDA(cudaGetLastError());
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

// Rows layout (every sketched index owns a contiguous row, e.g. a row-major n x d array sketched
// columnwise): the column kernels are fed through a transposed copy of a BLOCK of columns at a
// time (the reference also transposes for this direction, RFUT_Elemental.hpp:74-90), so the scratch
// is bounded (<= 4 GiB) instead of a second copy of the operand.
static size_t rows_block_cols(size_t rows, size_t m) {
    size_t cb = ((size_t)4 << 30) / (rows * sizeof(double));
    if (cb > 592) cb = 592;          // four rounds of the persistent kernels' CTAs
    if (cb < 1) cb = 1;
    return cb < m ? cb : m;
}
size_t fjlt_workspace_bytes(size_t N, size_t m, int fut, bool layout_rows, size_t S) {
    const size_t mb = layout_rows ? rows_block_cols(N, m) : m;      // columns in flight
    const size_t tr = layout_rows ? align256(N * mb * sizeof(double)) : 0;
    if (fut == FUT_DCT && dct_fused_ok(N) && !dct_legacy_forced())
        return tr + dct_fused_ws_bytes(N, mb, S) + 1024;
    if (fut == FUT_DCT && dct_fast_ok(N))
        return tr + dct_batch_cols(N, mb) * N * sizeof(double) +
               dct_table_bytes(N, S) + dct_sel_bytes(N, S, dct_batch_cols(N, mb)) + 1024;
    if (fut == FUT_DCT) return 4 * N * sizeof(double) + 256;
    if (fut == FUT_WHT && wht_fast_ok(N))         // sign words (+ transposed block)
        return tr + N / 8 + 256;
    return 256;
}

bool fjlt_shard_ok(int fut, size_t N, size_t offset, size_t nloc) {
    return fut == FUT_WHT && N >= (size_t)WSEG && (N & (N - 1)) == 0 && N <= ((size_t)1 << 28) &&
           offset % WSEG == 0 && nloc % WSEG == 0 && nloc > 0;
}

static void fjlt_apply_impl(int fut, bool layout_rows, const double *X, size_t ldx, size_t N, size_t m,
                            const double *D, const uint32_t *Dbits, const int32_t *samples, size_t S,
                            const int32_t *wht_slot_sample, const int32_t *wht_slot_out,
                            double *OUT, size_t ldo, void *workspace, size_t workspace_bytes,
                            cudaStream_t st, size_t offset, size_t nloc, const void *dct_plan,
                            bool out_rows);

void fjlt_apply(int fut, bool layout_rows, const double *X, size_t ldx, size_t N, size_t m,
                const double *D, const uint32_t *Dbits, const int32_t *samples, size_t S,
                const int32_t *wht_slot_sample, const int32_t *wht_slot_out,
                double *OUT, size_t ldo, void *workspace, size_t workspace_bytes,
                cudaStream_t st, size_t offset, size_t nloc, const void *dct_plan) {
    if (S == 0 || m == 0) return;
    if (workspace_bytes < fjlt_workspace_bytes(N, m, fut, layout_rows, S))
        throw error(SKB_ERR_ALLOC, "fjlt apply: workspace too small");
    const bool fast = (fut == FUT_DCT && dct_fast_ok(N)) || (fut == FUT_WHT && wht_fast_ok(N));
    if (layout_rows && fast) {
        // block of columns -> contiguous columns in scratch -> column kernel writing rows-layout output
        const size_t rows = fut == FUT_WHT ? nloc : N, cb = rows_block_cols(N, m);
        double *T = reinterpret_cast<double *>(workspace);
        char *rest = reinterpret_cast<char *>(workspace) + align256(N * cb * sizeof(double));
        const size_t rest_bytes = workspace_bytes - align256(N * cb * sizeof(double));
        for (size_t c0 = 0; c0 < m; c0 += cb) {
            const size_t mc = m - c0 < cb ? m - c0 : cb;
            {
                LaunchScope ls(KC_FJLT, st);
                launch_transpose_to_cols(X + c0, ldx, rows, mc, T, st);
                SKB_CUDA(cudaGetLastError());
            }
            fjlt_apply_impl(fut, false, T, rows, N, mc, D, Dbits, samples, S, wht_slot_sample, wht_slot_out,
                            OUT + c0, ldo, rest, rest_bytes, st, offset, nloc, dct_plan, true);
        }
        return;
    }
    fjlt_apply_impl(fut, layout_rows, X, ldx, N, m, D, Dbits, samples, S, wht_slot_sample, wht_slot_out, OUT, ldo,
                    workspace, workspace_bytes, st, offset, nloc, dct_plan, layout_rows);
}

// out_rows: OUT[i*ldo + j] (else OUT[i + j*ldo]); layout_rows is only seen by the general-N path here
static void fjlt_apply_impl(int fut, bool layout_rows, const double *X, size_t ldx, size_t N, size_t m,
                            const double *D, const uint32_t *Dbits, const int32_t *samples, size_t S,
                            const int32_t *wht_slot_sample, const int32_t *wht_slot_out,
                            double *OUT, size_t ldo, void *workspace, size_t workspace_bytes,
                            cudaStream_t st, size_t offset, size_t nloc, const void *dct_plan,
                            bool out_rows) {
    if (S == 0 || m == 0) return;
    if (nloc != N && !fjlt_shard_ok(fut, N, offset, nloc))
        throw error(SKB_ERR_INVALID_USAGE,
                    "FJLT: a shard of the sketched dimension needs the Walsh-Hadamard mode, N a power "
                    "of two >= 8192, and shard offset and length that are multiples of 8192");
    if (fut == FUT_WHT && (N & (N - 1)) != 0)
        throw error(SKB_ERR_INVALID_PARAM, "FJLT (WHT mode): N must be a power of two");
    if (fut == FUT_DCT && dct_fast_ok(N)) {
        char *ws = reinterpret_cast<char *>(workspace);
        const double *Xc = X;
        size_t ldc = ldx;
        const size_t oi_ = out_rows ? ldo : 1, oj_ = out_rows ? 1 : ldo;
        if (dct_plan && dct_fused_ok(N) && !dct_legacy_forced()) {
            // keep the 32-bit item counter of one launch far from wrapping
            const size_t tiles = N / 2 / FFT_POINTS,
                         cmax = ((size_t)1 << 30) / (2 * tiles + (S + DCT_POST_CHUNK - 1) / DCT_POST_CHUNK);
            for (size_t c0 = 0; c0 < m; c0 += cmax) {
                const size_t mc = m - c0 < cmax ? m - c0 : cmax;
                dct_fused_launch(Xc + c0 * ldc, ldc, N, mc, Dbits, samples, S, dct_plan, ws, OUT + c0 * oj_,
                                 oi_, oj_, st);
            }
            return;
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
        const size_t oi = oi_, oj = oj_;
        const double inscale = 1.0 / sqrt(2.0 * (double)N), outscale = sqrt((double)N / (double)S);
        // one radix-4 unit per thread per pass where possible (latency hiding by occupancy)
        auto threads_for = [](size_t units) { size_t t = 128; while (t < units && t < 1024) t *= 2; return (unsigned)t; };
        const unsigned thr1 = threads_for(pl.G1 * pl.N2 / 4), thr2 = threads_for(pl.G2 * pl.N1 / 4);
        for (size_t c0 = 0; c0 < m; c0 += B) {
            const size_t nb = m - c0 < B ? m - c0 : B;
            {
                LaunchScope ls(KC_FJLT, st);
                k_dct_pass1<<<dim3((unsigned)(pl.N1 / pl.G1), (unsigned)nb), thr1, smem1, st>>>(
                    Xc, ldc, c0, pl, Dbits, inscale, H);
            }
            {
                LaunchScope ls(KC_FJLT, st);
                k_dct_pass2<<<dim3((unsigned)(pl.N2 / pl.G2), (unsigned)nb), thr2, smem2, st>>>(pl, H);
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
        uint64_t *Dwht = reinterpret_cast<uint64_t *>(ws);      // N/64 words, rebuilt per apply (~2 us)
        {
            LaunchScope ls(KC_FJLT, st);
            const size_t nwords = N >> 6;
            k_wht_sign_words<<<(unsigned)((nwords + 255) / 256), 256, 0, st>>>(Dbits, nwords, Dwht);
            SKB_CUDA(cudaGetLastError());
        }
        const size_t oi = out_rows ? ldo : 1, oj = out_rows ? 1 : ldo;
        const double inscale = 1.0 / sqrt((double)N), outscale = sqrt((double)N / (double)S);
        const size_t smem = (size_t)WSMEM_BYTES;
        const bool al16 = (ldc % 2 == 0) && ((uintptr_t)Xc % 16 == 0);
        auto kern = al16 ? k_fjlt_wht_cols<true> : k_fjlt_wht_cols<false>;
        SKB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        unsigned grid = (unsigned)(m < fjlt_sms() ? m : fjlt_sms());
        if (g_wht_peer.ptrs && (S > (size_t)WG_THREADS * WSPT || out_rows || (S & 1) || ((uintptr_t)g_wht_peer.scratch & 15)))
            throw error(SKB_ERR_INVALID_USAGE, "FJLT peer output: needs an even s <= 8192 and column-major operands");
        const size_t per_pass = (size_t)WG_THREADS * WSPT;               // 8192 samples per pass
        for (size_t s0 = 0, p = 0; s0 < S; s0 += per_pass, p++) {
            LaunchScope ls(KC_FJLT, st);
            kern<<<grid, WTHREADS, smem, st>>>(Xc, ldc, nloc, m, Dwht, wht_slot_sample + p * per_pass,
                                               wht_slot_out + p * per_pass, offset >> WSEG_LOG,
                                               inscale, outscale, OUT, oi, oj, g_wht_peer);
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
