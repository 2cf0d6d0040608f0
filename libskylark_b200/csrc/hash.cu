// Hash sketches (CWT / MMT / WZT): Pi has one nonzero per input index.
//
// Reference loops replaced:
//   hash_transform_data_t::build            sketch/hash_transform_data.hpp:64-76
//   WZT_data_t::build                       sketch/WZT_data.hpp:110-126
//   dense apply (both directions)           sketch/hash_transform_Elemental.hpp:83-123
//   CSC -> dense apply (both directions)    sketch/hash_transform_Elemental.hpp:201-256
//
// B200 design (HBM-bound, no tensor cores):
//   * bucket/value tables are produced on the device from the counter stream (K0);
//   * for the layout in which every input index owns a contiguous row (the path a
//     C-ordered "columnwise" sketch takes), rows are grouped by bucket once per
//     transform (stable radix sort) so that the apply is a gather-side segmented
//     sum: each warp streams <=32 whole rows of one bucket with 128-bit loads and
//     writes one 64-column slice of one output row.  Rows inside a chunk are added in
//     ascending order with separate multiply and add, i.e. the reference's own rounding
//     sequence; a bucket of <= 32 rows is one chunk and is stored plainly (bit-identical to
//     the reference loop, identical from run to run).  A longer bucket spans several chunks
//     whose sums meet in red.f64: the result is then correct to rounding but its low bits
//     depend on the order in which the atomics land -- the same holds for the strided kernels
//     when the index is split over CTAs (ksplit > 1), for the CSC rowwise kernel and for the
//     stream-K flush of the dense GEMM.  Tests compare those with a tolerance, not bitwise.
//   * the strided layout accumulates in shared memory without atomics (lane-owned buckets,
//     K2p) and the CSC input per output column (shared atomics), both dump coalesced columns.
#include <algorithm>
#include <map>
#include <cstdlib>
#include <cub/block/block_radix_sort.cuh>
#include <cub/block/block_scan.cuh>
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>

#include "rng.cuh"
#include "skb_internal.h"

namespace skb {

// ---------------------------------------------------------------------------
// K0: tables
// ---------------------------------------------------------------------------
__global__ void k_hash_tables(uint64_t key, uint64_t idx_base, uint64_t val_base,
                              uint64_t pm_base, size_t first, size_t n, uint64_t S, int kind,
                              double p, int order, int32_t *__restrict__ bucket,
                              double *__restrict__ value) {
    size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    size_t g = first + t;
    {
        MicroUrng u(key, idx_base + g);
        bucket[t] = (int32_t)boost_uniform_int(u, S - 1);
    }
    MicroUrng u(key, val_base + g);
    double v;
    if (kind == HV_RADEMACHER) {
        v = boost_rademacher(u, order);
    } else if (kind == HV_CAUCHY) {
        v = boost_cauchy(u);
    } else {
        double e = boost_exponential(u);
        MicroUrng u2(key, pm_base + g);
        double pm = boost_rademacher(u2, order);
        v = __dmul_rn(pm, pow(__ddiv_rn(1.0, e), __ddiv_rn(1.0, p)));
    }
    value[t] = v;
}

void hash_build_tables(int seed, uint64_t idx_base, uint64_t val_base, uint64_t pm_base,
                       size_t first, size_t n, uint64_t S, int kind, double p, int order,
                       int32_t *bucket, double *value, cudaStream_t st) {
    if (n == 0) return;
    LaunchScope ls(KC_TABLES, st);
    unsigned blocks = (unsigned)((n + 255) / 256);
    k_hash_tables<<<blocks, 256, 0, st>>>(seed_to_key(seed), idx_base, val_base, pm_base,
                                          first, n, S, kind, p, order, bucket, value);
    SKB_CUDA(cudaGetLastError());
}

// ---------------------------------------------------------------------------
// Plan: rows grouped by bucket, cut in chunks of <= CHUNK rows
// ---------------------------------------------------------------------------
#ifndef SKB_K1_CHUNK
#define SKB_K1_CHUNK 32
#endif
static constexpr int CHUNK = SKB_K1_CHUNK;       // rows of one bucket a warp takes: 32 or 64

__global__ void k_iota(int32_t *a, size_t n) {
    size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t < n) a[t] = (int32_t)t;
}

__global__ void k_bucket_hist(const int32_t *__restrict__ bucket, size_t n, int32_t *hist) {
    size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t < n) atomicAdd(&hist[bucket[t]], 1);
}

__global__ void k_chunks_per_bucket(const int32_t *__restrict__ hist, int S, int32_t *nch) {
    int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b < S) nch[b] = (hist[b] + CHUNK - 1) / CHUNK;
}

// bucket_row0: exclusive scan of hist; chunk0: exclusive scan of nch
__global__ void k_fill_chunks(const int32_t *__restrict__ hist,
                              const int32_t *__restrict__ bucket_row0,
                              const int32_t *__restrict__ chunk0, int S,
                              int32_t *chunk_start, int32_t *chunk_bucket, int nchunks,
                              int32_t n_total) {
    int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b == 0) chunk_start[nchunks] = n_total;
    if (b >= S) return;
    int cnt = hist[b], r0 = bucket_row0[b], c0 = chunk0[b];
    int nc = (cnt + CHUNK - 1) / CHUNK;
    for (int c = 0; c < nc; c++) {
        chunk_start[c0 + c] = r0 + c * CHUNK;
        chunk_bucket[c0 + c] = nc > 1 ? ~b : b;
    }
}

// chunk_start[c+1] of the last chunk of a bucket must be the bucket end, which is
// the next bucket's start: true by construction because chunks are laid out in
// bucket order and every bucket's first chunk starts at bucket_row0.

void hash_free_plan(HashPlan *plan) {
    if (plan->rows) cudaFree(plan->rows);
    if (plan->chunk_start) cudaFree(plan->chunk_start);
    if (plan->chunk_bucket) cudaFree(plan->chunk_bucket);
    *plan = HashPlan();
}

void hash_build_plan(const int32_t *bucket, size_t n, int S, HashPlan *plan, cudaStream_t st) {
    hash_free_plan(plan);
    plan->n = n;
    if (n == 0) return;
    if (n > 0x7fffffffULL) throw error(SKB_ERR_INVALID_PARAM, "hash plan: n exceeds int32");
    int32_t *iota = nullptr, *keys_out = nullptr, *hist = nullptr, *row0 = nullptr,
            *nch = nullptr, *chunk0 = nullptr;
    void *tmp = nullptr;
    try {
        SKB_CUDA(cudaMalloc(&plan->rows, n * sizeof(int32_t)));
        SKB_CUDA(cudaMalloc(&iota, n * sizeof(int32_t)));
        SKB_CUDA(cudaMalloc(&keys_out, n * sizeof(int32_t)));
        SKB_CUDA(cudaMalloc(&hist, (size_t)(S + 1) * 4 * sizeof(int32_t)));
        row0 = hist + (S + 1); nch = row0 + (S + 1); chunk0 = nch + (S + 1);
        SKB_CUDA(cudaMemsetAsync(hist, 0, (size_t)(S + 1) * 4 * sizeof(int32_t), st));
        unsigned nb = (unsigned)((n + 255) / 256), sb = (unsigned)((S + 1 + 255) / 256);
        {
            LaunchScope ls(KC_TABLES, st);
            k_iota<<<nb, 256, 0, st>>>(iota, n);
        }
        {
            LaunchScope ls(KC_TABLES, st);
            k_bucket_hist<<<nb, 256, 0, st>>>(bucket, n, hist);
        }
        int end_bit = 1;
        while ((1LL << end_bit) < (long long)S && end_bit < 31) end_bit++;
        size_t tmp_bytes = 0, scan_bytes = 0;
        cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, bucket, keys_out, iota, plan->rows,
                                        (int)n, 0, end_bit, st);
        cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, hist, row0, S + 1, st);
        if (scan_bytes > tmp_bytes) tmp_bytes = scan_bytes;
        SKB_CUDA(cudaMalloc(&tmp, tmp_bytes ? tmp_bytes : 16));
        {
            LaunchScope ls(KC_TABLES, st);
            SKB_CUDA(cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, bucket, keys_out, iota,
                                                     plan->rows, (int)n, 0, end_bit, st));
        }
        {
            LaunchScope ls(KC_TABLES, st);
            SKB_CUDA(cub::DeviceScan::ExclusiveSum(tmp, tmp_bytes, hist, row0, S + 1, st));
        }
        {
            LaunchScope ls(KC_TABLES, st);
            k_chunks_per_bucket<<<sb, 256, 0, st>>>(hist, S, nch);
        }
        {
            LaunchScope ls(KC_TABLES, st);
            SKB_CUDA(cub::DeviceScan::ExclusiveSum(tmp, tmp_bytes, nch, chunk0, S + 1, st));
        }
        int32_t nchunks = 0;
        SKB_CUDA(cudaMemcpyAsync(&nchunks, chunk0 + S, sizeof(int32_t), cudaMemcpyDeviceToHost, st));
        SKB_CUDA(cudaStreamSynchronize(st));
        plan->nchunks = nchunks;
        SKB_CUDA(cudaMalloc(&plan->chunk_start, (size_t)(nchunks + 1) * sizeof(int32_t)));
        SKB_CUDA(cudaMalloc(&plan->chunk_bucket, (size_t)(nchunks + 1) * sizeof(int32_t)));
        {
            LaunchScope ls(KC_TABLES, st);
            k_fill_chunks<<<sb, 256, 0, st>>>(hist, row0, chunk0, S, plan->chunk_start,
                                              plan->chunk_bucket, nchunks, (int32_t)n);
        }
        SKB_CUDA(cudaGetLastError());
        SKB_CUDA(cudaStreamSynchronize(st));
    } catch (...) {
        cudaFree(iota); cudaFree(keys_out); cudaFree(hist); cudaFree(tmp);
        hash_free_plan(plan);
        throw;
    }
    cudaFree(iota); cudaFree(keys_out); cudaFree(hist); cudaFree(tmp);
}

// ---------------------------------------------------------------------------
// K1: contiguous-row layout.  One warp = one chunk (<=32 rows of one bucket) x one
// 64-column tile; lane l keeps (row id, value) of the chunk's l-th row and
// broadcasts them by shuffle, so the address chain has no dependent global load.
// ---------------------------------------------------------------------------
__device__ __forceinline__ double2 ldg_stream2(const double *p) {
    double2 v;
    asm volatile("ld.global.nc.L1::no_allocate.v2.f64 {%0, %1}, [%2];"
                 : "=d"(v.x), "=d"(v.y) : "l"(p));
    return v;
}
__device__ __forceinline__ double ldg_stream1(const double *p) {
    double v;
    asm volatile("ld.global.nc.L1::no_allocate.f64 %0, [%1];" : "=d"(v) : "l"(p));
    return v;
}

// 64 registers -> 4 CTAs (32 warps) per SM: measured 75% of HBM at C1 against 72% at 79 registers
template <int VEC>
__global__ void __launch_bounds__(256, 4)
k_hash_rows(const double *__restrict__ X, size_t ldx, int m,
            const int32_t *__restrict__ rows, const int32_t *__restrict__ chunk_start,
            const int32_t *__restrict__ chunk_bucket, const double *__restrict__ value,
            long long nitems, int ntiles, double *__restrict__ OUT, size_t ldo) {
    const int lane = threadIdx.x & 31;
    long long item = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (item >= nitems) return;
    const int chunk = (int)(item / ntiles), tile = (int)(item % ntiles);
    const int beg = chunk_start[chunk], cnt = chunk_start[chunk + 1] - beg;
    const int cb = chunk_bucket[chunk];
    const bool multi = cb < 0;
    const int b = multi ? ~cb : cb;

    int myrow = 0; double myval = 0.0;
    if (lane < cnt) { myrow = rows[beg + lane]; myval = value[myrow]; }
    int myrow1 = 0; double myval1 = 0.0;           // rows 32..63 of the chunk (CHUNK == 64)
    if (CHUNK > 32 && lane + 32 < cnt) { myrow1 = rows[beg + 32 + lane]; myval1 = value[myrow1]; }

    constexpr int TILE = 32 * VEC;
    const int col = tile * TILE + lane * VEC;
    const bool active = col < m;     // VEC == 2 implies m even, so col+1 < m too
    const double *xcol = X + col;
    double a0 = 0.0, a1 = 0.0;

    constexpr int U = 8;
    for (int r = 0; r < cnt; r += U) {
        double2 x[U]; double v[U];
#pragma unroll
        for (int i = 0; i < U; i++) {
            int k = __shfl_sync(0xffffffffu, (CHUNK > 32 && r >= 32) ? myrow1 : myrow, (r + i) & 31);
            v[i] = __shfl_sync(0xffffffffu, (CHUNK > 32 && r >= 32) ? myval1 : myval, (r + i) & 31);
            x[i].x = 0.0; x[i].y = 0.0;
            if (active && r + i < cnt) {
                const double *p = xcol + (size_t)k * ldx;
                if (VEC == 2) x[i] = ldg_stream2(p); else x[i].x = ldg_stream1(p);
            }
        }
#pragma unroll
        for (int i = 0; i < U; i++) {
            if (r + i < cnt) {   // keep the reference's rounding: product, then sum
                a0 = __dadd_rn(a0, __dmul_rn(v[i], x[i].x));
                if (VEC == 2) a1 = __dadd_rn(a1, __dmul_rn(v[i], x[i].y));
            }
        }
    }
    if (!active) return;
    double *o = OUT + (size_t)b * ldo + col;
    if (multi) {
        atomicAdd(o, a0);
        if (VEC == 2) atomicAdd(o + 1, a1);
    } else {
        if (VEC == 2) { double2 w; w.x = a0; w.y = a1; *reinterpret_cast<double2 *>(o) = w; }
        else *o = a0;
    }
}

static void zero_matrix(double *OUT, size_t ld, size_t rows_contig, size_t nvec, cudaStream_t st) {
    // nvec vectors of rows_contig contiguous doubles, stride ld
    if (nvec == 0 || rows_contig == 0) return;
    if (ld == rows_contig)
        SKB_CUDA(cudaMemsetAsync(OUT, 0, rows_contig * nvec * sizeof(double), st));
    else
        SKB_CUDA(cudaMemset2DAsync(OUT, ld * sizeof(double), 0, rows_contig * sizeof(double),
                                   nvec, st));
}

void hash_apply_rows(const double *X, size_t ldx, size_t m, const HashPlan &plan,
                     const double *value, double *OUT, size_t ldo, int S, cudaStream_t st) {
    zero_matrix(OUT, ldo, m, (size_t)S, st);
    if (plan.nchunks == 0 || m == 0) return;
    bool vec2 = (m % 2 == 0) && (ldx % 2 == 0) && (ldo % 2 == 0) &&
                ((uintptr_t)X % 16 == 0) && ((uintptr_t)OUT % 16 == 0);
    int tile = vec2 ? 64 : 32;
    int ntiles = (int)((m + tile - 1) / tile);
    long long nitems = (long long)plan.nchunks * ntiles;
    long long nblocks = (nitems * 32 + 255) / 256;
    if (nblocks > 0x7fffffffLL) throw error(SKB_ERR_INVALID_PARAM, "hash apply: grid too large");
    LaunchScope ls(KC_HASH, st);
    if (vec2)
        k_hash_rows<2><<<(unsigned)nblocks, 256, 0, st>>>(X, ldx, (int)m, plan.rows,
            plan.chunk_start, plan.chunk_bucket, value, nitems, ntiles, OUT, ldo);
    else
        k_hash_rows<1><<<(unsigned)nblocks, 256, 0, st>>>(X, ldx, (int)m, plan.rows,
            plan.chunk_start, plan.chunk_bucket, value, nitems, ntiles, OUT, ldo);
    SKB_CUDA(cudaGetLastError());
}

// ---------------------------------------------------------------------------
// K2: strided layout (every column of X is contiguous along the sketched index).
// A CTA owns CG columns and a slice of the sketched index; it accumulates into a
// shared S x CG tile and flushes it (atomics only if the index is split).
// ---------------------------------------------------------------------------
template <int CG, int KU>
__global__ void __launch_bounds__(256)
k_hash_cols_smem(const double *__restrict__ X, size_t ldx, size_t n, size_t m,
                 const int32_t *__restrict__ bucket, const double *__restrict__ value,
                 double *__restrict__ OUT, size_t ldo, int S, int ksplit) {
    extern __shared__ double acc[];   // [CG][S]
    const size_t j0 = (size_t)blockIdx.x * CG;
    const int ncol = (int)min((size_t)CG, m - j0);
    for (int i = threadIdx.x; i < S * CG; i += blockDim.x) acc[i] = 0.0;
    __syncthreads();
    const size_t per = (n + ksplit - 1) / ksplit;
    const size_t k0 = (size_t)blockIdx.y * per;
    const size_t k1 = min(n, k0 + per);
    // KU index steps per trip: KU*CG independent loads in flight per thread before the atomics
    for (size_t kb = k0 + threadIdx.x; kb < k1; kb += (size_t)blockDim.x * KU) {
        int h[KU]; double v[KU]; double x[KU][CG];
#pragma unroll
        for (int u = 0; u < KU; u++) {
            const size_t k = kb + (size_t)u * blockDim.x;
            const bool on = k < k1;
            h[u] = on ? bucket[k] : 0;
            v[u] = on ? value[k] : 0.0;
#pragma unroll
            for (int c = 0; c < CG; c++) x[u][c] = (on && c < ncol) ? ldg_stream1(X + k + (j0 + c) * ldx) : 0.0;
        }
#pragma unroll
        for (int u = 0; u < KU; u++) {
            if (kb + (size_t)u * blockDim.x < k1) {
#pragma unroll
                for (int c = 0; c < CG; c++)
                    if (c < ncol) atomicAdd(&acc[c * S + h[u]], __dmul_rn(v[u], x[u][c]));
            }
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < S * ncol; i += blockDim.x) {
        int c = i / S, h = i - c * S;
        double a = acc[c * S + h];
        double *o = OUT + h + (j0 + c) * ldo;
        if (ksplit == 1) *o = a;
        else if (a != 0.0) atomicAdd(o, a);
    }
}


// Fallback when S is too large for shared memory: L2 atomics.
__global__ void k_hash_cols_global(const double *__restrict__ X, size_t ldx, size_t n, size_t m,
                                   const int32_t *__restrict__ bucket,
                                   const double *__restrict__ value, double *__restrict__ OUT,
                                   size_t ldo) {
    size_t k = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    size_t j = blockIdx.y;
    if (k >= n) return;
    atomicAdd(OUT + bucket[k] + j * ldo, __dmul_rn(value[k], X[k + j * ldx]));
}

void hash_apply_cols(const double *X, size_t ldx, size_t n, size_t m, const int32_t *bucket,
                     const double *value, double *OUT, size_t ldo, int S, cudaStream_t st) {
    if (m == 0) return;
    constexpr int CG = 4, KU = 4;
    size_t smem = (size_t)S * CG * sizeof(double);
    if (smem <= 200 * 1024) {
        unsigned gx = (unsigned)((m + CG - 1) / CG);
        int ksplit = 1;
        // aim for the CTAs that fit per SM (by shared memory, <= 4); every split costs an
        // S x CG atomic flush
        long long per_sm = (long long)((220 * 1024) / (smem ? smem : 1));
        if (per_sm > 4) per_sm = 4;
        if (per_sm < 1) per_sm = 1;
        while ((long long)gx * ksplit < 148 * per_sm && (n / (ksplit * 2)) >= 4096) ksplit *= 2;
        if (ksplit > 1) zero_matrix(OUT, ldo, (size_t)S, m, st);
        SKB_CUDA(cudaFuncSetAttribute(k_hash_cols_smem<CG, KU>,
                                      cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        LaunchScope ls(KC_HASH, st);
        k_hash_cols_smem<CG, KU><<<dim3(gx, ksplit), 256, smem, st>>>(X, ldx, n, m, bucket, value,
                                                                   OUT, ldo, S, ksplit);
        SKB_CUDA(cudaGetLastError());
    } else {
        zero_matrix(OUT, ldo, (size_t)S, m, st);
        if (n == 0) return;
        if (m > 65535) throw error(SKB_ERR_INVALID_PARAM, "hash apply: too many columns");
        LaunchScope ls(KC_HASH, st);
        k_hash_cols_global<<<dim3((unsigned)((n + 255) / 256), (unsigned)m), 256, 0, st>>>(
            X, ldx, n, m, bucket, value, OUT, ldo);
        SKB_CUDA(cudaGetLastError());
    }
}

// ---------------------------------------------------------------------------
// K2p: strided layout without atomics (see HashColsPlan).  A CTA owns CG columns and a range of
// index tiles.  Tiles of 2048 indices x CG columns are staged in shared memory by cp.async
// (coalesced 16-byte pieces, double buffered).  Warp w owns
// the buckets [w * SW, (w + 1) * SW) and walks the rounds the plan queued for it (one contiguous
// stream per warp, read ahead across tile boundaries): in a round every lane takes a row of a
// DIFFERENT bucket, reads the CG values of that index from the tile and does load-add-store on the
// bucket's accumulators.  The plan also arranges each half warp of a round on 16 different bank pairs
// for both accesses, so a round costs 6 * CG shared-memory wavefronts.  Rows of a bucket are added
// in ascending order (the reference's rounding sequence, hash_transform_Elemental.hpp:83-100), the
// accumulators are written out once.
// History at C1 (100k x 500 -> 2000, column-major): 0.216 ms with shared-memory atomics; 0.172 ms
// with lane-owned buckets (44 % of the lane slots busy); 0.109 ms in this form.  What did NOT help:
// several batches of entry loads in flight per warp (waiting on the oldest also waits for the
// newer ones that share its scoreboard), and denser rounds without the bank-pair matching (as many
// wavefronts as before, in fewer instructions).
// ---------------------------------------------------------------------------
#ifndef SKB_HC_PF
#define SKB_HC_PF 0          // L2 prefetch distance in tiles (0 = none: measured 0.109 ms at C1, 4: 0.113, 8: 0.127)
#endif
constexpr int HC_K = 2048, HC_W = 8, HC_THREADS = HC_W * 32, HC_NBUF = 2, HC_PF = SKB_HC_PF;
#ifndef SKB_HC_NB
#define SKB_HC_NB 8
#endif
constexpr int HC_B = 2, HC_NB = SKB_HC_NB;      // rounds applied together; rounds per register set of plan entries      // rounds per batch; batches of plan entries in flight per warp

void hash_free_cols_plan(HashColsPlan *plan) {
    cudaFree(plan->entries); cudaFree(plan->vals); cudaFree(plan->off);
    *plan = HashColsPlan();
}

int hash_cols_plan_bucket_width(int S) { return (((S + HC_W - 1) / HC_W) + 31) / 32 * 32; }

// The plan itself: host arithmetic only (no CUDA), so the CPU test-suite can check its invariants
// (tests/test_host.py through sl_debug_hash_cols_plan).
void hash_schedule_cols_plan(const int32_t *hb, const double *hv, size_t n, int S, std::vector<uint32_t> &entries,
                             std::vector<double> &vals, std::vector<uint32_t> &off, bool *signs_out) {
    const int SW = hash_cols_plan_bucket_width(S);
    const int ntiles = (int)((n + HC_K - 1) / HC_K);
    bool signs = true;
    for (size_t k = 0; k < n && signs; k++) signs = hv[k] == 1.0 || hv[k] == -1.0;
    *signs_out = signs;
    entries.clear(); vals.clear();
    // Warp-major: the rounds of warp w for tiles 0, 1, ... follow each other, every (warp, tile) run
    // padded with empty rounds to whole batches of HC_B, so a warp reads ONE contiguous stream and its
    // prefetch runs across tile boundaries.  off[w * (ntiles + 1) + t] = first round of (w, t).
    const size_t nt = (size_t)ntiles;
    std::vector<std::vector<uint32_t>> we(HC_W);
    std::vector<std::vector<double>> wv(HC_W);
    off.assign((size_t)HC_W * (nt + 1), 0u);
    std::vector<size_t> wr(HC_W, 0);
    // A warp owns the SW buckets [w * SW, (w + 1) * SW); ANY lane may take any of its rows, as long as
    // the 32 rows of a round go to 32 different buckets (plain load-add-store, no atomics) and a bucket
    // sees its rows in ascending order.  Rounds are filled from the buckets with the most rows left
    // (that needs max(ceil(rows / 32), longest bucket) rounds, the minimum), and each half warp -- one
    // shared-memory wavefront of 8-byte accesses -- gets 16 different accumulator bank pairs
    // (bucket mod 16) where the rows allow it, then different tile bank pairs (k mod 16).
    // An empty slot adds element 0 of the tile into a spare accumulator of its lane (bucket SW + lane,
    // never written out): the round needs no branch.
    std::vector<std::vector<uint32_t>> fifo((size_t)HC_W * SW);
    std::vector<uint32_t> head((size_t)SW), cand;
    std::vector<char> in_group((size_t)SW);
    // a tile touches at most 2048 of the HC_W * SW buckets: only those are visited (the cost of a plan
    // stays O(n log) whatever S is)
    std::vector<int> touched;
    std::vector<std::vector<uint32_t>> live(HC_W);         // per warp: its buckets with rows in this tile
    for (size_t t = 0; t < nt; t++) {
        const size_t k0 = t * HC_K, k1 = std::min(n, k0 + HC_K);
        for (int h : touched) fifo[(size_t)h].clear();
        touched.clear();
        for (int w = 0; w < HC_W; w++) live[w].clear();
        for (size_t k = k0; k < k1; k++) {                  // hb < S <= HC_W * SW
            const int h = hb[k];
            if (fifo[(size_t)h].empty()) { touched.push_back(h); live[h / SW].push_back((uint32_t)(h % SW)); }
            fifo[(size_t)h].push_back((uint32_t)(k - k0));
        }
        for (int w = 0; w < HC_W; w++) {
            off[(size_t)w * (nt + 1) + t] = (uint32_t)wr[w];         // made absolute below
            std::vector<uint32_t> *fw = fifo.data() + (size_t)w * SW;
            std::sort(live[w].begin(), live[w].end());              // bucket order: ties below resolve as before
            size_t left = 0;
            size_t nr = 0;
            int degA[16] = {0}, degX[16] = {0};           // rows left per accumulator / tile bank pair
            for (uint32_t bk : live[w]) {
                head[bk] = 0; left += fw[bk].size(); in_group[bk] = 0;
                for (uint32_t kk : fw[bk]) { degA[bk & 15]++; degX[kk & 15]++; }
            }
            while (left) {
                // one round = two half warps; a half warp takes rows with 16 different accumulator bank
                // pairs AND 16 different tile bank pairs (a matching between the two), so each of its
                // six shared-memory accesses is one wavefront.  Rows whose bank pairs have the most
                // rows left go first -- those pairs decide how many rounds the tile needs.
                // (the HC_B rounds the kernel applies together also keep to different buckets: it reads
                // all their accumulators before it stores any)
                if (nr % HC_B == 0) for (uint32_t bk : live[w]) in_group[bk] = 0;
                cand.clear();
                for (uint32_t bk : live[w]) if (head[bk] < fw[bk].size() && !in_group[bk]) cand.push_back(bk);
                std::stable_sort(cand.begin(), cand.end(), [&](uint32_t x, uint32_t y) {
                    const int px = std::max(degA[x & 15], degX[fw[x][head[x]] & 15]);
                    const int py = std::max(degA[y & 15], degX[fw[y][head[y]] & 15]);
                    if (px != py) return px > py;
                    return fw[x].size() - head[x] > fw[y].size() - head[y];
                });
                int slot_b[32]; uint32_t slot_k[32];
                for (int l = 0; l < 32; l++) slot_b[l] = -1;
                int fill[2] = {0, 0};
                bool acc_used[2][16] = {{false}}, x_used[2][16] = {{false}};
                for (uint32_t bk : cand) {
                    if (fill[0] + fill[1] == 32) break;
                    const uint32_t kk = fw[bk][head[bk]];
                    const int c = (int)(bk & 15), xc = (int)(kk & 15);
                    int h = -1;
                    for (int hh = 0; hh < 2 && h < 0; hh++) {
                        const int g = fill[0] <= fill[1] ? hh : 1 - hh;            // emptier half first
                        if (fill[g] < 16 && !acc_used[g][c] && !x_used[g][xc]) h = g;
                    }
                    if (h < 0) continue;
                    head[bk]++; left--; degA[c]--; degX[xc]--; in_group[bk] = 1;
                    slot_b[h * 16 + fill[h]] = (int)bk; slot_k[h * 16 + fill[h]] = kk;
                    fill[h]++; acc_used[h][c] = true; x_used[h][xc] = true;
                }
                for (int h = 0; h < 2; h++) {
                    // idle lanes: the tile element of a busy lane of the half (same address = broadcast),
                    // or element 0, and a spare accumulator on a bank pair the half does not use
                    const uint32_t kfill = fill[h] ? slot_k[h * 16] : 0u;
                    int c = 0;
                    for (int i = 0; i < 16; i++) {
                        const int l = h * 16 + i;
                        if (slot_b[l] >= 0) {
                            const size_t k = k0 + slot_k[l];
                            uint32_t e = (slot_k[l] << 3) | ((uint32_t)slot_b[l] << 16);
                            if (signs && hv[k] < 0) e |= 0x80000000u;
                            we[w].push_back(e);
                            if (!signs) wv[w].push_back(hv[k]);
                        } else {
                            while (acc_used[h][c]) c++;
                            we[w].push_back((kfill << 3) | ((uint32_t)(SW + h * 16 + c) << 16));
                            if (!signs) wv[w].push_back(0.0);
                            c++;
                        }
                    }
                }
                nr++;
            }
            for (; nr % HC_B; nr++)
                for (int l = 0; l < 32; l++) {
                    we[w].push_back((uint32_t)(SW + l) << 16);
                    if (!signs) wv[w].push_back(0.0);
                }
            wr[w] += nr;
        }
    }
    size_t base = 0;
    for (int w = 0; w < HC_W; w++) {
        for (size_t t = 0; t < nt; t++) off[(size_t)w * (nt + 1) + t] += (uint32_t)base;
        off[(size_t)w * (nt + 1) + nt] = (uint32_t)(base + wr[w]);
        entries.insert(entries.end(), we[w].begin(), we[w].end());
        if (!signs) vals.insert(vals.end(), wv[w].begin(), wv[w].end());
        base += wr[w];
    }
    // the prefetch reads HC_Q batches past the round it works on
    for (int z = 0; z < 2 * HC_NB; z++)                     // the prefetch runs two sets past the last round
        for (int l = 0; l < 32; l++) entries.push_back((uint32_t)(SW + l) << 16);
    if (!signs) vals.insert(vals.end(), (size_t)2 * HC_NB * 32, 0.0);
}

void hash_build_cols_plan(const int32_t *bucket, const double *value, size_t n, int S, HashColsPlan *plan,
                          cudaStream_t st) {
    hash_free_cols_plan(plan);
    plan->n = n; plan->S = S;
    plan->SW = hash_cols_plan_bucket_width(S);
    plan->ntiles = (int)((n + HC_K - 1) / HC_K);
    if (n == 0) return;
    std::vector<int32_t> hb(n); std::vector<double> hv(n);
    SKB_CUDA(cudaMemcpyAsync(hb.data(), bucket, n * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
    SKB_CUDA(cudaMemcpyAsync(hv.data(), value, n * sizeof(double), cudaMemcpyDeviceToHost, st));
    SKB_CUDA(cudaStreamSynchronize(st));
    std::vector<uint32_t> entries, off;
    std::vector<double> vals;
    bool signs = true;
    hash_schedule_cols_plan(hb.data(), hv.data(), n, S, entries, vals, off, &signs);
    try {
        SKB_CUDA(cudaMalloc(&plan->entries, entries.size() * sizeof(uint32_t) + 256));
        SKB_CUDA(cudaMalloc(&plan->off, off.size() * sizeof(uint32_t)));
        SKB_CUDA(cudaMemcpy(plan->entries, entries.data(), entries.size() * sizeof(uint32_t), cudaMemcpyHostToDevice));
        SKB_CUDA(cudaMemcpy(plan->off, off.data(), off.size() * sizeof(uint32_t), cudaMemcpyHostToDevice));
        if (!signs) {
            SKB_CUDA(cudaMalloc(&plan->vals, vals.size() * sizeof(double) + 256));
            SKB_CUDA(cudaMemcpy(plan->vals, vals.data(), vals.size() * sizeof(double), cudaMemcpyHostToDevice));
        }
    } catch (...) { hash_free_cols_plan(plan); throw; }
}

__device__ __forceinline__ uint64_t hc_policy_evict_last() {
    uint64_t pol; asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol)); return pol;
}
__device__ __forceinline__ uint64_t hc_policy_evict_first() {
    uint64_t pol; asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol)); return pol;
}
__device__ __forceinline__ int ldg_i32_hint(const int32_t *p, uint64_t pol) {
    int v; asm volatile("ld.global.nc.L2::cache_hint.s32 %0, [%1], %2;" : "=r"(v) : "l"(p), "l"(pol)); return v;
}
__device__ __forceinline__ double ldg_f64_hint(const double *p, uint64_t pol) {
    double v; asm volatile("ld.global.nc.L2::cache_hint.f64 %0, [%1], %2;" : "=d"(v) : "l"(p), "l"(pol)); return v;
}
__device__ __forceinline__ unsigned ldg_u32_hint(const uint32_t *p, uint64_t pol) {
    unsigned v; asm volatile("ld.global.nc.L2::cache_hint.u32 %0, [%1], %2;" : "=r"(v) : "l"(p), "l"(pol)); return v;
}
__device__ __forceinline__ double hc_lds(unsigned a) {
    double v; asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(a)); return v;
}
__device__ __forceinline__ void hc_sts(unsigned a, double v) {
    asm volatile("st.shared.f64 [%0], %1;" ::"r"(a), "d"(v) : "memory");
}
// accumulators: [CG][HC_W][SW + 32] -- the buckets of warp w, then one spare per lane
__host__ __device__ inline int hc_acc_stride(int SW) { return HC_W * (SW + 32); }

template <int CG, bool SIGN>
__global__ void __launch_bounds__(HC_THREADS, CG == 4 ? 1 : 2)
k_hash_cols_plan(const double *__restrict__ X, size_t ldx, size_t n, size_t m,
                 const uint32_t *__restrict__ entries, const double *__restrict__ vals,
                 const uint32_t *__restrict__ off, int ntiles, int tiles_per_split,
                 double *__restrict__ OUT, size_t ldo, int S, int SW, int ksplit, int dbg) {
    extern __shared__ __align__(16) double hc_smem[];
    const int SPp = hc_acc_stride(SW);
    double *acc = hc_smem;                                  // [CG][SPp]
    double *tiles = hc_smem + (size_t)CG * SPp;             // [HC_NBUF][CG][HC_K]
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const size_t j0 = (size_t)blockIdx.x * CG;
    const int ncol = (int)min((size_t)CG, m - j0);
    const int t0 = blockIdx.y * tiles_per_split, t1 = min(ntiles, t0 + tiles_per_split);
    // columns past the last one are never staged: whatever their buffers hold goes into accumulators
    // that are not written out; start those from zeros all the same
    for (int i = tid; i < CG * SPp + HC_NBUF * CG * HC_K; i += HC_THREADS) hc_smem[i] = 0.0;
    __syncthreads();
    // the plan entries are re-read by every CTA while the operand streams through L2: ask for them to stay
    const uint64_t pol_last = hc_policy_evict_last();
    auto issue = [&](int t, int buf) {
        const size_t k0 = (size_t)t * HC_K;
        const int len = (int)min((size_t)HC_K, n - k0);
        if (dbg & 1) { asm volatile("cp.async.commit_group;" ::: "memory"); return; }
        if (len == HC_K) {                                  // whole tile: 4 pieces of 16 bytes per thread and column
#pragma unroll
            for (int c = 0; c < CG; c++) {
                if (c < ncol) {
                    const double *src = X + k0 + (j0 + c) * ldx + tid * 2;
                    const unsigned dst = (unsigned)__cvta_generic_to_shared(tiles + ((size_t)(buf * CG + c) * HC_K + tid * 2));
#pragma unroll
                    for (int i = 0; i < HC_K / (2 * HC_THREADS); i++)
                        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + i * HC_THREADS * 16),
                                     "l"(src + i * HC_THREADS * 2) : "memory");
                }
            }
        } else {
            for (int idx = tid; idx < CG * (HC_K / 2); idx += HC_THREADS) {
                const int c = idx / (HC_K / 2), p2 = (idx - c * (HC_K / 2)) * 2;
                if (c < ncol && p2 < len) {
                    const double *src = X + k0 + p2 + (j0 + c) * ldx;
                    const unsigned dst = (unsigned)__cvta_generic_to_shared(tiles + ((size_t)(buf * CG + c) * HC_K + p2));
                    if (p2 + 1 < len) asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
                    else asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(dst), "l"(src) : "memory");
                }
            }
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
        if (HC_PF > 0 && tid < ncol) {                      // keep L2 a few tiles ahead of the staging loads
            const size_t kp = (size_t)(t + HC_PF) * HC_K;
            if (t + HC_PF < t1 && kp + HC_K <= n)
                asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(X + kp + (j0 + tid) * ldx), "r"(HC_K * 8));
        }
    };
    // the round offsets of this warp's tiles come through a 64-entry ring in shared memory, refilled
    // 32 tiles ahead (a tile can be shorter than the L2 latency of a load issued when it starts)
    const uint32_t *woff = off + (size_t)warp * (ntiles + 1) + t0;
    volatile uint32_t *ring = reinterpret_cast<uint32_t *>(tiles + (size_t)HC_NBUF * CG * HC_K) + warp * 64;
    const int nti = t1 - t0;                                // offsets 0 .. nti are this CTA's
    if (nti > 0) {
        if (lane <= nti) ring[lane] = __ldg(woff + lane);
        if (lane + 32 <= nti) ring[lane + 32] = __ldg(woff + lane + 32);
    }
    __syncwarp();
    uint32_t r = nti > 0 ? ring[0] : 0u;
    // Plan entries come from L2 behind the tile traffic, more than a microsecond away.  Two register
    // sets of HC_NB rounds: while one is applied the other is in flight.  A set is refilled only AFTER
    // the first entry of the other set has been waited for (the refill addresses depend on it): the
    // hardware tracks loads with a handful of scoreboards, and waiting on an old load while newer ones
    // share its scoreboard would wait for those as well -- a queue of several batches in flight
    // measured no better than one.
    uint32_t ea[HC_NB], eb[HC_NB]; double va[HC_NB], vb[HC_NB];
    const uint32_t *ep = entries + (size_t)r * 32 + lane;   // next set to fetch
    const double *vp = vals + (size_t)r * 32 + lane;
    auto fetch = [&](uint32_t (&e)[HC_NB], double (&v)[HC_NB], uint32_t dep) {
        const uint32_t *ep_ = ep + dep; const double *vp_ = vp + dep;
#pragma unroll
        for (int i = 0; i < HC_NB; i++) {
            if (dbg & 4) e[i] = (((unsigned)(size_t)ep_ * 2654435761u + (unsigned)lane * 40503u + (unsigned)i * 977u) & 0x3FF8u) | (((unsigned)lane * 7u + (unsigned)i) & 0xffu) << 16;
            else e[i] = ldg_u32_hint(ep_ + i * 32, pol_last);
            v[i] = SIGN ? 0.0 : __ldg(vp_ + i * 32);
        }
        ep += HC_NB * 32; vp += HC_NB * 32;
    };
    const unsigned accb = (unsigned)__cvta_generic_to_shared(acc + warp * (SW + 32));   // + bucket of the warp * 8 bytes
    const unsigned tiles_s = (unsigned)__cvta_generic_to_shared(tiles);
    int t = t0 - 1;
    uint32_t rend = r;
    unsigned tile = tiles_s;
    bool live = nti > 0;
    if (live) {
        issue(t0, 0);
        fetch(ea, va, 0u);
    }
    // applies the HC_NB rounds of one set; a tile may end after any pair of rounds, so the hand-over to
    // the next tile (one barrier per tile per warp) sits in front of every pair
    auto apply_set = [&](uint32_t (&e_)[HC_NB], double (&v_)[HC_NB]) {
#pragma unroll
        for (int g = 0; g < HC_NB; g += HC_B) {
            while (live && r == rend) {                     // this warp has finished tile t
                if (++t == t1) { live = false; break; }
                asm volatile("cp.async.wait_group 0;" ::: "memory");
                asm volatile("bar.sync 0;" ::: "memory");   // tile t has landed; every warp is done with tile t - 1
                const int buf = (t - t0) & 1;
                if (t + 1 < t1) issue(t + 1, buf ^ 1);
                const int ti = t - t0;
                if (ti > 0 && (ti & 31) == 0 && ti + 32 + lane <= nti)
                    ring[(ti + 32 + lane) & 63] = __ldg(woff + ti + 32 + lane);
                __syncwarp();
                rend = ring[(ti + 1) & 63];
                tile = tiles_s + (unsigned)buf * (CG * HC_K * 8);
            }
            if (!live) return;
            if (dbg & 2) { r = rend; continue; }
            // the plan keeps the HC_B * 32 rows of a group on different buckets: all loads, then the
            // adds, then all stores
            unsigned aa[HC_B];
            double x[HC_B][CG], a[HC_B][CG];
#pragma unroll
            for (int i = 0; i < HC_B; i++) {
                const uint32_t e = e_[g + i];
                const unsigned xa = tile + (e & 0x3FF8u);                       // index k of the tile, column 0
                aa[i] = accb + ((e >> 13) & 0xFFF8u);                           // bucket of the warp: (e >> 16) & 0x1fff
#pragma unroll
                for (int c = 0; c < CG; c++) x[i][c] = hc_lds(xa + c * (HC_K * 8));
#pragma unroll
                for (int c = 0; c < CG; c++) a[i][c] = hc_lds(aa[i] + c * SPp * 8);
            }
#pragma unroll
            for (int i = 0; i < HC_B; i++) {
                const uint32_t e = e_[g + i];
#pragma unroll
                for (int c = 0; c < CG; c++) {              // product, then sum: the reference's rounding
                    const double p = SIGN ? __hiloint2double(__double2hiint(x[i][c]) ^ (int)(e & 0x80000000u), __double2loint(x[i][c]))
                                          : __dmul_rn(v_[g + i], x[i][c]);
                    a[i][c] = __dadd_rn(a[i][c], p);
                }
            }
#pragma unroll
            for (int i = 0; i < HC_B; i++)
#pragma unroll
                for (int c = 0; c < CG; c++) hc_sts(aa[i] + c * SPp * 8, a[i][c]);
            __syncwarp();                   // the next group's lanes may read what other lanes stored here
            r += HC_B;
        }
    };
    while (live) {
        fetch(eb, vb, ea[0] & 0x4000u);     // bits 14..15 of an entry are zero: a data dependence, not an offset
        apply_set(ea, va);
        if (!live) break;
        fetch(ea, va, eb[0] & 0x4000u);
        apply_set(eb, vb);
    }
    __syncthreads();
    for (int i = tid; i < S * ncol; i += HC_THREADS) {
        const int c = i / S, h = i - c * S, w = h / SW;
        const double a = acc[c * SPp + w * (SW + 32) + (h - w * SW)];
        double *o = OUT + h + (j0 + c) * ldo;
        if (ksplit == 1) *o = a;
        else if (a != 0.0) atomicAdd(o, a);
    }
}

bool hash_apply_cols_planned(const double *X, size_t ldx, size_t n, size_t m, const HashColsPlan &plan,
                             double *OUT, size_t ldo, int S, cudaStream_t st) {
    static const bool off_env = getenv("SKB_DEBUG_NO_COLS_PLAN") != nullptr;
    if (off_env || m == 0 || n == 0 || plan.n != n || plan.S != S || !plan.entries) return false;
    // fewer than 32 buckets per warp leave warps (S < 256: 8 warps x 32 buckets) without work while the
    // others serialise on their few buckets: the shared-atomics kernel is faster there (measured at
    // 100k x 500: S = 32: 0.26 ms against 0.46; 64: 0.22 / 0.27; 256: equal; 1000: 0.175 / 0.113)
    if (S < 256 || S > 32767 || (ldx & 1) || ((uintptr_t)X & 15)) return false;
    const int SPp = hc_acc_stride(plan.SW);
    auto smem_for = [&](int cg) { return ((size_t)cg * SPp + (size_t)HC_NBUF * cg * HC_K) * sizeof(double) + HC_W * 64 * sizeof(uint32_t); };
    // two columns per CTA and two CTAs per SM (one's barriers and staging overlap the other's rounds)
    // when that still fills the GPU: measured 0.172 ms against 0.209 ms with four columns at C1
    int CG = (m + 1) / 2 >= 148 && smem_for(2) <= 108 * 1024 ? 2 : 4;
    { static const char *e = getenv("SKB_HC_CG"); if (e && (atoi(e) == 1 || atoi(e) == 2 || atoi(e) == 4)) CG = atoi(e); }
    while (CG > 1 && smem_for(CG) > 216 * 1024) CG >>= 1;
    if (smem_for(CG) > 216 * 1024) return false;
    const unsigned gx = (unsigned)((m + CG - 1) / CG);
    int ksplit = 1;
    while ((long long)gx * ksplit < 120 && plan.ntiles / (ksplit * 2) >= 4) ksplit *= 2;
    const int tps = (plan.ntiles + ksplit - 1) / ksplit;
    if (ksplit > 1) zero_matrix(OUT, ldo, (size_t)S, m, st);
    const bool sign = plan.vals == nullptr;
    const size_t smem = smem_for(CG);
    // timing experiments only (results are wrong): 1 = no tile staging, 2 = no rounds, 4 = made-up entries
    static const int dbg = getenv("SKB_HC_DBG") ? atoi(getenv("SKB_HC_DBG")) : 0;
    LaunchScope ls(KC_HASH, st);
#define SKB_HC_LAUNCH(CGV, SG)                                                                               \
    do {                                                                                                    \
        auto kern = k_hash_cols_plan<CGV, SG>;                                                              \
        SKB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));       \
        kern<<<dim3(gx, (unsigned)ksplit), HC_THREADS, smem, st>>>(X, ldx, n, m, plan.entries, plan.vals,   \
                                                                   plan.off, plan.ntiles, tps, OUT, ldo, S, plan.SW, ksplit, dbg); \
    } while (0)
    if (CG == 4) { if (sign) SKB_HC_LAUNCH(4, true); else SKB_HC_LAUNCH(4, false); }
    else if (CG == 2) { if (sign) SKB_HC_LAUNCH(2, true); else SKB_HC_LAUNCH(2, false); }
    else { if (sign) SKB_HC_LAUNCH(1, true); else SKB_HC_LAUNCH(1, false); }
#undef SKB_HC_LAUNCH
#undef SKB_HC_LAUNCH
    SKB_CUDA(cudaGetLastError());
    return true;
}

// ---------------------------------------------------------------------------
// K3: CSC input, sketch over rows.  Persistent CTAs; each column is accumulated in
// a shared S-vector (shared atomics), then written as one coalesced dense column
// while the accumulator is re-zeroed.
// ---------------------------------------------------------------------------
template <bool VEC2>
__global__ void __launch_bounds__(512)
k_hash_csc_colwise_smem(const int32_t *__restrict__ indptr, const int32_t *__restrict__ indices,
                        const double *__restrict__ values, size_t ncols,
                        const int32_t *__restrict__ bucket, const double *__restrict__ value,
                        double *__restrict__ OUT, size_t ldo, int S) {
    extern __shared__ double acc[];   // [S]
    for (int i = threadIdx.x; i < S; i += blockDim.x) acc[i] = 0.0;
    __syncthreads();
    // output and nonzeros stream through L2 once (evict-first); the bucket / value tables are gathered at
    // random by every column and are asked to stay (see k_hash_csc_colwise_bulk)
    const uint64_t pol_first = hc_policy_evict_first(), pol_last = hc_policy_evict_last();
    // Three-deep software pipeline over the columns this CTA owns.  A nonzero needs three
    // dependent trips to L2/HBM (indptr -> index/value -> bucket/value tables); each trip of a
    // LATER column is issued before the current column's dense S-vector is written out and its
    // result is only touched one iteration later, so the whole chain hides behind the dumps.
    constexpr int PRE = 4;            // nonzeros per thread held in registers (4 * 512 per column)
    const size_t stride = gridDim.x;
    int hC[PRE]; double wC[PRE], xC[PRE];          // current column: bucket, table value, entry
    int rN[PRE]; double xN[PRE];                   // next column: row index, entry
    int q0 = 0, q1 = 0;                            // column after next: its indptr pair
    int e0 = 0, e1 = 0, t0 = 0, t1 = 0;            // the same pair, carried along with the column's stages
    auto load_ptr = [&](size_t col) {
        q0 = q1 = 0;
        if (col < ncols) { q0 = indptr[col]; q1 = indptr[col + 1]; }
    };
    auto load_entries = [&]() {                    // uses q0/q1
#pragma unroll
        for (int i = 0; i < PRE; i++) {
            const int p = q0 + (int)threadIdx.x + i * (int)blockDim.x;
            rN[i] = -1; xN[i] = 0.0;
            if (p < q1) { rN[i] = ldg_i32_hint(indices + p, pol_first); xN[i] = ldg_f64_hint(values + p, pol_first); }
        }
        e0 = q0; e1 = q1;
    };
    auto load_tables = [&]() {                     // uses rN/xN, fills hC/wC/xC
#pragma unroll
        for (int i = 0; i < PRE; i++) {
            hC[i] = -1; wC[i] = 0.0; xC[i] = xN[i];
            if (rN[i] >= 0) { hC[i] = ldg_i32_hint(bucket + rN[i], pol_last); wC[i] = ldg_f64_hint(value + rN[i], pol_last); }
        }
        t0 = e0; t1 = e1;
    };
    // prologue: column c fully staged, column c+stride at the entry stage, c+2*stride at indptr
    load_ptr(blockIdx.x); load_entries(); load_tables();
    load_ptr(blockIdx.x + stride); load_entries();
    load_ptr(blockIdx.x + 2 * stride);
    for (size_t col = blockIdx.x; col < ncols; col += stride) {
#pragma unroll
        for (int i = 0; i < PRE; i++)
            if (hC[i] >= 0) atomicAdd(&acc[hC[i]], __dmul_rn(wC[i], xC[i]));
        {   // columns with more than PRE * blockDim nonzeros (p0, p1 arrived three columns ago: a
            // fresh indptr load here would put an L2 round trip on every column's critical path)
            const int p0 = t0, p1 = t1;
            for (int p = p0 + PRE * (int)blockDim.x + (int)threadIdx.x; p < p1; p += blockDim.x) {
                const int row = ldg_i32_hint(indices + p, pol_first);
                atomicAdd(&acc[ldg_i32_hint(bucket + row, pol_last)],
                          __dmul_rn(ldg_f64_hint(value + row, pol_last), ldg_f64_hint(values + p, pol_first)));
            }
        }
        load_tables();                 // column col + stride   (rows arrived an iteration ago)
        load_entries();                // column col + 2*stride (indptr arrived an iteration ago)
        load_ptr(col + 3 * stride);
        __syncthreads();
        double *o = OUT + col * ldo;
        if (VEC2) {
            double2 *o2 = reinterpret_cast<double2 *>(o);
            double2 *a2 = reinterpret_cast<double2 *>(acc);
            const double2 z = make_double2(0.0, 0.0);
            for (int i = threadIdx.x; i < S / 2; i += blockDim.x) { __stcs(o2 + i, a2[i]); a2[i] = z; }
            if ((S & 1) && threadIdx.x == 0) { o[S - 1] = acc[S - 1]; acc[S - 1] = 0.0; }
        } else {
            for (int i = threadIdx.x; i < S; i += blockDim.x) { o[i] = acc[i]; acc[i] = 0.0; }
        }
        __syncthreads();
    }
}

// K3, bulk-store form.  Measured on the C4 shard: the loads + shared-memory sums of all columns take
// 1.5 ms, writing the 16 GB takes 2.1 - 2.6 ms (the write ceiling is 7.5 TB/s; 148 CTAs streaming whole
// columns reach 6.0 - 6.4, tools/colwrite_bw.cu), and a kernel with one S-vector per SM that dumps a
// column and then starts the next pays the SUM (3.7 ms, whether 512 threads store from registers or
// one bulk copy does).  Here a CTA holds only a SLICE of the S-vector (S / NH slots) and walks the NH
// slices of its column one after the other: sum the nonzeros whose bucket falls in the slice (they
// stay in registers for all slices, so nothing is gathered twice), hand the slice to the copy engine
// (cp.async.bulk.global.shared::cta: no registers, no LSU queue), wait until the engine has READ it,
// re-zero the slots it touched.  Several such CTAs share an SM, so one's copy streams while the
// others sum.  L2 hints keep the 15 MB bucket / value tables resident under the 16 GB write stream
// (output and nonzeros evict-first, tables evict-last): without them 3.9 ms, with them 3.4 at NH = 2
// in one CTA per SM.
template <int THREADS>
__global__ void __launch_bounds__(THREADS)
k_hash_csc_colwise_bulk(const int32_t *__restrict__ indptr, const int32_t *__restrict__ indices,
                        const double *__restrict__ values, size_t ncols,
                        const int32_t *__restrict__ bucket, const double *__restrict__ value,
                        double *__restrict__ OUT, size_t ldo, int S, int NH, int dbg) {
    extern __shared__ __align__(16) double acc[];   // [S / NH]; (S / NH) % 2 == 0, columns 16-byte aligned
    const int H = S / NH;
    for (int i = threadIdx.x; i < H; i += THREADS) acc[i] = 0.0;
    __syncthreads();
    constexpr int PRE = 2048 / THREADS;             // nonzeros per thread held in registers
    const size_t stride = gridDim.x;
    const unsigned acc_s = (unsigned)__cvta_generic_to_shared(acc);
    const uint64_t pol_first = hc_policy_evict_first(), pol_last = hc_policy_evict_last();
    int hC[PRE]; double wC[PRE], xC[PRE];
    int rN[PRE]; double xN[PRE];
    int q0 = 0, q1 = 0, e0 = 0, e1 = 0, t0 = 0, t1 = 0;
    auto load_ptr = [&](size_t col) {
        q0 = q1 = 0;
        if (col < ncols) { q0 = indptr[col]; q1 = indptr[col + 1]; }
    };
    auto load_entries = [&]() {
#pragma unroll
        for (int i = 0; i < PRE; i++) {
            const int p = q0 + (int)threadIdx.x + i * THREADS;
            rN[i] = -1; xN[i] = 0.0;
            if (p < q1) { rN[i] = ldg_i32_hint(indices + p, pol_first); xN[i] = ldg_f64_hint(values + p, pol_first); }
        }
        e0 = q0; e1 = q1;
    };
    auto load_tables = [&]() {
#pragma unroll
        for (int i = 0; i < PRE; i++) {
            hC[i] = -1; wC[i] = 0.0; xC[i] = xN[i];
            if (rN[i] >= 0) { hC[i] = ldg_i32_hint(bucket + rN[i], pol_last); wC[i] = ldg_f64_hint(value + rN[i], pol_last); }
        }
        t0 = e0; t1 = e1;
    };
    load_ptr(blockIdx.x); load_entries(); load_tables();
    load_ptr(blockIdx.x + stride); load_entries();
    load_ptr(blockIdx.x + 2 * stride);
    int hP[PRE];                       // buckets whose slots are still set in the buffer (previous slice)
#pragma unroll
    for (int i = 0; i < PRE; i++) hP[i] = -1;
    int pp0 = 0, pp1 = 0, plo = 0;     // the same for nonzeros beyond the registers: their range and slice
    for (size_t col = blockIdx.x; col < ncols; col += stride) {
        const int p0 = t0, p1 = t1;
        int hN[PRE]; double vN[PRE];
#pragma unroll
        for (int i = 0; i < PRE; i++) { hN[i] = hC[i]; vN[i] = __dmul_rn(wC[i], xC[i]); }
        load_tables();                 // column col + stride   (this column's registers are copied above)
        load_entries();                // column col + 2*stride
        load_ptr(col + 3 * stride);
        for (int sl = 0; sl < NH; sl++) {
            const int lo = sl * H, hi = lo + H;
            // the engine has read the previous slice out of the buffer
            if (threadIdx.x == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
            __syncthreads();
#pragma unroll
            for (int i = 0; i < PRE; i++) if (hP[i] >= plo && hP[i] < plo + H) acc[hP[i] - plo] = 0.0;
            for (int p = pp0 + PRE * THREADS + (int)threadIdx.x; p < pp1; p += THREADS) {
                const int h = bucket[indices[p]];
                if (h >= plo && h < plo + H) acc[h - plo] = 0.0;
            }
            __syncthreads();           // the buffer is all zeros again
#pragma unroll
            for (int i = 0; i < PRE; i++)
                if (hN[i] >= lo && hN[i] < hi && !(dbg & 2)) atomicAdd(&acc[hN[i] - lo], vN[i]);
            for (int p = p0 + PRE * THREADS + (int)threadIdx.x; p < p1; p += THREADS) {
                const int row = indices[p], h = bucket[row];
                if (h >= lo && h < hi) atomicAdd(&acc[h - lo], __dmul_rn(value[row], values[p]));
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");    // the sums, for the copy engine
            __syncthreads();
            if (threadIdx.x == 0 && !(dbg & 1)) {
                asm volatile("cp.async.bulk.global.shared::cta.bulk_group.L2::cache_hint [%0], [%1], %2, %3;" ::"l"(OUT + col * ldo + lo),
                             "r"(acc_s), "r"(H * 8), "l"(pol_first) : "memory");
                asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            }
#pragma unroll
            for (int i = 0; i < PRE; i++) hP[i] = hN[i];
            pp0 = p0; pp1 = p1; plo = lo;
        }
    }
    if (threadIdx.x == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

__global__ void k_hash_csc_colwise_global(const int32_t *__restrict__ indptr,
                                          const int32_t *__restrict__ indices,
                                          const double *__restrict__ values, size_t ncols,
                                          const int32_t *__restrict__ bucket,
                                          const double *__restrict__ value,
                                          double *__restrict__ OUT, size_t ldo) {
    const int lane = threadIdx.x & 31;
    size_t warp = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    size_t nwarps = ((size_t)gridDim.x * blockDim.x) >> 5;
    for (size_t col = warp; col < ncols; col += nwarps) {
        const int p0 = indptr[col], p1 = indptr[col + 1];
        for (int p = p0 + lane; p < p1; p += 32) {
            int row = indices[p];
            atomicAdd(OUT + col * ldo + bucket[row], __dmul_rn(value[row], values[p]));
        }
    }
}

static int g_csc_bulk = -1;              // -1: environment (SKB_CSC_BULK), else sl_set_option("csc_form")
void hash_set_csc_form(int bulk) { g_csc_bulk = bulk ? 1 : 0; }

void hash_apply_csc_columnwise(const int32_t *indptr, const int32_t *indices,
                               const double *values, size_t ncols, const int32_t *bucket,
                               const double *value, double *OUT, size_t ldo, int S,
                               cudaStream_t st, size_t nnz_hint) {
    if (ncols == 0) return;
    size_t smem = (size_t)S * sizeof(double);
    static int force_global = -1;
    if (force_global < 0) force_global = getenv("SKB_DEBUG_CSC_GLOBAL") ? 1 : 0;
    if (smem <= 220 * 1024 && !force_global) {
        const bool vec2 = (ldo % 2 == 0) && ((uintptr_t)OUT % 16 == 0);
        // Both forms carry the L2 hints that matter (3.70 -> 3.3 ms at the C4 shard); with them the plain
        // register dump (3.26 ms, any column length) and the bulk-store slices (3.30 ms, columns of up to
        // 2048 nonzeros kept in registers) measure the same, so the simpler one is the default and
        // sl_set_option("csc_form", "bulk") / SKB_CSC_BULK=1 selects the other.
        static const bool env_bulk = getenv("SKB_CSC_BULK") != nullptr;
        const bool want_bulk = g_csc_bulk < 0 ? env_bulk : g_csc_bulk == 1;
        const bool bulk = want_bulk && vec2 && S % 4 == 0 && nnz_hint > 0 && nnz_hint / ncols <= 1536;
        static const int dbg = getenv("SKB_CSC_DBG") ? atoi(getenv("SKB_CSC_DBG")) : 0;   // timing experiments only
        if (bulk) {
            // slices per column / threads per CTA: SKB_CSC_NH, SKB_CSC_THREADS (tuning)
            static const int nh_env = getenv("SKB_CSC_NH") ? atoi(getenv("SKB_CSC_NH")) : 0;
            static const int th_env = getenv("SKB_CSC_THREADS") ? atoi(getenv("SKB_CSC_THREADS")) : 0;
            int NH = nh_env > 0 ? nh_env : 2;
            while (NH > 1 && (S % (2 * NH) != 0)) NH--;
            const int threads = th_env == 512 ? 512 : 256;     // measured at the C4 shard: 3.30 ms (256) / 3.81 (512) at NH = 2
            const size_t sl_bytes = (size_t)(S / NH) * sizeof(double);
            int per_sm = (int)((220 * 1024) / (sl_bytes + 1024));
            const int cap = threads == 256 ? 8 : 2;          // register file: 62 registers x threads
            if (per_sm > cap) per_sm = cap;
            if (per_sm < 1) per_sm = 1;
            size_t grid = (size_t)148 * per_sm;
            if (grid > ncols) grid = ncols;
            LaunchScope ls(KC_HASH, st);
            if (threads == 256) {
                SKB_CUDA(cudaFuncSetAttribute(k_hash_csc_colwise_bulk<256>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sl_bytes));
                k_hash_csc_colwise_bulk<256><<<(unsigned)grid, 256, sl_bytes, st>>>(indptr, indices, values, ncols, bucket, value,
                                                                                    OUT, ldo, S, NH, dbg);
            } else {
                SKB_CUDA(cudaFuncSetAttribute(k_hash_csc_colwise_bulk<512>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sl_bytes));
                k_hash_csc_colwise_bulk<512><<<(unsigned)grid, 512, sl_bytes, st>>>(indptr, indices, values, ncols, bucket, value,
                                                                                    OUT, ldo, S, NH, dbg);
            }
            SKB_CUDA(cudaGetLastError());
            return;
        }
        auto kern = vec2 ? k_hash_csc_colwise_smem<true> : k_hash_csc_colwise_smem<false>;
        SKB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        int per_sm = smem <= 100 * 1024 ? 2 : 1;
        size_t grid = (size_t)148 * per_sm;
        if (grid > ncols) grid = ncols;
        LaunchScope ls(KC_HASH, st);
        kern<<<(unsigned)grid, 512, smem, st>>>(indptr, indices, values, ncols, bucket, value, OUT,
                                                ldo, S);
        SKB_CUDA(cudaGetLastError());
    } else {
        zero_matrix(OUT, ldo, (size_t)S, ncols, st);
        LaunchScope ls(KC_HASH, st);
        k_hash_csc_colwise_global<<<148 * 8, 256, 0, st>>>(indptr, indices, values, ncols, bucket,
                                                         value, OUT, ldo);
        SKB_CUDA(cudaGetLastError());
    }
}

// CSC input, sketch over columns: each sparse column lands in dense output column
// h(col); several input columns share one output column, hence L2 atomics.
__global__ void k_hash_csc_rowwise(const int32_t *__restrict__ indptr,
                                   const int32_t *__restrict__ indices,
                                   const double *__restrict__ values, size_t ncols,
                                   const int32_t *__restrict__ bucket,
                                   const double *__restrict__ value, double *__restrict__ OUT,
                                   size_t ldo) {
    const int lane = threadIdx.x & 31;
    size_t warp = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    size_t nwarps = ((size_t)gridDim.x * blockDim.x) >> 5;
    for (size_t col = warp; col < ncols; col += nwarps) {
        const int p0 = indptr[col], p1 = indptr[col + 1];
        const double v = value[col];
        double *o = OUT + (size_t)bucket[col] * ldo;
        for (int p = p0 + lane; p < p1; p += 32)
            atomicAdd(o + indices[p], __dmul_rn(v, values[p]));
    }
}

void hash_apply_csc_rowwise(const int32_t *indptr, const int32_t *indices,
                            const double *values, size_t nrows, size_t ncols,
                            const int32_t *bucket, const double *value, double *OUT,
                            size_t ldo, int S, cudaStream_t st) {
    zero_matrix(OUT, ldo, nrows, (size_t)S, st);
    if (ncols == 0) return;
    LaunchScope ls(KC_HASH, st);
    k_hash_csc_rowwise<<<148 * 8, 256, 0, st>>>(indptr, indices, values, ncols, bucket, value,
                                                OUT, ldo);
    SKB_CUDA(cudaGetLastError());
}

// ---------------------------------------------------------------------------
// K2s: sparse -> sparse (hash_transform_local_sparse.hpp:88-223).  Every stored entry becomes
// a (target column, target row) key and a scaled value; a stable radix sort groups equal
// keys while keeping the input order inside a group -- the order in which the reference
// accumulates -- and one thread per group adds its run left to right, so the sums round
// exactly as the reference's do.  Output columns come out row-sorted (the reference leaves
// them in first-touch order; same matrix).  Cancelled sums stay stored, as in the reference.
// ---------------------------------------------------------------------------
__global__ void k_sp_expand(const int32_t *__restrict__ indptr, const int32_t *__restrict__ indices,
                            const double *__restrict__ values, int ncols, int nnz, int rowwise,
                            const int32_t *__restrict__ bucket, const double *__restrict__ value,
                            uint64_t *__restrict__ keys, double *__restrict__ vals) {
    for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < nnz; e += gridDim.x * blockDim.x) {
        int lo = 0, hi = ncols;                  // column of entry e: last c with indptr[c] <= e
        while (hi - lo > 1) {
            int mid = (lo + hi) >> 1;
            if (indptr[mid] <= e) lo = mid; else hi = mid;
        }
        const int col = lo, row = indices[e];
        const int k = rowwise ? col : row;       // sketched index
        const uint32_t ocol = rowwise ? (uint32_t)bucket[k] : (uint32_t)col;
        const uint32_t orow = rowwise ? (uint32_t)row : (uint32_t)bucket[k];
        keys[e] = ((uint64_t)ocol << 32) | orow;
        vals[e] = __dmul_rn(values[e], value[k]);
    }
}

__global__ void k_sp_heads(const uint64_t *__restrict__ keys, int nnz, int32_t *__restrict__ head) {
    for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < nnz; e += gridDim.x * blockDim.x)
        head[e] = (e == 0 || keys[e] != keys[e - 1]) ? 1 : 0;
}

__global__ void k_sp_reduce(const uint64_t *__restrict__ keys, const double *__restrict__ vals,
                            const int32_t *__restrict__ pos, int nnz,
                            uint64_t *__restrict__ ukeys, int32_t *__restrict__ out_indices,
                            double *__restrict__ out_values) {
    for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < nnz; e += gridDim.x * blockDim.x) {
        const int r = pos[e] - 1;
        if (e != 0 && pos[e - 1] == pos[e]) continue;      // not the head of its run
        const uint64_t k = keys[e];
        double acc = vals[e];
        for (int q = e + 1; q < nnz && keys[q] == k; q++) acc = __dadd_rn(acc, vals[q]);
        ukeys[r] = k;
        out_indices[r] = (int32_t)(uint32_t)k;
        out_values[r] = acc;
    }
}

__global__ void k_sp_indptr(const uint64_t *__restrict__ ukeys, int nout, int ncols_out,
                            int32_t *__restrict__ out_indptr) {
    for (int c = blockIdx.x * blockDim.x + threadIdx.x; c <= ncols_out; c += gridDim.x * blockDim.x) {
        const uint64_t target = (uint64_t)(uint32_t)c << 32;
        int lo = 0, hi = nout;                    // first entry with key >= target
        while (lo < hi) {
            int mid = (lo + hi) >> 1;
            if (ukeys[mid] < target) lo = mid + 1; else hi = mid;
        }
        out_indptr[c] = lo;
    }
}

// Result arrays of the sparse path come from the device's stream-ordered pool (kept cached:
// no release threshold), so a steady stream of applies does not pay cudaMalloc/cudaFree.
template <class T>
static void sp_alloc(T **p, size_t bytes, cudaStream_t st) {
    static bool tuned[64] = {false};
    int dev = 0;
    SKB_CUDA(cudaGetDevice(&dev));
    if (dev >= 0 && dev < 64 && !tuned[dev]) {
        cudaMemPool_t pool;
        SKB_CUDA(cudaDeviceGetDefaultMemPool(&pool, dev));
        unsigned long long keep = ~0ULL;
        SKB_CUDA(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep));
        tuned[dev] = true;
    }
    SKB_CUDA(cudaMallocAsync((void **)p, bytes ? bytes : 16, st));
}
// cudaFree accepts pool allocations (no implicit synchronisation: callers are done with p)
void hash_sparse_free(void *p, cudaStream_t st) { if (p) { cudaStreamSynchronize(st); cudaFree(p); } }

// K2s fast path (columnwise, every column <= 10240 entries): one CTA sorts a column at a
// time in shared memory.  Key = bucket << PB | position in the column; a keys-only stable block
// radix sort over the bucket bits groups the buckets and the values are fetched through the position.  Equal
// buckets are then added left to right (input order, as the reference does) by the thread that
// holds the head of the run.  Results go to scratch at the column's input offset; a second
// kernel packs the columns once the output offsets are known.
#ifndef SKB_SPC_RADIX
#define SKB_SPC_RADIX 4
#endif
template <int THREADS, int ITEMS>
struct SpcTypes {
    static constexpr int CAP = THREADS * ITEMS;
    using Sort = cub::BlockRadixSort<uint32_t, THREADS, ITEMS, cub::NullType, SKB_SPC_RADIX>;
    using Scan = cub::BlockScan<int, THREADS>;
    union Tmp { typename Sort::TempStorage sort; typename Scan::TempStorage scan; };
    static constexpr size_t tmp_bytes = (sizeof(Tmp) + 15) / 16 * 16;
    static constexpr size_t smem_bytes = tmp_bytes + (size_t)CAP * 8 + ((size_t)CAP + 4) * 4;
};

// SEG: the entries of column `col` are not one contiguous run but the concatenation of `nseg`
// runs [vptr[s * ncols + col], vptr[s * ncols + col + 1]) -- the layout in which the partial
// sketches of nseg ranks arrive from an all-to-all (rank-major) -- taken in segment order; rows are
// their own buckets and values are taken as they are (a merge, not a sketch).
template <int THREADS, int ITEMS, bool SEG>
__global__ void __launch_bounds__(THREADS)
k_sp_col_sort(const int32_t *__restrict__ indptr, const int32_t *__restrict__ indices,
              const double *__restrict__ values, int ncols, const int32_t *__restrict__ bucket,
              const double *__restrict__ value, int pos_bits, int key_bits,
              int32_t *__restrict__ tmp_idx, double *__restrict__ tmp_val, int32_t *__restrict__ cnt,
              const int32_t *__restrict__ vptr, int nseg) {
    __shared__ int seg_pos[65], seg_src[64];
    using T = SpcTypes<THREADS, ITEMS>;
    constexpr int CAP = T::CAP;
    extern __shared__ __align__(16) unsigned char spc_smem[];
    typename T::Tmp &tmp = *reinterpret_cast<typename T::Tmp *>(spc_smem);
    double *svals = reinterpret_cast<double *>(spc_smem + T::tmp_bytes);
    uint32_t *skeys = reinterpret_cast<uint32_t *>(svals + CAP);
    const uint32_t pmask = (1u << pos_bits) - 1u;
    for (int col = blockIdx.x; col < ncols; col += gridDim.x) {
        const int c0 = indptr[col], cn = indptr[col + 1] - c0;
        if (SEG) {
            if (threadIdx.x == 0) {
                int p = 0;
                for (int sg = 0; sg < nseg; sg++) {
                    const int b = vptr[(size_t)sg * ncols + col];
                    seg_pos[sg] = p; seg_src[sg] = b;
                    p += vptr[(size_t)sg * ncols + col + 1] - b;
                }
                seg_pos[nseg] = p;
            }
            __syncthreads();
        }
        uint32_t keys[ITEMS];
        int sg = 0;
#pragma unroll
        for (int t = 0; t < ITEMS; t++) {                // blocked: thread holds ITEMS consecutive entries
            const int i = threadIdx.x * ITEMS + t;
            keys[t] = 0xffffffffu;
            if (i < cn) {
                if (SEG) {
                    while (i >= seg_pos[sg + 1]) sg++;
                    const int src = seg_src[sg] + (i - seg_pos[sg]);
                    keys[t] = ((uint32_t)indices[src] << pos_bits) | (uint32_t)i;
                    svals[i] = values[src];
                } else {
                    const int row = indices[c0 + i];
                    keys[t] = ((uint32_t)bucket[row] << pos_bits) | (uint32_t)i;
                    svals[i] = __dmul_rn(values[c0 + i], value[row]);
                }
            }
        }
        // the radix sort is stable and the entries start in column order, so only the bucket
        // bits need sorting (padding keys are all ones and stay behind)
        typename T::Sort(tmp.sort).Sort(keys, pos_bits, key_bits);   // blocked output
#pragma unroll
        for (int t = 0; t < ITEMS; t++) skeys[threadIdx.x * ITEMS + t] = keys[t];
        if (threadIdx.x == 0) skeys[CAP] = 0xffffffffu;
        __syncthreads();
        int nhead = 0;
        bool head[ITEMS];
#pragma unroll
        for (int t = 0; t < ITEMS; t++) {
            const int i = threadIdx.x * ITEMS + t;
            head[t] = i < cn && (i == 0 || (skeys[i - 1] >> pos_bits) != (keys[t] >> pos_bits));
            nhead += head[t];
        }
        int off, total;
        typename T::Scan(tmp.scan).ExclusiveSum(nhead, off, total);
#pragma unroll
        for (int t = 0; t < ITEMS; t++) {
            if (head[t]) {
                const uint32_t b = keys[t] >> pos_bits;
                double acc = svals[keys[t] & pmask];
                for (int q = threadIdx.x * ITEMS + t + 1; (skeys[q] >> pos_bits) == b && q < cn; q++)
                    acc = __dadd_rn(acc, svals[skeys[q] & pmask]);
                tmp_idx[c0 + off] = (int32_t)b;
                tmp_val[c0 + off] = acc;
                off++;
            }
        }
        if (threadIdx.x == 0) cnt[col] = total;
        __syncthreads();                               // shared arrays are reused by the next column
    }
}

template <int THREADS, int ITEMS>
static void launch_sp_col_sort(const int32_t *indptr, const int32_t *indices, const double *values,
                               int ncols, const int32_t *bucket, const double *value, int pos_bits,
                               int key_bits, int32_t *tmp_idx, double *tmp_val, int32_t *cnt, cudaStream_t st,
                               const int32_t *vptr = nullptr, int nseg = 0) {
    using T = SpcTypes<THREADS, ITEMS>;
    auto kern = vptr ? k_sp_col_sort<THREADS, ITEMS, true> : k_sp_col_sort<THREADS, ITEMS, false>;
    SKB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)T::smem_bytes));
    size_t per_sm = (size_t)(220 * 1024) / T::smem_bytes;
    if (per_sm * THREADS > 2048) per_sm = 2048 / THREADS;
    if (per_sm < 1) per_sm = 1;
    const unsigned grid = (unsigned)std::min<size_t>((size_t)ncols, 148 * per_sm);
    kern<<<grid, THREADS, T::smem_bytes, st>>>(indptr, indices, values, ncols, bucket, value, pos_bits,
                                               key_bits, tmp_idx, tmp_val, cnt, vptr, nseg);
    SKB_CUDA(cudaGetLastError());
}

__global__ void k_sp_col_pack(const int32_t *__restrict__ indptr, const int32_t *__restrict__ out_indptr,
                              int ncols, const int32_t *__restrict__ tmp_idx,
                              const double *__restrict__ tmp_val, int32_t *__restrict__ out_indices,
                              double *__restrict__ out_values) {
    const int lane = threadIdx.x & 31;
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarps = (gridDim.x * blockDim.x) >> 5;
    for (int col = warp; col < ncols; col += nwarps) {
        const int src = indptr[col], dst = out_indptr[col], n = out_indptr[col + 1] - dst;
        for (int i = lane; i < n; i += 32) {
            out_indices[dst + i] = tmp_idx[src + i];
            out_values[dst + i] = tmp_val[src + i];
        }
    }
}

__global__ void k_sp_max_col(const int32_t *__restrict__ indptr, int ncols, int32_t *out) {
    int m = 0;
    for (int c = blockIdx.x * blockDim.x + threadIdx.x; c < ncols; c += gridDim.x * blockDim.x)
        m = max(m, indptr[c + 1] - indptr[c]);
    m = __reduce_max_sync(0xffffffffu, m);
    if ((threadIdx.x & 31) == 0 && m > 0) atomicMax(out, m);
}

// Returns false (nothing done) when the shape does not qualify.
static bool sparse_columnwise_fast(const int32_t *indptr, const int32_t *indices, const double *values,
                                   size_t ncols, size_t nnz, const int32_t *bucket, const double *value,
                                   int S, int32_t **out_indptr, int32_t **out_indices,
                                   double **out_values, size_t *out_nnz, void **ws, size_t *ws_bytes,
                                   cudaStream_t st, const int32_t *vptr = nullptr, int nseg = 0) {
    static const bool off = getenv("SKB_DEBUG_SP_GENERIC") != nullptr;
    if (nnz == 0 || ncols == 0) return false;
    if (!vptr && (off || nnz / ncols < 256)) return false;   // short columns: the global sort wins
    auto al = [](size_t b) { return (b + 255) / 256 * 256; };
    size_t scan_bytes = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, (int32_t *)nullptr, (int32_t *)nullptr, (int)ncols + 1, st);
    const size_t need = al(nnz * 4) + al(nnz * 8) + al((ncols + 1) * 4) + al(scan_bytes) + 512;
    if (*ws_bytes < need) {
        SKB_CUDA(cudaStreamSynchronize(st));
        cudaFree(*ws); *ws = nullptr; *ws_bytes = 0;
        SKB_CUDA(cudaMalloc(ws, need));
        *ws_bytes = need;
    }
    char *w = (char *)*ws;
    int32_t *tmp_idx = (int32_t *)w; w += al(nnz * 4);
    double *tmp_val = (double *)w;   w += al(nnz * 8);
    int32_t *cnt = (int32_t *)w;     w += al((ncols + 1) * 4);
    void *scan_tmp = w;              w += al(scan_bytes);
    int32_t *dmax = (int32_t *)w;
    SKB_CUDA(cudaMemsetAsync(dmax, 0, sizeof(int32_t), st));
    {
        LaunchScope ls(KC_HASH, st);
        k_sp_max_col<<<(unsigned)std::min<size_t>((ncols + 255) / 256, 1184), 256, 0, st>>>(indptr, (int)ncols, dmax);
    }
    int32_t maxc = 0;
    SKB_CUDA(cudaMemcpyAsync(&maxc, dmax, sizeof(int32_t), cudaMemcpyDeviceToHost, st));
    SKB_CUDA(cudaStreamSynchronize(st));
    // capacity classes: (threads, entries per thread)
    const int cap = maxc <= 512 ? 512 : maxc <= 1024 ? 1024 : maxc <= 2048 ? 2048 : maxc <= 5120 ? 5120 :
                    maxc <= 10240 ? 10240 : 0;
    int pos_bits = 1, bucket_bits = 1;
    while ((1 << pos_bits) < (cap ? cap : 1)) pos_bits++;
    while ((1LL << bucket_bits) < (long long)S) bucket_bits++;
    if (!cap || pos_bits + bucket_bits > 32) return false;
    sp_alloc(out_indptr, (ncols + 1) * sizeof(int32_t), st);
    SKB_CUDA(cudaMemsetAsync(cnt + ncols, 0, sizeof(int32_t), st));
    {
        LaunchScope ls(KC_HASH, st);
        const int kb = pos_bits + bucket_bits, nc = (int)ncols;
        if (cap == 512) launch_sp_col_sort<256, 2>(indptr, indices, values, nc, bucket, value, pos_bits, kb, tmp_idx, tmp_val, cnt, st, vptr, nseg);
        else if (cap == 1024) launch_sp_col_sort<256, 4>(indptr, indices, values, nc, bucket, value, pos_bits, kb, tmp_idx, tmp_val, cnt, st, vptr, nseg);
        else if (cap == 2048) launch_sp_col_sort<256, 8>(indptr, indices, values, nc, bucket, value, pos_bits, kb, tmp_idx, tmp_val, cnt, st, vptr, nseg);
        else if (cap == 5120) launch_sp_col_sort<512, 10>(indptr, indices, values, nc, bucket, value, pos_bits, kb, tmp_idx, tmp_val, cnt, st, vptr, nseg);
        else launch_sp_col_sort<512, 20>(indptr, indices, values, nc, bucket, value, pos_bits, kb, tmp_idx, tmp_val, cnt, st, vptr, nseg);
    }
    {
        LaunchScope ls(KC_HASH, st);
        SKB_CUDA(cub::DeviceScan::ExclusiveSum(scan_tmp, scan_bytes, cnt, *out_indptr, (int)ncols + 1, st));
    }
    int32_t nout = 0;
    SKB_CUDA(cudaMemcpyAsync(&nout, *out_indptr + ncols, sizeof(int32_t), cudaMemcpyDeviceToHost, st));
    SKB_CUDA(cudaStreamSynchronize(st));
    sp_alloc(out_indices, (size_t)nout * 4, st);
    sp_alloc(out_values, (size_t)nout * 8, st);
    {
        LaunchScope ls(KC_HASH, st);
        k_sp_col_pack<<<148 * 8, 256, 0, st>>>(indptr, *out_indptr, (int)ncols, tmp_idx, tmp_val,
                                               *out_indices, *out_values);
        SKB_CUDA(cudaGetLastError());
    }
    *out_nnz = (size_t)nout;
    return true;
}

__global__ void k_seg_col_counts(const int32_t *__restrict__ vptr, int nseg, int ncols, int32_t *__restrict__ cnt) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j > ncols) return;
    int c = 0;
    if (j < ncols)
        for (int sg = 0; sg < nseg; sg++) c += vptr[(size_t)sg * ncols + j + 1] - vptr[(size_t)sg * ncols + j];
    cnt[j] = c;
}

// Merge of nseg rank-major CSC pieces (see k_sp_col_sort<.., SEG>): entries of equal (row, column)
// are added in segment order, rows of every column come out sorted.  colbase: caller-provided scratch
// of ncols + 1 ints.  Returns false when a merged column exceeds the shared-memory sort's capacity.
bool hash_merge_segments(const int32_t *vptr, const int32_t *indices, const double *values, int nseg,
                         size_t ncols, size_t nrows, size_t nnz, int32_t **out_indptr, int32_t **out_indices,
                         double **out_values, size_t *out_nnz, void **ws, size_t *ws_bytes, cudaStream_t st) {
    *out_indptr = nullptr; *out_indices = nullptr; *out_values = nullptr; *out_nnz = 0;
    if (nseg < 1 || nseg > 64 || nnz > 0x7fffffffULL || ncols > 0x7ffffffeULL)
        throw error(SKB_ERR_INVALID_PARAM, "sparse merge: bad sizes");
    static std::map<int, std::pair<int32_t *, size_t>> base_per_dev;
    int dev = 0;
    SKB_CUDA(cudaGetDevice(&dev));
    auto &cb = base_per_dev[dev];
    size_t scan_bytes = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, (int32_t *)nullptr, (int32_t *)nullptr, (int)ncols + 1, st);
    const size_t need = (ncols + 1) * 2 * sizeof(int32_t) + scan_bytes + 512;
    if (cb.second < need) {
        SKB_CUDA(cudaStreamSynchronize(st));
        cudaFree(cb.first); cb.first = nullptr; cb.second = 0;
        SKB_CUDA(cudaMalloc(&cb.first, need));
        cb.second = need;
    }
    int32_t *cnt = cb.first, *colbase = cb.first + (ncols + 1);
    void *scan_tmp = colbase + (ncols + 1);
    {
        LaunchScope ls(KC_HASH, st);
        k_seg_col_counts<<<(unsigned)((ncols + 1 + 255) / 256), 256, 0, st>>>(vptr, nseg, (int)ncols, cnt);
        SKB_CUDA(cub::DeviceScan::ExclusiveSum(scan_tmp, scan_bytes, cnt, colbase, (int)ncols + 1, st));
    }
    return sparse_columnwise_fast(colbase, indices, values, ncols, nnz, nullptr, nullptr, (int)nrows, out_indptr,
                                  out_indices, out_values, out_nnz, ws, ws_bytes, st, vptr, nseg);
}

void hash_apply_csc_sparse(const int32_t *indptr, const int32_t *indices, const double *values,
                           size_t nrows, size_t ncols, size_t nnz, bool rowwise,
                           const int32_t *bucket, const double *value, int S,
                           int32_t **out_indptr, int32_t **out_indices, double **out_values,
                           size_t *out_nnz, void **ws, size_t *ws_bytes, cudaStream_t st) {
    const size_t ncols_out = rowwise ? (size_t)S : ncols;
    const size_t nrows_out = rowwise ? nrows : (size_t)S;
    *out_indptr = nullptr; *out_indices = nullptr; *out_values = nullptr; *out_nnz = 0;
    if (nnz > 0x7fffffffULL || ncols > 0x7ffffffeULL || nrows > 0x7fffffffULL)
        throw error(SKB_ERR_INVALID_PARAM, "sparse apply: sizes exceed int32");
    if (!rowwise) {
        try {
            if (sparse_columnwise_fast(indptr, indices, values, ncols, nnz, bucket, value, S, out_indptr,
                                       out_indices, out_values, out_nnz, ws, ws_bytes, st))
                return;
        } catch (...) {
            hash_sparse_free(*out_indptr, st); hash_sparse_free(*out_indices, st); hash_sparse_free(*out_values, st);
            *out_indptr = nullptr; *out_indices = nullptr; *out_values = nullptr;
            throw;
        }
    }
    // temporaries live in the caller's grow-only workspace (no allocation in steady state)
    auto al = [](size_t b) { return (b + 255) / 256 * 256; };
    size_t sort_bytes = 0, scan_bytes = 0;
    if (nnz) {
        cub::DoubleBuffer<uint64_t> kq(nullptr, nullptr);
        cub::DoubleBuffer<double> vq(nullptr, nullptr);
        cub::DeviceRadixSort::SortPairs(nullptr, sort_bytes, kq, vq, (int)nnz, 0, 64, st);
        cub::DeviceScan::InclusiveSum(nullptr, scan_bytes, (int32_t *)nullptr, (int32_t *)nullptr,
                                      (int)nnz, st);
    }
    const size_t tmp_bytes = al(std::max(sort_bytes, scan_bytes));
    const size_t need = 4 * al(nnz * 8) + 2 * al(nnz * 4) + tmp_bytes + 256;
    if (*ws_bytes < need) {
        SKB_CUDA(cudaStreamSynchronize(st));
        cudaFree(*ws); *ws = nullptr; *ws_bytes = 0;
        SKB_CUDA(cudaMalloc(ws, need));
        *ws_bytes = need;
    }
    char *w = (char *)*ws;
    uint64_t *keys = (uint64_t *)w;  w += al(nnz * 8);
    uint64_t *keys2 = (uint64_t *)w; w += al(nnz * 8);
    double *vals = (double *)w;      w += al(nnz * 8);
    double *vals2 = (double *)w;     w += al(nnz * 8);
    int32_t *head = (int32_t *)w;    w += al(nnz * 4);
    int32_t *pos = (int32_t *)w;     w += al(nnz * 4);
    void *tmp = w;
    try {
        sp_alloc(out_indptr, (ncols_out + 1) * sizeof(int32_t), st);
        if (nnz == 0) {
            SKB_CUDA(cudaMemsetAsync(*out_indptr, 0, (ncols_out + 1) * sizeof(int32_t), st));
            sp_alloc(out_indices, 16, st);
            sp_alloc(out_values, 16, st);
            return;
        }
        const int n = (int)nnz;
        const unsigned nb = (unsigned)std::min<size_t>((nnz + 255) / 256, 148 * 32);
        {
            LaunchScope ls(KC_HASH, st);
            k_sp_expand<<<nb, 256, 0, st>>>(indptr, indices, values, (int)ncols, n, rowwise ? 1 : 0,
                                            bucket, value, keys, vals);
        }
        int row_bits = 1, col_bits = 1;
        while ((1ULL << row_bits) < nrows_out && row_bits < 32) row_bits++;
        while ((1ULL << col_bits) < ncols_out && col_bits < 32) col_bits++;
        // rows occupy the low word: sort its used bits, then the used bits of the high word
        cub::DoubleBuffer<uint64_t> kb(keys, keys2);
        cub::DoubleBuffer<double> vb(vals, vals2);
        {
            LaunchScope ls(KC_HASH, st);
            SKB_CUDA(cub::DeviceRadixSort::SortPairs(tmp, sort_bytes, kb, vb, n, 0, row_bits, st));
        }
        {
            LaunchScope ls(KC_HASH, st);
            SKB_CUDA(cub::DeviceRadixSort::SortPairs(tmp, sort_bytes, kb, vb, n, 32, 32 + col_bits, st));
        }
        {
            LaunchScope ls(KC_HASH, st);
            k_sp_heads<<<nb, 256, 0, st>>>(kb.Current(), n, head);
        }
        {
            LaunchScope ls(KC_HASH, st);
            SKB_CUDA(cub::DeviceScan::InclusiveSum(tmp, scan_bytes, head, pos, n, st));
        }
        int32_t nout = 0;
        SKB_CUDA(cudaMemcpyAsync(&nout, pos + (n - 1), sizeof(int32_t), cudaMemcpyDeviceToHost, st));
        SKB_CUDA(cudaStreamSynchronize(st));
        sp_alloc(out_indices, (size_t)nout * 4, st);
        sp_alloc(out_values, (size_t)nout * 8, st);
        uint64_t *ukeys = kb.Alternate();       // the other sort buffer is free by now
        {
            LaunchScope ls(KC_HASH, st);
            k_sp_reduce<<<nb, 256, 0, st>>>(kb.Current(), vb.Current(), pos, n, ukeys, *out_indices,
                                            *out_values);
        }
        {
            LaunchScope ls(KC_HASH, st);
            const unsigned cb = (unsigned)std::min<size_t>((ncols_out + 1 + 255) / 256, 148 * 32);
            k_sp_indptr<<<cb, 256, 0, st>>>(ukeys, nout, (int)ncols_out, *out_indptr);
        }
        SKB_CUDA(cudaGetLastError());
        *out_nnz = (size_t)nout;
    } catch (...) {
        hash_sparse_free(*out_indptr, st); hash_sparse_free(*out_indices, st); hash_sparse_free(*out_values, st);
        *out_indptr = nullptr; *out_indices = nullptr; *out_values = nullptr;
        throw;
    }
}

}  // namespace skb
