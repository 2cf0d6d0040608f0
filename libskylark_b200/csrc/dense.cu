// Dense sketches (JLT / CT) and random Fourier features (Gaussian / Laplacian RFT).
//
// Reference loops replaced:
//   realize_matrix_view                 sketch/dense_transform_data.hpp:79-101
//   local apply (Gemm)                  sketch/dense_transform_Elemental_local.hpp:79-111
//   row-sharded inner-panel apply       sketch/dense_transform_Elemental_coldist_star.hpp:200-229
//   RFT epilogue                        sketch/RFT_Elemental.hpp:94-117, 135-158
//
// B200 design.  S is never stored: entry S[i,k] = scale * sample(base + k*S + i) is
// regenerated from the Threefry2x64-13 counter stream.  A panel of KC input indices
// (KC*S variates, a contiguous range of the stream) is generated ONCE into an
// L2-resident scratch buffer by k_gen_panel and consumed by a stream-K FP64 tensor
// core GEMM (mma.sync m8n8k4 f64 = DMMA; sm_100 has no tcgen05 path for FP64).
// Generating inside the GEMM tile loop would repeat every Box-Muller d/128 times
// (8x at d = 1024); staging through L2 costs 16 bytes of L2 traffic per variate
// instead and keeps the FP64 pipe for the contraction.
//   * CTA tile 128(i) x 128(j) x 16(k), 8 warps (2 x 4), warp tile 64 x 32,
//     accumulators in registers (64 doubles / thread), 4-stage cp.async pipeline;
//   * shared tiles padded so that the 64-bit fragment loads are conflict free
//     (row stride = 4 mod 16 doubles);
//   * stream-K: the (tile, k-step) space of one launch is cut into gridDim equal
//     contiguous ranges, one per SM, so 256 tiles on 148 SMs lose nothing to wave
//     quantisation; partial tiles are flushed with red.global.add.f64;
//   * the RFT epilogue outscale*cos(scale_i*acc+shift_i) is fused into the flush
//     when a CTA owns the whole contraction of a tile (the rowwise feature-map case).
#include <algorithm>
#include <cstdio>
#include <cstdlib>

#include <cuda.h>          // CUtensorMap and its enums (the encoder itself is fetched at run time)

#include "rng.cuh"
#include "skb_internal.h"

namespace skb {

// ---------------------------------------------------------------------------
// Panel generation (K5a).  P[e] = scale * sample(base + first + e), e < count.
// ---------------------------------------------------------------------------
// 128-thread blocks with <= 80 registers: one such block fits beside the resident GEMM CTA
// of an SM (which holds 256 x 212 registers), so generation of panel p+1 overlaps GEMM p.
template <int DIST>
__global__ void __launch_bounds__(128)
k_gen_panel(uint64_t key, uint64_t stream0, size_t count, double scale, double *__restrict__ P) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    for (; e + stride < count; e += 2 * stride) {       // two independent variates in flight
        double v0 = DIST == DD_NORMAL ? normal_at(key, stream0 + e) : cauchy_at(key, stream0 + e);
        double v1 = DIST == DD_NORMAL ? normal_at(key, stream0 + e + stride)
                                      : cauchy_at(key, stream0 + e + stride);
        P[e] = __dmul_rn(scale, v0);
        P[e + stride] = __dmul_rn(scale, v1);
    }
    if (e < count) {
        double v = DIST == DD_NORMAL ? normal_at(key, stream0 + e) : cauchy_at(key, stream0 + e);
        P[e] = __dmul_rn(scale, v);
    }
}

void dense_generate_panel(const DenseParams &p, size_t k0, size_t kc, double *panel,
                          cudaStream_t st) {
    size_t count = kc * p.S;
    if (count == 0) return;
    static int skip_gen = -1;      // timing experiments only (results are garbage when set)
    if (skip_gen < 0) skip_gen = getenv("SKB_DEBUG_SKIP_GEN") ? 1 : 0;
    if (skip_gen) return;
    uint64_t stream0 = p.base + (uint64_t)k0 * p.S;
    static int gen_threads = 0;
    if (!gen_threads) { const char *e = getenv("SKB_GEN_THREADS"); gen_threads = e ? atoi(e) : 128; }
    size_t blocks = (count + 2 * gen_threads - 1) / (2 * gen_threads);
    size_t cap = (size_t)148 * 64 * (128 / gen_threads > 0 ? 128 / gen_threads : 1);
    if (blocks > cap) blocks = cap;
    LaunchScope ls(KC_GEN, st);
    if (p.dist == DD_NORMAL)
        k_gen_panel<DD_NORMAL><<<(unsigned)blocks, gen_threads, 0, st>>>(seed_to_key(p.seed), stream0,
                                                                 count, p.scale, panel);
    else
        k_gen_panel<DD_CAUCHY><<<(unsigned)blocks, gen_threads, 0, st>>>(seed_to_key(p.seed), stream0,
                                                                 count, p.scale, panel);
    SKB_CUDA(cudaGetLastError());
}

// K0 for tests: S[i0:i0+h, j0:j0+w] column-major
template <int DIST>
__global__ void k_export_block(uint64_t key, uint64_t base, size_t S, double scale, size_t i0,
                               size_t j0, size_t h, size_t w, double *__restrict__ out) {
    size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= h * w) return;
    size_t i = t % h, j = t / h;
    uint64_t idx = base + (uint64_t)(j0 + j) * S + (i0 + i);
    double v = DIST == DD_NORMAL ? normal_at(key, idx) : cauchy_at(key, idx);
    out[t] = __dmul_rn(scale, v);
}

void dense_export_block(const DenseParams &p, size_t i0, size_t j0, size_t h, size_t w,
                        double *out, cudaStream_t st) {
    if (h * w == 0) return;
    unsigned blocks = (unsigned)((h * w + 255) / 256);
    LaunchScope ls(KC_GEN, st);
    if (p.dist == DD_NORMAL)
        k_export_block<DD_NORMAL><<<blocks, 256, 0, st>>>(seed_to_key(p.seed), p.base, p.S,
                                                          p.scale, i0, j0, h, w, out);
    else
        k_export_block<DD_CAUCHY><<<blocks, 256, 0, st>>>(seed_to_key(p.seed), p.base, p.S,
                                                          p.scale, i0, j0, h, w, out);
    SKB_CUDA(cudaGetLastError());
}

// ---------------------------------------------------------------------------
// FP64 tensor-core GEMM (K5b): OUT[i,j] (+)= sum_kk P[kk*S + i] * X(kk, j)
// ---------------------------------------------------------------------------
#ifndef SKB_BK
#define SKB_BK 16
#endif
#ifndef SKB_STAGES
#define SKB_STAGES 4
#endif
static constexpr int BM = 128, BK = SKB_BK, STAGES = SKB_STAGES;
// Tile configuration per operand type (measured on B200):
//   fp64: 128 x 64 tile, 2 x 2 warps (64 x 32 warp tiles), TWO CTAs per SM.  Each CTA syncs all
//         its warps at every k-step; two independent CTAs per SM run out of phase, so one is
//         issuing DMMAs while the other sits at its barrier / fragment loads (C2: 292 vs 303 ms
//         for one 128 x 128 CTA of 8 warps).
//   fp32: 128 x 128 tile, 4 x 4 warps (32 x 32 warp tiles), one CTA per SM (C5 shard: 470 vs
//         611 ms for 2 x 4 warps): conversions and the fused cosine need more warps to cover.
template <typename TX> struct GemmCfg;
template <> struct GemmCfg<double> { static constexpr int BN = 64, WM = 2, WN = 2, CTAS = 2; };
template <> struct GemmCfg<float> { static constexpr int BN = 128, WM = 4, WN = 4, CTAS = 1; };
// Warp layout (WM x WN warps over the 128 x 128 tile) is a template parameter: measured on
// B200, 2 x 4 warps (64 x 32 warp tiles, 64 accumulators/thread) is fastest for fp64 operands
// (C2: 307 vs 334 ms), 4 x 4 warps (32 x 32 warp tiles) for fp32 operands with the fused
// cosine epilogue (C5 shard: 470 vs 611 ms) where conversions and the epilogue need covering.
static constexpr int LDP = BM + 4;   // doubles; = 4 mod 16 -> conflict-free 64-bit fragment loads

template <typename TX> struct XTile;
template <> struct XTile<double> {
    static constexpr int LD_ROWS = GemmCfg<double>::BN + 4;   // Xs[kk][j]
    static constexpr int LD_COLS = BK + 4;                    // Xs[j][kk]
};
template <> struct XTile<float> {
    static constexpr int LD_ROWS = GemmCfg<float>::BN + 8;    // = 8 mod 32 floats
    static constexpr int LD_COLS = BK + 4;                    // = 20 mod 32 floats
};

template <typename TX, bool ROWS> struct GemmSmem {
    static constexpr int P_ELEMS = BK * LDP;
    static constexpr int X_ELEMS = ROWS ? BK * XTile<TX>::LD_ROWS : GemmCfg<TX>::BN * XTile<TX>::LD_COLS;
    static constexpr size_t STAGE_BYTES = P_ELEMS * sizeof(double) +
                                          ((X_ELEMS * sizeof(TX) + 15) / 16) * 16;
    static constexpr size_t BYTES = STAGE_BYTES * STAGES;
};

template <int BYTES>
__device__ __forceinline__ void cp_async(void *smem, const void *gmem, int src_bytes) {
    unsigned s = (unsigned)__cvta_generic_to_shared(smem);
    if (BYTES == 16)
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(s), "l"(gmem), "r"(src_bytes));
    else if (BYTES == 8)
        asm volatile("cp.async.ca.shared.global [%0], [%1], 8, %2;" ::"r"(s), "l"(gmem), "r"(src_bytes));
    else
        asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(s), "l"(gmem), "r"(src_bytes));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;"); }
template <int N> __device__ __forceinline__ void cp_async_wait() {
    asm volatile("cp.async.wait_group %0;" ::"n"(N));
}

__device__ __forceinline__ void dmma884(double &c0, double &c1, double a, double b) {
    asm("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

struct GemmArgs {
    const double *P; size_t S; size_t kc;      // panel: kc rows of S
    const void *X; size_t ldx; size_t m;       // local input, kc sketched indices
    void *OUT; size_t ldo;                     // TX (direct) or double (atomic) output
    int ntm, ntn, nk;
    int direct_ok;        // full-K segments may store (and apply the epilogue) directly
    int whole_tiles;      // partition by whole tiles (no k split) instead of stream-K
    int out_is_acc;       // OUT points at a double accumulation buffer (atomic path for float)
    const double *scales, *shifts; double outscale; int has_epi;
};

// Copy a TR x TC tile from global (row stride gld elements) into shared (row stride sld)
// with CPB-byte cp.async, zero-filling beyond (row_limit, col_limit).  Every thread copies
// the same column offset of rows r0, r0 + RSTEP, ...: one division per call, then only
// pointer increments -- the copy issue sits on the critical path of every k-step.
template <typename T, int CPB, int NT, int TR, int TC>
__device__ __forceinline__ void load_tile(T *smem, int sld, const T *g, size_t gld,
                                          long long row_limit, long long col_limit) {
    constexpr int EPC = CPB / (int)sizeof(T);          // elements per copy
    constexpr int CPR = TC / EPC;                      // copies per row
    static_assert(NT % CPR == 0 || CPR % NT == 0, "thread count must tile the copies of a row");
    if (CPR <= NT) {
        constexpr int RSTEP = NT / (CPR <= NT ? CPR : NT), Q = (TR + RSTEP - 1) / RSTEP;
        const int r0 = threadIdx.x / CPR, col = (threadIdx.x % CPR) * EPC;
        const long long rem = col_limit - col;
        const int colbytes = rem <= 0 ? 0 : (rem >= EPC ? CPB : (int)rem * (int)sizeof(T));
        const T *src = g + (size_t)r0 * gld + col;
        T *dst = smem + r0 * sld + col;
#pragma unroll
        for (int q = 0; q < Q; q++) {
            const int r = r0 + q * RSTEP;
            if (TR % RSTEP == 0 || r < TR) {
                const int nb = r < row_limit ? colbytes : 0;
                cp_async<CPB>(dst, nb > 0 ? src : g, nb);
            }
            src += (size_t)RSTEP * gld;
            dst += RSTEP * sld;
        }
    } else {                                           // more copies per row than threads
        constexpr int PER = CPR / (CPR <= NT ? 1 : NT);
#pragma unroll 1
        for (int r = 0; r < TR; r++) {
#pragma unroll
            for (int p = 0; p < PER; p++) {
                const int col = (threadIdx.x + p * NT) * EPC;
                const long long rem = col_limit - col;
                int nb = 0;
                if (r < row_limit && rem > 0) nb = rem >= EPC ? CPB : (int)rem * (int)sizeof(T);
                cp_async<CPB>(smem + r * sld + col, nb > 0 ? g + (size_t)r * gld + col : g, nb);
            }
        }
    }
}

template <typename TX, bool ROWS, int CPB_P, int CPB_X, int WM, int WN>
__global__ void __launch_bounds__(32 * WM * WN, GemmCfg<TX>::CTAS) k_dense_gemm(GemmArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    using SM = GemmSmem<TX, ROWS>;
    constexpr int BN = GemmCfg<TX>::BN;
    constexpr int NT = 32 * WM * WN, WTM = BM / WM, WTN = BN / WN, MI = WTM / 8, NI = WTN / 8;
    constexpr int LDX = ROWS ? XTile<TX>::LD_ROWS : XTile<TX>::LD_COLS;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int wm = warp % WM, wn = warp / WM;
    const int g = lane >> 2, t4 = lane & 3;

    const long long total = (long long)a.ntm * a.ntn * a.nk;
    long long u, u_end;
    if (a.whole_tiles) {
        const long long nt = (long long)a.ntm * a.ntn;
        u = nt * blockIdx.x / gridDim.x * a.nk;
        u_end = nt * (blockIdx.x + 1) / gridDim.x * a.nk;
    } else {
        u = total * blockIdx.x / gridDim.x;
        u_end = total * (blockIdx.x + 1) / gridDim.x;
    }
    const TX *Xg = reinterpret_cast<const TX *>(a.X);

    while (u < u_end) {
        const int tile = (int)(u / a.nk);
        const int ks0 = (int)(u - (long long)tile * a.nk);
        const int ks1 = (int)min((long long)a.nk, ks0 + (u_end - u));
        const int tm = tile / a.ntn, tn = tile - tm * a.ntn;
        const size_t i0 = (size_t)tm * BM, j0 = (size_t)tn * BN;
        const long long i_lim = (long long)a.S - (long long)i0;   // valid i in tile
        const long long j_lim = (long long)a.m - (long long)j0;

        double acc[MI][NI][2];
#pragma unroll
        for (int mi = 0; mi < MI; mi++)
#pragma unroll
            for (int ni = 0; ni < NI; ni++) { acc[mi][ni][0] = 0.0; acc[mi][ni][1] = 0.0; }

        auto issue = [&](int ks, int slot) {
            unsigned char *base = smem_raw + (size_t)slot * SM::STAGE_BYTES;
            double *Ps = reinterpret_cast<double *>(base);
            TX *Xs = reinterpret_cast<TX *>(base + SM::P_ELEMS * sizeof(double));
            const long long kk0 = (long long)ks * BK;
            const long long k_lim = (long long)a.kc - kk0;          // valid kk in this step
            load_tile<double, CPB_P, NT, BK, BM>(Ps, LDP, a.P + (size_t)kk0 * a.S + i0, a.S, k_lim, i_lim);
            if (ROWS)
                load_tile<TX, CPB_X, NT, BK, BN>(Xs, LDX, Xg + (size_t)kk0 * a.ldx + j0, a.ldx, k_lim, j_lim);
            else
                load_tile<TX, CPB_X, NT, BN, BK>(Xs, LDX, Xg + (size_t)j0 * a.ldx + kk0, a.ldx, j_lim, k_lim);
        };

        const int nsteps = ks1 - ks0;
#pragma unroll
        for (int s = 0; s < STAGES - 1; s++) {
            if (s < nsteps) issue(ks0 + s, s);
            cp_async_commit();
        }
        for (int it = 0; it < nsteps; it++) {
            cp_async_wait<STAGES - 2>();
            __syncthreads();
            {
                int nx = it + STAGES - 1;
                if (nx < nsteps) issue(ks0 + nx, nx % STAGES);
                cp_async_commit();
            }
            const unsigned char *base = smem_raw + (size_t)(it % STAGES) * SM::STAGE_BYTES;
            const double *Ps = reinterpret_cast<const double *>(base);
            const TX *Xs = reinterpret_cast<const TX *>(base + SM::P_ELEMS * sizeof(double));
#pragma unroll
            for (int k4 = 0; k4 < BK / 4; k4++) {
                double af[MI], bf[NI];
                const double *pa = Ps + (k4 * 4 + t4) * LDP + wm * WTM + g;
#pragma unroll
                for (int mi = 0; mi < MI; mi++) af[mi] = pa[mi * 8];
#pragma unroll
                for (int ni = 0; ni < NI; ni++) {
                    int j = wn * WTN + ni * 8 + g;
                    bf[ni] = ROWS ? (double)Xs[(k4 * 4 + t4) * LDX + j]
                                  : (double)Xs[j * LDX + k4 * 4 + t4];
                }
#pragma unroll
                for (int mi = 0; mi < MI; mi++)
#pragma unroll
                    for (int ni = 0; ni < NI; ni++)
                        dmma884(acc[mi][ni][0], acc[mi][ni][1], af[mi], bf[ni]);
            }
        }
        cp_async_wait<0>();
        __syncthreads();

        // ---- flush -----------------------------------------------------------
        const bool full = (ks0 == 0 && ks1 == a.nk);
        const bool direct = full && a.direct_ok;
#pragma unroll
        for (int mi = 0; mi < MI; mi++) {
            const long long il = wm * WTM + mi * 8 + g;
            if (il >= i_lim) continue;
            const size_t i = i0 + il;
            double sc = 1.0, sh = 0.0;
            if (direct && a.has_epi) { sc = a.scales[i]; sh = a.shifts[i]; }
#pragma unroll
            for (int ni = 0; ni < NI; ni++) {
#pragma unroll
                for (int e = 0; e < 2; e++) {
                    const long long jl = wn * WTN + ni * 8 + t4 * 2 + e;
                    if (jl >= j_lim) continue;
                    const size_t j = j0 + jl;
                    const size_t off = ROWS ? i * a.ldo + j : i + j * a.ldo;
                    double v = acc[mi][ni][e];
                    if (direct) {
                        if (a.has_epi) v = __dmul_rn(a.outscale, cos(__dadd_rn(__dmul_rn(v, sc), sh)));
                        if (a.out_is_acc) reinterpret_cast<double *>(a.OUT)[off] = v;
                        else reinterpret_cast<TX *>(a.OUT)[off] = (TX)v;
                    } else {
                        if (a.out_is_acc || sizeof(TX) == 8)
                            atomicAdd(reinterpret_cast<double *>(a.OUT) + off, v);
                        else
                            atomicAdd(reinterpret_cast<float *>(a.OUT) + off, (float)v);
                    }
                }
            }
        }
        u += nsteps;
    }
}

// acc (double, same indexing as OUT) -> OUT with optional RFT epilogue
template <typename TX, bool ROWS>
__global__ void k_dense_finalize(const double *__restrict__ acc, TX *__restrict__ OUT, size_t S,
                                 size_t m, size_t ldo, const double *scales,
                                 const double *shifts, double outscale, int has_epi) {
    size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= S * m) return;
    size_t i, j;
    if (ROWS) { i = t / m; j = t - i * m; } else { j = t / S; i = t - j * S; }
    size_t off = ROWS ? i * ldo + j : i + j * ldo;
    double v = acc[off];
    if (has_epi) v = __dmul_rn(outscale, cos(__dadd_rn(__dmul_rn(v, scales[i]), shifts[i])));
    OUT[off] = (TX)v;
}

static size_t panel_mib() {
    static size_t v = 0;
    if (v == 0) {
        const char *e = getenv("SKB_PANEL_MIB");   // tuning knob; default chosen by measurement
        long x = e ? atol(e) : 0;
        v = x > 0 ? (size_t)x : 256;
    }
    return v;
}

static size_t panel_rows_for(size_t S) {
    size_t kc = (panel_mib() << 20) / (S * sizeof(double));
    kc = kc / BK * BK;
    if (kc < (size_t)BK) kc = BK;
    if (kc > 32768) kc = 32768;
    return kc;
}

size_t dense_workspace_bytes(const DenseParams &p, size_t nloc, bool layout_rows, bool f32,
                             size_t m, size_t ldo) {
    size_t kc = panel_rows_for(p.S);
    if (kc > nloc) kc = (nloc + BK - 1) / BK * BK;
    if (kc == 0) kc = BK;
    size_t panel = ((kc * p.S * sizeof(double) + 255) / 256) * 256;
    size_t npanels = nloc == 0 ? 0 : (nloc + kc - 1) / kc;
    // double accumulation image of OUT, used only by the float atomic path
    size_t accbuf = f32 ? (layout_rows ? p.S * ldo : m * ldo) * sizeof(double) : 0;
    return panel * (npanels > 1 ? 2 : 1) + accbuf + 256;
}

template <typename TX, bool ROWS>
static void launch_gemm(const GemmArgs &a, bool al16p, bool alx, unsigned grid, cudaStream_t st) {
    using SM = GemmSmem<TX, ROWS>;
    constexpr int XS = sizeof(TX) == 8 ? 8 : 4;   // small-copy size for X
    constexpr int WM = GemmCfg<TX>::WM, WN = GemmCfg<TX>::WN;
#define SKB_LAUNCH(CP, CX)                                                                       \
    do {                                                                                         \
        auto kern = k_dense_gemm<TX, ROWS, CP, CX, WM, WN>;                                      \
        SKB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,         \
                                      (int)SM::BYTES));                                          \
        kern<<<grid, 32 * WM * WN, SM::BYTES, st>>>(a);                                          \
    } while (0)
    LaunchScope ls(KC_GEMM, st);
    if (al16p && alx) SKB_LAUNCH(16, 16);
    else if (al16p) SKB_LAUNCH(16, XS);
    else if (alx) SKB_LAUNCH(8, 16);
    else SKB_LAUNCH(8, XS);
#undef SKB_LAUNCH
    SKB_CUDA(cudaGetLastError());
}

// Two internal streams per device: the GEMMs run on a high-priority stream, the panel
// generator on a low-priority one, so that generation of panel p+1 fills the FP64-pipe
// bubbles (and the idle integer pipe) of GEMM p instead of running ahead of it.
struct SideStreams {
    cudaStream_t gemm = nullptr, gen = nullptr;
    cudaEvent_t fork = nullptr, join = nullptr, gen_done[2] = {nullptr, nullptr},
                gemm_done[2] = {nullptr, nullptr};
};
static SideStreams &side_streams() {
    static SideStreams per_dev[64];
    int dev = 0;
    SKB_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) throw error(SKB_ERR_GENERIC, "device index out of range");
    SideStreams &s = per_dev[dev];
    if (!s.gemm) {
        int least = 0, greatest = 0;
        SKB_CUDA(cudaDeviceGetStreamPriorityRange(&least, &greatest));
        SKB_CUDA(cudaStreamCreateWithPriority(&s.gemm, cudaStreamNonBlocking, greatest));
        SKB_CUDA(cudaStreamCreateWithPriority(&s.gen, cudaStreamNonBlocking, least));
        cudaEvent_t *evs[] = {&s.fork, &s.join, &s.gen_done[0], &s.gen_done[1], &s.gemm_done[0],
                              &s.gemm_done[1]};
        for (cudaEvent_t *e : evs) SKB_CUDA(cudaEventCreateWithFlags(e, cudaEventDisableTiming));
    }
    return s;
}

template <typename TX, bool ROWS>
static void dense_apply_t(const DenseParams &p, const TX *X, size_t ldx, size_t m, size_t koff,
                          size_t nloc, TX *OUT, size_t ldo, const RftEpilogue &epi,
                          void *workspace, size_t workspace_bytes, cudaStream_t st,
                          bool accumulate) {
    const size_t S = p.S;
    if (accumulate && (sizeof(TX) != 8 || epi.shifts != nullptr))
        throw error(SKB_ERR_INVALID_USAGE, "dense apply: accumulation needs fp64 and no epilogue");
    if (S == 0 || m == 0) return;
    size_t kc_max = panel_rows_for(S);
    if (kc_max > nloc) kc_max = (nloc + BK - 1) / BK * BK;
    if (kc_max == 0) kc_max = BK;
    const size_t panel_bytes = ((kc_max * S * sizeof(double) + 255) / 256) * 256;
    if (workspace_bytes < panel_bytes)
        throw error(SKB_ERR_ALLOC, "dense apply: workspace too small");
    const size_t npanels = nloc == 0 ? 0 : (nloc + kc_max - 1) / kc_max;
    const size_t nbuf = npanels > 1 ? 2 : 1;
    if (workspace_bytes < panel_bytes * nbuf)
        throw error(SKB_ERR_ALLOC, "dense apply: workspace too small");
    double *panels[2] = {reinterpret_cast<double *>(workspace),
                         reinterpret_cast<double *>(reinterpret_cast<char *>(workspace) +
                                                    panel_bytes * (nbuf - 1))};
    double *accbuf = reinterpret_cast<double *>(reinterpret_cast<char *>(workspace) +
                                                panel_bytes * nbuf);

    constexpr int BN = GemmCfg<TX>::BN;
    const int ntm = (int)((S + BM - 1) / BM), ntn = (int)((m + BN - 1) / BN);
    const long long ntiles = (long long)ntm * ntn;
    const bool has_epi = epi.shifts != nullptr;
    const bool tile_mode = ntiles >= 4 * 148 * GemmCfg<TX>::CTAS;     // enough tiles: no k split
    const bool single = npanels == 1 && !accumulate;
    // direct stores (and the fused epilogue) need one CTA to own a tile's whole sum
    const bool direct_ok = single;
    const bool all_direct = single && tile_mode;
    const bool use_accbuf = sizeof(TX) == 4 && !all_direct;
    const bool need_zero = !all_direct && !accumulate;
    const bool fuse_epi = has_epi && all_direct;

    void *gemm_out = use_accbuf ? (void *)accbuf : (void *)OUT;
    // accbuf is laid out exactly like OUT (same ldo) so the flush code is shared
    if (use_accbuf && workspace_bytes < panel_bytes * nbuf + (ROWS ? S * ldo : m * ldo) * sizeof(double))
        throw error(SKB_ERR_ALLOC, "dense apply: workspace too small for accumulation buffer");
    // With several panels the work is forked onto the two internal streams and joined back.
    static int no_fork = -1;
    if (no_fork < 0) no_fork = getenv("SKB_DEBUG_NO_FORK") ? 1 : 0;
    const bool forked = npanels > 1 && !no_fork;
    cudaStream_t st_gemm = st, st_gen = st;
    SideStreams *ss = nullptr;
    if (forked) {
        ss = &side_streams();
        st_gemm = ss->gemm; st_gen = ss->gen;
        SKB_CUDA(cudaEventRecord(ss->fork, st));
        SKB_CUDA(cudaStreamWaitEvent(st_gemm, ss->fork, 0));
        SKB_CUDA(cudaStreamWaitEvent(st_gen, ss->fork, 0));
    }
    if (need_zero) {
        size_t contig = ROWS ? m : S, nvec = ROWS ? S : m;
        size_t esz = use_accbuf ? sizeof(double) : sizeof(TX);
        if (ldo == contig) SKB_CUDA(cudaMemsetAsync(gemm_out, 0, contig * nvec * esz, st_gemm));
        else SKB_CUDA(cudaMemset2DAsync(gemm_out, ldo * esz, 0, contig * esz, nvec, st_gemm));
    }

    const bool al16p = (S % 2 == 0);
    const int xe = 16 / (int)sizeof(TX);
    const bool alx = (ldx % xe == 0) && ((uintptr_t)X % 16 == 0);

    auto gen_panel = [&](size_t pi) {
        size_t k0 = pi * kc_max;
        size_t kc = nloc - k0 < kc_max ? nloc - k0 : kc_max;
        if (forked && pi >= 2)    // buffer pi%2 was last read by GEMM pi-2
            SKB_CUDA(cudaStreamWaitEvent(st_gen, ss->gemm_done[pi % 2], 0));
        dense_generate_panel(p, koff + k0, kc, panels[pi % 2], st_gen);
        if (forked) SKB_CUDA(cudaEventRecord(ss->gen_done[pi % 2], st_gen));
    };
    if (npanels > 0) gen_panel(0);
    for (size_t pi = 0; pi < npanels; pi++) {
        const size_t k0 = pi * kc_max;
        size_t kc = nloc - k0 < kc_max ? nloc - k0 : kc_max;
        if (pi + 1 < npanels) gen_panel(pi + 1);
        if (forked) SKB_CUDA(cudaStreamWaitEvent(st_gemm, ss->gen_done[pi % 2], 0));
        GemmArgs a;
        a.P = panels[pi % 2]; a.S = S; a.kc = kc;
        a.X = ROWS ? (const void *)(X + k0 * ldx) : (const void *)(X + k0);
        a.ldx = ldx; a.m = m; a.OUT = gemm_out; a.ldo = ldo;
        a.ntm = ntm; a.ntn = ntn; a.nk = (int)((kc + BK - 1) / BK);
        a.direct_ok = direct_ok ? 1 : 0;
        a.out_is_acc = use_accbuf ? 1 : 0;
        a.scales = epi.scales; a.shifts = epi.shifts; a.outscale = epi.outscale;
        a.has_epi = fuse_epi ? 1 : 0;
        if (!tile_mode && a.direct_ok) a.direct_ok = 0;   // stream-K ranges may split a tile
        a.whole_tiles = tile_mode ? 1 : 0;
        long long total = ntiles * a.nk;
        const long long ncta = 148LL * GemmCfg<TX>::CTAS;
        unsigned grid = tile_mode ? (unsigned)ncta : (unsigned)(total < ncta ? total : ncta);
        // alignment of the X slab for this panel
        bool alx_p = alx && (ROWS || (k0 % xe == 0));
        launch_gemm<TX, ROWS>(a, al16p, alx_p, grid, st_gemm);
        if (forked) SKB_CUDA(cudaEventRecord(ss->gemm_done[pi % 2], st_gemm));
    }
    if (nloc != 0 && (use_accbuf || (has_epi && !fuse_epi))) {
        const double *src = use_accbuf ? accbuf : reinterpret_cast<const double *>(OUT);
        size_t n = S * m;
        LaunchScope ls(KC_MISC, st_gemm);
        k_dense_finalize<TX, ROWS><<<(unsigned)((n + 255) / 256), 256, 0, st_gemm>>>(
            src, OUT, S, m, ldo, epi.scales, epi.shifts, epi.outscale, has_epi ? 1 : 0);
        SKB_CUDA(cudaGetLastError());
    }
    if (forked) {
        SKB_CUDA(cudaEventRecord(ss->join, st_gemm));
        SKB_CUDA(cudaStreamWaitEvent(st, ss->join, 0));
    }
}

// ---------------------------------------------------------------------------
// K5c: fp32 operands in the strided layout with the whole contraction in one panel -- the
// rowwise feature map of C5 (4M x 784 fp32 -> 16384).  The reference computes this case in
// single precision (W is cast to float, then sgemm: dense_transform_data.hpp:90-99); FP64 DMMA
// is far more precise than the 1e-5 bar asks and 10x slower than the fp32-class tensor path.
// Here every fp32 operand is split into two TF32 numbers (hi + lo, 21 mantissa bits together)
// and a product is three mma.sync.m16n8k8.tf32 (lo*hi, hi*lo, hi*hi) accumulated in fp32
// ("3xTF32"), which keeps fp32-level accuracy; the cosine epilogue stays in fp64 and is fused.
//   OUT[i + j*ldo] = f( sum_k Pf[k*S + i] * X[k + j*ldx] ),  i < S (features), j < m (points)
// CTA tile 128(i) x 128(j) x 32(k), 8 warps (2 x 4, warp tile 64 x 32), 3-stage cp.async
// pipeline, two CTAs per SM so that one CTA's epilogue overlaps the other's MMAs.
// ---------------------------------------------------------------------------
namespace tf32 {
constexpr int TBM = 128, TBN = 128, TBK = 32, TSTAGES = 3, TNT = 256;
constexpr int LDA = TBM + 8;      // floats, As[k][i]: = 8 mod 32 -> a fragment's 4 k-rows x 8 i hit 32 banks
constexpr int LDB = TBK + 4;      // floats, Bs[j][k]: = 4 mod 32 -> 8 j-rows x 4 k hit 32 banks
constexpr int STAGE_FLOATS = TBK * LDA + TBN * LDB;
constexpr size_t SMEM = (size_t)TSTAGES * STAGE_FLOATS * sizeof(float);
}  // namespace tf32

struct Tf32Args {
    const float *P; const float *X; float *OUT;
    size_t S, m, K, ldx, ldo;
    const double *scales, *shifts; double outscale; int has_epi;
    int ntm; long long ntiles;
};

__global__ void k_panel_to_f32(const double *__restrict__ P, float *__restrict__ Pf, size_t n) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) Pf[i] = (float)P[i];          // round to nearest, the reference's cast
}

// hi = x with the 13 low mantissa bits cleared (a TF32 number), lo = x - hi (exact in fp32).
// lo is passed as it is: the tensor core reads only the TF32 part of an operand, i.e. drops
// lo's own low bits (2^-21 relative to x).  Two instructions per element.
__device__ __forceinline__ void split_tf32(float x, unsigned &hi, unsigned &lo) {
    hi = __float_as_uint(x) & 0xffffe000u;
    lo = __float_as_uint(x - __uint_as_float(hi));
}
__device__ __forceinline__ void mma_tf32(float (&c)[4], const unsigned (&a)[4], const unsigned (&b)[2]) {
    asm("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}
__device__ __forceinline__ void mma_tf32_zero(float (&d)[4], const unsigned (&a)[4], const unsigned (&b)[2]) {
    asm("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%10,%10,%10,%10};"
        : "=f"(d[0]), "=f"(d[1]), "=f"(d[2]), "=f"(d[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]), "f"(0.f));
}

// The tensor core adds into its accumulator with truncation, so a long chain of accumulating
// MMAs shrinks the sum systematically (measured: 9e-6 relative over 294 MMAs).  Every k8 step
// therefore starts from a zero accumulator (3 MMAs: lo*hi, hi*lo, hi*hi) and its result is
// added to the running sum with an ordinary round-to-nearest FADD.
__global__ void __launch_bounds__(tf32::TNT, 2) k_dense_gemm_tf32x3(Tf32Args a) {
    using namespace tf32;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float *sm = reinterpret_cast<float *>(smem_raw);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int wm = warp & 1, wn = warp >> 1;              // 2 x 4 warps, warp tile 64(i) x 32(j)
    const int g = lane >> 2, t = lane & 3;
    const long long tile = blockIdx.x;
    const size_t i0 = (size_t)(tile % a.ntm) * TBM, j0 = (size_t)(tile / a.ntm) * TBN;   // i fastest: the X tile stays in L2
    const int nk = (int)((a.K + TBK - 1) / TBK);

    auto load_stage = [&](int st, int kb) {
        float *As = sm + (size_t)st * STAGE_FLOATS, *Bs = As + TBK * LDA;
        const size_t k0 = (size_t)kb * TBK;
#pragma unroll
        for (int c = 0; c < (TBK * TBM / 4) / TNT; c++) {          // 2 chunks of 16 B per thread
            const int q = tid + c * TNT, kk = q >> 5, ci = (q & 31) * 4;
            const bool ok = k0 + kk < a.K && i0 + ci < a.S;        // S, K are multiples of 4 here
            const float *src = ok ? a.P + (k0 + kk) * a.S + i0 + ci : a.P;
            cp_async<16>(As + kk * LDA + ci, src, ok ? 16 : 0);
        }
#pragma unroll
        for (int c = 0; c < (TBN * TBK / 4) / TNT; c++) {
            const int q = tid + c * TNT, jj = q >> 3, ck = (q & 7) * 4;
            const bool ok = j0 + jj < a.m && k0 + ck < a.K;
            const float *src = ok ? a.X + (j0 + jj) * a.ldx + k0 + ck : a.X;
            cp_async<16>(Bs + jj * LDB + ck, src, ok ? 16 : 0);
        }
        asm volatile("cp.async.commit_group;");
    };

    float acc[4][4][4];
#pragma unroll
    for (int mi = 0; mi < 4; mi++)
#pragma unroll
        for (int ni = 0; ni < 4; ni++)
#pragma unroll
            for (int e = 0; e < 4; e++) acc[mi][ni][e] = 0.f;

#pragma unroll
    for (int s = 0; s < TSTAGES - 1; s++) {
        if (s < nk) load_stage(s, s); else asm volatile("cp.async.commit_group;");
    }
    for (int kb = 0; kb < nk; kb++) {
        asm volatile("cp.async.wait_group %0;" ::"n"(TSTAGES - 2));
        __syncthreads();
        if (kb + TSTAGES - 1 < nk) load_stage((kb + TSTAGES - 1) % TSTAGES, kb + TSTAGES - 1);
        else asm volatile("cp.async.commit_group;");
        const float *As = sm + (size_t)(kb % TSTAGES) * STAGE_FLOATS, *Bs = As + TBK * LDA;
#pragma unroll
        for (int ks = 0; ks < TBK / 8; ks++) {
            unsigned bh[4][2], bl[4][2];
#pragma unroll
            for (int ni = 0; ni < 4; ni++) {
                const float *bp = Bs + (wn * 32 + ni * 8 + g) * LDB + ks * 8 + t;
                split_tf32(bp[0], bh[ni][0], bl[ni][0]);
                split_tf32(bp[4], bh[ni][1], bl[ni][1]);
            }
#pragma unroll
            for (int mi = 0; mi < 4; mi++) {
                const float *ap = As + (ks * 8 + t) * LDA + wm * 64 + mi * 16 + g;
                unsigned ah[4], al[4];
                split_tf32(ap[0], ah[0], al[0]);
                split_tf32(ap[8], ah[1], al[1]);
                split_tf32(ap[4 * LDA], ah[2], al[2]);
                split_tf32(ap[4 * LDA + 8], ah[3], al[3]);
#pragma unroll
                for (int ni = 0; ni < 4; ni++) {
                    float tmp[4];
                    mma_tf32_zero(tmp, al, bh[ni]);    // small terms first
                    mma_tf32(tmp, ah, bl[ni]);
                    mma_tf32(tmp, ah, bh[ni]);
#pragma unroll
                    for (int e = 0; e < 4; e++) acc[mi][ni][e] += tmp[e];
                }
            }
        }
    }
    asm volatile("cp.async.wait_group 0;");
    // epilogue: c0 (g, 2t) c1 (g, 2t+1) c2 (g+8, 2t) c3 (g+8, 2t+1) of every 16 x 8 block.
    // The cosine argument is formed and reduced to [-pi, pi] in fp64, the cosine itself is fp32.
#pragma unroll
    for (int mi = 0; mi < 4; mi++) {
#pragma unroll
        for (int h = 0; h < 2; h++) {
            const size_t i = i0 + wm * 64 + mi * 16 + g + h * 8;
            if (i >= a.S) continue;
            double sc = 1.0, sh = 0.0;
            if (a.has_epi) { sc = a.scales[i]; sh = a.shifts[i]; }
#pragma unroll
            for (int ni = 0; ni < 4; ni++) {
#pragma unroll
                for (int e = 0; e < 2; e++) {
                    const size_t j = j0 + wn * 32 + ni * 8 + 2 * t + e;
                    if (j >= a.m) continue;
                    float v = acc[mi][ni][h * 2 + e];
                    if (a.has_epi) {
                        const double arg = fma((double)v, sc, sh);
                        const double q = rint(arg * 0.15915494309189535);            // / (2 pi)
                        const double r = fma(q, -6.283185307179586, arg);             // |r| <= pi (+ 1e-15)
                        v = (float)(a.outscale * (double)cosf((float)r));
                    }
                    a.OUT[i + j * a.ldo] = v;
                }
            }
        }
    }
}

// ---------------------------------------------------------------------------
// K5d (form without TMA; the fallback when the driver offers no tensor-map encoder): the same
// 3xTF32 product on the 5th-generation tensor cores (tcgen05.mma.kind::tf32, accumulators in
// TMEM).  The tensor core reads shared memory itself, so operands are split BEFORE they get
// there.  A CTA owns a 128(i) x 128(j) tile: 4 producer warps load the raw fp32 rows (panel
// transposed to K-major by a pre-pass and padded with zeros to a multiple of 16 in k; X from the
// caller's buffer), split them in registers and store hi and lo tiles in the no-swizzle K-major
// core-matrix layout (element (r,k) at (r/8)*512 + (k/4)*128 + (r%8)*16 + (k%4)*4) of a 3-stage
// ring; one thread issues the MMAs -- hi*hi into one TMEM accumulator, lo*hi and hi*lo into a
// second one, so the long truncating chain carries only the large term once -- and the producer
// warps, done loading, read both accumulators back (tcgen05.ld), add them and apply the epilogue.
// Two CTAs fit an SM (96 KB of shared memory, 256 TMEM columns each): one's epilogue overlaps the
// other's MMAs.  The TMA forms further down are the ones normally used.
// Measured accumulator behaviour (tools/tcgen05_probe.cu): operands are truncated to TF32,
// a chain of 100 accumulating MMAs loses 1.6e-6 (relative, Frobenius), 32 lose 2.3e-7.
// ---------------------------------------------------------------------------
namespace tc {
constexpr int TM = 128, TN = 128, TK = 16, NSTAGE = 3;
constexpr int PRODUCERS = 128, THREADS = PRODUCERS + 32;
constexpr uint32_t LBO = 128, SBO = (TK / 4) * 128;            // bytes
constexpr uint32_t TILE_BYTES = TM * TK * 4;                   // 8 KB
constexpr uint32_t STAGE_BYTES = 4 * TILE_BYTES;               // P_hi, P_lo, X_hi, X_lo
constexpr size_t SMEM = (size_t)NSTAGE * STAGE_BYTES + 256;    // + barriers, TMEM base
constexpr uint32_t TMEM_COLS = 256;                            // main, cross
}  // namespace tc

// panel P[k*S + i] (fp64) -> Pt[i*Kp + k] (fp32, the reference's cast; K-major, zero padded in k)
__global__ void k_tc_transpose_panel(const double *__restrict__ P, size_t S, size_t K, size_t Kp,
                                     float *__restrict__ Pt, float *__restrict__ Pt_lo = nullptr) {
    __shared__ float tile[32][33];
    const size_t i0 = (size_t)blockIdx.x * 32, k0 = (size_t)blockIdx.y * 32;
    for (int r = threadIdx.y; r < 32; r += blockDim.y) {          // r = k, x = i (coalesced read)
        const size_t k = k0 + r, i = i0 + threadIdx.x;
        tile[r][threadIdx.x] = (k < K && i < S) ? (float)P[k * S + i] : 0.f;
    }
    __syncthreads();
    for (int r = threadIdx.y; r < 32; r += blockDim.y) {          // r = i, x = k (coalesced write)
        const size_t i = i0 + r, k = k0 + threadIdx.x;
        if (i < S && k < Kp) {
            const float x = tile[threadIdx.x][r];
            Pt[i * Kp + k] = x;
            if (Pt_lo) Pt_lo[i * Kp + k] = x - __uint_as_float(__float_as_uint(x) & 0xffffe000u);
        }
    }
}

__device__ __forceinline__ uint64_t tc_desc(uint32_t saddr) {
    return (uint64_t)((saddr >> 4) & 0x3fff) | ((uint64_t)(tc::LBO >> 4) << 16) |
           ((uint64_t)(tc::SBO >> 4) << 32) | ((uint64_t)1 << 46);      // version 1, no swizzle
}
__device__ __forceinline__ void tc_mma(uint32_t tmem, uint64_t ad, uint64_t bd, uint32_t idesc, uint32_t accum) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
                 ::"r"(tmem), "l"(ad), "l"(bd), "r"(idesc), "r"(accum) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    // only single threads and elected lanes wait here, so a plain try_wait loop (which suspends in
    // hardware for a bounded time) costs nothing elsewhere and adds no wake-up latency
    uint32_t done = 0;
    while (!done)
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
                     "selp.u32 %0, 1, 0, p;\n\t}" : "=r"(done) : "r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tc_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}

struct TcArgs {
    const float *Pt, *X; float *OUT;          // Pt[i*Kp + k] (panel, transposed, zero padded), X[j*ldx + k]
    size_t S, m, K, Kp, ldx, ldo;
    const double *scales, *shifts; double outscale; int has_epi;
    int ntm, ntn;
};

__global__ void __launch_bounds__(tc::THREADS, 2) k_dense_gemm_tc_tf32x3(TcArgs a) {
    using namespace tc;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const uint32_t sbase = (uint32_t)__cvta_generic_to_shared(smem_raw);
    const uint32_t bars = sbase + NSTAGE * STAGE_BYTES;        // full[3], empty[3], accfull: 8 B each
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(smem_raw + NSTAGE * STAGE_BYTES + 64);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    // tile order: groups of GI i-tiles, inside a group j outer / i inner, so the CTAs in flight share
    // GI panel tiles and a few X tiles (L2 working set of tens of MB instead of the whole panel)
    size_t ti, tj;
    {
        constexpr int GI = 16;
        const long long t = blockIdx.x, per = (long long)GI * a.ntn, grp = t / per, tp = t - grp * per;
        const int gi = (int)min((long long)GI, (long long)a.ntm - grp * GI);
        tj = (size_t)(tp / gi); ti = (size_t)(grp * GI + tp % gi);
    }
    const size_t i0 = ti * TM, j0 = tj * TN;
    const int nk = (int)(a.Kp / TK);
    auto full_bar = [&](int s) { return bars + 8u * s; };
    auto empty_bar = [&](int s) { return bars + 8u * (NSTAGE + s); };
    const uint32_t acc_bar = bars + 8u * (2 * NSTAGE);

    if (tid == 0) {
        for (int s = 0; s < NSTAGE; s++) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(full_bar(s)), "r"(PRODUCERS));
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(empty_bar(s)));
        }
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(acc_bar));
        asm volatile("fence.mbarrier_init.release.cluster;");
    }
    if (warp == 4) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                     ::"r"((uint32_t)__cvta_generic_to_shared(tmem_slot)), "n"(TMEM_COLS));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;");
    const uint32_t tmem = *tmem_slot;

    if (warp < 4) {
        // ===== producers: load raw fp32 once, split into hi / lo in registers, store both tiles.
        // A warp instruction reads 8 rows x 64 contiguous bytes (whole sectors): lane -> (row =
        // lane%8 + 8*warp + 32*c, 16-byte chunk = lane/8); the loads of k-block kb+1 are in flight
        // while k-block kb is split and stored =====
        // (8 lanes of a quarter warp = 8 consecutive rows of one chunk: their 16-byte stores fill one
        // 128-byte core matrix, conflict-free; the warp as a whole still reads 8 rows x 64 bytes)
        const int kc = lane >> 3;
        uint32_t drow[4]; const float *prow[4], *xrow[4]; bool okp[4], okx[4];
#pragma unroll
        for (int c = 0; c < 4; c++) {
            const int r = warp * 8 + (lane & 7) + 32 * c;
            drow[c] = (uint32_t)(r >> 3) * SBO + (uint32_t)(r & 7) * 16 + (uint32_t)kc * LBO;
            okp[c] = i0 + r < a.S; okx[c] = j0 + r < a.m;
            prow[c] = a.Pt + (okp[c] ? (i0 + r) * a.Kp : 0) + kc * 4;
            xrow[c] = a.X + (okx[c] ? (j0 + r) * a.ldx : 0) + kc * 4;
        }
        auto fetch = [&](int kb, float4 (&pv)[4], float4 (&xv)[4]) {
            const size_t kof = (size_t)kb * TK;
            const bool kin = kof + kc * 4 < a.K;          // K is a multiple of 4: a chunk is inside or outside
#pragma unroll
            for (int c = 0; c < 4; c++) {
                pv[c] = okp[c] ? __ldg(reinterpret_cast<const float4 *>(prow[c] + kof)) : make_float4(0.f, 0.f, 0.f, 0.f);
                xv[c] = (okx[c] && kin) ? __ldg(reinterpret_cast<const float4 *>(xrow[c] + kof)) : make_float4(0.f, 0.f, 0.f, 0.f);
            }
        };
        auto put = [&](uint32_t addr_hi, uint32_t addr_lo, float4 v) {
            float4 h, l;
            h.x = __uint_as_float(__float_as_uint(v.x) & 0xffffe000u); l.x = v.x - h.x;
            h.y = __uint_as_float(__float_as_uint(v.y) & 0xffffe000u); l.y = v.y - h.y;
            h.z = __uint_as_float(__float_as_uint(v.z) & 0xffffe000u); l.z = v.z - h.z;
            h.w = __uint_as_float(__float_as_uint(v.w) & 0xffffe000u); l.w = v.w - h.w;
            asm volatile("st.shared.v4.f32 [%0], {%1,%2,%3,%4};" ::"r"(addr_hi), "f"(h.x), "f"(h.y), "f"(h.z), "f"(h.w) : "memory");
            asm volatile("st.shared.v4.f32 [%0], {%1,%2,%3,%4};" ::"r"(addr_lo), "f"(l.x), "f"(l.y), "f"(l.z), "f"(l.w) : "memory");
        };
        auto stage_out = [&](int kb, float4 (&pv)[4], float4 (&xv)[4]) {
            const int s = kb % NSTAGE;
            if (kb >= NSTAGE) {          // one lane polls, the warp follows (the barrier unit is shared with the stores)
                if (lane == 0) mbar_wait(empty_bar(s), (uint32_t)((kb / NSTAGE - 1) & 1));
                __syncwarp();
            }
            const uint32_t dst = sbase + (uint32_t)s * STAGE_BYTES;
#pragma unroll
            for (int c = 0; c < 4; c++) {
                put(dst + drow[c], dst + TILE_BYTES + drow[c], pv[c]);
                put(dst + 2 * TILE_BYTES + drow[c], dst + 3 * TILE_BYTES + drow[c], xv[c]);
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");     // generic-proxy stores -> tensor core
            mbar_arrive(full_bar(s));
        };
        // two register sets: the loads of k-blocks kb+1 and kb+2 are in flight while kb is split and stored
        float4 pa[4], xa[4], pb[4], xb[4];
        fetch(0, pa, xa);
        if (nk > 1) fetch(1, pb, xb);
        for (int kb = 0; kb < nk; kb += 2) {
            stage_out(kb, pa, xa);
            if (kb + 2 < nk) fetch(kb + 2, pa, xa);
            if (kb + 1 < nk) {
                stage_out(kb + 1, pb, xb);
                if (kb + 3 < nk) fetch(kb + 3, pb, xb);
            }
        }
        // ===== epilogue: warp w owns TMEM lanes [32w, 32w+32) = rows i, columns = j =====
        if (lane == 0) mbar_wait(acc_bar, 0);
        __syncwarp();
        asm volatile("tcgen05.fence::after_thread_sync;");
        const size_t i = i0 + (size_t)tid;
        double sc = 1.0, sh = 0.0;
        if (a.has_epi && i < a.S) { sc = a.scales[i]; sh = a.shifts[i]; }
        for (int c0 = 0; c0 < TN; c0 += 8) {
            uint32_t mn[8], cr[8];
            const uint32_t taddr = tmem + ((uint32_t)(warp * 32) << 16) + (uint32_t)c0;
            asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                         : "=r"(mn[0]), "=r"(mn[1]), "=r"(mn[2]), "=r"(mn[3]), "=r"(mn[4]), "=r"(mn[5]), "=r"(mn[6]), "=r"(mn[7])
                         : "r"(taddr));
            asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                         : "=r"(cr[0]), "=r"(cr[1]), "=r"(cr[2]), "=r"(cr[3]), "=r"(cr[4]), "=r"(cr[5]), "=r"(cr[6]), "=r"(cr[7])
                         : "r"(taddr + (uint32_t)TN));
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
            if (i < a.S) {
#pragma unroll
                for (int e = 0; e < 8; e++) {
                    const size_t j = j0 + c0 + e;
                    if (j >= a.m) break;
                    float v = __uint_as_float(mn[e]) + __uint_as_float(cr[e]);
                    if (a.has_epi) {
                        const double arg = fma((double)v, sc, sh);
                        const double q = rint(arg * 0.15915494309189535);
                        const double rr = fma(q, -6.283185307179586, arg);
                        v = (float)(a.outscale * (double)cosf((float)rr));
                    }
                    a.OUT[i + j * a.ldo] = v;
                }
            }
        }
    } else if (lane == 0) {
        // ===== MMA issuer =====
        const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(TN >> 3) << 17) | ((uint32_t)(TM >> 4) << 24);
        const uint32_t t_main = tmem, t_cross = tmem + (uint32_t)TN;
        for (int kb = 0; kb < nk; kb++) {
            const int s = kb % NSTAGE;
            mbar_wait(full_bar(s), (uint32_t)((kb / NSTAGE) & 1));
            asm volatile("tcgen05.fence::after_thread_sync;");
            const uint32_t st0 = sbase + (uint32_t)s * STAGE_BYTES;
#pragma unroll
            for (int ks = 0; ks < TK / 8; ks++) {
                const uint32_t ko = (uint32_t)ks * 2 * LBO;
                const uint64_t phi = tc_desc(st0 + ko), plo = tc_desc(st0 + TILE_BYTES + ko);
                const uint64_t xhi = tc_desc(st0 + 2 * TILE_BYTES + ko), xlo = tc_desc(st0 + 3 * TILE_BYTES + ko);
                const uint32_t first = (kb | ks) == 0 ? 0u : 1u;
                tc_mma(t_cross, plo, xhi, idesc, first);
                tc_mma(t_cross, phi, xlo, idesc, 1u);
                tc_mma(t_main, phi, xhi, idesc, first);
            }
            tc_commit(empty_bar(s));          // frees the stage once these MMAs have read it
        }
        tc_commit(acc_bar);
    }
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    __syncwarp();
    if (warp == 4) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(tc::TMEM_COLS));
}

// Shapes the 3xTF32 path takes (fp32 operands, strided layout, no accumulation).
bool dense_tf32_path_ok(const DenseParams &p, const void *X, size_t ldx, size_t m, size_t nloc) {
    static const bool off = getenv("SKB_DEBUG_NO_TF32") != nullptr;
    const size_t S = p.S;
    if (off || S == 0 || m == 0 || nloc == 0) return false;
    if (S % 4 || nloc % 4 || ldx % 4 || ((uintptr_t)X % 16)) return false;
    if (nloc > panel_rows_for(S)) return false;
    const long long ntm = (long long)((S + tf32::TBM - 1) / tf32::TBM), ntn = (long long)((m + tf32::TBN - 1) / tf32::TBN);
    return ntm * ntn <= 0x7fffffffLL;
}
static size_t tc_workspace_bytes(const DenseParams &p, size_t nloc, size_t m);
static bool tc_enabled();
size_t dense_tf32_workspace_bytes(const DenseParams &p, size_t nloc, size_t m) {
    size_t need = ((nloc * p.S * sizeof(double) + 255) / 256) * 256 + ((nloc * p.S * sizeof(float) + 255) / 256) * 256 + 256;
    if (tc_enabled()) need = std::max(need, tc_workspace_bytes(p, nloc, m));
    return need;
}

// Returns false when the shape does not qualify for the 3xTF32 path.
static bool dense_apply_tf32x3(const DenseParams &p, const float *X, size_t ldx, size_t m, size_t koff,
                               size_t nloc, float *OUT, size_t ldo, const RftEpilogue &epi,
                               void *workspace, size_t workspace_bytes, cudaStream_t st) {
    const size_t S = p.S;
    if (!dense_tf32_path_ok(p, X, ldx, m, nloc)) return false;
    const size_t pbytes = ((nloc * S * sizeof(double) + 255) / 256) * 256;
    const size_t fbytes = ((nloc * S * sizeof(float) + 255) / 256) * 256;
    if (workspace_bytes < pbytes + fbytes) return false;
    const long long ntm = (long long)((S + tf32::TBM - 1) / tf32::TBM), ntn = (long long)((m + tf32::TBN - 1) / tf32::TBN);
    double *P = reinterpret_cast<double *>(workspace);
    float *Pf = reinterpret_cast<float *>(reinterpret_cast<char *>(workspace) + pbytes);
    dense_generate_panel(p, koff, nloc, P, st);
    {
        LaunchScope ls(KC_GEN, st);
        const size_t n = nloc * S;
        k_panel_to_f32<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(P, Pf, n);
        SKB_CUDA(cudaGetLastError());
    }
    Tf32Args a;
    a.P = Pf; a.X = X; a.OUT = OUT; a.S = S; a.m = m; a.K = nloc; a.ldx = ldx; a.ldo = ldo;
    a.scales = epi.scales; a.shifts = epi.shifts; a.outscale = epi.outscale;
    a.has_epi = epi.shifts != nullptr ? 1 : 0;
    a.ntm = (int)ntm; a.ntiles = ntm * ntn;
    SKB_CUDA(cudaFuncSetAttribute(k_dense_gemm_tf32x3, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tf32::SMEM));
    LaunchScope ls(KC_GEMM, st);
    k_dense_gemm_tf32x3<<<(unsigned)a.ntiles, tf32::TNT, tf32::SMEM, st>>>(a);
    SKB_CUDA(cudaGetLastError());
    return true;
}

// ---------------------------------------------------------------------------
// K5e: K5d with TMA.  The hardware truncates fp32 operands to TF32 (tools/tcgen05_probe.cu:
// 3.5e-8 against truncated operands, 2.9e-4 against rounded ones), so a raw fp32 tile IS the hi
// operand: one thread issues two cp.async.bulk.tensor loads per stage (16 x 128 boxes of the
// transposed panel and of X, 64-byte swizzle = the canonical K-major SWIZZLE_64B layout of the
// MMA), out-of-range rows and the k tail arrive as zeros.  The producer warps only compute the lo
// tiles (lo = x - trunc(x), element by element in the same swizzled positions): per stage 16 KB
// read and 16 KB written through the LSU instead of 16 KB of scattered global loads and 32 KB of
// stores.  Everything else (two TMEM accumulators, commit-based stage release, epilogue) as K5d.
// ---------------------------------------------------------------------------
typedef CUresult (*skb_encode_tiled_t)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                       const cuuint64_t *, const cuuint32_t *, const cuuint32_t *,
                                       CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                       CUtensorMapFloatOOBfill);
static skb_encode_tiled_t tma_encoder() {
    static skb_encode_tiled_t fn = nullptr;
    static bool tried = false;
    if (!tried) {
        tried = true;
        void *p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<skb_encode_tiled_t>(p);
        cudaGetLastError();
    }
    return fn;
}
// rows x cols fp32 matrix with row stride ld (elements), boxes of 16 columns x 128 rows
static bool tma_make_map(CUtensorMap *map, const float *base, size_t rows, size_t cols, size_t ld,
                         unsigned box_rows = 128) {
    skb_encode_tiled_t enc = tma_encoder();
    if (!enc) return false;
    const cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    const cuuint64_t strides[1] = {(cuuint64_t)ld * sizeof(float)};
    const cuuint32_t box[2] = {16, box_rows}, estr[2] = {1, 1};
    return enc(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float *>(base), dims, strides, box, estr,
               CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
               CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

namespace tm {
constexpr int TM = 128, TN = 128, TK = 16, NSTAGE = 3;
constexpr int PRODUCERS = 128, THREADS = PRODUCERS + 64;     // + MMA warp + TMA warp
constexpr uint32_t SBO = 8 * 64;                               // 8 rows of 64 bytes
constexpr uint32_t TILE_BYTES = TM * TK * 4;                   // 8 KB
constexpr uint32_t STAGE_BYTES = 4 * TILE_BYTES;               // P raw, P lo, X raw, X lo
constexpr size_t SMEM = (size_t)NSTAGE * STAGE_BYTES + 1024 + 256;   // + alignment slack + barriers
constexpr uint32_t TMEM_COLS = 256;
}  // namespace tm

__device__ __forceinline__ uint64_t tm_desc(uint32_t saddr) {       // K-major, SWIZZLE_64B
    return (uint64_t)((saddr >> 4) & 0x3fff) | ((uint64_t)1 << 16) | ((uint64_t)(tm::SBO >> 4) << 32) |
           ((uint64_t)1 << 46) | ((uint64_t)4 << 61);
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap *map, int c0, int c1, uint32_t bar) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                 ::"r"(dst), "l"(map), "r"(c0), "r"(c1), "r"(bar) : "memory");
}

struct TmArgs {
    float *OUT; size_t S, m, ldo;
    const double *scales, *shifts; double outscale; int has_epi;
    int ntm, ntn, nk;
};

__global__ void __launch_bounds__(tm::THREADS, 2)
k_dense_gemm_tc_tma(const __grid_constant__ CUtensorMap tmP, const __grid_constant__ CUtensorMap tmX, TmArgs a) {
    using namespace tm;
    extern __shared__ unsigned char smem_raw[];
    const uint32_t sbase = ((uint32_t)__cvta_generic_to_shared(smem_raw) + 1023u) & ~1023u;   // swizzle atoms want aligned tiles
    const uint32_t bars = sbase + NSTAGE * STAGE_BYTES;        // tma[3], full[3], empty[3], acc: 8 B each
    __shared__ uint32_t tmem_slot;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    size_t ti, tj;
    {
        constexpr int GI = 16;
        const long long t = blockIdx.x, per = (long long)GI * a.ntn, grp = t / per, tp = t - grp * per;
        const int gi = (int)min((long long)GI, (long long)a.ntm - grp * GI);
        tj = (size_t)(tp / gi); ti = (size_t)(grp * GI + tp % gi);
    }
    const size_t i0 = ti * TM, j0 = tj * TN;
    const int nk = a.nk;
    auto tma_bar = [&](int s) { return bars + 8u * s; };
    auto full_bar = [&](int s) { return bars + 8u * (NSTAGE + s); };
    auto empty_bar = [&](int s) { return bars + 8u * (2 * NSTAGE + s); };
    const uint32_t acc_bar = bars + 8u * (3 * NSTAGE);

    if (tid == 0) {
        for (int s = 0; s < NSTAGE; s++) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(tma_bar(s)));
            asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(full_bar(s)), "r"(PRODUCERS));
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(empty_bar(s)));
        }
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(acc_bar));
        asm volatile("fence.mbarrier_init.release.cluster;");
    }
    if (warp == 4) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                     ::"r"((uint32_t)__cvta_generic_to_shared(&tmem_slot)), "n"(TMEM_COLS));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;");
    const uint32_t tmem = tmem_slot;

    if (warp == 5) {
        if (lane == 0) {
            // ===== TMA issuer =====
            for (int kb = 0; kb < nk; kb++) {
                const int s = kb % NSTAGE;
                if (kb >= NSTAGE) mbar_wait(empty_bar(s), (uint32_t)((kb / NSTAGE - 1) & 1));
                const uint32_t st0 = sbase + (uint32_t)s * STAGE_BYTES;
                asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(tma_bar(s)), "r"(2 * TILE_BYTES) : "memory");
                tma_load_2d(st0, &tmP, kb * TK, (int)i0, tma_bar(s));
                tma_load_2d(st0 + 2 * TILE_BYTES, &tmX, kb * TK, (int)j0, tma_bar(s));
            }
        }
    } else if (warp == 4) {
        if (lane == 0) {
            // ===== MMA issuer =====
            const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(TN >> 3) << 17) | ((uint32_t)(TM >> 4) << 24);
            const uint32_t t_main = tmem, t_cross = tmem + (uint32_t)TN;
            for (int kb = 0; kb < nk; kb++) {
                const int s = kb % NSTAGE;
                mbar_wait(full_bar(s), (uint32_t)((kb / NSTAGE) & 1));
                asm volatile("tcgen05.fence::after_thread_sync;");
                const uint32_t st0 = sbase + (uint32_t)s * STAGE_BYTES;
#pragma unroll
                for (int ks = 0; ks < TK / 8; ks++) {
                    const uint32_t ko = (uint32_t)ks * 32;          // 8 floats inside the 64-byte swizzle row
                    const uint64_t phi = tm_desc(st0 + ko), plo = tm_desc(st0 + TILE_BYTES + ko);
                    const uint64_t xhi = tm_desc(st0 + 2 * TILE_BYTES + ko), xlo = tm_desc(st0 + 3 * TILE_BYTES + ko);
                    const uint32_t first = (kb | ks) == 0 ? 0u : 1u;
                    tc_mma(t_cross, plo, xhi, idesc, first);
                    tc_mma(t_cross, phi, xlo, idesc, 1u);
                    tc_mma(t_main, phi, xhi, idesc, first);
                }
                tc_commit(empty_bar(s));
            }
            tc_commit(acc_bar);
        }
    } else {
        // ===== lo producers: lo = x - trunc(x), same (swizzled) position in the lo tile =====
        for (int kb = 0; kb < nk; kb++) {
            const int s = kb % NSTAGE;
            if (lane == 0) mbar_wait(tma_bar(s), (uint32_t)((kb / NSTAGE) & 1));
            __syncwarp();
            const uint32_t st0 = sbase + (uint32_t)s * STAGE_BYTES;
#pragma unroll
            for (int tl = 0; tl < 2; tl++) {
                const uint32_t raw = st0 + (uint32_t)tl * 2 * TILE_BYTES, lo = raw + TILE_BYTES;
#pragma unroll
                for (int c = 0; c < 4; c++) {
                    const uint32_t off = (uint32_t)(tid + c * PRODUCERS) * 16;
                    float4 v, l;
                    asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(raw + off));
                    l.x = v.x - __uint_as_float(__float_as_uint(v.x) & 0xffffe000u);
                    l.y = v.y - __uint_as_float(__float_as_uint(v.y) & 0xffffe000u);
                    l.z = v.z - __uint_as_float(__float_as_uint(v.z) & 0xffffe000u);
                    l.w = v.w - __uint_as_float(__float_as_uint(v.w) & 0xffffe000u);
                    asm volatile("st.shared.v4.f32 [%0], {%1,%2,%3,%4};" ::"r"(lo + off), "f"(l.x), "f"(l.y), "f"(l.z), "f"(l.w) : "memory");
                }
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            mbar_arrive(full_bar(s));
        }
        // ===== epilogue: warp w owns TMEM lanes [32w, 32w+32) = rows i, columns = j =====
        if (lane == 0) mbar_wait(acc_bar, 0);
        __syncwarp();
        asm volatile("tcgen05.fence::after_thread_sync;");
        const size_t i = i0 + (size_t)tid;
        double sc = 1.0, sh = 0.0;
        if (a.has_epi && i < a.S) { sc = a.scales[i]; sh = a.shifts[i]; }
        const float osc = (float)a.outscale;
        for (int c0 = 0; c0 < TN; c0 += 8) {
            uint32_t mn[8], cr[8];
            const uint32_t taddr = tmem + ((uint32_t)(warp * 32) << 16) + (uint32_t)c0;
            asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                         : "=r"(mn[0]), "=r"(mn[1]), "=r"(mn[2]), "=r"(mn[3]), "=r"(mn[4]), "=r"(mn[5]), "=r"(mn[6]), "=r"(mn[7])
                         : "r"(taddr));
            asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                         : "=r"(cr[0]), "=r"(cr[1]), "=r"(cr[2]), "=r"(cr[3]), "=r"(cr[4]), "=r"(cr[5]), "=r"(cr[6]), "=r"(cr[7])
                         : "r"(taddr + (uint32_t)TN));
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
            if (i < a.S) {
#pragma unroll
                for (int e = 0; e < 8; e++) {
                    const size_t j = j0 + c0 + e;
                    if (j >= a.m) break;
                    float v = __uint_as_float(mn[e]) + __uint_as_float(cr[e]);
                    if (a.has_epi) {
                        const double arg = fma((double)v, sc, sh);
                        const double q = rint(arg * 0.15915494309189535);
                        v = osc * __cosf((float)fma(q, -6.283185307179586, arg));     // |r| <= pi: MUFU.COS is good to 4e-7 there
                    }
                    a.OUT[i + j * a.ldo] = v;
                }
            }
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    __syncwarp();
    if (warp == 4) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(tm::TMEM_COLS));
}

// ---------------------------------------------------------------------------
// K5d, persistent form: one CTA per SM walks the tile list.  Roles: warps 0-3 epilogue, warps
// 4-7 lo producers, warp 8 MMA issuer, warp 9 TMA issuer.  The operand ring (6 stages) runs
// across tile boundaries and TMEM holds two accumulator pairs, so the MMAs of tile t+1 start
// while the epilogue warps drain tile t.
// ---------------------------------------------------------------------------
namespace tp {
constexpr int TM = 128, TN = 128, TK = 16, NSTAGE = 6;
constexpr int EPI = 128, PRODUCERS = 128, THREADS = EPI + PRODUCERS + 64;
constexpr uint32_t TILE_BYTES = TM * TK * 4, STAGE_BYTES = 4 * TILE_BYTES;
constexpr size_t SMEM = (size_t)NSTAGE * STAGE_BYTES + 1024 + 256;
constexpr uint32_t TMEM_COLS = 512;            // 2 x (main 128 + cross 128)
constexpr int CHUNK = 16;                      // k-blocks per accumulation chunk: the truncating chain stays at 32 MMAs
}  // namespace tp

__device__ __forceinline__ void tp_tile(const TmArgs &a, long long t, size_t &i0, size_t &j0) {
    constexpr int GI = 16;
    const long long per = (long long)GI * a.ntn, grp = t / per, tp_ = t - grp * per;
    const int gi = (int)min((long long)GI, (long long)a.ntm - grp * GI);
    j0 = (size_t)(tp_ / gi) * tp::TN; i0 = (size_t)(grp * GI + tp_ % gi) * tp::TM;
}

__global__ void __launch_bounds__(tp::THREADS, 1)
k_dense_gemm_tc_tma_p(const __grid_constant__ CUtensorMap tmP, const __grid_constant__ CUtensorMap tmX, TmArgs a,
                      long long ntiles) {
    using namespace tp;
    extern __shared__ unsigned char smem_raw[];
    const uint32_t sbase = ((uint32_t)__cvta_generic_to_shared(smem_raw) + 1023u) & ~1023u;
    const uint32_t bars = sbase + NSTAGE * STAGE_BYTES;   // tma[6], full[6], empty[6], accfull[2], accempty[2]
    __shared__ uint32_t tmem_slot;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int nk = a.nk;
    auto tma_bar = [&](int s) { return bars + 8u * s; };
    auto full_bar = [&](int s) { return bars + 8u * (NSTAGE + s); };
    auto empty_bar = [&](int s) { return bars + 8u * (2 * NSTAGE + s); };
    auto accfull_bar = [&](int b) { return bars + 8u * (3 * NSTAGE + b); };
    auto accempty_bar = [&](int b) { return bars + 8u * (3 * NSTAGE + 2 + b); };

    if (tid == 0) {
        for (int s = 0; s < NSTAGE; s++) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(tma_bar(s)));
            asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(full_bar(s)), "r"(PRODUCERS));
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(empty_bar(s)));
        }
        for (int b = 0; b < 2; b++) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(accfull_bar(b)));
            asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(accempty_bar(b)), "r"(EPI));
        }
        asm volatile("fence.mbarrier_init.release.cluster;");
    }
    if (warp == 8) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                     ::"r"((uint32_t)__cvta_generic_to_shared(&tmem_slot)), "n"(TMEM_COLS));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;");
    const uint32_t tmem = tmem_slot;

    if (warp == 9) {
        if (lane == 0) {
            // ===== TMA issuer: g counts stages over all tiles of this CTA =====
            long long g = 0;
            for (long long t = blockIdx.x; t < ntiles; t += gridDim.x) {
                size_t i0, j0; tp_tile(a, t, i0, j0);
                for (int kb = 0; kb < nk; kb++, g++) {
                    const int s = (int)(g % NSTAGE);
                    if (g >= NSTAGE) mbar_wait(empty_bar(s), (uint32_t)((g / NSTAGE - 1) & 1));
                    const uint32_t st0 = sbase + (uint32_t)s * STAGE_BYTES;
                    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(tma_bar(s)), "r"(2 * TILE_BYTES) : "memory");
                    tma_load_2d(st0, &tmP, kb * TK, (int)i0, tma_bar(s));
                    tma_load_2d(st0 + 2 * TILE_BYTES, &tmX, kb * TK, (int)j0, tma_bar(s));
                }
            }
        }
    } else if (warp == 8) {
        if (lane == 0) {
            // ===== MMA issuer =====
            const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(TN >> 3) << 17) | ((uint32_t)(TM >> 4) << 24);
            long long g = 0; int u = 0;                 // u counts (tile, chunk) units; accumulator pair = u & 1
            for (long long t = blockIdx.x; t < ntiles; t += gridDim.x) {
                for (int k0 = 0; k0 < nk; k0 += CHUNK, u++) {
                    const int buf = u & 1;
                    if (u >= 2) mbar_wait(accempty_bar(buf), (uint32_t)(((u >> 1) - 1) & 1));   // epilogue drained this pair
                    asm volatile("tcgen05.fence::after_thread_sync;");
                    const uint32_t t_main = tmem + (uint32_t)buf * 256, t_cross = t_main + (uint32_t)TN;
                    const int k1 = min(nk, k0 + CHUNK);
                    for (int kb = k0; kb < k1; kb++, g++) {
                        const int s = (int)(g % NSTAGE);
                        mbar_wait(full_bar(s), (uint32_t)((g / NSTAGE) & 1));
                        asm volatile("tcgen05.fence::after_thread_sync;");
                        const uint32_t st0 = sbase + (uint32_t)s * STAGE_BYTES;
#pragma unroll
                        for (int ks = 0; ks < TK / 8; ks++) {
                            const uint32_t ko = (uint32_t)ks * 32;
                            const uint64_t phi = tm_desc(st0 + ko), plo = tm_desc(st0 + TILE_BYTES + ko);
                            const uint64_t xhi = tm_desc(st0 + 2 * TILE_BYTES + ko), xlo = tm_desc(st0 + 3 * TILE_BYTES + ko);
                            const uint32_t first = (kb == k0 && ks == 0) ? 0u : 1u;
                            tc_mma(t_cross, plo, xhi, idesc, first);
                            tc_mma(t_cross, phi, xlo, idesc, 1u);
                            tc_mma(t_main, phi, xhi, idesc, first);
                        }
                        tc_commit(empty_bar(s));
                    }
                    tc_commit(accfull_bar(buf));
                }
            }
        }
    } else if (warp >= 4) {
        // ===== lo producers (warps 4-7) =====
        const int pt = tid - EPI;
        long long g = 0;
        for (long long t = blockIdx.x; t < ntiles; t += gridDim.x) {
            for (int kb = 0; kb < nk; kb++, g++) {
                const int s = (int)(g % NSTAGE);
                if (lane == 0) mbar_wait(tma_bar(s), (uint32_t)((g / NSTAGE) & 1));
                __syncwarp();
                const uint32_t st0 = sbase + (uint32_t)s * STAGE_BYTES;
#pragma unroll
                for (int tl = 0; tl < 2; tl++) {
                    const uint32_t raw = st0 + (uint32_t)tl * 2 * TILE_BYTES, lo = raw + TILE_BYTES;
#pragma unroll
                    for (int c = 0; c < 4; c++) {
                        const uint32_t off = (uint32_t)(pt + c * PRODUCERS) * 16;
                        float4 v, l;
                        asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(raw + off));
                        l.x = v.x - __uint_as_float(__float_as_uint(v.x) & 0xffffe000u);
                        l.y = v.y - __uint_as_float(__float_as_uint(v.y) & 0xffffe000u);
                        l.z = v.z - __uint_as_float(__float_as_uint(v.z) & 0xffffe000u);
                        l.w = v.w - __uint_as_float(__float_as_uint(v.w) & 0xffffe000u);
                        asm volatile("st.shared.v4.f32 [%0], {%1,%2,%3,%4};" ::"r"(lo + off), "f"(l.x), "f"(l.y), "f"(l.z), "f"(l.w) : "memory");
                    }
                }
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                mbar_arrive(full_bar(s));
            }
        }
    } else {
        // ===== epilogue (warps 0-3): warp w owns TMEM lanes [32w, 32w+32) = rows i =====
        const float osc = (float)a.outscale;
        int u = 0;
        for (long long t = blockIdx.x; t < ntiles; t += gridDim.x) {
            size_t i0, j0; tp_tile(a, t, i0, j0);
            float acc[TN];                      // running sums in registers, added with round-to-nearest
#pragma unroll
            for (int c = 0; c < TN; c++) acc[c] = 0.f;
            for (int k0 = 0; k0 < nk; k0 += CHUNK, u++) {
                const int buf = u & 1;
                if (lane == 0) mbar_wait(accfull_bar(buf), (uint32_t)((u >> 1) & 1));
                __syncwarp();
                asm volatile("tcgen05.fence::after_thread_sync;");
                const uint32_t tb = tmem + (uint32_t)buf * 256 + ((uint32_t)(warp * 32) << 16);
#pragma unroll
                for (int c0 = 0; c0 < TN; c0 += 8) {
                    uint32_t mn[8], cr[8];
                    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                                 : "=r"(mn[0]), "=r"(mn[1]), "=r"(mn[2]), "=r"(mn[3]), "=r"(mn[4]), "=r"(mn[5]), "=r"(mn[6]), "=r"(mn[7])
                                 : "r"(tb + (uint32_t)c0));
                    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                                 : "=r"(cr[0]), "=r"(cr[1]), "=r"(cr[2]), "=r"(cr[3]), "=r"(cr[4]), "=r"(cr[5]), "=r"(cr[6]), "=r"(cr[7])
                                 : "r"(tb + (uint32_t)TN + (uint32_t)c0));
                    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
                    for (int e = 0; e < 8; e++) acc[c0 + e] += __uint_as_float(mn[e]) + __uint_as_float(cr[e]);
                }
                asm volatile("tcgen05.fence::before_thread_sync;");
                mbar_arrive(accempty_bar(buf));             // this accumulator pair may be overwritten
            }
            const size_t i = i0 + (size_t)tid;
            if (i < a.S) {
                double sc = 1.0, sh = 0.0;
                if (a.has_epi) { sc = a.scales[i]; sh = a.shifts[i]; }
#pragma unroll
                for (int c = 0; c < TN; c++) {
                    const size_t j = j0 + c;
                    if (j < a.m) {
                        float v = acc[c];
                        if (a.has_epi) {
                            const double arg = fma((double)v, sc, sh);
                            const double q = rint(arg * 0.15915494309189535);
                            v = osc * __cosf((float)fma(q, -6.283185307179586, arg));
                        }
                        a.OUT[i + j * a.ldo] = v;
                    }
                }
            }
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    __syncwarp();
    if (warp == 8) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(tp::TMEM_COLS));
}

// ---------------------------------------------------------------------------
// K5d, wide form (the default): 128(i) x 256(j) tiles.  Shared-memory bandwidth is what bounds
// the 128 x 128 forms: three MMAs per k8 each read both operands again (48 KB per k-block of 16)
// on top of the TMA writes (16 KB) and the lo pass (32 KB): 200 bytes per clock wanted, 128
// available.  With N = 256 a panel tile serves twice the work, the panel's lo tile comes
// precomputed through TMA (the panel is small and shared by every point tile), and the producer
// warps compute only X's lo tile: about 140 bytes per clock at the full MMA rate.  TMEM holds one
// accumulator pair (main 256 + cross 256 columns), so the epilogue of a tile is not overlapped
// (about a tenth of the tile time).  Stage = P raw 8 KB, P lo 8 KB, X raw 16 KB, X lo 16 KB; ring of 4.
// ---------------------------------------------------------------------------
namespace tw {
constexpr int TM = 128, TN = 256, TK = 16, NSTAGE = 4;
#ifndef SKB_TW_PRODUCERS
#define SKB_TW_PRODUCERS 128
#endif
constexpr int EPI = 128, PRODUCERS = SKB_TW_PRODUCERS, THREADS = EPI + PRODUCERS + 64;
constexpr int WARP_MMA = (EPI + PRODUCERS) / 32, WARP_TMA = WARP_MMA + 1;
constexpr uint32_t PT_BYTES = TM * TK * 4, XT_BYTES = TN * TK * 4;     // 8 KB, 16 KB
constexpr uint32_t STAGE_BYTES = 2 * PT_BYTES + 2 * XT_BYTES;          // 48 KB
constexpr uint32_t OFF_PLO = PT_BYTES, OFF_X = 2 * PT_BYTES, OFF_XLO = 2 * PT_BYTES + XT_BYTES;
constexpr size_t SMEM = (size_t)NSTAGE * STAGE_BYTES + 1024 + 256;
constexpr uint32_t TMEM_COLS = 512;            // main 256 + cross 256
}  // namespace tw

__device__ __forceinline__ void tw_tile(const TmArgs &a, long long t, size_t &i0, size_t &j0) {
    constexpr int GI = 16;
    const long long per = (long long)GI * a.ntn, grp = t / per, tp_ = t - grp * per;
    const int gi = (int)min((long long)GI, (long long)a.ntm - grp * GI);
    j0 = (size_t)(tp_ / gi) * tw::TN; i0 = (size_t)(grp * GI + tp_ % gi) * tw::TM;
}

__global__ void __launch_bounds__(tw::THREADS, 1)
k_dense_gemm_tc_tma_w(const __grid_constant__ CUtensorMap tmP, const __grid_constant__ CUtensorMap tmPlo,
                      const __grid_constant__ CUtensorMap tmX, TmArgs a,
                      long long ntiles) {
    using namespace tw;
    extern __shared__ unsigned char smem_raw[];
    const uint32_t sbase = ((uint32_t)__cvta_generic_to_shared(smem_raw) + 1023u) & ~1023u;
    const uint32_t bars = sbase + NSTAGE * STAGE_BYTES;   // tma[6], full[6], empty[6], accfull[2], accempty[2]
    __shared__ uint32_t tmem_slot;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int nk = a.nk;
    auto tma_bar = [&](int s) { return bars + 8u * s; };
    auto full_bar = [&](int s) { return bars + 8u * (NSTAGE + s); };
    auto empty_bar = [&](int s) { return bars + 8u * (2 * NSTAGE + s); };
    auto accfull_bar = [&](int b) { return bars + 8u * (3 * NSTAGE + b); };
    auto accempty_bar = [&](int b) { return bars + 8u * (3 * NSTAGE + 2 + b); };

    if (tid == 0) {
        for (int s = 0; s < NSTAGE; s++) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(tma_bar(s)));
            asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(full_bar(s)), "r"(PRODUCERS));
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(empty_bar(s)));
        }
        for (int b = 0; b < 2; b++) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(accfull_bar(b)));
            asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(accempty_bar(b)), "r"(EPI + PRODUCERS));
        }
        asm volatile("fence.mbarrier_init.release.cluster;");
    }
    if (warp == WARP_MMA) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                     ::"r"((uint32_t)__cvta_generic_to_shared(&tmem_slot)), "n"(TMEM_COLS));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;");
    const uint32_t tmem = tmem_slot;

    // epilogue of one tile for 128 of its 256 columns; warp w reads TMEM lanes [32(w%4), +32) = rows i
    auto drain = [&](long long t, int it, int cbeg) {
        const float osc = (float)a.outscale;
        size_t i0, j0; tw_tile(a, t, i0, j0);
        if (lane == 0) mbar_wait(accfull_bar(0), (uint32_t)(it & 1));
        __syncwarp();
        asm volatile("tcgen05.fence::after_thread_sync;");
        const size_t i = i0 + (size_t)((warp & 3) * 32 + lane);
        float scf = 1.f, shf = 0.f;
        if (a.has_epi && i < a.S) { scf = (float)a.scales[i]; shf = (float)a.shifts[i]; }
        const uint32_t tb = tmem + ((uint32_t)((warp & 3) * 32) << 16);
        for (int c0 = cbeg; c0 < cbeg + 128; c0 += 8) {
            uint32_t mn[8], cr[8];
            asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                         : "=r"(mn[0]), "=r"(mn[1]), "=r"(mn[2]), "=r"(mn[3]), "=r"(mn[4]), "=r"(mn[5]), "=r"(mn[6]), "=r"(mn[7])
                         : "r"(tb + (uint32_t)c0));
            asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                         : "=r"(cr[0]), "=r"(cr[1]), "=r"(cr[2]), "=r"(cr[3]), "=r"(cr[4]), "=r"(cr[5]), "=r"(cr[6]), "=r"(cr[7])
                         : "r"(tb + (uint32_t)TN + (uint32_t)c0));
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
            if (i < a.S) {
#pragma unroll
                for (int e = 0; e < 8; e++) {
                    const size_t j = j0 + c0 + e;
                    if (j >= a.m) break;
                    float v = __uint_as_float(mn[e]) + __uint_as_float(cr[e]);
                    if (a.has_epi) {       // fp32 throughout, like the accumulator: two-constant reduction to [-pi, pi]
                        const float arg = fmaf(v, scf, shf);
                        const float q = rintf(arg * 0.15915494f);
                        float r = fmaf(q, -6.2831855f, arg);
                        r = fmaf(q, 1.7484555e-7f, r);            // 2*pi = 6.2831855f - 1.7484555e-7
                        v = osc * __cosf(r);
                    }
                    a.OUT[i + j * a.ldo] = v;
                }
            }
        }
        asm volatile("tcgen05.fence::before_thread_sync;");
        mbar_arrive(accempty_bar(0));             // the accumulators may be overwritten once all 256 arrive
    };

    if (warp == WARP_TMA) {
        if (lane == 0) {
            // ===== TMA issuer: g counts stages over all tiles of this CTA =====
            long long g = 0;
            for (long long t = blockIdx.x; t < ntiles; t += gridDim.x) {
                size_t i0, j0; tw_tile(a, t, i0, j0);
                for (int kb = 0; kb < nk; kb++, g++) {
                    const int s = (int)(g % NSTAGE);
                    if (g >= NSTAGE) mbar_wait(empty_bar(s), (uint32_t)((g / NSTAGE - 1) & 1));
                    const uint32_t st0 = sbase + (uint32_t)s * STAGE_BYTES;
                    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(tma_bar(s)), "r"(2 * PT_BYTES + XT_BYTES) : "memory");
                    tma_load_2d(st0, &tmP, kb * TK, (int)i0, tma_bar(s));
                    tma_load_2d(st0 + OFF_PLO, &tmPlo, kb * TK, (int)i0, tma_bar(s));
                    tma_load_2d(st0 + OFF_X, &tmX, kb * TK, (int)j0, tma_bar(s));
                }
            }
        }
    } else if (warp == WARP_MMA) {
        if (lane == 0) {
            // ===== MMA issuer =====
            const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(TN >> 3) << 17) | ((uint32_t)(TM >> 4) << 24);
            long long g = 0; int it = 0;
            for (long long t = blockIdx.x; t < ntiles; t += gridDim.x, it++) {
                const int buf = 0;
                if (it >= 1) mbar_wait(accempty_bar(buf), (uint32_t)((it - 1) & 1));          // epilogue drained the accumulators
                asm volatile("tcgen05.fence::after_thread_sync;");
                const uint32_t t_main = tmem, t_cross = tmem + (uint32_t)TN;
                for (int kb = 0; kb < nk; kb++, g++) {
                    const int s = (int)(g % NSTAGE);
                    mbar_wait(full_bar(s), (uint32_t)((g / NSTAGE) & 1));
                    asm volatile("tcgen05.fence::after_thread_sync;");
                    const uint32_t st0 = sbase + (uint32_t)s * STAGE_BYTES;
#pragma unroll
                    for (int ks = 0; ks < TK / 8; ks++) {
                        const uint32_t ko = (uint32_t)ks * 32;
                        const uint64_t phi = tm_desc(st0 + ko), plo = tm_desc(st0 + OFF_PLO + ko);
                        const uint64_t xhi = tm_desc(st0 + OFF_X + ko), xlo = tm_desc(st0 + OFF_XLO + ko);
                        const uint32_t first = (kb | ks) == 0 ? 0u : 1u;
                        tc_mma(t_cross, plo, xhi, idesc, first);
                        tc_mma(t_cross, phi, xlo, idesc, 1u);
                        tc_mma(t_main, phi, xhi, idesc, first);
                    }
                    tc_commit(empty_bar(s));
                }
                tc_commit(accfull_bar(buf));
            }
        }
    } else if (warp >= 4) {
        // ===== lo producers (warps 4 .. WARP_MMA-1) =====
        const int pt = tid - EPI;
        long long g = 0; int pit = 0;
        for (long long t = blockIdx.x; t < ntiles; t += gridDim.x, pit++) {
            for (int kb = 0; kb < nk; kb++, g++) {
                const int s = (int)(g % NSTAGE);
                if (lane == 0) mbar_wait(tma_bar(s), (uint32_t)((g / NSTAGE) & 1));
                __syncwarp();
                const uint32_t st0 = sbase + (uint32_t)s * STAGE_BYTES;
                const uint32_t raw = st0 + OFF_X, lo = st0 + OFF_XLO;
#pragma unroll
                for (int c = 0; c < (int)(XT_BYTES / 16) / PRODUCERS; c++) {
                    const uint32_t off = (uint32_t)(pt + c * PRODUCERS) * 16;
                    float4 v, l;
                    asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(raw + off));
                    l.x = v.x - __uint_as_float(__float_as_uint(v.x) & 0xffffe000u);
                    l.y = v.y - __uint_as_float(__float_as_uint(v.y) & 0xffffe000u);
                    l.z = v.z - __uint_as_float(__float_as_uint(v.z) & 0xffffe000u);
                    l.w = v.w - __uint_as_float(__float_as_uint(v.w) & 0xffffe000u);
                    asm volatile("st.shared.v4.f32 [%0], {%1,%2,%3,%4};" ::"r"(lo + off), "f"(l.x), "f"(l.y), "f"(l.z), "f"(l.w) : "memory");
                }
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                mbar_arrive(full_bar(s));
            }
            drain(t, pit, 128);          // the MMAs of the next tile cannot start before the epilogue anyway
        }
    } else {
        // ===== epilogue (warps 0-3): columns [0, 128) of every tile; the lo producers take [128, 256) =====
        int it = 0;
        for (long long t = blockIdx.x; t < ntiles; t += gridDim.x, it++) drain(t, it, 0);
    }
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    __syncwarp();
    if (warp == WARP_MMA) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(tw::TMEM_COLS));
}

static bool tc_enabled() {      // SKB_DEBUG_NO_TCGEN05=1 falls back to the mma.sync kernel
    static const bool on = getenv("SKB_DEBUG_NO_TCGEN05") == nullptr;
    return on;
}
static size_t tc_kp(size_t K) { return (K + tc::TK - 1) / tc::TK * tc::TK; }
static size_t tc_workspace_bytes(const DenseParams &p, size_t nloc, size_t m) {
    auto al = [](size_t b) { return (b + 255) / 256 * 256; };
    (void)m;
    return al(nloc * p.S * sizeof(double)) + 2 * al(p.S * tc_kp(nloc) * 4) + 256;     // panel, Pt, Pt_lo
}

static bool dense_apply_tc_tf32x3(const DenseParams &p, const float *X, size_t ldx, size_t m, size_t koff,
                                  size_t nloc, float *OUT, size_t ldo, const RftEpilogue &epi,
                                  void *workspace, size_t workspace_bytes, cudaStream_t st) {
    // the accumulation chain of this kernel grows with k (one truncating add per 8): 3.4e-6 at
    // k = 784; longer contractions stay on the mma.sync kernel, whose error does not grow
    if (!tc_enabled() || nloc > 1024 || !dense_tf32_path_ok(p, X, ldx, m, nloc)) return false;
    if (workspace_bytes < tc_workspace_bytes(p, nloc, m)) return false;
    auto al = [](size_t b) { return (b + 255) / 256 * 256; };
    const size_t S = p.S, Kp = tc_kp(nloc);
    char *w = reinterpret_cast<char *>(workspace);
    double *P = reinterpret_cast<double *>(w); w += al(nloc * S * sizeof(double));
    float *Pt = reinterpret_cast<float *>(w);
    dense_generate_panel(p, koff, nloc, P, st);
    {
        LaunchScope ls(KC_GEN, st);
        dim3 grid((unsigned)((S + 31) / 32), (unsigned)((Kp + 31) / 32));
        k_tc_transpose_panel<<<grid, dim3(32, 8), 0, st>>>(P, S, nloc, Kp, Pt);
        SKB_CUDA(cudaGetLastError());
    }
    TcArgs a;
    a.Pt = Pt; a.X = X; a.OUT = OUT;
    a.S = S; a.m = m; a.K = nloc; a.Kp = Kp; a.ldx = ldx; a.ldo = ldo;
    a.scales = epi.scales; a.shifts = epi.shifts; a.outscale = epi.outscale;
    a.has_epi = epi.shifts != nullptr ? 1 : 0;
    const long long ntm = (long long)((S + tc::TM - 1) / tc::TM), ntn = (long long)((m + tc::TN - 1) / tc::TN);
    a.ntm = (int)ntm; a.ntn = (int)ntn;
    SKB_CUDA(cudaFuncSetAttribute(k_dense_gemm_tc_tf32x3, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tc::SMEM));
    LaunchScope ls(KC_GEMM, st);
    k_dense_gemm_tc_tf32x3<<<(unsigned)(ntm * ntn), tc::THREADS, tc::SMEM, st>>>(a);
    SKB_CUDA(cudaGetLastError());
    return true;
}

static bool dense_apply_tc_tma(const DenseParams &p, const float *X, size_t ldx, size_t m, size_t koff,
                               size_t nloc, float *OUT, size_t ldo, const RftEpilogue &epi,
                               void *workspace, size_t workspace_bytes, cudaStream_t st) {
    static const bool off = getenv("SKB_DEBUG_NO_TMA") != nullptr;
    if (off || !tc_enabled() || !dense_tf32_path_ok(p, X, ldx, m, nloc)) return false;
    if (workspace_bytes < tc_workspace_bytes(p, nloc, m) || !tma_encoder()) return false;
    auto al = [](size_t b) { return (b + 255) / 256 * 256; };
    const size_t S = p.S, Kp = tc_kp(nloc);
    char *w = reinterpret_cast<char *>(workspace);
    double *P = reinterpret_cast<double *>(w); w += al(nloc * S * sizeof(double));
    float *Pt = reinterpret_cast<float *>(w); w += al(S * Kp * 4);
    float *Pt_lo = reinterpret_cast<float *>(w);
    // form: wide = 128 x 256 tiles (default), persist / plain = the 128 x 128 kernels
    // the wide form keeps one accumulation chain per tile (error grows with k: 3.4e-6 at k = 784), the
    // persistent form drains its accumulators every 16 k-blocks (1.4e-6, independent of k): short
    // contractions take the faster one, long ones the chunked one
    static const char *form_env = getenv("SKB_TC_FORM");
    const bool wide = nloc <= 1024 && (!form_env || form_env[0] == 'w');
    const bool persist = nloc > 1024 || !form_env || form_env[0] != 'n';      // "nonpersistent"
    CUtensorMap tmP, tmPlo, tmX;
    if (!tma_make_map(&tmP, Pt, S, Kp, Kp) || !tma_make_map(&tmPlo, Pt_lo, S, Kp, Kp) ||
        !tma_make_map(&tmX, X, m, nloc, ldx, wide ? 256 : 128)) return false;
    dense_generate_panel(p, koff, nloc, P, st);
    {
        LaunchScope ls(KC_GEN, st);
        dim3 grid((unsigned)((S + 31) / 32), (unsigned)((Kp + 31) / 32));
        k_tc_transpose_panel<<<grid, dim3(32, 8), 0, st>>>(P, S, nloc, Kp, Pt, wide ? Pt_lo : nullptr);
        SKB_CUDA(cudaGetLastError());
    }
    TmArgs a;
    a.OUT = OUT; a.S = S; a.m = m; a.ldo = ldo;
    a.scales = epi.scales; a.shifts = epi.shifts; a.outscale = epi.outscale;
    a.has_epi = epi.shifts != nullptr ? 1 : 0;
    const int tn = wide ? tw::TN : tm::TN;
    const long long ntm = (long long)((S + tm::TM - 1) / tm::TM), ntn = (long long)((m + tn - 1) / tn);
    a.ntm = (int)ntm; a.ntn = (int)ntn; a.nk = (int)(Kp / tm::TK);
    const long long ntiles = ntm * ntn;
    LaunchScope ls(KC_GEMM, st);
    if (wide) {
        SKB_CUDA(cudaFuncSetAttribute(k_dense_gemm_tc_tma_w, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tw::SMEM));
        k_dense_gemm_tc_tma_w<<<(unsigned)std::min<long long>(ntiles, 148), tw::THREADS, tw::SMEM, st>>>(tmP, tmPlo, tmX, a, ntiles);
    } else if (persist) {
        SKB_CUDA(cudaFuncSetAttribute(k_dense_gemm_tc_tma_p, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tp::SMEM));
        k_dense_gemm_tc_tma_p<<<(unsigned)std::min<long long>(ntiles, 148), tp::THREADS, tp::SMEM, st>>>(tmP, tmX, a, ntiles);
    } else {
        SKB_CUDA(cudaFuncSetAttribute(k_dense_gemm_tc_tma, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tm::SMEM));
        k_dense_gemm_tc_tma<<<(unsigned)ntiles, tm::THREADS, tm::SMEM, st>>>(tmP, tmX, a);
    }
    SKB_CUDA(cudaGetLastError());
    return true;
}

void dense_apply(const DenseParams &p, bool layout_rows, bool f32, const void *X, size_t ldx,
                 size_t m, size_t koff, size_t nloc, void *OUT, size_t ldo,
                 const RftEpilogue &epi, void *workspace, size_t workspace_bytes,
                 cudaStream_t st, bool accumulate) {
    if (f32) {
        if (!layout_rows && !accumulate &&
            dense_apply_tc_tma(p, (const float *)X, ldx, m, koff, nloc, (float *)OUT, ldo, epi, workspace, workspace_bytes, st))
            return;
        if (!layout_rows && !accumulate &&
            dense_apply_tc_tf32x3(p, (const float *)X, ldx, m, koff, nloc, (float *)OUT, ldo, epi, workspace, workspace_bytes, st))
            return;
        if (!layout_rows && !accumulate &&
            dense_apply_tf32x3(p, (const float *)X, ldx, m, koff, nloc, (float *)OUT, ldo, epi, workspace, workspace_bytes, st))
            return;
        if (layout_rows) dense_apply_t<float, true>(p, (const float *)X, ldx, m, koff, nloc, (float *)OUT, ldo, epi, workspace, workspace_bytes, st, accumulate);
        else dense_apply_t<float, false>(p, (const float *)X, ldx, m, koff, nloc, (float *)OUT, ldo, epi, workspace, workspace_bytes, st, accumulate);
    } else {
        if (layout_rows) dense_apply_t<double, true>(p, (const double *)X, ldx, m, koff, nloc, (double *)OUT, ldo, epi, workspace, workspace_bytes, st, accumulate);
        else dense_apply_t<double, false>(p, (const double *)X, ldx, m, koff, nloc, (double *)OUT, ldo, epi, workspace, workspace_bytes, st, accumulate);
    }
}


// ---------------------------------------------------------------------------
// Allocation-free dense sketch (sl_dense_sketch_apply_noalloc, include/skylark_b200.h).
//
// For callers that are not allowed to allocate device memory while they run -- a
// real-time loop, a process whose memory pool is owned by another runtime and sized up
// front (a late cudaMalloc may fail there), a plugin host that forbids cudaMalloc after
// start-up.
// The generated panel of S lives in a fixed file-scope device array instead of a
// cudaMalloc'ed workspace.  Every GPU that loads this library carries these 32 MiB until
// the process exits.
// ---------------------------------------------------------------------------
static constexpr size_t NOALLOC_PANEL_DOUBLES = (size_t)4 << 20;        // 32 MiB
static __device__ __align__(256) double g_noalloc_panel[NOALLOC_PANEL_DOUBLES];

// In host code `g_noalloc_panel` is only a handle for the runtime, not a device address;
// the address each device gave the array is looked up once per device and kept here.  Two host
// threads making their first call on a device together both look it up and store the same
// value, so the cache needs no lock.
static double *g_noalloc_panel_addr[64] = {nullptr};

static double *noalloc_panel_for_current_device() {
    int dev = 0;
    SKB_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) throw error(SKB_ERR_GENERIC, "device index out of range");
    if (!g_noalloc_panel_addr[dev]) {
        double *ptr = nullptr;
        SKB_CUDA(cudaGetSymbolAddress((void **)&ptr, g_noalloc_panel));
        g_noalloc_panel_addr[dev] = ptr;
    }
    return g_noalloc_panel_addr[dev];
}

template <bool ROWS>
static void noalloc_apply(const DenseParams &p, const double *X, size_t ldx, size_t m, double *OUT,
                          size_t ldo, cudaStream_t st) {
    const size_t S = p.S, N = p.N;
    if (S == 0 || m == 0) return;
    size_t kc_max = NOALLOC_PANEL_DOUBLES / S / BK * BK;
    if (kc_max == 0)
        throw error(SKB_ERR_INVALID_PARAM,
                    "sl_dense_sketch_apply_noalloc: s = " + std::to_string(S) + " exceeds the "
                    "built-in panel (" + std::to_string(NOALLOC_PANEL_DOUBLES / BK) +
                    " is the largest sketch size it can hold); use sl_apply_sketch_transform");
    double *panel = noalloc_panel_for_current_device();
    // The panel keeps whatever the previous call left in it: every element the GEMM reads
    // (rows < kc) is rewritten by the generator first, rows beyond kc are never read.  The
    // output is accumulated with atomics, so it is the one thing that must start at zero.
    {
        size_t contig = ROWS ? m : S, nvec = ROWS ? S : m;
        if (ldo == contig) SKB_CUDA(cudaMemsetAsync(OUT, 0, contig * nvec * sizeof(double), st));
        else SKB_CUDA(cudaMemset2DAsync(OUT, ldo * sizeof(double), 0, contig * sizeof(double), nvec, st));
    }
    constexpr int BN = GemmCfg<double>::BN;
    const int ntm = (int)((S + BM - 1) / BM), ntn = (int)((m + BN - 1) / BN);
    const long long ntiles = (long long)ntm * ntn;
    const bool al16p = (S % 2 == 0);
    const bool alx = (ldx % 2 == 0) && ((uintptr_t)X % 16 == 0);
    const uint64_t key = seed_to_key(p.seed);
    for (size_t k0 = 0; k0 < N; k0 += kc_max) {
        const size_t kc = N - k0 < kc_max ? N - k0 : kc_max;
        const size_t count = kc * S;
        size_t blocks = (count + 255) / 256;
        if (blocks > (size_t)148 * 64) blocks = (size_t)148 * 64;
        profiler().launches[KC_GEN]++;
        if (p.dist == DD_NORMAL)
            k_gen_panel<DD_NORMAL><<<(unsigned)blocks, 128, 0, st>>>(key, p.base + (uint64_t)k0 * S, count, p.scale, panel);
        else
            k_gen_panel<DD_CAUCHY><<<(unsigned)blocks, 128, 0, st>>>(key, p.base + (uint64_t)k0 * S, count, p.scale, panel);
        SKB_CUDA(cudaGetLastError());
        GemmArgs a;
        a.P = panel; a.S = S; a.kc = kc;
        a.X = ROWS ? (const void *)(X + k0 * ldx) : (const void *)(X + k0);
        a.ldx = ldx; a.m = m; a.OUT = OUT; a.ldo = ldo;
        a.ntm = ntm; a.ntn = ntn; a.nk = (int)((kc + BK - 1) / BK);
        a.direct_ok = 0; a.whole_tiles = 0; a.out_is_acc = 0;
        a.scales = nullptr; a.shifts = nullptr; a.outscale = 1.0; a.has_epi = 0;
        const long long total = ntiles * a.nk;
        const long long ncta = 148LL * GemmCfg<double>::CTAS;
        const unsigned grid = (unsigned)(total < ncta ? total : ncta);
        const bool alx_p = alx && (ROWS || (k0 % 2 == 0));
        using SM = GemmSmem<double, ROWS>;
        profiler().launches[KC_GEMM]++;
#define SKB_NA_LAUNCH(CP, CX)                                                                      \
        do {                                                                                       \
            auto kern = k_dense_gemm<double, ROWS, CP, CX, GemmCfg<double>::WM, GemmCfg<double>::WN>; \
            SKB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,       \
                                          (int)SM::BYTES));                                        \
            kern<<<grid, 32 * GemmCfg<double>::WM * GemmCfg<double>::WN, SM::BYTES, st>>>(a);      \
        } while (0)
        if (al16p && alx_p) SKB_NA_LAUNCH(16, 16);
        else if (al16p) SKB_NA_LAUNCH(16, 8);
        else if (alx_p) SKB_NA_LAUNCH(8, 16);
        else SKB_NA_LAUNCH(8, 8);
#undef SKB_NA_LAUNCH
        SKB_CUDA(cudaGetLastError());
    }
}

}  // namespace skb

extern "C" {

// See include/skylark_b200.h for the contract.
int sl_dense_sketch_apply_noalloc(int seed, uint64_t counter_base, int dist, double scale,
                                  int64_t n, int64_t s, const double *A_dev, int64_t lda, int64_t m,
                                  double *SA_dev, int64_t ldsa, int dim, void *cuda_stream) {
    try {
        if (n <= 0 || s <= 0 || m < 0 || !A_dev || !SA_dev || (dist != 0 && dist != 1))
            throw skb::error(skb::SKB_ERR_INVALID_PARAM, "sl_dense_sketch_apply_noalloc: bad argument");
        if (dim != 1 && dim != 2)
            throw skb::error(skb::SKB_ERR_INVALID_PARAM, "sl_dense_sketch_apply_noalloc: dim must be 1 or 2");
        // column-major operands, as everywhere in this ABI
        const int64_t a_rows = dim == 1 ? n : m, sa_rows = dim == 1 ? s : m;
        if (lda < a_rows || ldsa < sa_rows)
            throw skb::error(skb::SKB_ERR_INVALID_PARAM, "sl_dense_sketch_apply_noalloc: leading dimension too small");
        // The built-in array that is used is the CURRENT device's; the operands must live there
        // too (a query, not an allocation).
        {
            int cur = 0;
            SKB_CUDA(cudaGetDevice(&cur));
            const void *ptrs[2] = {A_dev, SA_dev};
            for (const void *q : ptrs) {
                cudaPointerAttributes at;
                if (cudaPointerGetAttributes(&at, q) != cudaSuccess) { cudaGetLastError(); continue; }
                if (at.type == cudaMemoryTypeDevice && at.device != cur)
                    throw skb::error(skb::SKB_ERR_INVALID_USAGE,
                                     "sl_dense_sketch_apply_noalloc: A_dev / SA_dev belong to device " +
                                     std::to_string(at.device) + " but the current device is " +
                                     std::to_string(cur) + "; make the operands' device current first");
            }
        }
        skb::DenseParams p;
        p.seed = seed; p.base = counter_base; p.dist = dist; p.scale = scale;
        p.N = (size_t)n; p.S = (size_t)s;
        cudaStream_t st = (cudaStream_t)cuda_stream;
        if (dim == 2) skb::noalloc_apply<true>(p, A_dev, (size_t)lda, (size_t)m, SA_dev, (size_t)ldsa, st);
        else skb::noalloc_apply<false>(p, A_dev, (size_t)lda, (size_t)m, SA_dev, (size_t)ldsa, st);
    } catch (const skb::error &e) {
        skb::set_last_error(e.what());
        return e.code;
    } catch (const std::exception &e) {
        skb::set_last_error(e.what());
        return skb::SKB_ERR_GENERIC;
    }
    return 0;
}

}  // extern "C"
