// Internal (C++) interface between the C-ABI shim (capi.cu) and the kernel
// translation units.  Nothing here is exported; the public surface is
// include/skylark_b200.h.
#pragma once
#include <cstddef>
#include <cstdint>
#include <stdexcept>
#include <string>
#include <vector>
#include <cuda_runtime.h>

namespace skb {

// Error codes follow base/exception.hpp:9-24 of the reference.
enum {
    SKB_ERR_GENERIC = 100, SKB_ERR_MPI = 101, SKB_ERR_ELEMENTAL = 102,
    SKB_ERR_UNSUPPORTED_DIST = 103, SKB_ERR_COMBBLAS = 104, SKB_ERR_SKETCH = 105,
    SKB_ERR_NLA = 106, SKB_ERR_ALLOC = 107, SKB_ERR_RANDOM123 = 108,
    SKB_ERR_UNSUPPORTED_BASE = 109, SKB_ERR_IO = 110, SKB_ERR_INVALID_PARAM = 111,
    SKB_ERR_INVALID_USAGE = 112, SKB_ERR_ML = 113
};

struct error : public std::runtime_error {
    int code;
    error(int c, const std::string &m) : std::runtime_error(m), code(c) {}
};

#define SKB_CUDA(call)                                                              \
    do {                                                                            \
        cudaError_t e__ = (call);                                                   \
        if (e__ != cudaSuccess)                                                     \
            throw skb::error(e__ == cudaErrorMemoryAllocation ? skb::SKB_ERR_ALLOC  \
                                                              : skb::SKB_ERR_GENERIC, \
                             std::string("CUDA error: ") + cudaGetErrorString(e__) + \
                                 " at " __FILE__ ":" + std::to_string(__LINE__));   \
    } while (0)

// Kernel classes for launch counting / optional event timing.
enum KernelClass { KC_TABLES = 0, KC_HASH, KC_GEN, KC_GEMM, KC_FJLT, KC_MISC, KC_COUNT };

struct Profiler {
    bool enabled = false;
    unsigned long long launches[KC_COUNT] = {0};
    double ms[KC_COUNT] = {0};
};
Profiler &profiler();

// RAII bracket: counts one launch of class `kc`; when profiling is enabled it
// also times it with CUDA events on `stream` (resolved at profile read-out).
struct LaunchScope {
    LaunchScope(KernelClass kc, cudaStream_t stream);
    ~LaunchScope();
    KernelClass kc; cudaStream_t stream; cudaEvent_t e0, e1; bool timed;
};
void profiler_collect();   // synchronises and folds pending event pairs into ms[]
void set_last_error(const std::string &msg);   // message returned by sl_get_exception_info

// ---- hash transforms (hash.cu) ---------------------------------------------
enum HashValueKind { HV_RADEMACHER = 0, HV_CAUCHY = 1, HV_WZT = 2 };

// K0: bucket / value tables for stream indices [first, first+n).
void hash_build_tables(int seed, uint64_t idx_base, uint64_t val_base, uint64_t pm_base,
                       size_t first, size_t n, uint64_t S, int kind, double p, int order,
                       int32_t *bucket, double *value, cudaStream_t st);

// Rows of one bucket, split in chunks of <= 32 rows; built once per (transform, shard).
struct HashPlan {
    int32_t *rows = nullptr;         // n local row ids, stably sorted by bucket
    int32_t *chunk_start = nullptr;  // nchunks+1 offsets into rows
    int32_t *chunk_bucket = nullptr; // bucket id; bitwise-not (negative) when the bucket
                                     // spans several chunks (needs atomics)
    int nchunks = 0;
    size_t n = 0;
};
void hash_build_plan(const int32_t *bucket, size_t n, int S, HashPlan *plan, cudaStream_t st);
void hash_free_plan(HashPlan *plan);

// OUT[h(k)*ldo + j] += v(k) * X[k*ldx + j], j < m   (input index k has contiguous rows)
void hash_apply_rows(const double *X, size_t ldx, size_t m, const HashPlan &plan,
                     const double *value, double *OUT, size_t ldo, int S, cudaStream_t st);
// OUT[h(k) + j*ldo] += v(k) * X[k + j*ldx], j < m   (input index k strided by 1 inside a column)
void hash_apply_cols(const double *X, size_t ldx, size_t n, size_t m, const int32_t *bucket,
                     const double *value, double *OUT, size_t ldo, int S, cudaStream_t st);
// Plan of the strided layout (K2): sketched indices are cut in tiles of 2048; inside a tile the rows
// of warp w's buckets (w = bucket / SW) are arranged in rounds of 32 rows with 32 different buckets,
// each half warp on 16 different shared-memory bank pairs for the accumulators and for the tile, so
// a CTA accumulates in shared memory with plain loads and stores -- no atomics.  Built on the host
// once per (transform, shard) from the bucket / value tables.
struct HashColsPlan {
    uint32_t *entries = nullptr;   // rounds of 32 packed entries: k in tile << 3 | bucket in the warp's range << 16 | sign << 31
    double *vals = nullptr;        // value per entry; null when every value is +-1 (sign bit used)
    uint32_t *off = nullptr;       // first round of (warp, tile): 8 * (ntiles + 1) entries, warp-major streams
    int ntiles = 0, S = 0, SW = 0;
    size_t n = 0;
};
void hash_build_cols_plan(const int32_t *bucket, const double *value, size_t n, int S, HashColsPlan *plan,
                          cudaStream_t st);
void hash_free_cols_plan(HashColsPlan *plan);
// the host-only part of the plan (no CUDA): rounds of 32 packed entries per warp, warp-major, and the
// first round of every (warp, tile)
int hash_cols_plan_bucket_width(int S);
void hash_schedule_cols_plan(const int32_t *bucket_host, const double *value_host, size_t n, int S,
                             std::vector<uint32_t> &entries, std::vector<double> &vals, std::vector<uint32_t> &off,
                             bool *all_signs);
// false = shape or alignment outside the planned kernel's range (caller falls back to hash_apply_cols)
bool hash_apply_cols_planned(const double *X, size_t ldx, size_t n, size_t m, const HashColsPlan &plan,
                             double *OUT, size_t ldo, int S, cudaStream_t st);
// CSC input (rows are the sketched index): OUT[h(row) + col*ldo] += v(row)*val
void hash_apply_csc_columnwise(const int32_t *indptr, const int32_t *indices,
                               const double *values, size_t ncols, const int32_t *bucket,
                               const double *value, double *OUT, size_t ldo, int S,
                               cudaStream_t st, size_t nnz_hint = 0);      // nnz of the operand: picks the kernel form
void hash_set_csc_form(int bulk);        // 0: register dump (default), 1: bulk-store slices
// CSC input (columns are the sketched index): OUT[row + h(col)*ldo] += v(col)*val
void hash_apply_csc_rowwise(const int32_t *indptr, const int32_t *indices,
                            const double *values, size_t nrows, size_t ncols,
                            const int32_t *bucket, const double *value, double *OUT,
                            size_t ldo, int S, cudaStream_t st);

// CSC input -> CSC output (device arrays allocated here from the stream-ordered pool, owned by
// the caller, released with hash_sparse_free):
// columnwise gives S x ncols, rowwise nrows x S; rows inside a column come out sorted.
// Temporaries live in the grow-only workspace (*ws, *ws_bytes).  Stream-ordered: the result
// arrays are complete once work queued on st so far has finished.
void hash_apply_csc_sparse(const int32_t *indptr, const int32_t *indices, const double *values,
                           size_t nrows, size_t ncols, size_t nnz, bool rowwise,
                           const int32_t *bucket, const double *value, int S,
                           int32_t **out_indptr, int32_t **out_indices, double **out_values,
                           size_t *out_nnz, void **ws, size_t *ws_bytes, cudaStream_t st);
void hash_sparse_free(void *p, cudaStream_t st);
// Merge of nseg rank-major CSC pieces: column j = concatenation over s of entries
// [vptr[s*ncols + j], vptr[s*ncols + j + 1]); equal (row, column) entries are added in segment
// order, rows come out sorted.  false = a merged column is too long for the shared-memory sort.
bool hash_merge_segments(const int32_t *vptr, const int32_t *indices, const double *values, int nseg,
                         size_t ncols, size_t nrows, size_t nnz, int32_t **out_indptr, int32_t **out_indices,
                         double **out_values, size_t *out_nnz, void **ws, size_t *ws_bytes, cudaStream_t st);

// ---- dense transforms (dense.cu) -------------------------------------------
enum DenseDist { DD_NORMAL = 0, DD_CAUCHY = 1 };
struct DenseParams {
    int seed; uint64_t base; int dist; double scale;   // S[i,k] = scale*sample(base + k*S + i)
    size_t N, S;
};
struct RftEpilogue {          // out = outscale * cos(scales[i]*acc + shifts[i]); null = none
    const double *scales = nullptr, *shifts = nullptr; double outscale = 1.0;
};
// Panel P[kk*S + i] = S[i, k0+kk] for kk < kc  (i fastest, as the counter stream runs)
void dense_generate_panel(const DenseParams &p, size_t k0, size_t kc, double *panel,
                          cudaStream_t st);
// Workspace the apply needs (bytes) for a given shape.
size_t dense_workspace_bytes(const DenseParams &p, size_t nloc, bool layout_rows, bool f32,
                             size_t m, size_t ldo);
// fp32 operands in the strided layout whose contraction fits one panel take the 3xTF32 kernel,
// which needs only the panel (fp64 + fp32 copy) as workspace.
bool dense_tf32_path_ok(const DenseParams &p, const void *X, size_t ldx, size_t m, size_t nloc);
size_t dense_tf32_workspace_bytes(const DenseParams &p, size_t nloc, size_t m);
// layout_rows = true : X[k*ldx + j] (k-th input index is a contiguous row of m), OUT[i*ldo + j]
// layout_rows = false: X[k + j*ldx],                                             OUT[i + j*ldo]
// Sketches input indices [koff, koff+nloc) of the transform (row shard); accumulate=false
// overwrites OUT.  f32 selects float X / OUT (S is still generated in double).
void dense_apply(const DenseParams &p, bool layout_rows, bool f32, const void *X, size_t ldx,
                 size_t m, size_t koff, size_t nloc, void *OUT, size_t ldo,
                 const RftEpilogue &epi, void *workspace, size_t workspace_bytes,
                 cudaStream_t st, bool accumulate = false);   // accumulate: OUT += (fp64, no epilogue)
// K0 for tests: dense block S[i0:i0+h, j0:j0+w] column-major into out
void dense_export_block(const DenseParams &p, size_t i0, size_t j0, size_t h, size_t w,
                        double *out, cudaStream_t st);

// ---- FJLT (fjlt.cu) ----------------------------------------------------------
enum FutKind { FUT_DCT = 0, FUT_WHT = 1 };
void fjlt_build_tables(int seed, uint64_t base, size_t N, size_t S, int order, double *D,
                       int32_t *samples, uint32_t *Dbits, cudaStream_t st);
size_t fjlt_workspace_bytes(size_t N, size_t m, int fut, bool layout_rows, size_t S);
// layout_rows as above.  OUT[i, j] = sqrt(N/S) * FUT(scale * D .* X[:, j])[samples[i]]
void fjlt_apply(int fut, bool layout_rows, const double *X, size_t ldx, size_t N, size_t m,
                const double *D, const uint32_t *Dbits, const int32_t *samples, size_t S,
                const int32_t *wht_slot_sample, const int32_t *wht_slot_out,
                double *OUT, size_t ldo, void *workspace, size_t workspace_bytes,
                cudaStream_t st, size_t offset, size_t nloc,    // X holds rows [offset, offset+nloc)
                const void *dct_plan = nullptr);                // prebuilt tables of the fused DCT form
// Per-transform tables of the fused DCT-II kernel (twiddles, post factors, the bitmap of needed
// bins and its ranks): bytes (0 when N is outside the fused form's range) and the build.
size_t fjlt_dct_plan_bytes(size_t N, size_t S);
void fjlt_dct_plan_build(size_t N, size_t S, const int32_t *samples, void *buf, cudaStream_t st);
bool fjlt_shard_ok(int fut, size_t N, size_t offset, size_t nloc);
void fjlt_set_sm_reserve(int sms);      // SMs the persistent FJLT kernels leave free
// Peer-output mode of the WHT kernel for the calling thread's next applies (nullptr = off); see fjlt.cu.
void fjlt_set_peer_out(double *const *ptrs, double *scratch, int cols_per_owner, int rank, int S);
// Host-side layout of the sampled gather of the WHT kernel (see fjlt.cu).
void fjlt_build_wht_slots(const int32_t *samples, size_t S, std::vector<int32_t> &slot_sample,
                          std::vector<int32_t> &slot_out);

}  // namespace skb
