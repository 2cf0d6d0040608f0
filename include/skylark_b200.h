/*
 * skylark_b200.h -- C ABI of libcskylark_b200.so, the B200-native replacement for the
 * sketch-apply subset of libSkylark's libcskylark.so.
 *
 * Part 1 keeps the reference's entry points (names, argument order and meaning, error
 * codes) so that python-skylark's lib.py / sketch.py bind to it unchanged; each
 * declaration cites the reference declaration it replaces (paths are relative to the
 * reference checkout).  Part 2 holds additive extensions (explicit contexts, device
 * buffers, streams, row shards, table export for tests, launch accounting).
 *
 * All matrices follow the reference convention: dense = column-major with leading
 * dimension, sparse = CSC with int32 indices; buffers are attached, never copied or
 * owned by the wrapper.  Return value 0 = success, otherwise a code from
 * base/exception.hpp:9-24 (100..113); sl_get_exception_info() gives the message.
 * Every compute entry point needs a CUDA device: there is no CPU fallback.
 */
#ifndef SKYLARK_B200_H
#define SKYLARK_B200_H

#include <stdbool.h>
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct sl_context_t sl_context_t;
typedef struct sl_sketch_transform_t sl_sketch_transform_t;

#define SL_COLUMNWISE 1 /* capi/basec.hpp:12 */
#define SL_ROWWISE 2    /* capi/basec.hpp:13 */

/* ------------------------------------------------------------------ Part 1 ---- */

/* capi/basec.hpp:22, capi/csketch.cpp:74-376: space separated ("T","In","Out") tuples. */
char *sl_supported_sketch_transforms(void);
/* capi/basec.hpp:28 */
const char *sl_strerror(const int errorcode);
/* capi/basec.hpp:33-34 */
bool sl_has_elemental(void);
bool sl_has_combblas(void);

/* capi/basec.hpp:43-44, capi/cbase.cpp:37-44 */
int sl_create_default_context(int seed, sl_context_t **ctxt);
/* capi/basec.hpp:52-53, capi/cbase.cpp:46-53: the communicator is ignored there too. */
int sl_create_context(int seed, void *comm_ignored, sl_context_t **ctxt);
/* capi/basec.hpp:58 */
int sl_free_context(sl_context_t *ctxt);

/* capi/sketchc.hpp:21-23, capi/csketch.cpp:379-545.  Variadic tail: CT `double C`;
 * WZT `double p`; GaussianRFT / LaplacianRFT `double sigma`; UST `int replace` (the reference has
 * UST_t, sketch/UST_data.hpp:19-115, but does not register it in its C API: an extension, with
 * ("UST","Matrix","Matrix") advertised by sl_supported_sketch_transforms).  Unknown type -> 111. */
int sl_create_sketch_transform(sl_context_t *ctxt, char *type, int n, int s,
                               sl_sketch_transform_t **sketch, ...);
/* capi/sketchc.hpp:29-30 */
int sl_deserialize_sketch_transform(const char *data, sl_sketch_transform_t **sketch);
/* capi/sketchc.hpp:36-37; *data is malloc'ed, caller frees with free(). */
int sl_serialize_sketch_transform(const sl_sketch_transform_t *sketch, char **data);
/* capi/sketchc.hpp:42-43 */
int sl_free_sketch_transform(sl_sketch_transform_t *S);
/* capi/sketchc.hpp:53-56, capi/csketch.cpp:614-1958.  input_type/output_type are
 * "Matrix" or "SparseMatrix"; A/SA come from the sl_wrap_* functions; dim is
 * SL_COLUMNWISE (SA = S*A) or SL_ROWWISE (SA = A*S^T).  An unsupported combination
 * returns 109 (the reference silently returns 0, capi/csketch.cpp:1957).
 * Host buffers (the reference's only kind) are copied to the device and the result back inside the
 * call, which returns when SA is complete.  Pageable buffers above 64 MiB are staged through pinned
 * memory by a few worker threads of the library (SKB_COPY_THREADS, default min(8, cores / 2); 0 = one
 * plain cudaMemcpyAsync); the threads only read A and write SA, and only during the call. */
int sl_apply_sketch_transform(sl_sketch_transform_t *S, char *input_type, void *A,
                              char *output_type, void *SA, int dim);

/* capi/basec.hpp:62-64, capi/cbase.cpp:63-75: host buffer, column-major m x n, ld = m. */
int sl_wrap_raw_matrix(double *data, int m, int n, void **A);
int sl_free_raw_matrix_wrap(void *A_);
/* capi/basec.hpp:66-83, capi/cbase.cpp:77-121: host CSC, int32 indptr[n_cols+1]/indices.
 * When a wrapper is the SA of a sparse -> sparse CWT/MMT/WZT apply
 * (sketch/hash_transform_local_sparse.hpp:88-223) the library attaches new arrays it owns and
 * raises struct_updated; read nnz/height/width, then copy out with sl_raw_sp_matrix_data before
 * freeing the wrapper.  For device wrappers (sl_wrap_device_sp_matrix) the three destinations
 * of sl_raw_sp_matrix_data are device pointers. */
int sl_wrap_raw_sp_matrix(int *indptr, int *ind, double *data, int nnz, int n_rows,
                          int n_cols, void **A);
int sl_free_raw_sp_matrix_wrap(void *A_);
int sl_raw_sp_matrix_struct_updated(void *A_, bool *struct_updated);
int sl_raw_sp_matrix_reset_update_flag(void *A_);
int sl_raw_sp_matrix_nnz(void *A_, int *nnz);
int sl_raw_sp_matrix_height(void *A_, int *height);
int sl_raw_sp_matrix_width(void *A_, int *width);
int sl_raw_sp_matrix_data(void *A_, int32_t *indptr, int32_t *indices, double *values);

/* capi/basec.hpp:85-87; *info is malloc'ed (the reference mixes new[] with free()). */
void sl_get_exception_info(char **info);
void sl_print_exception_trace(void);

/* ------------------------------------------------------------------ Part 2 ---- */

/* context_t(seed, counter) (base/context.hpp:25-27) and its accessors (:166). */
int sl_create_context_at(int seed, uint64_t counter, sl_context_t **ctxt);
int sl_context_get_seed(sl_context_t *ctxt, int *seed);
int sl_context_get_counter(sl_context_t *ctxt, uint64_t *counter);
/* context_t::allocate_random_samples_array's bookkeeping (base/context.hpp:94-101): reserves `size`
 * stream indices, returns the first; 108 when the stream would be exhausted. */
int sl_context_allocate(sl_context_t *ctxt, uint64_t size, uint64_t *base);

#define SL_DTYPE_F64 0
#define SL_DTYPE_F32 1
/* Device-resident column-major matrix (m x n, leading dimension ld, elements of dtype)
 * on CUDA device `device`; the counterpart of sl_wrap_raw_matrix for torch tensors.
 * Free with sl_free_raw_matrix_wrap. */
int sl_wrap_device_matrix(void *dptr, int64_t m, int64_t n, int64_t ld, int dtype, int device,
                          void **A);
/* Device-resident CSC.  Free with sl_free_raw_sp_matrix_wrap. */
int sl_wrap_device_sp_matrix(int32_t *indptr, int32_t *ind, double *data, int64_t nnz,
                             int64_t n_rows, int64_t n_cols, int device, void **A);

/* CUDA stream (cudaStream_t) used by subsequent calls of the calling thread; NULL = the
 * legacy default stream.  Device-matrix applies are asynchronous on that stream. */
int sl_set_stream(void *cuda_stream);

/* Row-sharded apply: A holds sketched indices [index_offset, index_offset + local) of
 * the transform's N; SA receives this shard's partial sum (to be summed across shards,
 * cf. sketch/hash_transform_Elemental.hpp:543-561 and
 * sketch/dense_transform_Elemental_coldist_star.hpp:200-229).  RFT cannot be sharded this
 * way (the cosine is not additive); FJLT can in the Walsh-Hadamard mode only, with N a power
 * of two >= 8192 and 8192-aligned shards (in DCT mode shard the other dimension, as the
 * reference re-distributes to, sketch/FJLT_Elemental.hpp:149-150). */
int sl_apply_sketch_transform_shard(sl_sketch_transform_t *S, char *input_type, void *A,
                                    char *output_type, void *SA, int dim,
                                    int64_t index_offset);

/* Row-sharded FJLT (Walsh-Hadamard mode) with the reduce-scatter fused into the kernel -- what
 * sketch/FJLT_Elemental.hpp:144-169 does with two redistributions, for a group of GPUs that can store
 * into each other's memory.  A: this rank's rows [index_offset, ...) of the n x d matrix (device,
 * column-major, d a multiple of `world`).  peer_ptrs_dev: DEVICE array of `world` pointers, entry r =
 * the staging buffer of rank r as mapped into this process, each world * (d/world) * s doubles laid out
 * [source rank][local column][sample].  The kernel writes the partial sketch of column j into slab `rank`
 * of the buffer of rank j / (d/world) (plain stores over NVLink, no TMA); after all ranks' kernels have
 * finished (a group barrier, the caller's job) every rank adds its `world` slabs.  scratch_dev: 148 * s
 * doubles of local device memory.  Needs an even s <= 8192.  dist.fjlt_wht_fused_reduce_scatter drives it. */
int sl_fjlt_apply_shard_peer(sl_sketch_transform_t *S, void *A, int64_t index_offset, void *peer_ptrs_dev,
                             int world, int rank, void *scratch_dev);

/* Options: "boost_order" = "160" (default; Boost >= 1.57 Rademacher draw order) | "155";
 * "fjlt_fut" = "dct" (default; what the reference computes) | "wht";
 * "sm_reserve" = SMs the persistent FJLT kernels leave free for a collective running beside them (0..100);
 * "csc_form" = "dump" (default) | "bulk": kernel form of the CSC -> dense hash sketch (same results). */
int sl_set_option(const char *key, const char *value);

/* Testing hook, host arithmetic only (no device needed): the plan of the strided hash kernel for the given
 * bucket / value tables (n entries, s buckets).  entries = rounds of 32 packed words
 * (index in its tile of 2048 << 3 | bucket inside the warp's range << 16 | sign << 31), one stream per warp
 * (8 warps), off[w * (ntiles + 1) + t] = first round of warp w in tile t; bucket_width = buckets per warp.
 * Both arrays are malloc'ed; the caller frees them. */
int sl_debug_hash_cols_plan(const int32_t *bucket, const double *value, int64_t n, int s, uint32_t **entries,
                            int64_t *n_entries, uint32_t **off, int64_t *n_off, int *bucket_width);

/* Generated tables, copied to host memory (tests compare them bit for bit):
 * what = "row_idx" (int64, hash), "row_value" (double, hash), "D" (double, FJLT),
 * "samples" (int64, FJLT and UST), "shifts" (double, RFT), "scales" (double, RFT);
 * elements [first, first+count). */
int sl_sketch_export(sl_sketch_transform_t *S, const char *what, int64_t first, int64_t count,
                     void *host_out);
/* Dense S block S[i0:i0+h, j0:j0+w], column-major doubles, to host memory
 * (sketch/dense_transform_data.hpp:79-101). */
int sl_sketch_export_dense(sl_sketch_transform_t *S, int64_t i0, int64_t j0, int64_t h,
                           int64_t w, double *host_out);
/* Raw MicroURNG draws of stream indices [base, base+count): first[i] = X1, second[i] = X0
 * (base/randgen.hpp:109-114), generated on the device. */
int sl_rng_export_raw(int seed, uint64_t base, int64_t count, uint64_t *first,
                      uint64_t *second);

/* Variates of stream indices [base, base+count), generated on the device with the semantics of
 * base/randgen.hpp:98-115 + Boost.Random: kind 0 = Rademacher
 * (utility/distributions.hpp:40-52, current boost_order), 1 = normal(0,1), 2 = cauchy(0,1),
 * 3 = uniform_real(a, b), 4 = uniform_int(0, a) (as doubles).  Used by the composite transforms of the Python layer (Fastfood, PPT),
 * whose tables are several such arrays (sketch/FRFT_data.hpp:91-120, sketch/PPT_data.hpp:83-99). */
int sl_rng_export(int kind, int seed, uint64_t base, int64_t count, double a, double b, double *host_out);

/* Launch accounting.  Classes: 0 tables, 1 hash apply, 2 dense panel generation,
 * 3 dense GEMM, 4 FJLT, 5 misc.  With timing enabled every launch is bracketed by CUDA
 * events on its stream; sl_profile_read synchronises and returns the totals. */
#define SL_KERNEL_CLASSES 6
int sl_profile_enable(int timing_on);
int sl_profile_reset(void);
int sl_profile_read(unsigned long long launches[SL_KERNEL_CLASSES],
                    double ms[SL_KERNEL_CLASSES]);

/* Dense sketch (JLT / CT arithmetic) that allocates nothing.  No counterpart in the
 * reference.  It exists for callers that must not allocate device memory at the moment they
 * sketch: a latency-critical loop that cannot afford an allocator call, a process whose
 * GPU memory is handed out up front by another allocator (so a later cudaMalloc may simply
 * fail), or a host application that forbids cudaMalloc once it is running.  sl_apply_sketch_transform keeps
 * a cudaMalloc'ed, growing workspace per transform; this call instead stages the generated
 * panel of S in a fixed 32 MiB array that is part of the library image on every GPU that
 * loads it (held until the process exits), and makes no cudaMalloc / cudaFree / stream or
 * event creation of its own.
 *   SA = scale * G * A        (dim = SL_COLUMNWISE; A is n x m, SA is s x m)
 *   SA = scale * A * G^T      (dim = SL_ROWWISE;    A is m x n, SA is m x s)
 * with G[i,k] = sample(counter_base + k*s + i), dist 0 = normal, 1 = Cauchy -- i.e. the S of
 * a JLT created from a context at `counter_base` with scale sqrt(1/s), or a CT with C/s.
 * A_dev / SA_dev are column-major fp64 device buffers with leading dimensions lda / ldsa;
 * the work is queued on `cuda_stream` (NULL = default stream) and the call returns without
 * synchronising.
 * Limits and rules:
 *   - s must be at most 262144 (16 panel rows have to fit in the 4 Mi-double array); larger s
 *     is refused with code 111 and a message.  n and m are not limited.
 *   - The array used is the one of the CURRENT device: A_dev, SA_dev and the stream must belong
 *     to it (operands of another device are refused with code 112).
 *   - The first call on each device looks up the array's device address (and loads the
 *     kernels); make that first call outside any stream capture.
 *   - All calls on a device share the one array.  Nothing in it needs clearing between
 *     calls (each call rewrites what it reads; SA is zeroed by the call itself), but two
 *     calls that run on the GPU at the same time -- different streams, or different host
 *     threads -- overwrite each other's panel and give wrong results.  Serialise them. */
int sl_dense_sketch_apply_noalloc(int seed, uint64_t counter_base, int dist, double scale,
                                  int64_t n, int64_t s, const double *A_dev, int64_t lda,
                                  int64_t m, double *SA_dev, int64_t ldsa, int dim,
                                  void *cuda_stream);

/* Pinned host memory for the host-buffer path (cudaHostAlloc / cudaFreeHost). */
int sl_host_alloc_pinned(size_t bytes, void **ptr);
int sl_host_free_pinned(void *ptr);
int sl_device_count(int *count);
/* No counterpart in the reference (its distributed hash apply reduces dense partials,
 * sketch/hash_transform_Elemental.hpp:543-561).  A, SA: device sparse wrappers.  SA receives
 * A with entries of equal (row, column) added in storage order and the rows of every column
 * sorted -- the merge step after ranks have exchanged sparse partial sketches. */
int sl_sp_matrix_coalesce(void *A, void *SA);
/* The same merge straight from the receive buffers of an all-to-all, without re-ordering them first:
 * the partial sketches of `nseg` ranks lie one after the other (rank-major); entries of column j sent by
 * rank r are indices/values[vptr[r * n_cols + j] .. vptr[r * n_cols + j + 1]) (device arrays, vptr has
 * nseg * n_cols + 1 entries).  SA (device sparse wrapper) receives the n_rows x n_cols sum; equal
 * (row, column) entries are added in rank order.  109 when a merged column has more than 10240 entries. */
int sl_sp_merge_segments(const int32_t *vptr_dev, const int32_t *indices_dev, const double *values_dev, int nseg,
                         int64_t n_rows, int64_t n_cols, int64_t nnz, void *SA);

#ifdef __cplusplus
}
#endif
#endif /* SKYLARK_B200_H */
