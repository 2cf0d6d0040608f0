/*
 * skylark_oracle.c -- CPU restatement of libSkylark's sketch-apply hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under libskylark_b200/ may include, link
 * or call this file; only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs use it, as the checker or the timed
 * CPU baseline, never as the product path.
 *
 * PARITY UNPINNED: the reference's own tests hold no golden vector for this
 * path and the reference cannot be built here
 * (Elemental, Boost, Random123, MPI, FFTW are absent).  The arithmetic of the
 * two un-vendored third-party dependencies is restated from their published
 * algorithms:
 *   - Random123 1.08  (conda/random123/meta.yaml:3-8): Threefry2x64 with 13
 *     rounds and r123::MicroURNG, as used by base/randgen.hpp:27-30,109-114.
 *     Pinned here against Random123's published 20-round known-answer vectors.
 *   - Boost.Random 1.60.0 (conda/boost/meta.yaml:3; 1.55.0 on Travis,
 *     script/bootstrap-travis-deps:6-10): uniform_int_distribution,
 *     discrete_distribution, uniform_01, normal_distribution (Box-Muller),
 *     cauchy_distribution, exponential_distribution,
 *     uniform_real_distribution.
 *
 * Every function cites the reference file:line it follows (paths relative to
 * the reference checkout).  Compile with -ffp-contract=off so that products
 * and sums round separately, as the reference's -O3 x86-64 build does.
 */
#include <math.h>
#include <stddef.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#ifdef _OPENMP
#include <omp.h>
#endif

/* ------------------------------------------------------------------------ */
/* Random123: Threefry2x64_R<R>  (Random123/threefry.h; used at            */
/* base/randgen.hpp:27 with R = 13)                                         */
/* ------------------------------------------------------------------------ */

static inline uint64_t rotl64(uint64_t x, int r) {
    return (x << r) | (x >> (64 - r));
}

void sko_threefry2x64(int rounds, const uint64_t ctr[2], const uint64_t key[2],
                      uint64_t out[2]) {
    static const int ROT[8] = {16, 42, 12, 31, 16, 32, 24, 21};
    uint64_t ks[3];
    ks[0] = key[0];
    ks[1] = key[1];
    ks[2] = 0x1BD11BDAA9FC1A22ULL ^ key[0] ^ key[1];
    uint64_t x0 = ctr[0] + ks[0];
    uint64_t x1 = ctr[1] + ks[1];
    for (int r = 0; r < rounds; r++) {
        x0 += x1;
        x1 = rotl64(x1, ROT[r % 8]);
        x1 ^= x0;
        if ((r + 1) % 4 == 0) {
            int q = (r + 1) / 4;
            x0 += ks[q % 3];
            x1 += ks[(q + 1) % 3] + (uint64_t)q;
        }
    }
    out[0] = x0;
    out[1] = x1;
}

/* r123::MicroURNG<Threefry2x64_R<13>> as constructed at
 * base/randgen.hpp:109-113: ctr = {base+index, 0}, key = {seed, 0}.  Block n
 * places n in the high 32 bits of the last counter word; words are handed out
 * last-first (rdata[--last_elem]). */
typedef struct {
    uint64_t c0, k0;
    uint64_t n;
    int last;
    uint64_t r[2];
} urng_t;

static inline void urng_init(urng_t *u, int seed, uint64_t stream_index) {
    u->c0 = stream_index;
    u->k0 = (uint64_t)(int64_t)seed; /* static_cast<uint64_t>(int), randgen.hpp:37 */
    u->n = 0;
    u->last = 0;
}

static inline uint64_t urng_next(urng_t *u) {
    if (u->last == 0) {
        uint64_t c[2] = {u->c0, u->n << 32};
        uint64_t k[2] = {u->k0, 0};
        sko_threefry2x64(13, c, k, u->r);
        u->n++;
        u->last = 2;
    }
    return u->r[--u->last];
}

/* ------------------------------------------------------------------------ */
/* Boost.Random pieces (64-bit engine with min 0, max 2^64-1)               */
/* ------------------------------------------------------------------------ */

/* boost::random::uniform_01<double> (new_uniform_01::operator()) */
static inline double boost_u01(urng_t *u) {
    for (;;) {
        double r = (double)urng_next(u) * 0x1p-64;
        if (r < 1.0) return r;
    }
}

/* boost::random::detail::generate_uniform_int, brange = 2^64-1 > range */
static inline uint64_t boost_uniform_int(urng_t *u, uint64_t range) {
    if (range == 0) return 0;
    const uint64_t brange = ~(uint64_t)0;
    uint64_t bucket = brange / (range + 1);
    if (brange % (range + 1) == range) bucket++;
    for (;;) {
        uint64_t r = urng_next(u) / bucket;
        if (r <= range) return r;
    }
}

/* utility/distributions.hpp:40-52: discrete_distribution{0.5,0,0.5} - 1.
 * Alias table for these weights: [0]=(0.5,->2) [1]=(0,->0) [2]=(1,-).
 * order 160: Boost >= 1.57 draws the index first, then uniform_01.
 * order 155: Boost <= 1.55 draws uniform_01 first, then the index. */
static inline double boost_rademacher(urng_t *u, int order) {
    static const double first[3] = {0.5, 0.0, 1.0};
    static const int second[3] = {2, 0, 2};
    uint64_t r;
    double t;
    if (order == 155) {
        t = boost_u01(u);
        r = boost_uniform_int(u, 2);
    } else {
        r = boost_uniform_int(u, 2);
        t = boost_u01(u);
    }
    int res = (t < first[r]) ? (int)r : second[r];
    return (double)res - 1.0;
}

#define BOOST_PI 3.14159265358979323846

/* ------------------------------------------------------------------------ */
/* The C library behind Boost's distributions.                              */
/* Boost.Random calls std::log / std::cos / std::tan / std::pow.  The libm  */
/* the reference was built against at its pinned versions (glibc <= 2.27:   */
/* the IBM Accurate Mathematical Library) rounds these functions correctly; */
/* the glibc on this box (2.39) is within 0.52 ulp but not correctly        */
/* rounded.  Mode 1 (default) therefore evaluates them in extended          */
/* precision and rounds once: x87 long double first and, when that result   */
/* sits too close to a rounding boundary to decide, __float128.  Mode 0     */
/* calls this box's libm directly (what the reference would cost here: the  */
/* timed CPU baseline uses it).                                             */
/* ------------------------------------------------------------------------ */
#include <quadmath.h>
static int g_libm_cr = 1;
void sko_set_libm(int correctly_rounded) { g_libm_cr = correctly_rounded ? 1 : 0; }
int sko_get_libm(void) { return g_libm_cr; }

static inline int ld_ambiguous(long double v) {
    union { long double ld; struct { uint64_t m; uint16_t se; } p; } u;
    u.ld = v;
    int low = (int)(u.p.m & 0x7FF);            /* the 11 bits below a double's last place */
    int dist = low - 0x400;
    if (dist < 0) dist = -dist;
    return dist <= 16;                         /* libm's long double routines: a few ulp_ld */
}
static inline double m_log(double x) {
    if (!g_libm_cr) return log(x);
    long double v = logl((long double)x);
    return ld_ambiguous(v) ? (double)logq((__float128)x) : (double)v;
}
static inline double m_cos(double x) {
    if (!g_libm_cr) return cos(x);
    long double v = cosl((long double)x);
    return ld_ambiguous(v) ? (double)cosq((__float128)x) : (double)v;
}
static inline double m_tan(double x) {
    if (!g_libm_cr) return tan(x);
    long double v = tanl((long double)x);
    return ld_ambiguous(v) ? (double)tanq((__float128)x) : (double)v;
}
static inline double m_pow(double x, double y) {
    if (!g_libm_cr) return pow(x, y);
    long double v = powl((long double)x, (long double)y);
    return ld_ambiguous(v) ? (double)powq((__float128)x, (__float128)y) : (double)v;
}

/* boost::random::normal_distribution<double> (<= 1.63): Box-Muller, always the
 * cosine branch because randgen.hpp:113 clones a reset distribution. */
static inline double boost_normal(urng_t *u) {
    double r1 = boost_u01(u);
    double r2 = boost_u01(u);
    double rho = sqrt(-2.0 * m_log(1.0 - r2));
    return rho * m_cos(2.0 * BOOST_PI * r1);
}

/* boost::random::cauchy_distribution<double>(0,1) */
static inline double boost_cauchy(urng_t *u) {
    return m_tan(BOOST_PI * (boost_u01(u) - 0.5));
}

/* boost::random::exponential_distribution<double>(1) (<= 1.63) */
static inline double boost_exponential(urng_t *u) {
    return -1.0 * m_log(1.0 - boost_u01(u));
}

/* boost::random::uniform_real_distribution<double>(a,b) */
static inline double boost_uniform_real(urng_t *u, double a, double b) {
    for (;;) {
        double num = (double)urng_next(u);
        double r = num / 0x1p64 * (b - a) + a;
        if (r < b) return r;
    }
}

/* ------------------------------------------------------------------------ */
/* random_samples_array_t::operator[]  (base/randgen.hpp:98-115)            */
/* One sample per stream index; all exported array fillers below take the   */
/* array's base counter and produce elements [0, count).                    */
/* ------------------------------------------------------------------------ */

void sko_raw_draws(int seed, uint64_t base, size_t count, uint64_t *first,
                   uint64_t *second) {
    for (size_t i = 0; i < count; i++) {
        urng_t u;
        urng_init(&u, seed, base + i);
        first[i] = urng_next(&u);
        second[i] = urng_next(&u);
    }
}

void sko_uniform_int(int seed, uint64_t base, size_t count, uint64_t hi,
                     uint64_t *out) {
#pragma omp parallel for if (count > 100000)
    for (size_t i = 0; i < count; i++) {
        urng_t u;
        urng_init(&u, seed, base + i);
        out[i] = boost_uniform_int(&u, hi);
    }
}

void sko_rademacher(int seed, uint64_t base, size_t count, int order,
                    double *out) {
#pragma omp parallel for if (count > 100000)
    for (size_t i = 0; i < count; i++) {
        urng_t u;
        urng_init(&u, seed, base + i);
        out[i] = boost_rademacher(&u, order);
    }
}

void sko_normal(int seed, uint64_t base, size_t count, double *out) {
#pragma omp parallel for if (count > 100000)
    for (size_t i = 0; i < count; i++) {
        urng_t u;
        urng_init(&u, seed, base + i);
        out[i] = boost_normal(&u);
    }
}

void sko_cauchy(int seed, uint64_t base, size_t count, double *out) {
#pragma omp parallel for if (count > 100000)
    for (size_t i = 0; i < count; i++) {
        urng_t u;
        urng_init(&u, seed, base + i);
        out[i] = boost_cauchy(&u);
    }
}

void sko_exponential(int seed, uint64_t base, size_t count, double *out) {
#pragma omp parallel for if (count > 100000)
    for (size_t i = 0; i < count; i++) {
        urng_t u;
        urng_init(&u, seed, base + i);
        out[i] = boost_exponential(&u);
    }
}

void sko_uniform_real(int seed, uint64_t base, size_t count, double a, double b,
                      double *out) {
    for (size_t i = 0; i < count; i++) {
        urng_t u;
        urng_init(&u, seed, base + i);
        out[i] = boost_uniform_real(&u, a, b);
    }
}

/* UST_data_t::build without replacement (sketch/UST_data.hpp:89-98): inside-out shuffle,
 * i-th draw = context.random_value(uniform_int_distribution<int>(0, i)) at stream index base + i
 * (base/context.hpp:131-136); the first S entries of the shuffled range are the samples. */
void sko_ust_noreplace(int seed, uint64_t base, size_t N, size_t S, int64_t *out) {
    int64_t *work = (int64_t *)malloc((N ? N : 1) * sizeof(int64_t));
    for (size_t i = 0; i < N; i++) {
        urng_t u;
        urng_init(&u, seed, base + i);
        size_t j = (size_t)boost_uniform_int(&u, (uint64_t)i);
        work[i] = work[j];
        work[j] = (int64_t)i;
    }
    for (size_t i = 0; i < S; i++) out[i] = work[i];
    free(work);
}

/* context_t::random_int via random_array_t (base/randgen.hpp:171-187) */
int sko_random_int(int seed, uint64_t base) {
    urng_t u;
    urng_init(&u, seed, base);
    return (int)urng_next(&u);
}

/* WZT_data_t::build (sketch/WZT_data.hpp:110-126):
 * value[i] = pm[i] * pow(1/e[i], 1/p), e from [val_base..), pm from [pm_base..) */
void sko_wzt_values(int seed, uint64_t val_base, uint64_t pm_base, size_t count,
                    double p, int order, double *out) {
    for (size_t i = 0; i < count; i++) {
        urng_t u;
        urng_init(&u, seed, val_base + i);
        double e = boost_exponential(&u);
        urng_init(&u, seed, pm_base + i);
        double pm = boost_rademacher(&u, order);
        out[i] = pm * m_pow(1.0 / e, 1.0 / p);
    }
}

/* ------------------------------------------------------------------------ */
/* Hash transforms                                                          */
/* ------------------------------------------------------------------------ */

/* hash_transform_t<El::Matrix,El::Matrix>::apply_impl, columnwise
 * (sketch/hash_transform_Elemental.hpp:83-100).  A is N x w column-major with
 * leading dimension lda; SA is S x w column-major. */
void sko_hash_apply_columnwise(const double *A, size_t N, size_t w, size_t lda,
                               const uint64_t *row_idx, const double *row_value,
                               double *SA, size_t S, size_t ldsa) {
    for (size_t j = 0; j < w; j++) memset(SA + j * ldsa, 0, S * sizeof(double));
    for (size_t r = 0; r < N; r++) {
        size_t nr = (size_t)row_idx[r];
        double sc = row_value[r];
        for (size_t c = 0; c < w; c++) {
            double v = sc * A[r + c * lda];
            SA[nr + c * ldsa] += v;
        }
    }
}

/* same, rowwise (sketch/hash_transform_Elemental.hpp:106-123).  A is h x N
 * column-major; SA is h x S column-major. */
void sko_hash_apply_rowwise(const double *A, size_t h, size_t N, size_t lda,
                            const uint64_t *row_idx, const double *row_value,
                            double *SA, size_t S, size_t ldsa) {
    for (size_t j = 0; j < S; j++) memset(SA + j * ldsa, 0, h * sizeof(double));
    for (size_t c = 0; c < N; c++) {
        size_t nc = (size_t)row_idx[c];
        double sc = row_value[c];
        for (size_t r = 0; r < h; r++) {
            double v = sc * A[r + c * lda];
            SA[r + nc * ldsa] += v;
        }
    }
}

/* hash_transform_t<sparse_matrix_t,El::Matrix>::apply_impl, columnwise
 * (sketch/hash_transform_Elemental.hpp:201-225): CSC A (N x w). */
void sko_hash_apply_csc_columnwise(const int32_t *indptr, const int32_t *indices,
                                   const double *values, size_t w,
                                   const uint64_t *row_idx,
                                   const double *row_value, double *SA, size_t S,
                                   size_t ldsa) {
    for (size_t j = 0; j < w; j++) memset(SA + j * ldsa, 0, S * sizeof(double));
#pragma omp parallel for
    for (long col = 0; col < (long)w; col++) {
        for (int32_t j = indptr[col]; j < indptr[col + 1]; j++) {
            int32_t row = indices[j];
            double val = values[j];
            SA[(size_t)col * ldsa + row_idx[row]] += row_value[row] * val;
        }
    }
}

/* same, rowwise (sketch/hash_transform_Elemental.hpp:231-256): CSC A (h x N),
 * SA is h x S column-major. */
void sko_hash_apply_csc_rowwise(const int32_t *indptr, const int32_t *indices,
                                const double *values, size_t h, size_t N,
                                const uint64_t *row_idx, const double *row_value,
                                double *SA, size_t S, size_t ldsa) {
    for (size_t j = 0; j < S; j++) memset(SA + j * ldsa, 0, h * sizeof(double));
    for (size_t col = 0; col < N; col++) {
        for (int32_t j = indptr[col]; j < indptr[col + 1]; j++) {
            int32_t row = indices[j];
            double val = values[j];
            SA[row_idx[col] * ldsa + (size_t)row] += row_value[col] * val;
        }
    }
}

/* hash_transform_t<sparse_matrix_t,sparse_matrix_t>::apply_impl
 * (sketch/hash_transform_local_sparse.hpp:88-223), both directions.  Output arrays must hold
 * nnz(A) entries; returns the number of stored entries.  Rows inside an output column appear in
 * first-touch order and cancelled sums stay stored, as in the reference. */
size_t sko_hash_apply_csc_sparse(const int32_t *indptr, const int32_t *indices,
                                 const double *values, size_t h, size_t w, int rowwise,
                                 const uint64_t *row_idx, const double *row_value, size_t S,
                                 int32_t *out_indptr, int32_t *out_indices, double *out_values) {
    size_t nnz = 0;
    out_indptr[0] = 0;
    if (!rowwise) {
        long *idx_map = (long *)malloc(S * sizeof(long));
        for (size_t i = 0; i < S; i++) idx_map[i] = -1;
        for (size_t col = 0; col < w; col++) {
            for (int32_t idx = indptr[col]; idx < indptr[col + 1]; idx++) {
                size_t row = (size_t)indices[idx];
                double val = values[idx] * row_value[row];
                row = (size_t)row_idx[row];
                if (idx_map[row] == -1) {
                    idx_map[row] = (long)nnz; out_indices[nnz] = (int32_t)row; out_values[nnz] = val; nnz++;
                } else {
                    out_values[idx_map[row]] += val;
                }
            }
            out_indptr[col + 1] = (int32_t)nnz;
            for (size_t i = (size_t)out_indptr[col]; i < nnz; i++) idx_map[out_indices[i]] = -1;
        }
        free(idx_map);
    } else {
        /* inv_mapping: input columns of each target column, in ascending order */
        size_t *cnt = (size_t *)calloc(S + 1, sizeof(size_t));
        for (size_t c = 0; c < w; c++) cnt[row_idx[c] + 1]++;
        for (size_t b = 0; b < S; b++) cnt[b + 1] += cnt[b];
        size_t *fill = (size_t *)malloc(S * sizeof(size_t));
        size_t *cols = (size_t *)malloc((w ? w : 1) * sizeof(size_t));
        for (size_t b = 0; b < S; b++) fill[b] = cnt[b];
        for (size_t c = 0; c < w; c++) cols[fill[row_idx[c]]++] = c;
        long *idx_map = (long *)malloc((h ? h : 1) * sizeof(long));
        for (size_t i = 0; i < h; i++) idx_map[i] = -1;
        for (size_t b = 0; b < S; b++) {
            for (size_t q = cnt[b]; q < cnt[b + 1]; q++) {
                size_t col = cols[q];
                for (int32_t idx = indptr[col]; idx < indptr[col + 1]; idx++) {
                    size_t row = (size_t)indices[idx];
                    double val = values[idx] * row_value[col];
                    if (idx_map[row] == -1) {
                        idx_map[row] = (long)nnz; out_indices[nnz] = (int32_t)row; out_values[nnz] = val; nnz++;
                    } else {
                        out_values[idx_map[row]] += val;
                    }
                }
            }
            out_indptr[b + 1] = (int32_t)nnz;
            for (size_t i = (size_t)out_indptr[b]; i < nnz; i++) idx_map[out_indices[i]] = -1;
        }
        free(cnt); free(fill); free(cols); free(idx_map);
    }
    return nnz;
}

/* ------------------------------------------------------------------------ */
/* Dense transforms                                                         */
/* ------------------------------------------------------------------------ */

/* dense_transform_data_t::realize_matrix_view
 * (sketch/dense_transform_data.hpp:79-101): R is height x width column-major,
 * R[i_loc, j_loc] = scale * entries[(j + j_loc*row_stride)*S + (i + i_loc*col_stride)].
 * dist: 0 normal (JLT_data.hpp:29, RFT_data.hpp:118), 1 cauchy (CT_data.hpp:37). */
void sko_realize_dense(int seed, uint64_t base, int dist, double scale, size_t S,
                       size_t i, size_t j, size_t height, size_t width,
                       size_t col_stride, size_t row_stride, double *R) {
#pragma omp parallel for
    for (long j_loc = 0; j_loc < (long)width; j_loc++) {
        size_t j_glob = j + (size_t)j_loc * row_stride;
        for (size_t i_loc = 0; i_loc < height; i_loc++) {
            size_t i_glob = i + i_loc * col_stride;
            uint64_t idx = base + (uint64_t)j_glob * S + i_glob;
            urng_t u;
            urng_init(&u, seed, idx);
            double sample = dist == 0 ? boost_normal(&u) : boost_cauchy(&u);
            R[(size_t)j_loc * height + i_loc] = scale * sample;
        }
    }
}

/* same, cast to float (dense_transform_data.hpp:96-98 with T = float) */
void sko_realize_dense_f32(int seed, uint64_t base, int dist, double scale,
                           size_t S, size_t i, size_t j, size_t height,
                           size_t width, float *R) {
#pragma omp parallel for
    for (long j_loc = 0; j_loc < (long)width; j_loc++) {
        for (size_t i_loc = 0; i_loc < height; i_loc++) {
            uint64_t idx = base + (uint64_t)(j + j_loc) * S + (i + i_loc);
            urng_t u;
            urng_init(&u, seed, idx);
            double sample = dist == 0 ? boost_normal(&u) : boost_cauchy(&u);
            R[(size_t)j_loc * height + i_loc] = (float)(scale * sample);
        }
    }
}

/* base::Gemm(NORMAL, NORMAL, 1, R, A, 0, SA) as called by
 * dense_transform_Elemental_local.hpp:97-111 -- plain triple loop, column-major.
 * C (m x n) = op(A) * op(B); ta/tb = 0 normal, 1 transpose. */
void sko_gemm(int ta, int tb, size_t m, size_t n, size_t k, const double *A,
              size_t lda, const double *B, size_t ldb, double *C, size_t ldc) {
#pragma omp parallel for
    for (long j = 0; j < (long)n; j++) {
        double *c = C + (size_t)j * ldc;
        for (size_t i = 0; i < m; i++) c[i] = 0.0;
        for (size_t p = 0; p < k; p++) {
            double b = tb ? B[(size_t)j + p * ldb] : B[p + (size_t)j * ldb];
            if (!ta) {
                const double *a = A + p * lda;
                for (size_t i = 0; i < m; i++) c[i] += a[i] * b;
            } else {
                for (size_t i = 0; i < m; i++) c[i] += A[p + i * lda] * b;
            }
        }
    }
}

/* RFT_t::apply_impl epilogue (sketch/RFT_Elemental.hpp:94-117 columnwise,
 * :135-158 rowwise): x = outscale * cos(x * scales[f] + shifts[f]); the feature
 * index f is the row for columnwise and the column for rowwise. */
void sko_rft_epilogue(double *SA, size_t rows, size_t cols, size_t ld,
                      int rowwise, const double *scales, const double *shifts,
                      double outscale) {
#pragma omp parallel for
    for (long j = 0; j < (long)cols; j++)
        for (size_t i = 0; i < rows; i++) {
            size_t f = rowwise ? (size_t)j : i;
            double x = SA[i + (size_t)j * ld];
            x *= scales[f];
            x += shifts[f];
            x = cos(x);
            x = outscale * x;
            SA[i + (size_t)j * ld] = x;
        }
}

void sko_rft_epilogue_f32(float *SA, size_t rows, size_t cols, size_t ld,
                          int rowwise, const double *scales,
                          const double *shifts, double outscale) {
    /* value_type is float when the matrices are float (RFT_Elemental.hpp:62) */
#pragma omp parallel for
    for (long j = 0; j < (long)cols; j++)
        for (size_t i = 0; i < rows; i++) {
            size_t f = rowwise ? (size_t)j : i;
            float x = SA[i + (size_t)j * ld];
            x *= scales[f];
            x += shifts[f];
            x = cosf(x);
            x = outscale * x;
            SA[i + (size_t)j * ld] = x;
        }
}

/* ------------------------------------------------------------------------ */
/* FJLT                                                                     */
/* ------------------------------------------------------------------------ */

/* FFTW REDFT10 definition (utility/fft/fftw_futs.h:117-126):
 * Y[k] = 2 * sum_j X[j] cos(pi (j + 1/2) k / N).  O(N^2), small N only; used
 * to check the FFT-based path the Python oracle takes for large N. */
void sko_dct2_naive(const double *x, size_t N, double *y) {
    for (size_t k = 0; k < N; k++) {
        long double acc = 0.0L;
        for (size_t j = 0; j < N; j++)
            acc += (long double)x[j] *
                   cosl(3.14159265358979323846264338327950288L *
                        ((long double)j + 0.5L) * (long double)k / (long double)N);
        y[k] = (double)(2.0L * acc);
    }
}

/* RFUT_t::apply_impl_vdist scaling (sketch/RFUT_Elemental.hpp:173-180):
 * TA[i,j] = scale * D[i] * A[i,j] with scale = 1/sqrt(2N) (fftw_futs.h:44-46). */
void sko_rfut_scale(const double *A, size_t N, size_t w, size_t lda,
                    const double *D, double scale, double *TA) {
    for (size_t j = 0; j < w; j++)
        for (size_t i = 0; i < N; i++)
            TA[i + j * N] = scale * D[i] * A[i + j * lda];
}

/* SpiralWHT wht_apply restated as the textbook in-place butterfly in natural
 * (Hadamard) order; WHT_t scale 1/sqrt(N) (sketch/FUT.hpp:25-113).  Used only
 * by the WHT mode of FJLT, which the reference does not have. */
void sko_wht_inplace(double *x, size_t N) {
    for (size_t h = 1; h < N; h <<= 1)
        for (size_t i = 0; i < N; i += (h << 1))
            for (size_t j = i; j < i + h; j++) {
                double a = x[j], b = x[j + h];
                x[j] = a + b;
                x[j + h] = a - b;
            }
}

/* FJLT_t::apply_impl_vdist gather (sketch/FJLT_Elemental.hpp:157-166):
 * SA[i,j] = sqrt(N/S) * TA[samples[i], j]. */
void sko_fjlt_gather(const double *TA, size_t N, size_t w, const uint64_t *samples,
                     size_t S, double scale, double *SA, size_t ldsa) {
    for (size_t j = 0; j < w; j++)
        for (size_t i = 0; i < S; i++)
            SA[i + j * ldsa] = scale * TA[samples[i] + j * N];
}

void sko_set_num_threads(int n) {
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}

int sko_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
