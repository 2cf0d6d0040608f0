// C ABI of libcskylark_b200.so (declared in include/skylark_b200.h).
//
// Mirrors the sketch subset of the reference's capi/{cbase,csketch}.cpp: contexts with
// seed + counter semantics (base/context.hpp:19-183), sketch data objects whose random
// state is a set of counter-stream bases (sketch/*_data.hpp), raw-matrix wrappers, JSON
// (de)serialisation that replays the RNG, the error-code convention of
// base/exception.hpp:9-24.  All arithmetic runs on the GPU; there is no CPU fallback.
#include <cmath>
#include <condition_variable>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../include/skylark_b200.h"
#include "rng.cuh"
#include "skb_internal.h"

using namespace skb;

// ---------------------------------------------------------------------------
// process-wide state
// ---------------------------------------------------------------------------
static std::string g_last_error;
static thread_local cudaStream_t g_stream = nullptr;
static int g_boost_order = SKB_ORDER_160;
static int g_fjlt_fut = FUT_DCT;

namespace skb {
void set_last_error(const std::string &msg) { g_last_error = msg; }
Profiler &profiler() { static Profiler p; return p; }
struct PendingPair { KernelClass kc; cudaEvent_t e0, e1; };
static std::vector<PendingPair> g_pending;

LaunchScope::LaunchScope(KernelClass kc_, cudaStream_t st) : kc(kc_), stream(st), timed(false) {
    Profiler &p = profiler();
    p.launches[kc]++;
    if (p.enabled) {
        if (cudaEventCreate(&e0) == cudaSuccess && cudaEventCreate(&e1) == cudaSuccess) {
            cudaEventRecord(e0, stream);
            timed = true;
        }
    }
}
LaunchScope::~LaunchScope() {
    if (timed) {
        cudaEventRecord(e1, stream);
        g_pending.push_back({kc, e0, e1});
    }
}
void profiler_collect() {
    Profiler &p = profiler();
    for (auto &pp : g_pending) {
        cudaEventSynchronize(pp.e1);
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, pp.e0, pp.e1) == cudaSuccess) p.ms[pp.kc] += ms;
        cudaEventDestroy(pp.e0);
        cudaEventDestroy(pp.e1);
    }
    g_pending.clear();
}
}  // namespace skb

static const char *const k_errmsg[] = {   // base/exception.hpp:9-24
    "Skylark failure",
    "Skylark failed while communicating (MPI)",
    "Skylark failed in a call to Elemental",
    "Skylark failed due to a unsupported matrix distribution",
    "Skylark failed in a call to CombBLAS",
    "Skylark failed in sketching layer",
    "Skylark failed in nla operation",
    "Skylark failed when allocating memory in a sketch",
    "Skylark failed in a call into the Random123 layer",
    "Skylark failed due to a unsupported base operation",
    "Skylark failed in I/O calls",
    "Skylark failed because invalid parameter was passed",
    "Skylark failed because base classes were used incorrectly",
    "Skylark failed in ml operation"};

#define SL_TRY try {
#define SL_CATCH                                                                   \
    }                                                                              \
    catch (const skb::error &e) { g_last_error = e.what(); return e.code; }        \
    catch (const std::bad_alloc &) { g_last_error = "out of host memory"; return SKB_ERR_ALLOC; } \
    catch (const std::exception &e) { g_last_error = e.what(); return SKB_ERR_GENERIC; } \
    return 0;

static void require_device() {
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) {
        cudaGetLastError();
        throw error(SKB_ERR_GENERIC, "libcskylark_b200: no CUDA device available; this library has "
                                     "no CPU path");
    }
}

struct DeviceGuard {
    int prev = -1; bool changed = false;
    explicit DeviceGuard(int dev) {
        if (dev < 0) return;
        SKB_CUDA(cudaGetDevice(&prev));
        if (prev != dev) { SKB_CUDA(cudaSetDevice(dev)); changed = true; }
    }
    ~DeviceGuard() { if (changed) cudaSetDevice(prev); }
};

// ---------------------------------------------------------------------------
// objects
// ---------------------------------------------------------------------------
struct sl_context_t {
    int seed; uint64_t counter;
    uint64_t allocate(uint64_t size) {   // base/context.hpp:94-101, randgen.hpp:64-72
        if (size != 0 && UINT64_MAX - counter < size - 1)
            throw error(SKB_ERR_RANDOM123, "The random number generator stream cannot supply the "
                                           "requested number of samples without being exhausted");
        uint64_t b = counter; counter += size; return b;
    }
};

enum MatKind { MK_DENSE = 0, MK_SPARSE = 1 };
struct sl_matrix_t {
    MatKind kind; bool on_device; int device; int dtype;
    void *data; int64_t m, n, ld;                       // dense
    int32_t *indptr, *indices; double *values; int64_t nnz;   // sparse (m x n CSC)
    bool struct_updated;
    struct SpOwned *own = nullptr;   // arrays a sparse result was attached with (wrapper-owned)
};
// Storage behind a sparse SA after an apply replaced its structure (the reference's
// sparse_matrix_t::attach(..., ownership=true), base/sparse_matrix.hpp); freed with the wrapper.
struct SpOwned {
    std::vector<int32_t> h_indptr, h_indices; std::vector<double> h_values;
    int32_t *d_indptr = nullptr, *d_indices = nullptr; double *d_values = nullptr; int device = -1;
    void release_device() {
        if (!d_indptr && !d_indices && !d_values) return;
        int prev = -1; cudaGetDevice(&prev); if (device >= 0) cudaSetDevice(device);
        cudaFree(d_indptr); cudaFree(d_indices); cudaFree(d_values);   // pool memory: goes back to the pool
        d_indptr = d_indices = nullptr; d_values = nullptr;
        if (prev >= 0) cudaSetDevice(prev);
    }
    ~SpOwned() { release_device(); }
};

enum TType { T_JLT, T_CT, T_CWT, T_MMT, T_WZT, T_GRFT, T_LRFT, T_FJLT, T_UST };

struct DevState {
    int32_t *bucket = nullptr; double *value = nullptr;
    std::map<std::pair<size_t, size_t>, HashPlan> plans;
    std::map<std::pair<size_t, size_t>, HashColsPlan> cplans;
    double *shifts = nullptr, *scales = nullptr;
    double *D = nullptr; int32_t *samples = nullptr; uint32_t *Dbits = nullptr;
    int32_t *wht_slot_sample = nullptr, *wht_slot_out = nullptr;
    void *dct_plan = nullptr;
    void *ws = nullptr; size_t ws_bytes = 0;
};

struct sl_sketch_transform_t {
    TType type; std::string name;
    int N, S;
    int seed; uint64_t counter0;        // creation context
    double param = 0.0;                 // C | p | sigma
    int order, fut;
    // stream bases (Appendix B of SURVEY.md; *_data.hpp build() functions)
    uint64_t base0 = 0, base1 = 0, base2 = 0;
    double scale = 1.0, outscale = 1.0;
    std::map<int, DevState> dev;

    ~sl_sketch_transform_t() {
        for (auto &kv : dev) {
            int prev = -1; cudaGetDevice(&prev); cudaSetDevice(kv.first);
            DevState &d = kv.second;
            cudaFree(d.bucket); cudaFree(d.value); cudaFree(d.shifts); cudaFree(d.scales);
            cudaFree(d.D); cudaFree(d.samples); cudaFree(d.Dbits); cudaFree(d.wht_slot_sample);
            cudaFree(d.wht_slot_out); cudaFree(d.dct_plan); cudaFree(d.ws);
            for (auto &p : d.plans) hash_free_plan(&p.second);
            for (auto &p : d.cplans) hash_free_cols_plan(&p.second);
            if (prev >= 0) cudaSetDevice(prev);
        }
    }
    bool is_hash() const { return type == T_CWT || type == T_MMT || type == T_WZT; }
    bool is_dense() const { return type == T_JLT || type == T_CT; }
    bool is_rft() const { return type == T_GRFT || type == T_LRFT; }
};

static bool str2type(const char *s, TType *t) {   // capi/csketch.cpp:48-67 (sketch subset)
    static const struct { const char *n; TType t; } tab[] = {
        {"JLT", T_JLT}, {"CT", T_CT}, {"CWT", T_CWT}, {"MMT", T_MMT}, {"WZT", T_WZT},
        {"GaussianRFT", T_GRFT}, {"LaplacianRFT", T_LRFT}, {"FJLT", T_FJLT}, {"UST", T_UST}};
    for (auto &e : tab) if (strcmp(s, e.n) == 0) { *t = e.t; return true; }
    return false;
}

// Builds the random state the way the reference's *_data_t constructors do and
// advances the context (counter rules: SURVEY.md Appendix B).
static sl_sketch_transform_t *make_transform(TType type, const char *name, int n, int s,
                                             double param, sl_context_t &ctx) {
    if (n <= 0 || s <= 0) throw error(SKB_ERR_INVALID_PARAM, "sketch dimensions must be positive");
    std::unique_ptr<sl_sketch_transform_t> t(new sl_sketch_transform_t());
    t->type = type; t->name = name; t->N = n; t->S = s;
    t->seed = ctx.seed; t->counter0 = ctx.counter; t->param = param;
    t->order = g_boost_order; t->fut = g_fjlt_fut;
    const uint64_t N = (uint64_t)n, S = (uint64_t)s;
    switch (type) {
    case T_CWT: case T_MMT:           // hash_transform_data.hpp:64-76
        t->base0 = ctx.allocate(N); t->base1 = ctx.allocate(N); break;
    case T_WZT:                       // WZT_data.hpp:49-52, 110-126
        if (param < 1 || 2 < param)
            throw error(SKB_ERR_SKETCH, "WZT parameter p has to be in (1, 2)");
        t->base0 = ctx.allocate(N); t->base1 = ctx.allocate(N); t->base2 = ctx.allocate(N); break;
    case T_JLT:                       // JLT_data.hpp:27-32
        t->scale = sqrt(1.0 / (double)s); t->base0 = ctx.allocate(N * S); break;
    case T_CT:                        // CT_data.hpp:35-41
        t->scale = param / (double)s; t->base0 = ctx.allocate(N * S); break;
    case T_GRFT: case T_LRFT:         // RFT_data.hpp:71-89, 115-123
        t->scale = 1.0 / param; t->outscale = sqrt(2.0 / s);
        t->base0 = ctx.allocate(N * S); t->base1 = ctx.allocate(S); break;
    case T_FJLT:                      // FJLT_data.hpp:79-86 + RFUT_data.hpp:27-31 (overlap kept)
        t->base0 = ctx.counter; ctx.allocate(S); break;
    case T_UST:                       // UST_data.hpp:81-102: param != 0 = with replacement
        if (param == 0.0 && s > n)
            throw error(SKB_ERR_INVALID_PARAM, "UST without replacement needs s <= n");
        t->base0 = ctx.allocate(param != 0.0 ? S : N); break;
    }
    return t.release();
}

static DevState &dev_state(sl_sketch_transform_t *t, int device) { return t->dev[device]; }

static void ensure_ws(DevState &d, size_t bytes) {
    if (d.ws_bytes >= bytes) return;
    if (d.ws) { SKB_CUDA(cudaStreamSynchronize(g_stream)); cudaFree(d.ws); d.ws = nullptr; d.ws_bytes = 0; }
    SKB_CUDA(cudaMalloc(&d.ws, bytes));
    d.ws_bytes = bytes;
}

// Device tables are cached per (transform, device).  They are built on the calling thread's
// stream and the build is waited for once, so that a later apply on ANY stream may read them; the
// DevState pointers are published only after every allocation and launch has succeeded.
static void ensure_hash_tables(sl_sketch_transform_t *t, DevState &d) {
    if (d.bucket) return;
    int32_t *bucket = nullptr; double *value = nullptr;
    try {
        SKB_CUDA(cudaMalloc(&bucket, (size_t)t->N * sizeof(int32_t)));
        SKB_CUDA(cudaMalloc(&value, (size_t)t->N * sizeof(double)));
        int kind = t->type == T_CWT ? HV_RADEMACHER : t->type == T_MMT ? HV_CAUCHY : HV_WZT;
        hash_build_tables(t->seed, t->base0, t->base1, t->base2, 0, (size_t)t->N, (uint64_t)t->S, kind,
                          t->param, t->order, bucket, value, g_stream);
        SKB_CUDA(cudaStreamSynchronize(g_stream));
    } catch (...) { cudaFree(bucket); cudaFree(value); throw; }
    d.value = value; d.bucket = bucket;
}

static void ensure_rft_tables(sl_sketch_transform_t *t, DevState &d) {
    if (d.shifts) return;
    // shifts ~ uniform_real(0, 2*pi) (RFT_data.hpp:79-82); scales = 1 (:86).  Integer-exact
    // arithmetic, so generating on the host and copying is identical to the device.
    std::vector<double> sh((size_t)t->S), sc((size_t)t->S, 1.0);
    const double pi = 3.141592653589793238462643383279502884;   // boost::math::constants::pi
    for (int i = 0; i < t->S; i++) {
        MicroUrng u(seed_to_key(t->seed), t->base1 + (uint64_t)i);
        sh[i] = boost_uniform_real(u, 0.0, 2 * pi);
    }
    SKB_CUDA(cudaMalloc(&d.shifts, sh.size() * sizeof(double)));
    SKB_CUDA(cudaMalloc(&d.scales, sc.size() * sizeof(double)));
    SKB_CUDA(cudaMemcpy(d.shifts, sh.data(), sh.size() * sizeof(double), cudaMemcpyHostToDevice));
    SKB_CUDA(cudaMemcpy(d.scales, sc.data(), sc.size() * sizeof(double), cudaMemcpyHostToDevice));
}

static void ensure_fjlt_tables(sl_sketch_transform_t *t, DevState &d) {
    if (d.D) return;
    double *D = nullptr; int32_t *samples = nullptr, *ss = nullptr, *so = nullptr; uint32_t *Dbits = nullptr;
    void *plan = nullptr;
    try {
        SKB_CUDA(cudaMalloc(&D, (size_t)t->N * sizeof(double)));
        SKB_CUDA(cudaMalloc(&samples, (size_t)t->S * sizeof(int32_t)));
        SKB_CUDA(cudaMalloc(&Dbits, ((size_t)t->N / 32 + 1) * sizeof(uint32_t)));
        fjlt_build_tables(t->seed, t->base0, (size_t)t->N, (size_t)t->S, t->order, D, samples, Dbits, g_stream);
        if (t->fut == FUT_DCT) {
            const size_t pb = fjlt_dct_plan_bytes((size_t)t->N, (size_t)t->S);
            if (pb) {
                SKB_CUDA(cudaMalloc(&plan, pb));
                fjlt_dct_plan_build((size_t)t->N, (size_t)t->S, samples, plan, g_stream);
            }
        }
        SKB_CUDA(cudaStreamSynchronize(g_stream));
        if (t->fut == FUT_WHT) {     // gather layout of the WHT kernel, computed once from the samples
            std::vector<int32_t> hs((size_t)t->S), slot_sample, slot_out;
            SKB_CUDA(cudaMemcpy(hs.data(), samples, hs.size() * sizeof(int32_t), cudaMemcpyDeviceToHost));
            fjlt_build_wht_slots(hs.data(), hs.size(), slot_sample, slot_out);
            SKB_CUDA(cudaMalloc(&ss, slot_sample.size() * sizeof(int32_t) + 16));
            SKB_CUDA(cudaMalloc(&so, slot_out.size() * sizeof(int32_t) + 16));
            SKB_CUDA(cudaMemcpy(ss, slot_sample.data(), slot_sample.size() * sizeof(int32_t), cudaMemcpyHostToDevice));
            SKB_CUDA(cudaMemcpy(so, slot_out.data(), slot_out.size() * sizeof(int32_t), cudaMemcpyHostToDevice));
        }
    } catch (...) { cudaFree(D); cudaFree(samples); cudaFree(Dbits); cudaFree(ss); cudaFree(so); cudaFree(plan); throw; }
    d.samples = samples; d.Dbits = Dbits; d.wht_slot_sample = ss; d.wht_slot_out = so; d.dct_plan = plan; d.D = D;
}

// UST_data_t::build (sketch/UST_data.hpp:81-102).  With replacement: samples[i] =
// uniform_int(0, N-1) at stream index base + i.  Without: the inside-out shuffle
//   for i in [0, N): j = uniform_int<int>(0, i) at stream index base + i; work[i] = work[j]; work[j] = i
// whose first S entries are the samples -- a sequential recurrence, so it runs on the host once per
// transform (integer arithmetic only: identical on host and device) and the S indices are copied.
static void ensure_ust_tables(sl_sketch_transform_t *t, DevState &d) {
    if (d.samples) return;
    std::vector<int32_t> smp((size_t)t->S);
    const uint64_t key = seed_to_key(t->seed);
    if (t->param != 0.0) {
        for (int i = 0; i < t->S; i++) {
            MicroUrng u(key, t->base0 + (uint64_t)i);
            smp[(size_t)i] = (int32_t)boost_uniform_int(u, (uint64_t)t->N - 1);
        }
    } else {
        std::vector<int32_t> work((size_t)t->N);
        for (int i = 0; i < t->N; i++) {
            MicroUrng u(key, t->base0 + (uint64_t)i);
            const int j = (int)boost_uniform_int(u, (uint64_t)i);
            work[(size_t)i] = work[(size_t)j];
            work[(size_t)j] = i;
        }
        for (int i = 0; i < t->S; i++) smp[(size_t)i] = work[(size_t)i];
    }
    int32_t *samples = nullptr;
    try {
        SKB_CUDA(cudaMalloc(&samples, smp.size() * sizeof(int32_t) + 16));
        SKB_CUDA(cudaMemcpy(samples, smp.data(), smp.size() * sizeof(int32_t), cudaMemcpyHostToDevice));
    } catch (...) { cudaFree(samples); throw; }
    d.samples = samples;
}

// UST_t::apply (sketch/UST_Elemental.hpp: sa[j*ldsa + i] = a[j*lda + samples[i]] columnwise,
// sa[j*ldsa + i] = a[samples[j]*lda + i] rowwise).  X holds sketched indices [offset, offset+nloc);
// samples outside that range give zeros, so the partial results of a partition add up.
// layout_rows: X[k*ldx + j], OUT[i*ldo + j]  (contiguous rows are copied whole);
// otherwise:   X[k + j*ldx], OUT[i + j*ldo].
__global__ void k_ust_gather(const double *__restrict__ X, size_t ldx, size_t m, const int32_t *__restrict__ samples,
                             size_t S, int64_t offset, int64_t nloc, double *__restrict__ OUT, size_t ldo,
                             int layout_rows) {
    const size_t total = S * m;
    for (size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (size_t)gridDim.x * blockDim.x) {
        size_t i, j;
        if (layout_rows) { i = t / m; j = t - i * m; } else { j = t / S; i = t - j * S; }
        const int64_t k = (int64_t)samples[i] - offset;
        const bool in = k >= 0 && k < nloc;
        const double v = in ? (layout_rows ? X[(size_t)k * ldx + j] : X[(size_t)k + j * ldx]) : 0.0;
        if (layout_rows) OUT[i * ldo + j] = v; else OUT[i + j * ldo] = v;
    }
}

static DenseParams dense_params(const sl_sketch_transform_t *t) {
    DenseParams p;
    p.seed = t->seed; p.base = t->base0; p.scale = t->scale;
    p.dist = (t->type == T_JLT || t->type == T_GRFT) ? DD_NORMAL : DD_CAUCHY;
    p.N = (size_t)t->N; p.S = (size_t)t->S;
    return p;
}

// ---------------------------------------------------------------------------
// Host -> device copy of a PAGEABLE buffer (numpy arrays, what the reference's callers pass).  One
// cudaMemcpyAsync from pageable memory runs at the driver's single-threaded staging rate (11.6 GB/s
// measured on the C1 operand).  Here a few worker threads copy 32 MiB chunks into a ring of pinned
// buffers and every chunk goes out as its own asynchronous copy, so the host-side memcpy of chunk
// c + 1 overlaps the transfer of chunk c.  Stream-ordered like cudaMemcpyAsync; the caller's buffer
// has been read completely when the function returns.  SKB_COPY_THREADS = 0 restores the plain copy.
// ---------------------------------------------------------------------------
namespace {
struct CopyPool {
    struct Task { char *dst; const char *src; size_t n; };
    std::vector<std::thread> workers;
    std::mutex mu;
    std::condition_variable cv, done_cv;
    std::vector<Task> q;
    size_t pending = 0;
    explicit CopyPool(int n) {
        for (int i = 0; i < n; i++)
            workers.emplace_back([this] {
                for (;;) {
                    Task t;
                    {
                        std::unique_lock<std::mutex> lk(mu);
                        cv.wait(lk, [this] { return !q.empty(); });
                        t = q.back(); q.pop_back();
                    }
                    memcpy(t.dst, t.src, t.n);
                    {
                        std::lock_guard<std::mutex> lk(mu);
                        if (--pending == 0) done_cv.notify_all();
                    }
                }
            });
        for (auto &w : workers) w.detach();        // the pool lives as long as the process
    }
    void run(const std::vector<Task> &tasks) {               // returns when every task has been copied
        if (tasks.empty()) return;
        {
            std::lock_guard<std::mutex> lk(mu);
            for (const Task &t : tasks) { q.push_back(t); pending++; }
        }
        cv.notify_all();
        std::unique_lock<std::mutex> lk(mu);
        done_cv.wait(lk, [this] { return pending == 0; });
    }
    void copy(char *dst, const char *src, size_t n) {        // returns when all of [src, src + n) is in dst
        const size_t parts = workers.size(), piece = ((n + parts - 1) / parts + 4095) / 4096 * 4096;
        {
            std::lock_guard<std::mutex> lk(mu);
            for (size_t o = 0; o < n; o += piece) { q.push_back({dst + o, src + o, std::min(piece, n - o)}); pending++; }
        }
        cv.notify_all();
        std::unique_lock<std::mutex> lk(mu);
        done_cv.wait(lk, [this] { return pending == 0; });
    }
};
struct PinnedRing {
    static constexpr int NBUF = 3;
    static constexpr size_t CHUNK = (size_t)32 << 20;
    char *buf[NBUF] = {nullptr, nullptr, nullptr};
    cudaEvent_t sent[NBUF] = {nullptr, nullptr, nullptr};
    bool in_flight[NBUF] = {false, false, false};
};
}  // namespace

static int copy_threads() {
    static int n = -1;
    if (n < 0) {
        const char *e = getenv("SKB_COPY_THREADS");
        const unsigned hw = std::thread::hardware_concurrency();
        n = (e && *e) ? atoi(e) : (int)std::min(8u, std::max(2u, hw / 2));
        if (n < 0) n = 0;
        if (n > 32) n = 32;
    }
    return n;
}

static bool host_pointer_is_pageable(const void *p) {
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, p) == cudaSuccess) return at.type == cudaMemoryTypeUnregistered;
    cudaGetLastError();
    return true;
}
static CopyPool *copy_pool() {
    static CopyPool *pool = new CopyPool(copy_threads());     // never destroyed: no join at exit
    return pool;
}
static PinnedRing *pinned_ring(int device) {                  // nullptr: no pinned memory to be had
    static PinnedRing rings[64];
    PinnedRing &r = rings[device];
    if (!r.buf[0]) {
        for (int b = 0; b < PinnedRing::NBUF; b++) {
            if (cudaHostAlloc((void **)&r.buf[b], PinnedRing::CHUNK, cudaHostAllocDefault) != cudaSuccess) {
                cudaGetLastError();
                for (int z = 0; z < b; z++) { cudaFreeHost(r.buf[z]); r.buf[z] = nullptr; }
                r.buf[0] = nullptr;
                return nullptr;
            }
            SKB_CUDA(cudaEventCreateWithFlags(&r.sent[b], cudaEventDisableTiming));
        }
    }
    return &r;
}
// `height` pieces of `width` bytes, `spitch` apart in pageable host memory, packed on the device
// (what cudaMemcpy2DAsync with dpitch == width does)
static void h2d_2d_from_pageable(void *dst_dev, const void *src_host, size_t spitch, size_t width, size_t height,
                                 int device, cudaStream_t st) {
    PinnedRing *r = nullptr;
    if (host_pointer_is_pageable(src_host) && copy_threads() >= 2 && width * height >= 2 * PinnedRing::CHUNK &&
        width <= PinnedRing::CHUNK && device >= 0 && device < 64)
        r = pinned_ring(device);
    if (!r) {
        SKB_CUDA(cudaMemcpy2DAsync(dst_dev, width, src_host, spitch, width, height, cudaMemcpyHostToDevice, st));
        return;
    }
    const size_t per = std::max<size_t>(1, PinnedRing::CHUNK / width);       // pieces per chunk
    std::vector<CopyPool::Task> tasks;
    size_t c = 0;
    for (size_t h0 = 0; h0 < height; h0 += per, c++) {
        const int b = (int)(c % PinnedRing::NBUF);
        const size_t nh = std::min(per, height - h0);
        if (r->in_flight[b]) SKB_CUDA(cudaEventSynchronize(r->sent[b]));
        tasks.clear();
        // pieces of at least 1 MiB per task: whole pieces grouped, or one piece cut
        const size_t group = std::max<size_t>(1, ((size_t)1 << 20) / width);
        for (size_t i = 0; i < nh; i += group)
            for (size_t j = i; j < std::min(nh, i + group); j++)
                tasks.push_back({r->buf[b] + j * width, (const char *)src_host + (h0 + j) * spitch, width});
        copy_pool()->run(tasks);
        SKB_CUDA(cudaMemcpyAsync((char *)dst_dev + h0 * width, r->buf[b], nh * width, cudaMemcpyHostToDevice, st));
        SKB_CUDA(cudaEventRecord(r->sent[b], st));
        r->in_flight[b] = true;
    }
}

static void h2d_from_pageable(void *dst_dev, const void *src_host, size_t bytes, int device, cudaStream_t st) {
    PinnedRing *rp = nullptr;        // pinned / registered memory goes out directly at PCIe speed
    if (host_pointer_is_pageable(src_host) && copy_threads() >= 2 && bytes >= 2 * PinnedRing::CHUNK &&
        device >= 0 && device < 64)
        rp = pinned_ring(device);
    if (!rp) {
        SKB_CUDA(cudaMemcpyAsync(dst_dev, src_host, bytes, cudaMemcpyHostToDevice, st));
        return;
    }
    PinnedRing &r = *rp;
    CopyPool *pool = copy_pool();
    size_t c = 0;
    for (size_t o = 0; o < bytes; o += PinnedRing::CHUNK, c++) {
        const int b = (int)(c % PinnedRing::NBUF);
        const size_t n = std::min(PinnedRing::CHUNK, bytes - o);
        if (r.in_flight[b]) SKB_CUDA(cudaEventSynchronize(r.sent[b]));     // its last transfer has left the buffer
        pool->copy(r.buf[b], (const char *)src_host + o, n);
        SKB_CUDA(cudaMemcpyAsync((char *)dst_dev + o, r.buf[b], n, cudaMemcpyHostToDevice, st));
        SKB_CUDA(cudaEventRecord(r.sent[b], st));
        r.in_flight[b] = true;
    }
}

// The mirror image for results (feature maps return far more than they take: 500k x 16384 floats from
// 500k x 784): chunks come back into the pinned ring and the worker threads move chunk c into the
// caller's pageable buffer while chunk c + 1 is on the wire.  Synchronous: the result is in dst_host
// when the function returns.
static void d2h_to_pageable(void *dst_host, const void *src_dev, size_t bytes, int device, cudaStream_t st) {
    PinnedRing *rp = nullptr;
    if (host_pointer_is_pageable(dst_host) && copy_threads() >= 2 && bytes >= 2 * PinnedRing::CHUNK &&
        device >= 0 && device < 64)
        rp = pinned_ring(device);
    if (!rp) {
        SKB_CUDA(cudaMemcpyAsync(dst_host, src_dev, bytes, cudaMemcpyDeviceToHost, st));
        SKB_CUDA(cudaStreamSynchronize(st));
        return;
    }
    PinnedRing &r = *rp;
    CopyPool *pool = copy_pool();
    for (int b = 0; b < PinnedRing::NBUF; b++)
        if (r.in_flight[b]) { SKB_CUDA(cudaEventSynchronize(r.sent[b])); r.in_flight[b] = false; }
    const size_t nchunks = (bytes + PinnedRing::CHUNK - 1) / PinnedRing::CHUNK;
    auto issue = [&](size_t c) {
        const int b = (int)(c % PinnedRing::NBUF);
        const size_t o = c * PinnedRing::CHUNK, n = std::min(PinnedRing::CHUNK, bytes - o);
        SKB_CUDA(cudaMemcpyAsync(r.buf[b], (const char *)src_dev + o, n, cudaMemcpyDeviceToHost, st));
        SKB_CUDA(cudaEventRecord(r.sent[b], st));
    };
    for (size_t c = 0; c < nchunks && c + 1 < (size_t)PinnedRing::NBUF; c++) issue(c);     // NBUF - 1 chunks ahead
    for (size_t c = 0; c < nchunks; c++) {
        const int b = (int)(c % PinnedRing::NBUF);
        if (c + PinnedRing::NBUF - 1 < nchunks) issue(c + PinnedRing::NBUF - 1);           // into the buffer drained last turn
        const size_t o = c * PinnedRing::CHUNK, n = std::min(PinnedRing::CHUNK, bytes - o);
        SKB_CUDA(cudaEventSynchronize(r.sent[b]));
        pool->copy((char *)dst_host + o, r.buf[b], n);
    }
}

// ---------------------------------------------------------------------------
// apply on device-resident operands
// ---------------------------------------------------------------------------
static void apply_device(sl_sketch_transform_t *t, const sl_matrix_t *A, const sl_matrix_t *SA,
                         int dim, int64_t offset, bool shard, bool accumulate);

// Dense transform (JLT / CT / RFT) of a CSC operand -- the (T, SparseMatrix, Matrix) combinations of
// capi/csketch.cpp:80,100,177,197.  Blocks of the sketched index are expanded to dense panels in
// scratch and go through the dense kernels, panel results accumulating in SA; the work is that of
// the dense sketch of the same shape (the reference realises all of S and multiplies, also O(s)
// per stored entry plus the generation of S).
__global__ void k_csc_expand_rows(const int32_t *__restrict__ indptr, const int32_t *__restrict__ indices,
                                  const double *__restrict__ values, size_t ncols, int64_t k0, int64_t kc,
                                  double *__restrict__ X) {       // X[(row-k0) + col*kc]
    const int lane = threadIdx.x & 31;
    size_t warp = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarps = ((size_t)gridDim.x * blockDim.x) >> 5;
    for (size_t j = warp; j < ncols; j += nwarps)
        for (int p = indptr[j] + lane; p < indptr[j + 1]; p += 32) {
            const int64_t r = indices[p];
            if (r >= k0 && r < k0 + kc) atomicAdd(X + (r - k0) + j * kc, values[p]);
        }
}
__global__ void k_csc_expand_cols(const int32_t *__restrict__ indptr, const int32_t *__restrict__ indices,
                                  const double *__restrict__ values, int64_t k0, int64_t kc, size_t ld,
                                  double *__restrict__ X) {       // X[(col-k0)*ld + row]
    const int lane = threadIdx.x & 31;
    size_t warp = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarps = ((size_t)gridDim.x * blockDim.x) >> 5;
    for (size_t kk = warp; kk < (size_t)kc; kk += nwarps)
        for (int p = indptr[k0 + kk] + lane; p < indptr[k0 + kk + 1]; p += 32)
            atomicAdd(X + kk * ld + indices[p], values[p]);
}

static void apply_dense_on_sparse(sl_sketch_transform_t *t, const sl_matrix_t *A, const sl_matrix_t *SA,
                                  int dim, int64_t offset, bool shard) {
    const bool colwise = dim == SL_COLUMNWISE, rft = t->is_rft();
    const int64_t sk = colwise ? A->m : A->n, other = colwise ? A->n : A->m;
    if (offset < 0 || offset + sk > t->N || (!shard && sk != t->N))
        throw error(SKB_ERR_INVALID_PARAM, "apply: sketched dimension of A does not match N");
    if (SA->kind != MK_DENSE || SA->dtype != SL_DTYPE_F64)
        throw error(SKB_ERR_UNSUPPORTED_BASE, "apply: a dense transform of a sparse matrix needs a float64 dense SA");
    const int64_t sa_sk = colwise ? SA->m : SA->n, sa_other = colwise ? SA->n : SA->m;
    if (sa_sk != t->S || sa_other != other) throw error(SKB_ERR_INVALID_PARAM, "apply: SA has the wrong shape");
    if (sk == 0 || other == 0) return;
    cudaStream_t st = g_stream;
    static std::map<int, std::pair<void *, size_t>> scratch_per_dev;      // grow-only dense panel per device
    auto &sc = scratch_per_dev[A->device];
    const size_t budget = (size_t)1 << 30;
    int64_t kc = (int64_t)(budget / ((size_t)other * 8)) / 16 * 16;
    if (kc < 16) kc = 16;
    if (kc > sk) kc = sk;
    if (rft) {      // the cosine is not additive over panels: one panel or nothing
        if ((size_t)sk * (size_t)other * 8 > ((size_t)16 << 30))
            throw error(SKB_ERR_ALLOC, "apply: RFT of a sparse matrix expands it to dense; this one is too large");
        kc = sk;
    }
    const size_t need = (size_t)kc * (size_t)other * 8;
    if (sc.second < need) {
        SKB_CUDA(cudaStreamSynchronize(st));
        cudaFree(sc.first); sc.first = nullptr; sc.second = 0;
        SKB_CUDA(cudaMalloc(&sc.first, need));
        sc.second = need;
    }
    if (!rft)       // panels accumulate
        SKB_CUDA(cudaMemset2DAsync(SA->data, (size_t)SA->ld * 8, 0, (size_t)SA->m * 8, (size_t)SA->n, st));
    for (int64_t k0 = 0; k0 < sk; k0 += kc) {
        const int64_t kn = sk - k0 < kc ? sk - k0 : kc;
        SKB_CUDA(cudaMemsetAsync(sc.first, 0, (size_t)kn * (size_t)other * 8, st));
        sl_matrix_t dA = *A;
        dA.kind = MK_DENSE; dA.data = sc.first; dA.dtype = SL_DTYPE_F64;
        {
            LaunchScope ls(KC_MISC, st);
            if (colwise) {
                dA.m = kn; dA.n = other; dA.ld = kn;
                k_csc_expand_rows<<<148 * 8, 256, 0, st>>>(A->indptr, A->indices, A->values, (size_t)A->n, k0, kn,
                                                           (double *)sc.first);
            } else {
                dA.m = other; dA.n = kn; dA.ld = other;
                k_csc_expand_cols<<<148 * 8, 256, 0, st>>>(A->indptr, A->indices, A->values, k0, kn, (size_t)other,
                                                           (double *)sc.first);
            }
            SKB_CUDA(cudaGetLastError());
        }
        apply_device(t, &dA, SA, dim, offset + k0, true, !rft);
    }
}

static void apply_device(sl_sketch_transform_t *t, const sl_matrix_t *A, const sl_matrix_t *SA,
                         int dim, int64_t offset, bool shard, bool accumulate) {
    const bool colwise = dim == SL_COLUMNWISE;
    const int64_t sk = colwise ? A->m : A->n;        // sketched extent present in A
    const int64_t other = colwise ? A->n : A->m;
    if (offset < 0 || offset + sk > t->N || (!shard && sk != t->N))
        throw error(SKB_ERR_INVALID_PARAM, "apply: sketched dimension of A does not match N");
    if (SA->kind != MK_DENSE) throw error(SKB_ERR_UNSUPPORTED_BASE, "apply: sparse output goes through apply_sparse_out");
    const int64_t sa_sk = colwise ? SA->m : SA->n, sa_other = colwise ? SA->n : SA->m;
    if (sa_sk != t->S || sa_other != other)
        throw error(SKB_ERR_INVALID_PARAM, "apply: SA has the wrong shape");
    if (A->dtype != SA->dtype) throw error(SKB_ERR_INVALID_PARAM, "apply: A and SA differ in dtype");
    const bool f32 = A->dtype == SL_DTYPE_F32;
    if (f32 && !(t->is_dense() || t->is_rft()))
        throw error(SKB_ERR_UNSUPPORTED_BASE, "apply: float32 is supported for JLT/CT/RFT only");
    DevState &d = dev_state(t, A->device);
    cudaStream_t st = g_stream;
    const bool layout_rows = !colwise;   // rowwise on column-major == contiguous rows per index

    if (accumulate && !t->is_dense())
        throw error(SKB_ERR_INVALID_USAGE, "apply: accumulation is implemented for JLT/CT only");
    if (A->kind == MK_SPARSE) {
        if (t->is_dense() || t->is_rft()) { apply_dense_on_sparse(t, A, SA, dim, offset, shard); return; }
        if (!t->is_hash()) throw error(SKB_ERR_UNSUPPORTED_BASE, "apply: FJLT and UST take dense operands only");
        ensure_hash_tables(t, d);
        const int32_t *b = d.bucket + offset; const double *v = d.value + offset;
        if (colwise)
            hash_apply_csc_columnwise(A->indptr, A->indices, A->values, (size_t)A->n, b, v,
                                      (double *)SA->data, (size_t)SA->ld, t->S, st, (size_t)A->nnz);
        else
            hash_apply_csc_rowwise(A->indptr, A->indices, A->values, (size_t)A->m, (size_t)A->n, b,
                                   v, (double *)SA->data, (size_t)SA->ld, t->S, st);
        return;
    }
    if (t->is_hash()) {
        ensure_hash_tables(t, d);
        if (layout_rows) {
            auto key = std::make_pair((size_t)offset, (size_t)sk);
            auto it = d.plans.find(key);
            if (it == d.plans.end()) {
                HashPlan plan;
                hash_build_plan(d.bucket + offset, (size_t)sk, t->S, &plan, st);
                it = d.plans.emplace(key, plan).first;
            }
            hash_apply_rows((const double *)A->data, (size_t)A->ld, (size_t)other, it->second,
                            d.value + offset, (double *)SA->data, (size_t)SA->ld, t->S, st);
        } else {
            // strided layout: the atomics-free planned kernel when the shape allows, for operands
            // large enough to pay for building the plan once (host side, O(n))
            bool done = false;
            if ((size_t)sk * (size_t)other >= ((size_t)1 << 20) && t->S >= 256 && t->S <= 23552) {
                auto key = std::make_pair((size_t)offset, (size_t)sk);
                auto it = d.cplans.find(key);
                if (it == d.cplans.end()) {
                    HashColsPlan plan;
                    hash_build_cols_plan(d.bucket + offset, d.value + offset, (size_t)sk, t->S, &plan, st);
                    it = d.cplans.emplace(key, plan).first;
                }
                done = hash_apply_cols_planned((const double *)A->data, (size_t)A->ld, (size_t)sk, (size_t)other,
                                               it->second, (double *)SA->data, (size_t)SA->ld, t->S, st);
            }
            if (!done)
                hash_apply_cols((const double *)A->data, (size_t)A->ld, (size_t)sk, (size_t)other,
                                d.bucket + offset, d.value + offset, (double *)SA->data,
                                (size_t)SA->ld, t->S, st);
        }
        return;
    }
    if (t->is_dense() || t->is_rft()) {
        DenseParams p = dense_params(t);
        RftEpilogue epi;
        if (t->is_rft()) {
            if (shard && sk != t->N)
                throw error(SKB_ERR_INVALID_USAGE, "apply: RFT cannot be applied to a shard of the "
                                                   "sketched dimension (the cosine is not additive)");
            ensure_rft_tables(t, d);
            epi.scales = d.scales; epi.shifts = d.shifts; epi.outscale = t->outscale;
        }
        size_t need;
        if (f32 && !layout_rows && !accumulate &&
            dense_tf32_path_ok(p, A->data, (size_t)A->ld, (size_t)other, (size_t)sk))
            need = dense_tf32_workspace_bytes(p, (size_t)sk, (size_t)other);      // no fp64 image of SA needed
        else
            need = dense_workspace_bytes(p, (size_t)sk, layout_rows, f32, (size_t)other, (size_t)SA->ld);
        ensure_ws(d, need);
        dense_apply(p, layout_rows, f32, A->data, (size_t)A->ld, (size_t)other, (size_t)offset,
                    (size_t)sk, SA->data, (size_t)SA->ld, epi, d.ws, d.ws_bytes, st, accumulate);
        return;
    }
    if (t->type == T_UST) {
        ensure_ust_tables(t, d);
        const size_t total = (size_t)t->S * (size_t)other;
        if (total == 0) return;
        size_t blocks = (total + 255) / 256;
        if (blocks > (size_t)148 * 32) blocks = (size_t)148 * 32;
        LaunchScope ls(KC_MISC, st);
        k_ust_gather<<<(unsigned)blocks, 256, 0, st>>>((const double *)A->data, (size_t)A->ld, (size_t)other, d.samples,
                                                       (size_t)t->S, offset, sk, (double *)SA->data, (size_t)SA->ld,
                                                       layout_rows ? 1 : 0);
        SKB_CUDA(cudaGetLastError());
        return;
    }
    // FJLT.  Row shards are possible in the Walsh-Hadamard mode only (H_N = H_P (x) H_{N/P}:
    // the partial sketches of whole 8192-row segments add up); the DCT does not split that way
    // (the reference re-distributes instead, FJLT_Elemental.hpp:149-150) -- shard its columns.
    if (sk != t->N && !fjlt_shard_ok(t->fut, (size_t)t->N, (size_t)offset, (size_t)sk))
        throw error(SKB_ERR_INVALID_USAGE,
                    "apply: FJLT over a shard of the sketched dimension needs fjlt_fut=wht, N a power of "
                    "two >= 8192 and 8192-aligned shards; in DCT mode shard the other dimension");
    ensure_fjlt_tables(t, d);
    size_t need = fjlt_workspace_bytes((size_t)t->N, (size_t)other, t->fut, layout_rows, (size_t)t->S);
    ensure_ws(d, need);
    fjlt_apply(t->fut, layout_rows, (const double *)A->data, (size_t)A->ld, (size_t)t->N,
               (size_t)other, d.D, d.Dbits, d.samples, (size_t)t->S, d.wht_slot_sample, d.wht_slot_out,
               (double *)SA->data, (size_t)SA->ld,
               d.ws, d.ws_bytes, st, (size_t)offset, (size_t)sk, d.dct_plan);
}

// Hands a freshly computed device CSC to the wrapper SA: kept on the device for device wrappers,
// copied into wrapper-owned host arrays otherwise.
static void attach_sparse_result(sl_matrix_t *SA, int32_t *oi, int32_t *ox, double *ov, size_t onnz,
                                 int64_t om, int64_t on, int device) {
    if (!SA->own) SA->own = new SpOwned();
    SpOwned &o = *SA->own;
    o.release_device();
    if (SA->on_device) {
        o.d_indptr = oi; o.d_indices = ox; o.d_values = ov; o.device = device;
        SA->indptr = oi; SA->indices = ox; SA->values = ov;
    } else {
        try {
            o.h_indptr.resize((size_t)on + 1); o.h_indices.resize(onnz); o.h_values.resize(onnz);
            SKB_CUDA(cudaMemcpyAsync(o.h_indptr.data(), oi, ((size_t)on + 1) * 4, cudaMemcpyDeviceToHost, g_stream));
            if (onnz) {
                SKB_CUDA(cudaMemcpyAsync(o.h_indices.data(), ox, onnz * 4, cudaMemcpyDeviceToHost, g_stream));
                SKB_CUDA(cudaMemcpyAsync(o.h_values.data(), ov, onnz * 8, cudaMemcpyDeviceToHost, g_stream));
            }
            SKB_CUDA(cudaStreamSynchronize(g_stream));
        } catch (...) { hash_sparse_free(oi, g_stream); hash_sparse_free(ox, g_stream); hash_sparse_free(ov, g_stream); throw; }
        hash_sparse_free(oi, g_stream); hash_sparse_free(ox, g_stream); hash_sparse_free(ov, g_stream);
        SA->indptr = o.h_indptr.data(); SA->indices = o.h_indices.data(); SA->values = o.h_values.data();
    }
    SA->m = om; SA->n = on; SA->nnz = (int64_t)onnz; SA->struct_updated = true;
}

// Sparse in, sparse out (CWT/MMT/WZT; sketch/hash_transform_local_sparse.hpp:88-223).  A is
// device resident; the result replaces SA's structure, on the device for device wrappers and
// in wrapper-owned host arrays otherwise (read back through sl_raw_sp_matrix_*).
static void apply_sparse_out(sl_sketch_transform_t *t, const sl_matrix_t *A, sl_matrix_t *SA,
                             int dim, int64_t offset, bool shard, int device) {
    const bool colwise = dim == SL_COLUMNWISE;
    if (!t->is_hash() || A->kind != MK_SPARSE)
        throw error(SKB_ERR_UNSUPPORTED_BASE, "apply: sparse output needs CWT/MMT/WZT and a sparse input");
    const int64_t sk = colwise ? A->m : A->n;
    if (offset < 0 || offset + sk > t->N || (!shard && sk != t->N))
        throw error(SKB_ERR_INVALID_PARAM, "apply: sketched dimension of A does not match N");
    DevState &d = dev_state(t, device);
    ensure_hash_tables(t, d);
    int32_t *oi = nullptr, *ox = nullptr; double *ov = nullptr; size_t onnz = 0;
    hash_apply_csc_sparse(A->indptr, A->indices, A->values, (size_t)A->m, (size_t)A->n,
                          (size_t)A->nnz, !colwise, d.bucket + offset, d.value + offset, t->S,
                          &oi, &ox, &ov, &onnz, &d.ws, &d.ws_bytes, g_stream);
    attach_sparse_result(SA, oi, ox, ov, onnz, colwise ? t->S : A->m, colwise ? A->n : t->S, device);
}

// Host operands of a dense sketch: the sketched dimension is cut in chunks; chunk c+1 is
// copied to the device (second stream, double-buffered staging) while chunk c is sketched
// into the device-resident SA with global input indices -- the same additivity the
// row-sharded multi-GPU path uses.  Staging buffers persist per device (grow-only).
struct HostStage {
    void *buf[2] = {nullptr, nullptr}; size_t cap = 0;
    void *out = nullptr; size_t out_cap = 0;
    cudaStream_t copy = nullptr;
    cudaEvent_t copied[2] = {nullptr, nullptr}, consumed[2] = {nullptr, nullptr}, start = nullptr;
};
static HostStage &host_stage(int dev) {
    static HostStage per_dev[64];
    if (dev < 0 || dev >= 64) throw error(SKB_ERR_GENERIC, "device index out of range");
    HostStage &h = per_dev[dev];
    if (!h.copy) {
        SKB_CUDA(cudaStreamCreateWithFlags(&h.copy, cudaStreamNonBlocking));
        cudaEvent_t *evs[] = {&h.copied[0], &h.copied[1], &h.consumed[0], &h.consumed[1], &h.start};
        for (cudaEvent_t *e : evs) SKB_CUDA(cudaEventCreateWithFlags(e, cudaEventDisableTiming));
    }
    return h;
}

static void apply_host_dense_pipelined(sl_sketch_transform_t *t, sl_matrix_t *A, sl_matrix_t *SA,
                                       int dim, int64_t offset, bool shard, int device) {
    const bool colwise = dim == SL_COLUMNWISE;
    const int64_t sk = colwise ? A->m : A->n, other = colwise ? A->n : A->m;
    if (offset < 0 || offset + sk > t->N || (!shard && sk != t->N))
        throw error(SKB_ERR_INVALID_PARAM, "apply: sketched dimension of A does not match N");
    const int64_t sa_sk = colwise ? SA->m : SA->n, sa_other = colwise ? SA->n : SA->m;
    if (sa_sk != t->S || sa_other != other) throw error(SKB_ERR_INVALID_PARAM, "apply: SA has the wrong shape");
    cudaStream_t st = g_stream;
    HostStage &h = host_stage(device);
    const size_t chunk_bytes_target = (size_t)256 << 20;
    int64_t krows = (int64_t)(chunk_bytes_target / ((size_t)other * 8));
    krows = krows / 16 * 16;
    if (krows < 16) krows = 16;
    if (krows > sk) krows = sk;
    const size_t need = (size_t)krows * (size_t)other * 8;
    if (h.cap < need) {
        SKB_CUDA(cudaStreamSynchronize(st));
        SKB_CUDA(cudaStreamSynchronize(h.copy));
        for (int b = 0; b < 2; b++) { cudaFree(h.buf[b]); h.buf[b] = nullptr; }
        h.cap = 0;
        for (int b = 0; b < 2; b++) SKB_CUDA(cudaMalloc(&h.buf[b], need));
        h.cap = need;
    }
    const size_t obytes = (size_t)SA->ld * (size_t)SA->n * 8;
    if (h.out_cap < obytes) {
        SKB_CUDA(cudaStreamSynchronize(st));
        cudaFree(h.out); h.out = nullptr; h.out_cap = 0;
        SKB_CUDA(cudaMalloc(&h.out, obytes ? obytes : 16));
        h.out_cap = obytes;
    }
    sl_matrix_t dSA = *SA;
    dSA.on_device = true; dSA.device = device; dSA.data = h.out;
    SKB_CUDA(cudaMemsetAsync(h.out, 0, obytes, st));
    SKB_CUDA(cudaEventRecord(h.start, st));
    SKB_CUDA(cudaStreamWaitEvent(h.copy, h.start, 0));
    int c = 0;
    for (int64_t k0 = 0; k0 < sk; k0 += krows, c++) {
        const int64_t kc = sk - k0 < krows ? sk - k0 : krows;
        const int b = c & 1;
        if (c >= 2) SKB_CUDA(cudaStreamWaitEvent(h.copy, h.consumed[b], 0));
        sl_matrix_t dA = *A;
        dA.on_device = true; dA.device = device; dA.data = h.buf[b];
        if (colwise) {     // A is sk x other column-major: rows [k0, k0+kc) of every column
            dA.m = kc; dA.n = other; dA.ld = kc;
            h2d_2d_from_pageable(h.buf[b], (const double *)A->data + k0, (size_t)A->ld * 8, (size_t)kc * 8,
                                 (size_t)other, device, h.copy);
        } else {           // A is other x sk column-major: columns [k0, k0+kc) are contiguous
            dA.m = other; dA.n = kc; dA.ld = A->ld;
            h2d_from_pageable(h.buf[b], (const double *)A->data + (size_t)k0 * A->ld,
                              (size_t)kc * (size_t)A->ld * 8, device, h.copy);
        }
        SKB_CUDA(cudaEventRecord(h.copied[b], h.copy));
        SKB_CUDA(cudaStreamWaitEvent(st, h.copied[b], 0));
        apply_device(t, &dA, &dSA, dim, offset + k0, true, true);
        SKB_CUDA(cudaEventRecord(h.consumed[b], st));
    }
    SKB_CUDA(cudaMemcpyAsync(SA->data, h.out, obytes, cudaMemcpyDeviceToHost, st));
    SKB_CUDA(cudaStreamSynchronize(st));
}

// Host operands: stage through device memory (copies inside the call, synchronous).
static void apply_any(sl_sketch_transform_t *t, sl_matrix_t *A, sl_matrix_t *SA, int dim,
                      int64_t offset, bool shard) {
    require_device();
    if (dim != SL_COLUMNWISE && dim != SL_ROWWISE)
        throw error(SKB_ERR_INVALID_PARAM, "apply: dim must be SL_COLUMNWISE or SL_ROWWISE");
    if (A->on_device != SA->on_device)
        throw error(SKB_ERR_INVALID_USAGE, "apply: A and SA must both be host or both be device buffers");
    if (A->on_device) {
        if (A->device != SA->device) throw error(SKB_ERR_INVALID_USAGE, "apply: A and SA are on different devices");
        DeviceGuard g(A->device);
        if (SA->kind == MK_SPARSE) apply_sparse_out(t, A, SA, dim, offset, shard, A->device);
        else apply_device(t, A, SA, dim, offset, shard, false);
        return;
    }
    if (SA->kind == MK_SPARSE && A->kind != MK_SPARSE)
        throw error(SKB_ERR_UNSUPPORTED_BASE, "apply: sparse output needs CWT/MMT/WZT and a sparse input");
    int device = 0;
    SKB_CUDA(cudaGetDevice(&device));
    cudaStream_t st = g_stream;
    if (t->is_dense() && A->kind == MK_DENSE && A->dtype == SL_DTYPE_F64 && SA->dtype == SL_DTYPE_F64 &&
        (size_t)A->m * (size_t)A->n * 8 > ((size_t)64 << 20)) {
        apply_host_dense_pipelined(t, A, SA, dim, offset, shard, device);
        return;
    }
    // Staging buffers come from a grow-only pool per device: a host-operand apply costs its copies
    // and its kernels, not four cudaMalloc / cudaFree round trips (each a device synchronisation).
    sl_matrix_t dA = *A, dSA = *SA;
    dA.on_device = dSA.on_device = true; dA.device = dSA.device = device;
    struct Slot { void *p = nullptr; size_t cap = 0; };
    static Slot pool[64][4];
    if (device < 0 || device >= 64) throw error(SKB_ERR_GENERIC, "device index out of range");
    auto stage = [&](int slot, size_t bytes) -> void * {
        Slot &sl = pool[device][slot];
        if (sl.cap < bytes || !sl.p) {
            SKB_CUDA(cudaStreamSynchronize(st));
            cudaFree(sl.p); sl.p = nullptr; sl.cap = 0;
            const size_t want = bytes + bytes / 8 + 256;       // some headroom for the next, slightly larger call
            SKB_CUDA(cudaMalloc(&sl.p, want));
            sl.cap = want;
        }
        return sl.p;
    };
    try {
        const size_t esz = A->dtype == SL_DTYPE_F32 ? 4 : 8;
        if (A->kind == MK_DENSE) {
            size_t bytes = (size_t)A->ld * (size_t)A->n * esz;
            dA.data = stage(0, bytes);
            h2d_from_pageable(dA.data, A->data, bytes, device, st);
        } else {
            size_t pb = (size_t)(A->n + 1) * 4, ib = (size_t)A->nnz * 4, vb = (size_t)A->nnz * 8;
            dA.indptr = (int32_t *)stage(0, pb); dA.indices = (int32_t *)stage(1, ib); dA.values = (double *)stage(2, vb);
            SKB_CUDA(cudaMemcpyAsync(dA.indptr, A->indptr, pb, cudaMemcpyHostToDevice, st));
            h2d_from_pageable(dA.indices, A->indices, ib, device, st);
            h2d_from_pageable(dA.values, A->values, vb, device, st);
        }
        if (SA->kind == MK_SPARSE) {
            apply_sparse_out(t, &dA, SA, dim, offset, shard, device);
            return;
        }
        size_t obytes = (size_t)SA->ld * (size_t)SA->n * (SA->dtype == SL_DTYPE_F32 ? 4 : 8);
        dSA.data = stage(3, obytes);
        apply_device(t, &dA, &dSA, dim, offset, shard, false);
        d2h_to_pageable(SA->data, dSA.data, obytes, device, st);
    } catch (...) {
        cudaStreamSynchronize(st);
        throw;
    }
}

// ---------------------------------------------------------------------------
// tiny JSON (boost::property_tree::write_json style: every value is a string)
// ---------------------------------------------------------------------------
static void json_skip(const char *&p) { while (*p == ' ' || *p == '\n' || *p == '\t' || *p == '\r') p++; }
static std::string json_string(const char *&p) {
    std::string s;
    if (*p != '"') throw error(SKB_ERR_IO, "malformed JSON: expected string");
    p++;
    while (*p && *p != '"') { if (*p == '\\' && p[1]) p++; s.push_back(*p++); }
    if (*p != '"') throw error(SKB_ERR_IO, "malformed JSON: unterminated string");
    p++;
    return s;
}
static void json_object(const char *&p, const std::string &prefix, std::map<std::string, std::string> &out) {
    json_skip(p);
    if (*p != '{') throw error(SKB_ERR_IO, "malformed JSON: expected object");
    p++;
    for (;;) {
        json_skip(p);
        if (*p == '}') { p++; return; }
        std::string key = json_string(p);
        json_skip(p);
        if (*p != ':') throw error(SKB_ERR_IO, "malformed JSON: expected ':'");
        p++;
        json_skip(p);
        if (*p == '{') json_object(p, prefix + key + ".", out);
        else if (*p == '"') out[prefix + key] = json_string(p);
        else {
            std::string v;
            while (*p && *p != ',' && *p != '}' && *p != ' ' && *p != '\n' && *p != '\r' && *p != '\t') v.push_back(*p++);
            out[prefix + key] = v;
        }
        json_skip(p);
        if (*p == ',') { p++; continue; }
        if (*p == '}') { p++; return; }
        throw error(SKB_ERR_IO, "malformed JSON: expected ',' or '}'");
    }
}
static const std::string &json_get(const std::map<std::string, std::string> &m, const char *k) {
    auto it = m.find(k);
    if (it == m.end()) throw error(SKB_ERR_IO, std::string("serialized sketch lacks key ") + k);
    return it->second;
}

static char *dup_cstr(const std::string &s) {
    char *r = (char *)malloc(s.size() + 1);
    if (r) memcpy(r, s.c_str(), s.size() + 1);
    return r;
}

static sl_matrix_t *as_matrix(void *p) {
    if (!p) throw error(SKB_ERR_INVALID_PARAM, "null matrix handle");
    return (sl_matrix_t *)p;
}

static void check_type_string(const char *s, const sl_matrix_t *M) {
    if (!s) throw error(SKB_ERR_INVALID_PARAM, "null matrix type string");
    bool dense = strcmp(s, "Matrix") == 0, sparse = strcmp(s, "SparseMatrix") == 0;
    if (!dense && !sparse)
        throw error(SKB_ERR_UNSUPPORTED_DIST, std::string("unsupported matrix type: ") + s);
    if ((dense && M->kind != MK_DENSE) || (sparse && M->kind != MK_SPARSE))
        throw error(SKB_ERR_INVALID_PARAM, "matrix type string does not match the wrapped object");
}

// ---------------------------------------------------------------------------
// exported functions
// ---------------------------------------------------------------------------
extern "C" {

char *sl_supported_sketch_transforms(void) {
    static char s[] =
        "(\"JLT\",\"Matrix\",\"Matrix\") (\"JLT\",\"SparseMatrix\",\"Matrix\") "
        "(\"CT\",\"Matrix\",\"Matrix\") (\"CT\",\"SparseMatrix\",\"Matrix\") "
        "(\"FJLT\",\"Matrix\",\"Matrix\") (\"UST\",\"Matrix\",\"Matrix\") "
        "(\"CWT\",\"Matrix\",\"Matrix\") (\"CWT\",\"SparseMatrix\",\"Matrix\") "
        "(\"CWT\",\"SparseMatrix\",\"SparseMatrix\") "
        "(\"MMT\",\"Matrix\",\"Matrix\") (\"MMT\",\"SparseMatrix\",\"Matrix\") "
        "(\"MMT\",\"SparseMatrix\",\"SparseMatrix\") "
        "(\"WZT\",\"Matrix\",\"Matrix\") (\"WZT\",\"SparseMatrix\",\"Matrix\") "
        "(\"WZT\",\"SparseMatrix\",\"SparseMatrix\") "
        "(\"GaussianRFT\",\"Matrix\",\"Matrix\") (\"GaussianRFT\",\"SparseMatrix\",\"Matrix\") "
        "(\"LaplacianRFT\",\"Matrix\",\"Matrix\") (\"LaplacianRFT\",\"SparseMatrix\",\"Matrix\") ";
    return s;
}

const char *sl_strerror(const int errorcode) {
    if (errorcode < 100 || errorcode > 113) return "Unknown error code";
    return k_errmsg[errorcode - 100];
}

bool sl_has_elemental(void) { return false; }
bool sl_has_combblas(void) { return false; }

int sl_create_default_context(int seed, sl_context_t **ctxt) {
    SL_TRY
    *ctxt = new sl_context_t{seed, 0};
    SL_CATCH
}
int sl_create_context(int seed, void *, sl_context_t **ctxt) {
    return sl_create_default_context(seed, ctxt);
}
int sl_create_context_at(int seed, uint64_t counter, sl_context_t **ctxt) {
    SL_TRY
    *ctxt = new sl_context_t{seed, counter};
    SL_CATCH
}
int sl_free_context(sl_context_t *ctxt) { delete ctxt; return 0; }
int sl_context_get_seed(sl_context_t *ctxt, int *seed) { *seed = ctxt->seed; return 0; }
int sl_context_get_counter(sl_context_t *ctxt, uint64_t *counter) { *counter = ctxt->counter; return 0; }
int sl_context_allocate(sl_context_t *ctxt, uint64_t size, uint64_t *base) {
    SL_TRY
    if (!ctxt || !base) throw error(SKB_ERR_INVALID_PARAM, "null argument");
    *base = ctxt->allocate(size);
    SL_CATCH
}

int sl_create_sketch_transform(sl_context_t *ctxt, char *type_, int n, int s,
                               sl_sketch_transform_t **sketch, ...) {
    TType type;
    if (!type_ || !str2type(type_, &type)) {
        g_last_error = std::string("unknown sketch transform type: ") + (type_ ? type_ : "(null)");
        return SKB_ERR_INVALID_PARAM;      // capi/csketch.cpp:385-386
    }
    double p1 = 0.0;
    if (type == T_CT || type == T_WZT || type == T_GRFT || type == T_LRFT) {
        va_list argp;
        va_start(argp, sketch);
        p1 = va_arg(argp, double);
        va_end(argp);
    } else if (type == T_UST) {       // `int replace` (UST_data_t::params_t, sketch/UST_data.hpp:27-33)
        va_list argp;
        va_start(argp, sketch);
        p1 = va_arg(argp, int) ? 1.0 : 0.0;
        va_end(argp);
    }
    SL_TRY
    if (!ctxt) throw error(SKB_ERR_INVALID_PARAM, "null context");
    *sketch = make_transform(type, type_, n, s, p1, *ctxt);
    SL_CATCH
}

int sl_serialize_sketch_transform(const sl_sketch_transform_t *t, char **data) {
    SL_TRY
    char buf[1024];
    std::string extra;
    if (t->type == T_CT) { snprintf(buf, sizeof buf, ",\n    \"C\": \"%.17g\"", t->param); extra = buf; }
    if (t->type == T_WZT) { snprintf(buf, sizeof buf, ",\n    \"P\": \"%.17g\"", t->param); extra = buf; }
    if (t->is_rft()) { snprintf(buf, sizeof buf, ",\n    \"sigma\": \"%.17g\"", t->param); extra = buf; }
    if (t->type == T_UST) extra = std::string(",\n    \"replace\": \"") + (t->param != 0.0 ? "true" : "false") + "\"";
    snprintf(buf, sizeof buf,
             "{\n    \"skylark_object_type\": \"sketch\",\n    \"sketch_type\": \"%s\",\n"
             "    \"skylark_version\": \"0.20\",\n    \"N\": \"%d\",\n    \"S\": \"%d\",\n"
             "    \"creation_context\": {\n        \"skylark_object_type\": \"context\",\n"
             "        \"skylark_version\": \"0.20\",\n        \"seed\": \"%d\",\n"
             "        \"counter\": \"%llu\"\n    }%s\n}\n",
             t->name.c_str(), t->N, t->S, t->seed, (unsigned long long)t->counter0, extra.c_str());
    *data = dup_cstr(buf);
    if (!*data) throw std::bad_alloc();
    SL_CATCH
}

int sl_deserialize_sketch_transform(const char *data, sl_sketch_transform_t **sketch) {
    SL_TRY
    if (!data) throw error(SKB_ERR_INVALID_PARAM, "null JSON");
    std::map<std::string, std::string> kv;
    const char *p = data;
    json_object(p, "", kv);
    const std::string &name = json_get(kv, "sketch_type");
    TType type;
    if (!str2type(name.c_str(), &type))
        throw error(SKB_ERR_SKETCH, "cannot deserialize sketch type " + name);   // sketch_add.hpp:15-54
    int n = (int)strtod(json_get(kv, "N").c_str(), nullptr);
    int s = (int)strtod(json_get(kv, "S").c_str(), nullptr);
    sl_context_t ctx;
    ctx.seed = (int)strtol(json_get(kv, "creation_context.seed").c_str(), nullptr, 10);
    ctx.counter = strtoull(json_get(kv, "creation_context.counter").c_str(), nullptr, 10);
    double p1 = 0.0;
    if (type == T_CT) p1 = strtod(json_get(kv, "C").c_str(), nullptr);
    if (type == T_WZT) p1 = strtod(json_get(kv, "P").c_str(), nullptr);
    if (type == T_GRFT || type == T_LRFT) p1 = strtod(json_get(kv, "sigma").c_str(), nullptr);
    if (type == T_UST) { const std::string &r = json_get(kv, "replace"); p1 = (r == "true" || r == "1") ? 1.0 : 0.0; }
    *sketch = make_transform(type, name.c_str(), n, s, p1, ctx);   // replays the RNG (CWT_data.hpp:48-52)
    SL_CATCH
}

int sl_free_sketch_transform(sl_sketch_transform_t *S) { delete S; return 0; }

int sl_apply_sketch_transform(sl_sketch_transform_t *S, char *input_type, void *A,
                              char *output_type, void *SA, int dim) {
    SL_TRY
    if (!S) throw error(SKB_ERR_INVALID_PARAM, "null sketch transform");
    sl_matrix_t *a = as_matrix(A), *sa = as_matrix(SA);
    check_type_string(input_type, a);
    check_type_string(output_type, sa);
    apply_any(S, a, sa, dim, 0, false);
    SL_CATCH
}

int sl_apply_sketch_transform_shard(sl_sketch_transform_t *S, char *input_type, void *A,
                                    char *output_type, void *SA, int dim, int64_t index_offset) {
    SL_TRY
    if (!S) throw error(SKB_ERR_INVALID_PARAM, "null sketch transform");
    sl_matrix_t *a = as_matrix(A), *sa = as_matrix(SA);
    check_type_string(input_type, a);
    check_type_string(output_type, sa);
    apply_any(S, a, sa, dim, index_offset, true);
    SL_CATCH
}

int sl_wrap_raw_matrix(double *data, int m, int n, void **A) {
    SL_TRY
    sl_matrix_t *M = new sl_matrix_t();
    M->kind = MK_DENSE; M->on_device = false; M->device = -1; M->dtype = SL_DTYPE_F64;
    M->data = data; M->m = m; M->n = n; M->ld = m;
    *A = M;
    SL_CATCH
}
int sl_free_raw_matrix_wrap(void *A_) { delete (sl_matrix_t *)A_; return 0; }

int sl_wrap_device_matrix(void *dptr, int64_t m, int64_t n, int64_t ld, int dtype, int device,
                          void **A) {
    SL_TRY
    if (m < 0 || n < 0 || ld < m || (dtype != SL_DTYPE_F64 && dtype != SL_DTYPE_F32))
        throw error(SKB_ERR_INVALID_PARAM, "sl_wrap_device_matrix: bad shape or dtype");
    sl_matrix_t *M = new sl_matrix_t();
    M->kind = MK_DENSE; M->on_device = true; M->device = device; M->dtype = dtype;
    M->data = dptr; M->m = m; M->n = n; M->ld = ld;
    *A = M;
    SL_CATCH
}

int sl_wrap_raw_sp_matrix(int *indptr, int *ind, double *data, int nnz, int n_rows, int n_cols,
                          void **A) {
    SL_TRY
    sl_matrix_t *M = new sl_matrix_t();
    M->kind = MK_SPARSE; M->on_device = false; M->device = -1; M->dtype = SL_DTYPE_F64;
    M->indptr = indptr; M->indices = ind; M->values = data; M->nnz = nnz; M->m = n_rows; M->n = n_cols;
    M->struct_updated = false;
    *A = M;
    SL_CATCH
}
int sl_wrap_device_sp_matrix(int32_t *indptr, int32_t *ind, double *data, int64_t nnz,
                             int64_t n_rows, int64_t n_cols, int device, void **A) {
    SL_TRY
    sl_matrix_t *M = new sl_matrix_t();
    M->kind = MK_SPARSE; M->on_device = true; M->device = device; M->dtype = SL_DTYPE_F64;
    M->indptr = indptr; M->indices = ind; M->values = data; M->nnz = nnz; M->m = n_rows; M->n = n_cols;
    M->struct_updated = false;
    *A = M;
    SL_CATCH
}
int sl_free_raw_sp_matrix_wrap(void *A_) {
    sl_matrix_t *M = (sl_matrix_t *)A_;
    if (M) delete M->own;
    delete M;
    return 0;
}
int sl_raw_sp_matrix_struct_updated(void *A_, bool *u) { *u = ((sl_matrix_t *)A_)->struct_updated; return 0; }
int sl_raw_sp_matrix_reset_update_flag(void *A_) { ((sl_matrix_t *)A_)->struct_updated = false; return 0; }
int sl_raw_sp_matrix_nnz(void *A_, int *nnz) { *nnz = (int)((sl_matrix_t *)A_)->nnz; return 0; }
int sl_raw_sp_matrix_height(void *A_, int *h) { *h = (int)((sl_matrix_t *)A_)->m; return 0; }
int sl_raw_sp_matrix_width(void *A_, int *w) { *w = (int)((sl_matrix_t *)A_)->n; return 0; }
int sl_raw_sp_matrix_data(void *A_, int32_t *indptr, int32_t *indices, double *values) {
    SL_TRY
    sl_matrix_t *M = as_matrix(A_);
    if (M->on_device) {     // device wrapper: the three destinations are device pointers too
        DeviceGuard g(M->device);
        SKB_CUDA(cudaMemcpyAsync(indptr, M->indptr, (size_t)(M->n + 1) * sizeof(int32_t), cudaMemcpyDeviceToDevice, g_stream));
        SKB_CUDA(cudaMemcpyAsync(indices, M->indices, (size_t)M->nnz * sizeof(int32_t), cudaMemcpyDeviceToDevice, g_stream));
        SKB_CUDA(cudaMemcpyAsync(values, M->values, (size_t)M->nnz * sizeof(double), cudaMemcpyDeviceToDevice, g_stream));
        SKB_CUDA(cudaStreamSynchronize(g_stream));
        return 0;
    }
    memcpy(indptr, M->indptr, (size_t)(M->n + 1) * sizeof(int32_t));
    memcpy(indices, M->indices, (size_t)M->nnz * sizeof(int32_t));
    memcpy(values, M->values, (size_t)M->nnz * sizeof(double));
    SL_CATCH
}

void sl_get_exception_info(char **info) { *info = dup_cstr(g_last_error); }
void sl_print_exception_trace(void) { fprintf(stderr, "%s\n", g_last_error.c_str()); }

int sl_set_stream(void *cuda_stream) { g_stream = (cudaStream_t)cuda_stream; return 0; }

int sl_set_option(const char *key, const char *value) {
    SL_TRY
    if (!key || !value) throw error(SKB_ERR_INVALID_PARAM, "null option");
    if (strcmp(key, "boost_order") == 0) {
        if (strcmp(value, "160") == 0) g_boost_order = SKB_ORDER_160;
        else if (strcmp(value, "155") == 0) g_boost_order = SKB_ORDER_155;
        else throw error(SKB_ERR_INVALID_PARAM, "boost_order must be 160 or 155");
    } else if (strcmp(key, "fjlt_fut") == 0) {
        if (strcmp(value, "dct") == 0) g_fjlt_fut = FUT_DCT;
        else if (strcmp(value, "wht") == 0) g_fjlt_fut = FUT_WHT;
        else throw error(SKB_ERR_INVALID_PARAM, "fjlt_fut must be dct or wht");
    } else if (strcmp(key, "sm_reserve") == 0) {
        // SMs the persistent FJLT kernels leave free (so that a collective running beside them,
        // dist.apply_distributed_overlapped, finds room for its CTAs); 0 = use the whole GPU
        const int v = atoi(value);
        if (v < 0 || v > 100) throw error(SKB_ERR_INVALID_PARAM, "sm_reserve must be in [0, 100]");
        fjlt_set_sm_reserve(v);
    } else if (strcmp(key, "csc_form") == 0) {
        // kernel form of the CSC -> dense hash sketch: "dump" (default) or "bulk" (slices of the
        // accumulator through the copy engine; same speed at C4, kept selectable for A/B and tests)
        if (strcmp(value, "dump") == 0) hash_set_csc_form(0);
        else if (strcmp(value, "bulk") == 0) hash_set_csc_form(1);
        else throw error(SKB_ERR_INVALID_PARAM, "csc_form must be dump or bulk");
    } else throw error(SKB_ERR_INVALID_PARAM, std::string("unknown option ") + key);
    SL_CATCH
}

int sl_debug_hash_cols_plan(const int32_t *bucket, const double *value, int64_t n, int s, uint32_t **entries,
                            int64_t *n_entries, uint32_t **off, int64_t *n_off, int *bucket_width) {
    SL_TRY
    if (!bucket || !value || n < 0 || s < 1 || !entries || !n_entries || !off || !n_off || !bucket_width)
        throw error(SKB_ERR_INVALID_PARAM, "sl_debug_hash_cols_plan: bad argument");
    std::vector<uint32_t> e, o;
    std::vector<double> v;
    bool signs = true;
    hash_schedule_cols_plan(bucket, value, (size_t)n, s, e, v, o, &signs);
    *entries = (uint32_t *)malloc(e.size() * sizeof(uint32_t) + 1);
    *off = (uint32_t *)malloc(o.size() * sizeof(uint32_t) + 1);
    if (!*entries || !*off) { free(*entries); free(*off); throw error(SKB_ERR_ALLOC, "out of host memory"); }
    memcpy(*entries, e.data(), e.size() * sizeof(uint32_t));
    memcpy(*off, o.data(), o.size() * sizeof(uint32_t));
    *n_entries = (int64_t)e.size(); *n_off = (int64_t)o.size();
    *bucket_width = hash_cols_plan_bucket_width(s);
    SL_CATCH
}

int sl_sketch_export(sl_sketch_transform_t *t, const char *what, int64_t first, int64_t count,
                     void *host_out) {
    SL_TRY
    require_device();
    if (!t || !what || first < 0 || count < 0) throw error(SKB_ERR_INVALID_PARAM, "sl_sketch_export: bad argument");
    int device = 0;
    SKB_CUDA(cudaGetDevice(&device));
    DevState &d = dev_state(t, device);
    cudaStream_t st = g_stream;
    auto copy_f64 = [&](const double *src, int64_t limit) {
        if (first + count > limit) throw error(SKB_ERR_RANDOM123, "Index is out of bounds");
        SKB_CUDA(cudaMemcpyAsync(host_out, src + first, (size_t)count * 8, cudaMemcpyDeviceToHost, st));
        SKB_CUDA(cudaStreamSynchronize(st));
    };
    auto copy_i32_as_i64 = [&](const int32_t *src, int64_t limit) {
        if (first + count > limit) throw error(SKB_ERR_RANDOM123, "Index is out of bounds");
        std::vector<int32_t> tmp((size_t)count);
        SKB_CUDA(cudaMemcpyAsync(tmp.data(), src + first, (size_t)count * 4, cudaMemcpyDeviceToHost, st));
        SKB_CUDA(cudaStreamSynchronize(st));
        for (int64_t i = 0; i < count; i++) ((int64_t *)host_out)[i] = tmp[(size_t)i];
    };
    if (strcmp(what, "row_idx") == 0 && t->is_hash()) { ensure_hash_tables(t, d); copy_i32_as_i64(d.bucket, t->N); }
    else if (strcmp(what, "row_value") == 0 && t->is_hash()) { ensure_hash_tables(t, d); copy_f64(d.value, t->N); }
    else if (strcmp(what, "D") == 0 && t->type == T_FJLT) { ensure_fjlt_tables(t, d); copy_f64(d.D, t->N); }
    else if (strcmp(what, "samples") == 0 && t->type == T_FJLT) { ensure_fjlt_tables(t, d); copy_i32_as_i64(d.samples, t->S); }
    else if (strcmp(what, "samples") == 0 && t->type == T_UST) { ensure_ust_tables(t, d); copy_i32_as_i64(d.samples, t->S); }
    else if (strcmp(what, "shifts") == 0 && t->is_rft()) { ensure_rft_tables(t, d); copy_f64(d.shifts, t->S); }
    else if (strcmp(what, "scales") == 0 && t->is_rft()) { ensure_rft_tables(t, d); copy_f64(d.scales, t->S); }
    else throw error(SKB_ERR_INVALID_PARAM, std::string("sl_sketch_export: no table '") + what + "' for " + t->name);
    SL_CATCH
}

int sl_sketch_export_dense(sl_sketch_transform_t *t, int64_t i0, int64_t j0, int64_t h, int64_t w,
                           double *host_out) {
    SL_TRY
    require_device();
    if (!t || !(t->is_dense() || t->is_rft())) throw error(SKB_ERR_INVALID_PARAM, "sl_sketch_export_dense: not a dense transform");
    if (i0 < 0 || j0 < 0 || h < 0 || w < 0 || i0 + h > t->S || j0 + w > t->N)
        throw error(SKB_ERR_RANDOM123, "Index is out of bounds");
    double *buf = nullptr;
    size_t bytes = (size_t)h * (size_t)w * 8;
    SKB_CUDA(cudaMalloc(&buf, bytes ? bytes : 16));
    try {
        dense_export_block(dense_params(t), (size_t)i0, (size_t)j0, (size_t)h, (size_t)w, buf, g_stream);
        SKB_CUDA(cudaMemcpyAsync(host_out, buf, bytes, cudaMemcpyDeviceToHost, g_stream));
        SKB_CUDA(cudaStreamSynchronize(g_stream));
    } catch (...) { cudaFree(buf); throw; }
    cudaFree(buf);
    SL_CATCH
}

}  // extern "C"

// variates of stream indices [base, base + count) for the composite transforms of the Python layer
__global__ void k_rng_export(int kind, uint64_t key, uint64_t base, size_t count, double a, double b, int order,
                             double *out) {
    size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= count) return;
    MicroUrng u(key, base + t);
    double v;
    if (kind == 0) v = boost_rademacher(u, order);
    else if (kind == 1) v = boost_normal(u);
    else if (kind == 2) v = boost_cauchy(u);
    else if (kind == 3) v = boost_uniform_real(u, a, b);
    else v = (double)boost_uniform_int(u, (uint64_t)a);       // kind 4: uniform_int(0, a), a < 2^53
    out[t] = v;
}

__global__ void k_raw_draws(uint64_t key, uint64_t base, size_t count, uint64_t *first, uint64_t *second) {
    size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= count) return;
    MicroUrng u(key, base + t);
    first[t] = u.next();
    second[t] = u.next();
}

extern "C" {

int sl_rng_export_raw(int seed, uint64_t base, int64_t count, uint64_t *first, uint64_t *second) {
    SL_TRY
    require_device();
    if (count < 0) throw error(SKB_ERR_INVALID_PARAM, "negative count");
    if (count == 0) return 0;
    uint64_t *buf = nullptr;
    SKB_CUDA(cudaMalloc(&buf, (size_t)count * 16));
    try {
        {
            LaunchScope ls(KC_TABLES, g_stream);
            k_raw_draws<<<(unsigned)((count + 255) / 256), 256, 0, g_stream>>>(
                seed_to_key(seed), base, (size_t)count, buf, buf + count);
        }
        SKB_CUDA(cudaGetLastError());
        SKB_CUDA(cudaMemcpyAsync(first, buf, (size_t)count * 8, cudaMemcpyDeviceToHost, g_stream));
        SKB_CUDA(cudaMemcpyAsync(second, buf + count, (size_t)count * 8, cudaMemcpyDeviceToHost, g_stream));
        SKB_CUDA(cudaStreamSynchronize(g_stream));
    } catch (...) { cudaFree(buf); throw; }
    cudaFree(buf);
    SL_CATCH
}

// See include/skylark_b200.h.
int sl_fjlt_apply_shard_peer(sl_sketch_transform_t *S, void *A_, int64_t index_offset, void *peer_ptrs_dev,
                             int world, int rank, void *scratch_dev) {
    SL_TRY
    require_device();
    if (!S || !A_ || !peer_ptrs_dev || !scratch_dev || world < 1 || rank < 0 || rank >= world)
        throw error(SKB_ERR_INVALID_PARAM, "sl_fjlt_apply_shard_peer: bad argument");
    sl_matrix_t *A = as_matrix(A_);
    if (S->type != T_FJLT || S->fut != FUT_WHT || A->kind != MK_DENSE || !A->on_device || A->dtype != SL_DTYPE_F64)
        throw error(SKB_ERR_UNSUPPORTED_BASE, "sl_fjlt_apply_shard_peer: needs an FJLT in Walsh-Hadamard mode and a "
                                              "float64 device matrix");
    if (A->n % world != 0)
        throw error(SKB_ERR_INVALID_PARAM, "sl_fjlt_apply_shard_peer: the number of columns must be a multiple of the group size");
    DeviceGuard g(A->device);
    // a dummy SA wrapper of the right shape: the kernel sends its columns to the owners instead
    sl_matrix_t SA = *A;
    SA.m = S->S; SA.n = A->n; SA.ld = S->S; SA.data = scratch_dev;
    fjlt_set_peer_out(reinterpret_cast<double *const *>(peer_ptrs_dev), reinterpret_cast<double *>(scratch_dev),
                      (int)(A->n / world), rank, S->S);
    try {
        apply_device(S, A, &SA, SL_COLUMNWISE, index_offset, true, false);
    } catch (...) { fjlt_set_peer_out(nullptr, nullptr, 0, 0, 0); throw; }
    fjlt_set_peer_out(nullptr, nullptr, 0, 0, 0);
    SL_CATCH
}

int sl_rng_export(int kind, int seed, uint64_t base, int64_t count, double a, double b, double *host_out) {
    SL_TRY
    require_device();
    if (count < 0 || kind < 0 || kind > 4) throw error(SKB_ERR_INVALID_PARAM, "sl_rng_export: bad argument");
    if (count == 0) return 0;
    double *buf = nullptr;
    SKB_CUDA(cudaMalloc(&buf, (size_t)count * 8));
    try {
        {
            LaunchScope ls(KC_TABLES, g_stream);
            k_rng_export<<<(unsigned)((count + 255) / 256), 256, 0, g_stream>>>(kind, seed_to_key(seed), base, (size_t)count,
                                                                               a, b, g_boost_order, buf);
        }
        SKB_CUDA(cudaGetLastError());
        SKB_CUDA(cudaMemcpyAsync(host_out, buf, (size_t)count * 8, cudaMemcpyDeviceToHost, g_stream));
        SKB_CUDA(cudaStreamSynchronize(g_stream));
    } catch (...) { cudaFree(buf); throw; }
    cudaFree(buf);
    SL_CATCH
}

int sl_profile_enable(int on) { profiler().enabled = on != 0; return 0; }
int sl_profile_reset(void) {
    profiler_collect();
    Profiler &p = profiler();
    for (int i = 0; i < KC_COUNT; i++) { p.launches[i] = 0; p.ms[i] = 0; }
    return 0;
}
int sl_profile_read(unsigned long long launches[SL_KERNEL_CLASSES], double ms[SL_KERNEL_CLASSES]) {
    profiler_collect();
    Profiler &p = profiler();
    for (int i = 0; i < KC_COUNT; i++) { launches[i] = p.launches[i]; ms[i] = p.ms[i]; }
    return 0;
}

int sl_host_alloc_pinned(size_t bytes, void **ptr) {
    SL_TRY
    require_device();
    SKB_CUDA(cudaHostAlloc(ptr, bytes ? bytes : 16, cudaHostAllocDefault));
    SL_CATCH
}
int sl_host_free_pinned(void *ptr) { cudaFreeHost(ptr); return 0; }
// Sums entries of a device CSC that share (row, column), in storage order, and sorts the rows of
// every column: the merge step after sparse partial sketches have been exchanged between ranks.
__global__ void k_iota_ones(int32_t *idx, double *one, size_t n) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) { idx[i] = (int32_t)i; one[i] = 1.0; }
}
int sl_sp_matrix_coalesce(void *A_, void *SA_) {
    SL_TRY
    require_device();
    sl_matrix_t *A = as_matrix(A_), *SA = as_matrix(SA_);
    if (A->kind != MK_SPARSE || SA->kind != MK_SPARSE || !A->on_device || !SA->on_device)
        throw error(SKB_ERR_INVALID_USAGE, "sl_sp_matrix_coalesce: both operands must be device sparse matrices");
    DeviceGuard g(A->device);
    static std::map<int, std::pair<void *, size_t>> ws_per_dev;      // grow-only scratch per device
    auto &ws = ws_per_dev[A->device];
    int32_t *idx = nullptr; double *one = nullptr;
    const size_t m = (size_t)A->m;
    SKB_CUDA(cudaMallocAsync((void **)&idx, (m + 1) * sizeof(int32_t), g_stream));
    SKB_CUDA(cudaMallocAsync((void **)&one, (m + 1) * sizeof(double), g_stream));
    int32_t *oi = nullptr, *ox = nullptr; double *ov = nullptr; size_t onnz = 0;
    try {
        if (m) k_iota_ones<<<(unsigned)((m + 255) / 256), 256, 0, g_stream>>>(idx, one, m);
        hash_apply_csc_sparse(A->indptr, A->indices, A->values, m, (size_t)A->n, (size_t)A->nnz, false,
                              idx, one, (int)(m ? m : 1), &oi, &ox, &ov, &onnz, &ws.first, &ws.second, g_stream);
    } catch (...) { cudaFreeAsync(idx, g_stream); cudaFreeAsync(one, g_stream); throw; }
    cudaFreeAsync(idx, g_stream); cudaFreeAsync(one, g_stream);
    attach_sparse_result(SA, oi, ox, ov, onnz, A->m, A->n, A->device);
    SL_CATCH
}
// See include/skylark_b200.h.
int sl_sp_merge_segments(const int32_t *vptr_dev, const int32_t *indices_dev, const double *values_dev, int nseg,
                         int64_t n_rows, int64_t n_cols, int64_t nnz, void *SA_) {
    SL_TRY
    require_device();
    sl_matrix_t *SA = as_matrix(SA_);
    if (SA->kind != MK_SPARSE || !SA->on_device)
        throw error(SKB_ERR_INVALID_USAGE, "sl_sp_merge_segments: SA must be a device sparse matrix");
    if (!vptr_dev || nseg < 1 || n_rows < 0 || n_cols < 0 || nnz < 0)
        throw error(SKB_ERR_INVALID_PARAM, "sl_sp_merge_segments: bad argument");
    DeviceGuard g(SA->device);
    static std::map<int, std::pair<void *, size_t>> ws_per_dev;      // grow-only scratch per device
    auto &ws = ws_per_dev[SA->device];
    int32_t *oi = nullptr, *ox = nullptr; double *ov = nullptr; size_t onnz = 0;
    if (!hash_merge_segments(vptr_dev, indices_dev, values_dev, nseg, (size_t)n_cols, (size_t)n_rows, (size_t)nnz,
                             &oi, &ox, &ov, &onnz, &ws.first, &ws.second, g_stream))
        throw error(SKB_ERR_UNSUPPORTED_BASE, "sl_sp_merge_segments: a merged column exceeds 10240 entries (or the "
                                              "input is empty); re-order and use sl_sp_matrix_coalesce");
    attach_sparse_result(SA, oi, ox, ov, onnz, n_rows, n_cols, SA->device);
    SL_CATCH
}
int sl_device_count(int *count) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); n = 0; }
    *count = n;
    return 0;
}

}  // extern "C"
