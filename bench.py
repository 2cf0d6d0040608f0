#!/usr/bin/env python
"""Benchmark of the sketch-apply hot path (contract: see the task statement / DESIGN.md).

Workload (N = 1): BASELINE.json configs[1] -- JLT Gaussian columnwise sketch of a dense fp64
1,000,000 x 1024 matrix to s = 4096 on one B200.  A "step" is one apply: 2*s*n*d flop.
N > 1 is weak scaling: every rank holds its own 1,000,000-row shard of an (N*1e6) x 1024
matrix, sketches it with global row indices and the 33.5 MB partial sketches are summed with
an NCCL all-reduce inside the step (the exchange step of a row-sharded columnwise sketch).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

`--impl reference` times the reference's CPU algorithm for the same path (the oracle port:
realise S entry by entry from the counter stream, then BLAS dgemm) on the host cores.
"""
import argparse
import ctypes
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

N_ROWS, D_COLS, S_ROWS = 1_000_000, 1024, 4096
SKETCH_SEED = 123834      # python-skylark/skylark/examples/example_sketch.py:53
DATA_SEED = 0
METRIC = "JLT columnwise sketch throughput (2*s*n*d flop per apply)"
UNIT = "TFLOP/s"
WORKLOAD = "C2: JLT fp64 1000000x1024 -> s=4096, columnwise, row-major A"


CPU_SAMPLE_ROWS = 48 * 8192   # ~10-15 s of host work on the 16-core box (0.3 TFLOP/s measured)


def flops_per_step(n_rows):
    return 2.0 * S_ROWS * n_rows * D_COLS


# ---------------------------------------------------------------------------
# reference arm / cpu_baseline: the oracle port on the host cores
# ---------------------------------------------------------------------------
def cpu_reference_sample(n_sample, repeats=1):
    """Times the reference's CPU path (dense_transform_data.hpp:79-101 realisation +
    dense_transform_Elemental_local.hpp:97-111 Gemm) on the first n_sample rows."""
    import numpy as np
    from oracle import oracle as o
    o.lib()
    try:
        import torch
        torch.set_num_threads(_host_cores())
    except Exception:
        torch = None
    try:
        o.lib().sko_set_num_threads(_host_cores())
    except Exception:
        pass
    o.set_libm(False)       # timed as the reference would run here: this box's libm, not the checker's
    rng = np.random.default_rng(DATA_SEED)
    A = rng.standard_normal((n_sample, D_COLS))
    ctx = o.Context(SKETCH_SEED)
    t = o.JLT(N_ROWS, S_ROWS, ctx)
    times = []
    panel = 8192        # rows of A per panel, so S is never materialised whole (the reference
                        # blocks the same way, dense_transform_Elemental_coldist_star.hpp:200-229)
    for _ in range(repeats):
        t0 = time.perf_counter()
        SA = np.zeros((S_ROWS, D_COLS))
        for k0 in range(0, n_sample, panel):
            kc = min(panel, n_sample - k0)
            R = o.realize_dense(t.seed, t.base, 0, t.scale, S_ROWS, 0, k0, S_ROWS, kc)
            SA += R @ A[k0:k0 + kc]
        times.append(time.perf_counter() - t0)
    del SA
    dt = statistics.median(times)
    cores = _host_cores()
    return {"seconds": dt, "tflops": flops_per_step(n_sample) / dt / 1e12, "cores": cores,
            "threads": o.num_threads(),
            "sample": "first %d of %d rows (time is linear in rows): S realised entry by entry "
                      "(Threefry2x64-13 + Box-Muller, OpenMP) then numpy dgemm" % (n_sample, N_ROWS)}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    n_sample = CPU_SAMPLE_ROWS
    for _ in range(args.warmup if args.warmup < 2 else 1):
        cpu_reference_sample(1024)
    res = [cpu_reference_sample(n_sample) for _ in range(max(1, min(args.steps, 3)))]
    dt = statistics.median(r["seconds"] for r in res)
    val = flops_per_step(n_sample) / dt / 1e12
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus,
            "steps": len(res), "warmup": 1, "ms_per_step": dt * 1e3, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": WORKLOAD, "note": "reference CPU algorithm (oracle port; the "
                       "Elemental/MPI build is unavailable on this box), bounded sample"},
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": res[0]["cores"],
                             "threads": res[0]["threads"], "kind": "port", "sample": res[0]["sample"]},
            "threads": res[0]["threads"],
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))
    return 0


# ---------------------------------------------------------------------------
# clocks sampler
# ---------------------------------------------------------------------------
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu = gpu_index
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q,
                 "--format=csv,noheader,nounits", "-lms", "200"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        for r in self.rows:
            try:
                sm.append(float(r[1])); mx.append(float(r[2]))
                for name, v in zip(names, r[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            except Exception:
                pass
        return {"sm_mhz": statistics.median(sm) if sm else None,
                "sm_max_mhz": max(mx) if mx else None, "reasons": sorted(reasons),
                "samples": len(sm)}


# ---------------------------------------------------------------------------
# ours
# ---------------------------------------------------------------------------
def measured_hbm_gbs():
    """(GB/s, source): the driver-measured copy bandwidth of this pool's B200s, else the
    fallback the profiling recipe states."""
    try:
        v = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
        return v, "MEASURED_PEAKS.json hbm_gbs (driver-measured copy, read+write)"
    except Exception:
        return 6500.0, "fallback: MEASURED_PEAKS.json absent"


def committed_dram_traffic(kernel):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of `kernel`, from the committed
    summary of an `ncu --set full` capture of this bench command (profiles/dram_traffic.json,
    written by tools/ncu_summary.py); None when no capture of the current kernel is committed."""
    try:
        rec = json.load(open(os.path.join(ROOT, "profiles", "dram_traffic.json")))[kernel]
        return rec["dram_bytes_per_launch"], rec.get("source")
    except Exception:
        return None, None


def measure_fp64_gemm_peak(torch):
    """cuBLAS DGEMM 8192^3 through torch.matmul, best of 5 -- the FP64 counterpart of the
    driver's bf16 figure in MEASURED_PEAKS.json (which has no FP64 entry)."""
    n = 8192
    a = torch.randn(n, n, dtype=torch.float64, device="cuda")
    b = torch.randn(n, n, dtype=torch.float64, device="cuda")
    torch.matmul(a, b)
    torch.cuda.synchronize()
    best = float("inf")
    for _ in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); torch.matmul(a, b); e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    del a, b
    return 2.0 * n ** 3 / (best * 1e-3) / 1e12


def _time_apply(torch, S, A, SA, dim, reps, flush):
    S.apply(A, SA, dim)
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        if flush is not None:
            flush.zero_()            # evict L2 between timed runs
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); S.apply(A, SA, dim); e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return statistics.median(ts)


def cpu_other_configs():
    """The reference's CPU algorithm (oracle port) for configs 1, 3, 4, 5 on bounded samples, so
    that every GPU number in `other_configs` has the CPU path timed beside it."""
    import numpy as np
    import scipy.sparse
    from oracle import oracle as o
    out = {}
    o.set_libm(False)
    rng = np.random.default_rng(DATA_SEED)
    cores = _host_cores()

    def best(fn, reps=2):
        ts = []
        for _ in range(reps):
            t0 = time.perf_counter(); fn(); ts.append(time.perf_counter() - t0)
        return min(ts)

    # C1 at full size: the serial scatter loop of hash_transform_Elemental.hpp:106-123
    n, d, s = 100_000, 500, 2000
    t = o.CWT(n, s, o.Context(SKETCH_SEED))
    A = rng.standard_normal((n, d))
    # a C-ordered n x d array reaches the reference as the column-major d x n matrix with the
    # direction flipped (lib.py:113-123): the rowwise loop, which walks memory contiguously
    dt = best(lambda: t.apply(A.T, "rowwise"))
    out["c1_cwt"] = {"seconds": dt, "value": (n * d * 8 + s * d * 8) / dt / 1e9, "unit": "GB/s", "cores": 1,
                     "sample": "full size (100000x500 -> 2000), the reference's serial rowwise loop on "
                               "the transposed view"}
    # C3: sign-scale + per-column DCT-II (scipy/pocketfft = REDFT10) + gather on 16 of the 2048 columns
    n, s, dc = 2 ** 20, 8192, 16
    f = o.FJLT(n, s, o.Context(SKETCH_SEED))
    A = rng.standard_normal((n, dc))
    dt = best(lambda: f.apply(A, "columnwise", fut="dct"), 1)
    out["c3_fjlt_dct"] = {"seconds": dt, "value": (n * dc * 8 + s * dc * 8) / dt / 1e9, "unit": "GB/s",
                          "cores": 1, "sample": "16 of 2048 columns (time is linear in columns)"}
    # C4: CSC scatter loop (OpenMP over columns) on 2000 of the 100000 columns of one shard
    n_loc, dcols, s, nnz_col = 1_250_000, 2000, 20_000, 1250
    t = o.CWT(n_loc, s, o.Context(SKETCH_SEED))
    indptr = (np.arange(dcols + 1, dtype=np.int64) * nnz_col).astype(np.int32)
    indices = rng.integers(0, n_loc, dcols * nnz_col, dtype=np.int32)
    values = rng.standard_normal(dcols * nnz_col)
    dt = best(lambda: t.apply_csc(indptr, indices, values, (n_loc, dcols), "columnwise"))
    out["c4_cwt_csc"] = {"seconds": dt, "value": (dcols * nnz_col * 12 + s * dcols * 8) / dt / 1e9,
                         "unit": "GB/s", "cores": o.num_threads(),
                         "sample": "2000 of 100000 columns of one shard (time is linear in columns)"}
    # C5: realise W in double, cast to float, sgemm, float cosine pass on 2048 of the 500000 points
    m, n, s = 2048, 784, 16384
    r = o.GaussianRFT(n, s, 10.0, o.Context(SKETCH_SEED))
    X = rng.random((m, n), dtype=np.float32)
    dt = best(lambda: r.apply(X, "rowwise", dtype=np.float32), 1)
    out["c5_rft"] = {"seconds": dt, "value": 2.0 * m * n * s / dt / 1e12, "unit": "TFLOP/s", "cores": cores,
                     "sample": "2048 of 500000 points; includes realising the 784x16384 W, as every "
                               "reference apply does (dense_transform_Elemental_local.hpp:101-102)"}
    return out


def other_configs(torch, lib, sketch, hbm_gbs, fp64_peak):
    """The remaining BASELINE.json configs on one GPU (device-resident, CUDA events, median of
    5, L2 flushed between runs); reported beside the main line, not as the bench value."""
    out = {}
    dev = "cuda"
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def guarded(name, fn):
        sampler = ClockSampler(torch.cuda.current_device())
        sampler.start()
        try:
            out[name] = fn()
        except Exception as e:      # never lose the main line
            out[name] = {"error": str(e)[:200]}
        out[name]["clocks"] = sampler.stop()
        torch.cuda.empty_cache()

    def c1(order):
        n, d, s = 100_000, 500, 2000
        A = torch.randn(n, d, dtype=torch.float64, device=dev)
        if order == "F":
            A = A.t().contiguous().t()
        S = sketch.CWT(n, s, lib.Context(SKETCH_SEED))
        SA = S.apply(A, None, "columnwise")
        ms = _time_apply(torch, S, A, SA, "columnwise", 5, flush)
        gbs = (n * d * 8 + s * d * 8) / (ms * 1e-3) / 1e9
        return {"workload": "C1: CWT fp64 100000x500 -> 2000, %s-major A" % ("row" if order == "C" else "column"),
                "ms": ms, "value": gbs, "unit": "GB/s", "frac_of_measured_hbm": gbs / hbm_gbs}

    def c3(fut, d):
        n, s = 2 ** 20, 8192
        lib.set_option("fjlt_fut", fut)
        try:
            S = sketch.FJLT(n, s, lib.Context(SKETCH_SEED))
        finally:
            lib.set_option("fjlt_fut", "dct")
        A = torch.randn(d, n, dtype=torch.float64, device=dev).t()     # column-major n x d
        SA = S.apply(A, None, "columnwise")
        ms = _time_apply(torch, S, A, SA, "columnwise", 3, None)
        gbs = (n * d * 8 + s * d * 8) / (ms * 1e-3) / 1e9
        return {"workload": "C3: FJLT(%s) fp64 2^20x%d -> 8192, column-major A, 1 GPU" % (fut, d),
                "ms": ms, "value": gbs, "unit": "GB/s", "frac_of_measured_hbm": gbs / hbm_gbs}

    def c4():
        # one GPU's shard of the 8-GPU config: 1.25e6 of the 1e7 rows, 1e5 columns, 1.25e8 nnz
        n_total, n_loc, d, s, nnz_col = 10_000_000, 1_250_000, 100_000, 20_000, 1250
        g = torch.Generator(device=dev).manual_seed(DATA_SEED)
        indptr = (torch.arange(d + 1, device=dev, dtype=torch.int64) * nnz_col).to(torch.int32)
        indices = torch.randint(0, n_loc, (d * nnz_col,), device=dev, dtype=torch.int32, generator=g)
        values = torch.randn(d * nnz_col, dtype=torch.float64, device=dev, generator=g)
        A = lib.DeviceCSC(indptr, indices, values, (n_loc, d))
        S = sketch.CWT(n_total, s, lib.Context(SKETCH_SEED))
        SA = torch.empty(d, s, dtype=torch.float64, device=dev).t()     # column-major s x d, 16 GB
        S.apply_shard(A, SA, "columnwise", 0)
        torch.cuda.synchronize()
        ts = []
        for _ in range(3):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); S.apply_shard(A, SA, "columnwise", 0); e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        ms = statistics.median(ts)
        nbytes = d * nnz_col * 12 + (d + 1) * 4 + s * d * 8
        gbs = nbytes / (ms * 1e-3) / 1e9
        return {"workload": "C4 shard: CWT CSC 1.25e6(of 1e7)x1e5, 1.25e8 nnz -> 20000 (dense 16 GB partial)",
                "ms": ms, "value": gbs, "unit": "GB/s", "frac_of_measured_hbm": gbs / hbm_gbs}

    def c5():
        m, n, s = 500_000, 784, 16384      # one GPU's share of the 4M points
        S = sketch.GaussianRFT(n, s, 10.0, context=lib.Context(SKETCH_SEED))
        A = torch.rand(m, n, dtype=torch.float32, device=dev)
        SA = torch.empty(m, s, dtype=torch.float32, device=dev)          # 32.8 GB
        ms = _time_apply(torch, S, A, SA, "rowwise", 2, None)
        tf = 2.0 * m * n * s / (ms * 1e-3) / 1e12
        # fp32 operands take the 3xTF32 path on the 5th-generation tensor cores (tcgen05.mma.kind::tf32,
        # three MMAs per product, fp32-level accuracy).  Ceiling: the TF32 rate of that instruction
        # measured on this pool with tools/tcgen05_probe.cu (875 TFLOP/s over 148 SMs) divided by three.
        # For reference: 26.9 TFLOP/s on the FP64-DMMA path, 55 on the mma.sync 3xTF32 kernel.
        tf32_eff_peak = 875.0 / 3.0
        return {"workload": "C5 shard: GaussianRFT fp32 500000x784 -> 16384 rowwise, fused cos, 3xTF32 (tcgen05 + TMA)",
                "ms": ms, "value": tf, "unit": "TFLOP/s", "frac_of_3xtf32_tcgen05_peak": tf / tf32_eff_peak,
                "ratio_to_measured_fp64_gemm": tf / fp64_peak}

    def c1_host():
        # the reference's own calling convention: numpy operands through sl_apply_sketch_transform,
        # host <-> device copies inside the call (pageable host memory, staging from the pool)
        import numpy as np
        n, d, s = 100_000, 500, 2000
        A = np.random.default_rng(DATA_SEED).standard_normal((n, d))
        SA = np.empty((s, d))
        S = sketch.CWT(n, s, lib.Context(SKETCH_SEED))
        S.apply(A, SA, "columnwise")
        ts = []
        for _ in range(5):
            t0 = time.perf_counter(); S.apply(A, SA, "columnwise"); ts.append(time.perf_counter() - t0)
        ms = statistics.median(ts) * 1e3
        At = torch.from_numpy(A)
        dA = torch.empty_like(At, device=dev)
        cs = []
        for _ in range(5):
            torch.cuda.synchronize(); t0 = time.perf_counter(); dA.copy_(At); torch.cuda.synchronize()
            cs.append(time.perf_counter() - t0)
        copy_ms = statistics.median(cs) * 1e3
        return {"workload": "C1 end to end on HOST numpy operands (400 MB in, 8 MB out) through sl_apply_sketch_transform",
                "ms": ms, "h2d_copy_alone_ms": copy_ms, "ratio_to_copy": ms / copy_ms,
                "value": (n * d * 8 + s * d * 8) / (ms * 1e-3) / 1e9, "unit": "GB/s"}

    guarded("c1_cwt_rowmajor", lambda: c1("C"))
    guarded("c1_cwt_host_operands", c1_host)
    guarded("c1_cwt_colmajor", lambda: c1("F"))
    guarded("c3_fjlt_wht", lambda: c3("wht", 2048))
    guarded("c3_fjlt_dct", lambda: c3("dct", 2048))
    guarded("c4_cwt_csc_shard", c4)
    guarded("c5_rft_shard", c5)
    try:
        out["cpu_baseline"] = cpu_other_configs()
    except Exception as e:
        out["cpu_baseline"] = {"error": str(e)[:200]}
    return out


def multi_gpu_configs(torch, dist, lib, sketch, rank, world, hbm_gbs):
    """BASELINE.json configs 3-5 on `world` GPUs (all ranks run this in lockstep; device time via
    CUDA events, max over ranks).  C3: strong scaling of the fixed 2^20 x 2048 problem -- Walsh-
    Hadamard mode on row shards with the reduction overlapped block by block, DCT mode (what the
    reference computes) through the reference's rows -> columns all-to-all and on column shards.
    C4: 1.25e6 rows per rank of a (1.25e6*world) x 1e5 CSC matrix + reduce-scatter of the 16 GB
    dense partial, or a sparse exchange, or column shards.  C5: 5e5 points per rank, no collective.
    Every exchanged result is checked OUTSIDE the timed region against a single-GPU recompute of a
    column slice on rank 0 ("distributed == local", tests/unit/DenseSketchApplyElementalTest.cpp:28-106);
    the relative Frobenius error is printed as `check_rel_fro` and must be < 1e-12."""
    from libskylark_b200 import dist as sd
    dev = "cuda"
    out = {}
    TOL = 1e-12

    def timed(fn, reps):
        fn()
        torch.cuda.synchronize(); dist.barrier()
        ts = []
        for _ in range(reps):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); fn(); e1.record()
            torch.cuda.synchronize()
            t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ts.append(float(t.item()))
            dist.barrier()
        return statistics.median(ts)

    def rank0_time(fn, reps):
        """single-GPU time of the whole problem on rank 0 (the others wait): the T1 of a speed-up"""
        ms = 0.0
        if rank == 0:
            fn(); torch.cuda.synchronize()
            ts = []
            for _ in range(reps):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(); fn(); e1.record(); torch.cuda.synchronize()
                ts.append(e0.elapsed_time(e1))
            ms = statistics.median(ts)
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.broadcast(t, 0)
        return float(t.item())

    def bcast_check(v):
        t = torch.tensor([v if v is not None else -1.0], dtype=torch.float64, device=dev)
        dist.broadcast(t, 0)
        v = float(t.item())
        if not (0 <= v < TOL):
            raise RuntimeError("multi-GPU result differs from the single-GPU recompute: %g" % v)
        return v

    def rel(a, b):
        return float((a - b).norm() / b.norm())

    def guarded(name, fn):
        try:
            out[name] = fn()
        except Exception as e:      # keep the bench line if one leg fails
            out[name] = {"error": repr(e)[:300]}
        torch.cuda.empty_cache()
        dist.barrier()

    # ---- C3: fixed 2^20 x 2048 problem, every rank generates the SAME matrix and keeps its rows
    n, d, s = 2 ** 20, 2048, 8192
    CHK = 8                                             # columns recomputed on one GPU
    nloc = n // world

    def c3_data():
        g = torch.Generator(device=dev).manual_seed(DATA_SEED + 3)
        full = torch.randn(d, n, dtype=torch.float64, device=dev, generator=g)     # column-major n x d
        shard = full[:, rank * nloc:(rank + 1) * nloc].contiguous().t()             # rows of this rank
        chk = full[:CHK].contiguous().t()                                           # n x CHK, all rows
        return full, shard, chk

    def c3_wht():
        if n % (8192 * world):
            return {"skipped": "needs 8192-aligned row shards"}
        lib.set_option("fjlt_fut", "wht")
        try:
            S = sketch.FJLT(n, s, lib.Context(SKETCH_SEED))
        finally:
            lib.set_option("fjlt_fut", "dct")
        full, A, chk = c3_data()
        t1 = rank0_time(lambda: S.apply(full.t(), None, "columnwise"), 3)
        del full
        torch.cuda.empty_cache()
        SA = torch.empty(d, s, dtype=torch.float64, device=dev).t()
        res = {}
        for key, red in (("allreduce", "allreduce"), ("reduce_scatter", "reduce_scatter")):
            ms = timed(lambda: sd.apply_distributed_overlapped(S, A, SA, "columnwise", rank * nloc,
                                                               reduce=red, blocks=4), 5)
            got = sd.apply_distributed_overlapped(S, A, SA, "columnwise", rank * nloc, reduce=red, blocks=4)
            err = None
            if rank == 0:       # allreduce: SA whole; reduce_scatter: rank 0 owns the first d/(4*world) columns of block 0
                want = S.apply(chk, None, "columnwise")
                err = rel(got[:, :CHK], want)
            res[key] = {"ms": ms, "check_rel_fro": bcast_check(err)}
        ms_plain = timed(lambda: sd.apply_distributed(S, A, SA, "columnwise", rank * nloc, reduce="allreduce"), 3)
        ms_plain_rs = timed(lambda: sd.apply_distributed(S, A, SA, "columnwise", rank * nloc,
                                                         reduce="reduce_scatter"), 3)
        # the overlapped forms again as CUDA graphs: at ~1 ms per step the host cannot enqueue four
        # kernels and four collectives fast enough, a replayed graph can
        graph_ms = {}
        for key, red in (("allreduce", "allreduce"), ("reduce_scatter", "reduce_scatter"), ("plain_rs", "reduce_scatter")):
            try:
                if key == "plain_rs":
                    fn = lambda: sd.apply_distributed(S, A, SA, "columnwise", rank * nloc, reduce="reduce_scatter")
                else:
                    fn = lambda: sd.apply_distributed_overlapped(S, A, SA, "columnwise", rank * nloc, reduce=red, blocks=4)
                side = torch.cuda.Stream()
                side.wait_stream(torch.cuda.current_stream())
                with torch.cuda.stream(side):
                    fn()
                torch.cuda.current_stream().wait_stream(side)
                torch.cuda.synchronize(); dist.barrier()
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g):
                    keep = fn()
                graph_ms[key] = timed(g.replay, 5)
                g.replay()
                torch.cuda.synchronize()
                err = rel(keep[:, :CHK], S.apply(chk, None, "columnwise")) if rank == 0 else None
                bcast_check(err)
                del g, keep
            except Exception as e:      # capture is an optimisation, not a requirement
                graph_ms[key] = None
                graph_ms[key + "_error"] = repr(e)[:160]
                torch.cuda.synchronize()
        fused_ms = None
        try:        # the reduce-scatter fused into the kernel (peer stores), result = s x d/world per rank
            fn = lambda: sd.fjlt_wht_fused_reduce_scatter(S, A, rank * nloc)
            fused_ms = timed(fn, 5)
            got = fn()
            err = rel(got[:, :CHK], S.apply(chk, None, "columnwise")) if rank == 0 else None
            bcast_check(err)
            # as a CUDA graph of TWO steps (the two staging sets alternate, so a graph must hold both)
            try:
                side = torch.cuda.Stream()
                side.wait_stream(torch.cuda.current_stream())
                with torch.cuda.stream(side):
                    fn(); fn()
                torch.cuda.current_stream().wait_stream(side)
                torch.cuda.synchronize(); dist.barrier()
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g):
                    fn()
                    keep = fn()
                graph_ms["fused"] = timed(g.replay, 5) / 2.0
                g.replay()
                torch.cuda.synchronize()
                err = rel(keep[:, :CHK], S.apply(chk, None, "columnwise")) if rank == 0 else None
                bcast_check(err)
                del g, keep
            except Exception as e:
                graph_ms["fused_graph_error"] = repr(e)[:160]
                torch.cuda.synchronize()
        except Exception as e:
            graph_ms["fused_error"] = repr(e)[:200]
            fused_ms = None
            torch.cuda.synchronize()
        variants = {"overlapped all-reduce (4 column blocks)": res["allreduce"]["ms"],
                    "overlapped reduce-scatter (4 column blocks)": res["reduce_scatter"]["ms"],
                    "one kernel, then one all-reduce": ms_plain,
                    "one kernel, then one reduce-scatter": ms_plain_rs}
        if fused_ms:
            variants["reduce-scatter fused into the kernel (peer stores)"] = fused_ms
        if graph_ms.get("allreduce"):
            variants["overlapped all-reduce, CUDA graph"] = graph_ms["allreduce"]
        if graph_ms.get("reduce_scatter"):
            variants["overlapped reduce-scatter, CUDA graph"] = graph_ms["reduce_scatter"]
        if graph_ms.get("fused"):
            variants["reduce-scatter fused into the kernel, CUDA graph of two steps"] = graph_ms["fused"]
        if graph_ms.get("plain_rs"):
            variants["one kernel, then one reduce-scatter, CUDA graph"] = graph_ms["plain_rs"]
        best = min(variants.values())
        return {"workload": "C3: FJLT(wht) fp64 2^20x2048 -> 8192 row-sharded over %d GPUs (strong scaling); "
                            "kernel of column block b+1 overlapped with the NCCL reduction of block b" % world,
                "ms": best, "variant": min(variants, key=variants.get), "ms_by_variant": variants,
                "graph_capture": {k: v for k, v in graph_ms.items() if k.endswith("_error")} or "ok",
                "ms_one_gpu_same_run": t1, "speedup_vs_one_gpu": t1 / best,
                "scaling_efficiency": t1 / best / world,
                "check_rel_fro": max(res["allreduce"]["check_rel_fro"], res["reduce_scatter"]["check_rel_fro"]),
                "value": (n * d * 8 + s * d * 8) / (best * 1e-3) / 1e9, "unit": "GB/s", "scaling": "strong"}

    def c3_dct():
        S = sketch.FJLT(n, s, lib.Context(SKETCH_SEED))
        full, A, chk = c3_data()
        t1 = rank0_time(lambda: S.apply(full.t(), None, "columnwise"), 3)
        dloc = d // world
        Acols = full[rank * dloc:(rank + 1) * dloc].contiguous().t()                # n x d/world, all rows
        del full
        torch.cuda.empty_cache()
        # (i) the reference's route: rows -> columns all-to-all, local transform, all-gather
        ms_a2a = timed(lambda: sd.fjlt_apply_row_sharded(S, A, gather=True), 3)
        got = sd.fjlt_apply_row_sharded(S, A, gather=True)
        err = rel(got[:, :CHK], S.apply(chk, None, "columnwise")) if rank == 0 else None
        e1 = bcast_check(err)
        # (ii) the data already column-sharded: no exchange before the transform; all-gather after
        ms_col = timed(lambda: sd.gather_columns(S.apply(Acols, None, "columnwise"), d), 3)
        got = sd.gather_columns(S.apply(Acols, None, "columnwise"), d)
        err = rel(got[:, :CHK], S.apply(chk, None, "columnwise")) if rank == 0 else None
        e2 = bcast_check(err)
        return {"workload": "C3: FJLT(dct, the reference's transform) fp64 2^20x2048 -> 8192 over %d GPUs (strong "
                            "scaling): row shards re-sharded to columns by all-to-all (FJLT_Elemental.hpp:149-150), "
                            "local DCT, all-gather; and the same from column shards" % world,
                "ms": ms_a2a, "ms_column_sharded": ms_col, "ms_one_gpu_same_run": t1,
                "speedup_vs_one_gpu": t1 / ms_a2a, "speedup_vs_one_gpu_column_sharded": t1 / ms_col,
                "check_rel_fro": max(e1, e2),
                "value": (n * d * 8 + s * d * 8) / (ms_a2a * 1e-3) / 1e9, "unit": "GB/s", "scaling": "strong"}

    guarded("c3_fjlt_wht", c3_wht)
    guarded("c3_fjlt_dct", c3_dct)

    # ---- C4
    n_loc, d, s, nnz_col = 1_250_000, 100_000, 20_000, 1250
    if d % world == 0:
        g = torch.Generator(device=dev).manual_seed(DATA_SEED + rank)
        indptr = (torch.arange(d + 1, device=dev, dtype=torch.int64) * nnz_col).to(torch.int32)
        indices = torch.randint(0, n_loc, (d * nnz_col,), device=dev, dtype=torch.int32, generator=g)
        values = torch.randn(d * nnz_col, dtype=torch.float64, device=dev, generator=g)
        A = lib.DeviceCSC(indptr, indices, values, (n_loc, d))
        S = sketch.CWT(n_loc * world, s, lib.Context(SKETCH_SEED))

        def c4_slice_reference():
            """columns [0, CHK) of the whole matrix gathered on every rank (global row ids), sketched
            by ONE GPU: what rank 0 compares its share of the exchanged result with"""
            k = CHK * nnz_col
            idx = [torch.empty(k, dtype=torch.int32, device=dev) for _ in range(world)]
            val = [torch.empty(k, dtype=torch.float64, device=dev) for _ in range(world)]
            dist.all_gather(idx, indices[:k].contiguous())
            dist.all_gather(val, values[:k].contiguous())
            if rank != 0:
                return None
            cols = []
            for c in range(CHK):
                cols.append((torch.cat([idx[r][c * nnz_col:(c + 1) * nnz_col] + r * n_loc for r in range(world)]),
                             torch.cat([val[r][c * nnz_col:(c + 1) * nnz_col] for r in range(world)])))
            ip = (torch.arange(CHK + 1, device=dev, dtype=torch.int64) * nnz_col * world).to(torch.int32)
            Ac = lib.DeviceCSC(ip, torch.cat([c[0] for c in cols]), torch.cat([c[1] for c in cols]),
                               (n_loc * world, CHK))
            SAc = torch.empty(CHK, s, dtype=torch.float64, device=dev).t()
            S.apply(Ac, SAc, "columnwise")
            return SAc

        def c4_dense():
            SA = torch.empty(d, s, dtype=torch.float64, device=dev).t()          # column-major s x d
            kernel_ms = []

            def step():
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                S.apply_shard(A, SA, "columnwise", rank * n_loc)
                e1.record()
                res = sd.reduce_partial(SA, None, "reduce_scatter")     # each rank keeps d/world columns
                torch.cuda.synchronize()
                kernel_ms.append(e0.elapsed_time(e1))
                return res
            ms = timed(step, 2)
            got = step()
            want = c4_slice_reference()
            err = rel(got[:, :CHK], want) if rank == 0 else None
            nbytes = world * (d * nnz_col * 12 + (d + 1) * 4 + s * d * 8)
            return {"workload": "C4: CWT CSC %dx100000, %.2e nnz -> 20000 row-sharded over %d GPUs, "
                                "reduce-scatter of the 16 GB partials" % (n_loc * world, 1.25e8 * world, world),
                    "ms": ms, "kernel_ms_rank0": statistics.median(kernel_ms), "check_rel_fro": bcast_check(err),
                    "value": nbytes / (ms * 1e-3) / 1e9, "unit": "GB/s (all GPUs, kernel bytes / "
                    "end-to-end time incl. collective)", "scaling": "weak"}

        def c4_sparse():
            def step():
                res = sd.apply_distributed_sparse(S, A, rank * n_loc, None)
                torch.cuda.synchronize()
                return res
            ms = timed(step, 2)
            res, (c0, nc) = step()
            want = c4_slice_reference()
            err = None
            if rank == 0:
                ip = res.indptr[:CHK + 1].to(torch.int64)
                dense = torch.zeros(CHK, s, dtype=torch.float64, device=dev)
                colid = torch.repeat_interleave(torch.arange(CHK, device=dev), ip[1:] - ip[:-1])
                e = int(ip[-1])
                dense[colid, res.indices[:e].to(torch.int64)] = res.values[:e]
                err = rel(dense.t(), want)
            return {"workload": "C4: same matrix, sparse partial sketches (CSC -> CSC), column-block all-to-all, "
                                "merge on the device; result = d/world columns per GPU as CSC",
                    "ms": ms, "check_rel_fro": bcast_check(err),
                    "value": world * (d * nnz_col * 12 + (d + 1) * 4) / (ms * 1e-3) / 1e9,
                    "unit": "GB/s (all GPUs, input bytes / end-to-end time incl. exchange and merge)", "scaling": "weak"}

        guarded("c4_cwt_csc", c4_dense)
        guarded("c4_cwt_csc_sparse_exchange", c4_sparse)
        del A, S, indptr, indices, values
        torch.cuda.empty_cache()

        def c4_columns():
            # the partition that needs no exchange at all: shard the COLUMNS of A (every GPU holds all
            # 1e7 rows of d/world columns, the same 1.25e8 nnz) -> each GPU owns its columns of SA outright
            dc, nnz_c = d // world, nnz_col * world
            n_all = n_loc * world
            g = torch.Generator(device=dev).manual_seed(DATA_SEED + 100 + rank)
            indptr = (torch.arange(dc + 1, device=dev, dtype=torch.int64) * nnz_c).to(torch.int32)
            indices = torch.randint(0, n_all, (dc * nnz_c,), device=dev, dtype=torch.int32, generator=g)
            values = torch.randn(dc * nnz_c, dtype=torch.float64, device=dev, generator=g)
            A = lib.DeviceCSC(indptr, indices, values, (n_all, dc))
            S = sketch.CWT(n_all, s, lib.Context(SKETCH_SEED))
            SA = torch.empty(dc, s, dtype=torch.float64, device=dev).t()
            ms = timed(lambda: S.apply(A, SA, "columnwise"), 3)
            return {"workload": "C4: same 1e7 x 1e5 matrix, COLUMNS sharded over %d GPUs (all rows of %d columns each): "
                                "no collective, each GPU owns its %d columns of SA (dense)" % (world, dc, dc),
                    "ms": ms, "value": world * (dc * nnz_c * 12 + (dc + 1) * 4 + s * dc * 8) / (ms * 1e-3) / 1e9,
                    "unit": "GB/s (all GPUs)", "scaling": "strong"}
        guarded("c4_cwt_csc_column_sharded", c4_columns)

    # ---- C5
    def c5():
        m, n, s = 500_000, 784, 16384
        S = sketch.GaussianRFT(n, s, 10.0, context=lib.Context(SKETCH_SEED))
        A = torch.rand(m, n, dtype=torch.float32, device=dev)
        SA = torch.empty(m, s, dtype=torch.float32, device=dev)
        ms = timed(lambda: S.apply(A, SA, "rowwise"), 2)
        # no exchange to check; the rows of every rank against the FP64 path of the same transform
        want = S.apply(A[:256].double(), None, "rowwise")
        err = rel(SA[:256].double(), want)
        t = torch.tensor([err], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        if not float(t.item()) < 1e-5:
            raise RuntimeError("C5 fp32 features differ from the fp64 path: %g" % float(t.item()))
        return {"workload": "C5: GaussianRFT fp32 %dx784 -> 16384 rowwise, %d points per GPU, no collective"
                            % (m * world, m), "ms": ms, "check_rel_fro_vs_fp64_path": float(t.item()),
                "value": 2.0 * m * world * n * s / (ms * 1e-3) / 1e12, "unit": "TFLOP/s", "scaling": "weak"}
    guarded("c5_rft", c5)
    return out


def run_ours(args):
    import torch
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; this package has no CPU path")
    torch.cuda.set_device(local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    from libskylark_b200 import lib, sketch
    from libskylark_b200.dist import reduce_partial

    n_total = N_ROWS * world
    ctx = lib.Context(SKETCH_SEED)
    S = sketch.JLT(n_total, S_ROWS, ctx)
    offset = rank * N_ROWS

    gen = torch.Generator(device="cuda").manual_seed(DATA_SEED + rank)
    A = torch.randn(N_ROWS, D_COLS, dtype=torch.float64, device="cuda", generator=gen)
    SA = torch.empty(S_ROWS, D_COLS, dtype=torch.float64, device="cuda")

    def step():
        S.apply_shard(A, SA, "columnwise", offset)
        if world > 1:
            reduce_partial(SA, None, "allreduce")

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        step()
    barrier()

    lib.profile_reset()
    lib.profile_enable(True)
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        step()
    e1.record()
    barrier()
    elapsed_ms = e0.elapsed_time(e1)
    clocks = sampler.stop() if rank == 0 else None
    prof = lib.profile_read()
    lib.profile_enable(False)
    if world > 1:
        t = torch.tensor([elapsed_ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        elapsed_ms = float(t.item())
    ms_per_step = elapsed_ms / args.steps
    value = flops_per_step(N_ROWS) * world / (ms_per_step * 1e-3) / 1e12

    # ---- end-to-end: host buffers through the C ABI, copies inside the timed region ----
    e2e_steps = max(1, min(args.steps, 3))
    h2d = N_ROWS * D_COLS * 8
    d2h = S_ROWS * D_COLS * 8
    L = lib.lib
    hA, hSA = ctypes.c_void_p(), ctypes.c_void_p()
    lib.callsl("sl_host_alloc_pinned", h2d, ctypes.byref(hA))
    lib.callsl("sl_host_alloc_pinned", d2h, ctypes.byref(hSA))
    hA_t = torch.frombuffer((ctypes.c_char * h2d).from_address(hA.value), dtype=torch.float64)
    hA_t.view(N_ROWS, D_COLS).copy_(A)
    torch.cuda.synchronize()
    Aw, SAw = ctypes.c_void_p(), ctypes.c_void_p()
    # row-major n x d == column-major d x n: wrap the transpose, flip the direction
    # (python-skylark/skylark/lib.py:113-123)
    lib.callsl("sl_wrap_raw_matrix", ctypes.cast(hA, ctypes.POINTER(ctypes.c_double)), D_COLS,
               N_ROWS, ctypes.byref(Aw))
    lib.callsl("sl_wrap_raw_matrix", ctypes.cast(hSA, ctypes.POINTER(ctypes.c_double)), D_COLS,
               S_ROWS, ctypes.byref(SAw))
    lib.callsl("sl_set_stream", ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))

    def e2e_step():
        lib.callsl("sl_apply_sketch_transform_shard", S._obj, "Matrix", Aw, "Matrix", SAw, 2,
                   offset)
        if world > 1:     # partial sketches are summed on the device, result read back
            part = torch.frombuffer((ctypes.c_char * d2h).from_address(hSA.value),
                                    dtype=torch.float64).cuda(non_blocking=True)
            dist.all_reduce(part)
            part.cpu()

    e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_step()
    barrier()
    e2e_s = (time.perf_counter() - t0) / e2e_steps
    if world > 1:
        t = torch.tensor([e2e_s], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t.item())
    e2e_val = flops_per_step(N_ROWS) * world / e2e_s / 1e12
    lib.callsl("sl_free_raw_matrix_wrap", Aw)
    lib.callsl("sl_free_raw_matrix_wrap", SAw)
    del hA_t
    L.sl_host_free_pinned(hA)
    L.sl_host_free_pinned(hSA)

    multi = None
    main_check = None
    if world > 1:
        # "distributed == local": columns [0, 8) of the all-reduced sketch against ONE GPU sketching
        # the same columns of the whole (world * 1e6)-row matrix (outside the timed region)
        step()
        sl = A[:, :8].contiguous()
        allsl = torch.empty(world * N_ROWS, 8, dtype=torch.float64, device="cuda")
        dist.all_gather_into_tensor(allsl, sl)
        err = -1.0
        if rank == 0:
            want = S.apply(allsl, None, "columnwise")
            err = float((SA[:, :8] - want).norm() / want.norm())
        t = torch.tensor([err], dtype=torch.float64, device="cuda")
        dist.broadcast(t, 0)
        main_check = float(t.item())
        if not (0 <= main_check < 1e-12):
            raise SystemExit("bench.py: the all-reduced sketch differs from the single-GPU recompute: %g" % main_check)
        del allsl, sl
        del A, SA
        torch.cuda.empty_cache()
        try:
            multi = multi_gpu_configs(torch, dist, lib, sketch, rank, world, measured_hbm_gbs()[0])
        except Exception as e:      # keep the main line
            multi = {"error": str(e)[:300]}
    if rank != 0:
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return 0

    # ---- roofline of the dominant kernel (the FP64 tensor-core GEMM) ----
    gemm_launches, gemm_ms = prof["dense_gemm"]
    gen_launches, gen_ms = prof["dense_gen"]
    launches = sum(v[0] for v in prof.values())
    flops_per_launch = flops_per_step(N_ROWS) * args.steps / max(1, gemm_launches)
    peak = measure_fp64_gemm_peak(torch)
    traffic, traffic_src = committed_dram_traffic("k_dense_gemm")
    achieved = flops_per_launch / (gemm_ms / max(1, gemm_launches) * 1e-3) / 1e12 if gemm_ms else None
    roofline = {"bound": "tensor", "kernel": "k_dense_gemm (FP64 DMMA m8n8k4)",
                "achieved": achieved, "peak": peak, "unit": "TFLOP/s",
                "frac": (achieved / peak) if achieved else None,
                "traffic": traffic, "traffic_unit": "DRAM bytes per launch (dram__bytes_read.sum + "
                "dram__bytes_write.sum, ncu --set full): " + str(traffic_src),
                "algorithmic_operand_bytes_per_launch": 8192 * S_ROWS * 8 + 8192 * D_COLS * 8 + S_ROWS * D_COLS * 8,
                "peak_source": "cuBLAS DGEMM 8192^3 via torch.matmul fp64, best of 5, measured in "
                               "this run (MEASURED_PEAKS.json has no FP64 entry)",
                "flops_per_launch": flops_per_launch,
                "kernel_ms_per_step": {"dense_gemm": gemm_ms / args.steps,
                                       "dense_gen": gen_ms / args.steps},
                "launches_per_step": {"dense_gemm": gemm_launches / args.steps,
                                      "dense_gen": gen_launches / args.steps}}
    roofline["hbm_gbs_measured"], roofline["hbm_gbs_source"] = measured_hbm_gbs()

    cpu = cpu_reference_sample(CPU_SAMPLE_ROWS) if world == 1 else None
    extra = None
    if world == 1:
        del A, SA, S
        torch.cuda.empty_cache()
        extra = other_configs(torch, lib, sketch, roofline["hbm_gbs_measured"], peak)
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": WORKLOAD, "rows_per_gpu": N_ROWS, "d": D_COLS, "s": S_ROWS,
                       "sketch_seed": SKETCH_SEED, "partition": "row shards, allreduce of the s x d "
                       "partial" if world > 1 else "single GPU",
                       "l2": "inputs (8.19 GB per GPU) are larger than the 126 MB L2"},
            "e2e": {"value": e2e_val, "unit": UNIT, "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": d2h, "ms_per_step": e2e_s * 1e3, "steps": e2e_steps,
                    "path": "sl_apply_sketch_transform_shard on pinned HOST buffers (C ABI)"},
            "gpu_launches": launches, "clocks": clocks, "roofline": roofline}
    if main_check is not None:
        line["check_rel_fro_vs_one_gpu"] = main_check
    if cpu is not None:
        line["cpu_baseline"] = {"value": cpu["tflops"], "unit": UNIT, "cores": cpu["cores"],
                                "kind": "port", "sample": cpu["sample"], "seconds": cpu["seconds"]}
    if extra is not None:
        line["other_configs"] = extra
    if multi is not None:
        line["other_configs"] = multi
    print(json.dumps(line))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def _host_cores():
    return len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)


def _pin_cpu_threads():
    """The CPU legs use every host core this process may run on.  torchrun exports
    OMP_NUM_THREADS=1 to its workers, which would time the reference arm on one core: set the
    OpenMP / BLAS thread counts explicitly, before numpy or the oracle library are loaded."""
    n = str(_host_cores())
    for k in ("OMP_NUM_THREADS", "OPENBLAS_NUM_THREADS", "MKL_NUM_THREADS"):
        os.environ[k] = n
    return int(n)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    args = ap.parse_args()
    if args.impl == "reference":
        if int(os.environ.get("RANK", "0")) != 0:
            return 0
        _pin_cpu_threads()
        return run_reference(args)
    if int(os.environ.get("WORLD_SIZE", "1")) == 1:
        _pin_cpu_threads()          # cpu_baseline legs run on rank 0 of a single-GPU run only
    return run_ours(args)


if __name__ == "__main__":
    sys.exit(main())
