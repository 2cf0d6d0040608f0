"""Host-side logic that needs no GPU: the C-ABI library loads and exports every declared
symbol, contexts / counters / JSON follow the reference's rules, errors use its codes."""
import ctypes
import json
import os
import pickle
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "golden.json")))


def test_library_exports_every_declared_symbol():
    from libskylark_b200 import _build
    hdr = open(os.path.join(ROOT, "include", "skylark_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    names = set(re.findall(r"\b(sl_[a-z0-9_]+)\s*\(", hdr))
    assert len(names) >= 40
    L = ctypes.CDLL(_build.SO)
    missing = [n for n in sorted(names) if not hasattr(L, n)]
    assert not missing, missing


def test_supported_list_is_evaluable(sk):
    combos = sk.SUPPORTED_SKETCH_TRANSFORMS
    assert ("CWT", "Matrix", "Matrix") in combos and ("CWT", "SparseMatrix", "Matrix") in combos
    assert ("JLT", "Matrix", "Matrix") in combos and ("FJLT", "Matrix", "Matrix") in combos
    assert all(len(c) == 3 for c in combos)


def test_global_context_counter_rules(sk):
    from libskylark_b200 import lib
    sk.initialize(3)
    assert lib.global_counter() == 0
    sk.CWT(10, 6); assert lib.global_counter() == 20
    sk.MMT(7, 3); assert lib.global_counter() == 34
    sk.WZT(5, 4, 1.5); assert lib.global_counter() == 49
    sk.JLT(4, 3); assert lib.global_counter() == 61
    sk.CT(4, 3, 2.0); assert lib.global_counter() == 73
    sk.GaussianRFT(3, 5, 2.0); assert lib.global_counter() == 93
    sk.FJLT(16, 4); assert lib.global_counter() == 97      # advances by S only
    sk.LaplacianRFT(3, 5, 2.0); assert lib.global_counter() == 117


def test_explicit_context_and_positional_form(sk):
    from libskylark_b200 import lib
    c = lib.Context(7, 100)
    assert (c.seed, c.counter) == (7, 100)
    S = sk.CWT(10, 6, c)                  # BASELINE.json form: T(n, s, context)
    assert c.counter == 120
    d = S.serialize()
    assert d["creation_context"]["seed"] == "7" and d["creation_context"]["counter"] == "100"
    sk.JLT(4, 3, context=c)
    assert c.counter == 132
    assert sk.RFT is sk.GaussianRFT and sk.FastJLT is sk.JLT and sk.CountSketch is sk.CWT


def test_serialize_roundtrip_and_pickle(sk):
    for make in (lambda: sk.CWT(10, 6), lambda: sk.MMT(10, 6), lambda: sk.WZT(10, 6, 1.25),
                 lambda: sk.JLT(10, 6), lambda: sk.CT(10, 6, 2.5), lambda: sk.FJLT(16, 6),
                 lambda: sk.GaussianRFT(10, 6, 3.0), lambda: sk.LaplacianRFT(10, 6, 0.5)):
        S = make()
        d = S.serialize()
        assert d["skylark_object_type"] == "sketch" and d["skylark_version"] == "0.20"
        S2 = sk.deserialize_sketch(d)
        assert type(S2) is type(S)
        assert S2.serialize() == d
        S3 = pickle.loads(pickle.dumps(S))
        assert S3.serialize() == d
        assert (S3.getindim(), S3.getsketchdim()) == (S.getindim(), S.getsketchdim())
    # numbers instead of boost's strings are accepted too
    d = sk.CT(10, 6, 2.5).serialize()
    d2 = dict(d, N=10, S=6, C=2.5, creation_context=dict(d["creation_context"], seed=123834, counter=int(d["creation_context"]["counter"])))
    assert sk.deserialize_sketch(d2).serialize() == d


def test_error_codes(sk):
    from libskylark_b200 import errors, lib
    with pytest.raises(errors.LowerLayerError) as e:
        sk._SketchTransform("NoSuchTransform", 4, 2)
    assert e.value.code == 111                      # capi/csketch.cpp:385-386
    with pytest.raises(errors.LowerLayerError) as e:
        sk.WZT(10, 5, 3.0)
    assert e.value.code == 105                      # WZT_data.hpp:49-52 sketch_exception
    with pytest.raises(errors.LowerLayerError) as e:
        sk.CWT(0, 5)
    assert e.value.code == 111
    assert lib.lib.sl_strerror(108).decode() == "Skylark failed in a call into the Random123 layer"
    with pytest.raises(errors.UnsupportedError):
        sk.CWT(10, 5, forceppy=True)
    with pytest.raises(errors.UnsupportedError):
        sk.CWT(10, 5, defouttype="NoSuchType")
    c = lib.Context(1, 2**64 - 5)
    with pytest.raises(errors.LowerLayerError) as e:
        sk.CWT(10, 5, c)
    assert e.value.code == 108                      # stream exhausted, randgen.hpp:64-72


def test_reference_names_outside_the_path_fail_clearly(sk):
    from libskylark_b200 import errors
    for name in ("SJLT", "SparseJLT", "MaternRFT", "FastMaternRFT", "NURST", "GaussianQRFT", "ExpSemigroupRLT"):
        with pytest.raises(errors.UnsupportedError, match="outside the GPU sketch-apply path"):
            getattr(sk, name)(10, 5)


def test_ust_counters_and_json_on_the_host(sk):
    # UST_data.hpp:81-102 / SURVEY App. B: c+S with replacement, c+N without (no device needed
    # until the samples are asked for)
    from libskylark_b200 import lib
    c = lib.Context(9, 100)
    u = sk.UST(50, 7, True, context=c)
    assert c.counter == 107 and sk.URST is sk.UST and sk.UniformSampler is sk.UST
    v = sk.UST(50, 7, False, context=c)
    assert c.counter == 157
    d = v.serialize()
    assert d["sketch_type"] == "UST" and d["replace"] == "false" and d["creation_context"]["counter"] == "107"
    assert u.serialize()["replace"] == "true"
    w = sk.deserialize_sketch(d)
    assert w.serialize() == d


def test_adapter_checks(sk):
    from libskylark_b200 import errors
    S = sk.CWT(10, 5)
    with pytest.raises(errors.UnsupportedError):
        S.apply(np.zeros((10, 3), dtype=np.float32), None, 0)
    with pytest.raises(errors.UnsupportedError):
        S.apply(np.zeros((20, 3))[::2], None, 0)       # views refused, lib.py:103-104
    with pytest.raises(errors.DimensionMistmatchError):
        S.apply(np.zeros((11, 3)), None, 0)
    with pytest.raises(errors.DimensionMistmatchError):
        S.apply(np.zeros((10, 3)), np.zeros((4, 3)), 0)
    with pytest.raises(errors.UnsupportedError):
        S.apply(np.zeros((10, 3)), np.zeros((5, 3), order="F"), 0)
    with pytest.raises(ValueError):
        S.apply(np.zeros((10, 3)), None, "diagonal")
    with pytest.raises(errors.InvalidObjectError):
        S.apply([[1.0]], None, 0)


def test_compute_fails_loudly_without_gpu(sk):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from libskylark_b200 import errors
    S = sk.CWT(10, 5)
    with pytest.raises(errors.LowerLayerError, match="no CUDA device"):
        S.apply(np.zeros((10, 3)), None, 0)
    with pytest.raises(errors.LowerLayerError, match="no CUDA device"):
        S.row_idx()


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, "libskylark_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in src.lower(), os.path.join(dirpath, f)


def test_shard_range():
    from libskylark_b200.dist import shard_range
    for n, w in ((10, 3), (8, 8), (5, 8), (1000003, 8)):
        parts = [shard_range(n, r, w) for r in range(w)]
        assert parts[0][0] == 0 and sum(c for _, c in parts) == n
        for (o1, c1), (o2, _) in zip(parts, parts[1:]):
            assert o1 + c1 == o2


@pytest.mark.parametrize("n,s,valued", [(5000, 256, False), (9001, 2000, False), (4096, 513, True), (300, 20000, False)])
def test_strided_hash_plan_invariants(n, s, valued):
    """The host-built plan of the atomics-free strided hash kernel (hash.cu: hash_schedule_cols_plan), checked
    without a device: every index appears exactly once, in its own tile and in the stream of the warp that
    owns its bucket; the 32 rows of a round -- and of the two rounds the kernel applies together -- go to
    different buckets (plain load-add-store is then race-free); a bucket sees its rows in ascending order
    (the reference's summation order, hash_transform_Elemental.hpp:83-100); idle slots point at the spare
    accumulators; and each half warp of a round sits on 16 different shared-memory bank pairs for the
    accumulators (a performance property, asserted because the kernel's 6-wavefront round depends on it)."""
    from libskylark_b200 import _build
    L = ctypes.CDLL(_build.SO)
    rng = np.random.default_rng(n + s)
    bucket = rng.integers(0, s, size=n, dtype=np.int32)
    value = (rng.standard_normal(n) if valued else rng.choice([-1.0, 1.0], size=n)).astype(np.float64)
    ent, off = ctypes.POINTER(ctypes.c_uint32)(), ctypes.POINTER(ctypes.c_uint32)()
    n_ent, n_off, sw = ctypes.c_int64(), ctypes.c_int64(), ctypes.c_int()
    L.sl_debug_hash_cols_plan.argtypes = (ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64, ctypes.c_int,
                                          ctypes.POINTER(ctypes.POINTER(ctypes.c_uint32)), ctypes.POINTER(ctypes.c_int64),
                                          ctypes.POINTER(ctypes.POINTER(ctypes.c_uint32)), ctypes.POINTER(ctypes.c_int64),
                                          ctypes.POINTER(ctypes.c_int))
    rc = L.sl_debug_hash_cols_plan(bucket.ctypes.data, value.ctypes.data, n, s, ctypes.byref(ent), ctypes.byref(n_ent),
                                   ctypes.byref(off), ctypes.byref(n_off), ctypes.byref(sw))
    assert rc == 0
    libc = ctypes.CDLL(None)
    libc.free.argtypes = (ctypes.c_void_p,)
    try:
        E = np.ctypeslib.as_array(ent, shape=(n_ent.value,)).copy().reshape(-1, 32)
        O = np.ctypeslib.as_array(off, shape=(n_off.value,)).copy()
    finally:
        libc.free(ent); libc.free(off)
    SW, W, K = sw.value, 8, 2048
    assert SW % 32 == 0 and W * SW >= s
    ntiles = (n + K - 1) // K
    O = O.reshape(W, ntiles + 1)
    assert E.shape[0] >= int(O[-1, -1]) + 2                      # the kernel's prefetch may run past the last round
    seen = np.zeros(n, dtype=np.int64)
    last_row = {}
    for w in range(W):
        assert np.all(np.diff(O[w]) >= 0) and np.all(np.diff(O[w]) % 2 == 0)       # whole pairs of rounds
        if w:
            assert O[w, 0] == O[w - 1, -1]                                         # streams follow each other
        for t in range(ntiles):
            rounds = E[O[w, t]:O[w, t + 1]]
            for r0 in range(0, len(rounds), 2):
                pair_buckets = []
                for r in rounds[r0:r0 + 2]:
                    k = (r >> 3) & 0x7FF
                    b = (r >> 16) & 0x1FFF
                    real = b < SW
                    assert np.all(b[~real] >= SW) and np.all(b[~real] < SW + 32)        # spare accumulators
                    g = t * K + k[real].astype(np.int64)                                # global index
                    assert np.all(g < n)
                    assert np.array_equal(bucket[g], w * SW + b[real].astype(np.int64))  # right warp, right bucket
                    sign = (r[real] >> 31).astype(bool)
                    if not valued:
                        assert np.array_equal(sign, value[g] < 0)
                    seen[g] += 1
                    assert len(set(b[real].tolist())) == int(real.sum())                # 32 different buckets
                    pair_buckets += b[real].tolist()
                    for half in (slice(0, 16), slice(16, 32)):                          # one wavefront per half warp
                        assert len(set((b[half] % 16).tolist())) == 16
                        kk = k[half][real[half]]
                        assert len(set((kk % 16).tolist())) == len(kk)
                    for gi, bi in zip(g.tolist(), b[real].tolist()):                    # ascending rows per bucket
                        key = (w, bi)
                        assert last_row.get(key, -1) < gi
                        last_row[key] = gi
                assert len(set(pair_buckets)) == len(pair_buckets)                      # the pair applied together
    assert np.all(seen == 1)
