"""The CPU oracle against published KATs, the committed golden fixtures and the structural
properties the reference's own tests check: apply == explicit Pi*A
(tests/unit/SparseSketchApplyElementalTest.cpp:47-161), serialize/replay
(tests/unit/SerializationTest.cpp:13-87), counter rules."""
import json
import os

import numpy as np
import pytest
import scipy.sparse

GOLDEN = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "golden.json")))


def test_threefry_published_kat(oracle):
    for e in GOLDEN["threefry2x64_20_published"]:
        out = oracle.threefry2x64(20, [int(x, 16) for x in e["ctr"]], [int(x, 16) for x in e["key"]])
        assert ["%016x" % v for v in out] == e["out"]


def test_threefry_13_rounds_fixture(oracle):
    for e in GOLDEN["threefry2x64_13_derived"]:
        out = oracle.threefry2x64(13, [int(x, 16) for x in e["ctr"]], [int(x, 16) for x in e["key"]])
        assert ["%016x" % v for v in out] == e["out"]


def test_survey_vectors(oracle):
    # SURVEY.md Appendix A.6, produced by an independent restatement during the survey
    a, _ = oracle.raw_draws(0, 0, 3)
    assert ["%016x" % int(x) for x in a] == ["e91f9fee4b7a6fb5", "5b048c394b60b9f2", "1344b97241c1649a"]
    a, _ = oracle.raw_draws(123834, 0, 3)
    assert ["%016x" % int(x) for x in a] == ["258a7c04a8bb648a", "ca2331bf0906bd56", "9e6cd017c0c49cfe"]
    t = oracle.CWT(10, 6, oracle.Context(0))
    assert t.row_idx.tolist() == [5, 2, 0, 2, 3, 4, 3, 2, 2, 4]
    assert t.row_value.tolist() == [1, -1, 1, -1, 1, 1, 1, 1, 1, 1]
    t = oracle.CWT(10, 6, oracle.Context(0), order=155)
    assert t.row_value.tolist() == [1, -1, 1, -1, -1, -1, 1, 1, -1, 1]
    t = oracle.CWT(10, 6, oracle.Context(123834))
    assert t.row_idx.tolist() == [0, 4, 3, 3, 0, 0, 0, 0, 4, 1]
    assert t.row_value.tolist() == [1, 1, -1, 1, -1, -1, 1, 1, 1, -1]
    assert oracle.normal(0, 0, 3).tolist() == [2.0260583387707065, -0.7534839669473243, 2.064557991338214]
    assert oracle.normal(123834, 0, 3).tolist() == [0.8345286007232288, 0.408487308275462, -1.7218389514988737]
    assert oracle.cauchy(0, 0, 3).tolist() == [3.467985194452353, -0.4878003059151096, -4.149938550812701]


def test_golden_streams(oracle):
    for c in GOLDEN["streams"]:
        a, b = oracle.raw_draws(c["seed"], c["base"], 8)
        assert ["%016x" % int(x) for x in a] == c["first"]
        assert ["%016x" % int(x) for x in b] == c["second"]
        assert [int(x) for x in oracle.uniform_int(c["seed"], c["base"], 8, 1999)] == c["uniform_int_S1999"]
        assert oracle.rademacher(c["seed"], c["base"], 8, 160).tolist() == c["rademacher_160"]
        assert oracle.rademacher(c["seed"], c["base"], 8, 155).tolist() == c["rademacher_155"]
        assert [x.hex() for x in oracle.normal(c["seed"], c["base"], 8).tolist()] == c["normal"]
        assert [x.hex() for x in oracle.cauchy(c["seed"], c["base"], 8).tolist()] == c["cauchy"]


def test_counter_rules(oracle):
    c = oracle.Context(3)
    oracle.CWT(10, 6, c); assert c.counter == 20
    oracle.MMT(7, 3, c); assert c.counter == 34
    oracle.WZT(5, 4, 1.5, c); assert c.counter == 49
    oracle.JLT(4, 3, c); assert c.counter == 61
    oracle.CT(4, 3, 2.0, c); assert c.counter == 73
    oracle.GaussianRFT(3, 5, 2.0, c); assert c.counter == 73 + 15 + 5
    before = c.counter
    f = oracle.FJLT(16, 4, c)
    assert c.counter == before + 4            # advances by S only
    # D and samples share stream indices [c, ...)
    assert np.array_equal(f.D, oracle.rademacher(3, before, 16))
    assert np.array_equal(f.samples, oracle.uniform_int(3, before, 4, 15))
    with pytest.raises(ValueError):
        oracle.WZT(5, 4, 2.5, c)


def test_uniform_int_and_sign_statistics(oracle):
    idx = oracle.uniform_int(11, 0, 200000, 9)
    assert idx.min() == 0 and idx.max() == 9
    assert abs(np.bincount(idx.astype(int)) / 200000 - 0.1).max() < 0.005
    for order in (160, 155):
        s = oracle.rademacher(11, 0, 200000, order)
        assert set(np.unique(s)) == {-1.0, 1.0}
        assert abs(s.mean()) < 0.01
    z = oracle.normal(11, 0, 200000)
    assert abs(z.mean()) < 0.01 and abs(z.std() - 1) < 0.01


@pytest.mark.parametrize("T", ["CWT", "MMT", "WZT"])
def test_hash_apply_equals_explicit_product(oracle, T):
    rng = np.random.default_rng(0)
    c = oracle.Context(5)
    t = {"CWT": lambda: oracle.CWT(50, 7, c), "MMT": lambda: oracle.MMT(50, 7, c),
         "WZT": lambda: oracle.WZT(50, 7, 1.3, c)}[T]()
    P = t.explicit()
    A = rng.standard_normal((50, 9))
    np.testing.assert_allclose(t.apply(A, "columnwise"), P @ A, rtol=1e-13, atol=1e-13)
    B = rng.standard_normal((9, 50))
    np.testing.assert_allclose(t.apply(B, "rowwise"), B @ P.T, rtol=1e-13, atol=1e-13)
    Asp = scipy.sparse.random(50, 9, 0.3, format="csc", random_state=1)
    SA = t.apply_csc(Asp.indptr, Asp.indices, Asp.data, Asp.shape, "columnwise")
    np.testing.assert_allclose(SA, P @ Asp.toarray(), rtol=1e-13, atol=1e-13)
    Bsp = scipy.sparse.random(9, 50, 0.3, format="csc", random_state=2)
    SB = t.apply_csc(Bsp.indptr, Bsp.indices, Bsp.data, Bsp.shape, "rowwise")
    np.testing.assert_allclose(SB, Bsp.toarray() @ P.T, rtol=1e-13, atol=1e-13)


def test_dense_realize_views(oracle):
    c = oracle.Context(9)
    j = oracle.JLT(12, 5, c)
    R = j.realize()
    assert R.shape == (5, 12)
    # element (i, k) = scale * normal(base + k*S + i)   (dense_transform_data.hpp:90-99)
    z = oracle.normal(9, j.base, 60).reshape(12, 5).T
    assert np.array_equal(R, j.scale * z)
    sub = oracle.realize_dense(9, j.base, 0, j.scale, 5, 1, 2, 2, 3, col_stride=2, row_stride=3)
    assert np.array_equal(sub, R[1::2, 2::3][:2, :3])
    ct = oracle.CT(12, 5, 3.0, c)
    assert np.array_equal(ct.realize(), (3.0 / 5) * oracle.cauchy(9, ct.base, 60).reshape(12, 5).T)


def test_rft_matches_definition(oracle):
    rng = np.random.default_rng(1)
    c = oracle.Context(2)
    r = oracle.GaussianRFT(6, 8, 2.5, c)
    A = rng.standard_normal((6, 4))
    W = r.realize()
    want = np.sqrt(2.0 / 8) * np.cos(W @ A + r.shifts[:, None])
    np.testing.assert_allclose(r.apply(A, "columnwise"), want, rtol=1e-14, atol=1e-14)
    B = rng.standard_normal((4, 6))
    want = np.sqrt(2.0 / 8) * np.cos(B @ W.T + r.shifts[None, :])
    np.testing.assert_allclose(r.apply(B, "rowwise"), want, rtol=1e-14, atol=1e-14)
    assert (r.shifts >= 0).all() and (r.shifts < 2 * np.pi).all()


def test_fjlt_dct_definition(oracle):
    rng = np.random.default_rng(2)
    x = rng.standard_normal(24)
    np.testing.assert_allclose(oracle.dct2(x), oracle.dct2_naive(x), rtol=1e-12, atol=1e-12)
    c = oracle.Context(4)
    f = oracle.FJLT(24, 5, c)
    A = rng.standard_normal((24, 3))
    want = np.empty((5, 3))
    for j in range(3):
        y = oracle.dct2_naive(f.D * A[:, j] / np.sqrt(48.0))
        want[:, j] = np.sqrt(24 / 5) * y[f.samples.astype(int)]
    np.testing.assert_allclose(f.apply(A, "columnwise"), want, rtol=1e-12, atol=1e-12)
    np.testing.assert_allclose(f.apply(A.T.copy(), "rowwise"), want.T, rtol=1e-12, atol=1e-12)


def test_wht(oracle):
    x = np.arange(8.0)
    H = np.array([[(-1) ** bin(i & j).count("1") for j in range(8)] for i in range(8)], float)
    np.testing.assert_allclose(oracle.wht(x), H @ x)


def test_golden_applies(oracle):
    # end results of every transform on the fixed input, created in one sequence from one context
    import scipy.sparse
    from util import golden_input, unhex, rel_fro
    g = GOLDEN["applies"]
    n, d = g["n"], g["d"]
    A, Asp = golden_input(n, d)
    c = oracle.Context(g["seed"])
    cw = oracle.CWT(n, 5, c)
    assert np.array_equal(cw.apply(A, "columnwise"), unhex(g["cwt_columnwise"], (5, d)))
    assert np.array_equal(cw.apply(np.ascontiguousarray(A.T), "rowwise"), unhex(g["cwt_rowwise_of_transpose"], (d, 5)))
    sp = cw.apply_csc_sparse(scipy.sparse.csc_matrix(Asp), "columnwise")
    sp.sort_indices()
    assert sp.indptr.tolist() == g["cwt_sparse_columnwise"]["indptr"]
    assert sp.indices.tolist() == g["cwt_sparse_columnwise"]["indices"]
    assert [x.hex() for x in sp.data.tolist()] == g["cwt_sparse_columnwise"]["data"]
    rest = [("mmt_columnwise", oracle.MMT(n, 4, c), 4, {}), ("wzt_columnwise", oracle.WZT(n, 4, 1.5, c), 4, {}),
            ("jlt_columnwise", oracle.JLT(n, 6, c), 6, {}), ("ct_columnwise", oracle.CT(n, 6, 2.0, c), 6, {}),
            ("gaussian_rft_columnwise", oracle.GaussianRFT(n, 7, 3.0, c), 7, {}),
            ("laplacian_rft_columnwise", oracle.LaplacianRFT(n, 7, 3.0, c), 7, {})]
    for key, t, s, kw in rest:
        assert rel_fro(t.apply(A, "columnwise", **kw), unhex(g[key], (s, d))) < 1e-13, key
    fj = oracle.FJLT(n, 8, c)
    assert rel_fro(fj.apply(A, "columnwise", fut="dct"), unhex(g["fjlt_dct_columnwise"], (8, d))) < 1e-13
    assert rel_fro(fj.apply(A, "columnwise", fut="wht"), unhex(g["fjlt_wht_columnwise"], (8, d))) < 1e-13
    assert c.counter == g["counter_after"]
