"""UST (uniform sampling, sketch/UST_data.hpp:81-102 + UST_Elemental.hpp) on the GPU against the
oracle: sample indices bit-exact in both modes, apply exact (it is a gather), counter rules of
SURVEY App. B (c+S with replacement, c+N without), shards add up, serialise -> replay."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("replace", [True, False])
def test_ust_samples_and_apply(sk, oracle, replace):
    import torch
    from libskylark_b200 import lib
    n, s, d = 5000, 700, 37
    c, oc = lib.Context(321, 55), oracle.Context(321, 55)
    U, O = sk.UST(n, s, replace, context=c), oracle.UST(n, s, replace, oc)
    assert c.counter == oc.counter == 55 + (s if replace else n)
    assert np.array_equal(U.samples(), O.samples)
    if not replace:
        assert len(set(U.samples().tolist())) == s
    A = np.random.default_rng(1).standard_normal((n, d))
    assert np.array_equal(U.apply(A.copy(), None, "columnwise"), O.apply(A, "columnwise"))
    assert np.array_equal(U.apply(np.asfortranarray(A), None, 0), O.apply(A, 0))
    B = np.ascontiguousarray(A.T)
    assert np.array_equal(U.apply(B, None, "rowwise"), O.apply(B, "rowwise"))
    assert np.array_equal(U.apply(np.asfortranarray(B), None, 1), O.apply(B, 1))
    At = torch.from_numpy(A).cuda()
    parts = sum(U.apply_shard(At[a:b], None, 0, a) for a, b in ((0, 1234), (1234, 4000), (4000, 5000)))
    assert np.array_equal(parts.cpu().numpy(), O.apply(A, 0))
    U2 = sk.deserialize_sketch(U.serialize())
    assert U.serialize()["replace"] == ("true" if replace else "false")
    assert np.array_equal(U2.samples(), O.samples)


def test_ust_rejects_oversampling_without_replacement(sk):
    from libskylark_b200 import errors, lib
    with pytest.raises(errors.LowerLayerError):
        sk.UST(10, 11, False, context=lib.Context(1))
    assert sk.UST(10, 11, True, context=lib.Context(1)).samples().max() < 10
