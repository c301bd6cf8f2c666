"""uavmvs_reduce_noalloc (K9 with its scratch in a static __device__ array, no allocation inside the call) against the
existing uavmvs_reduce, through the C ABI.  The entry point has no counterpart in the reference,
so the checker is the existing entry point, whose own parity with the oracle is in
test_parity_gpu.py.  Kept in a file that pytest collects LAST.

Tolerance: min / max identical; sum within 1e-6 relative (the gate the other K9 comparisons use, test_sharded_gpu.py:
uavmvs_reduce adds its per-CTA totals with atomics, in no fixed order).  The test makes no assumption about what earlier
tests left in the shared scratch array: it passes alone in a fresh process and after the rest of the suite.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _pair(api, values, out3):
    """(no-alloc result, existing entry point's result) for one host vector."""
    a = api.Array.create(values.shape[0])
    a.upload(values)
    api.reduce_noalloc(a, out3)                                   # asynchronous on the default stream ...
    got = out3.download()[:3].copy()                              # ... which the download (cudaMemcpy) is ordered after
    ref = np.array(api.reduce(a), np.float32)
    return got, ref


def _check(got, ref):
    assert got[1] == ref[1] and got[2] == ref[2]                  # min, max: identical
    assert abs(float(got[0]) - float(ref[0])) <= 1e-6 * abs(float(ref[0])) + 1e-12      # sum: 1e-6 relative


def test_reduce_noalloc_equals_reduce_on_two_calls_in_a_row():
    from uavmvs_b200 import api
    assert api.device_count() >= 1
    rng = np.random.default_rng(21)
    out3 = api.Array.create(3)
    # first call: 1M values (489 partials); second call, right after it, on a SHORTER vector: the scratch still holds the
    # first call's partials beyond the second call's, and they must not leak into the result
    big = rng.uniform(0.0, 1.2, 1_000_000).astype(np.float32)
    small = rng.uniform(0.5, 2.0, 5000).astype(np.float32)
    small[1234] = -3.25
    got_big, ref_big = _pair(api, big, out3)
    got_small, ref_small = _pair(api, small, out3)
    _check(got_big, ref_big)
    _check(got_small, ref_small)
    assert got_small[1] == np.float32(-3.25)
    # also against numpy in double
    assert abs(float(got_big[0]) - float(big.astype(np.float64).sum())) <= 1e-6 * float(big.astype(np.float64).sum())
    assert got_big[1] == big.min() and got_big[2] == big.max()
    # fixed order of additions: a repeated call gives the same bits
    again, _ = _pair(api, big, out3)
    assert np.array_equal(again.view(np.uint32), got_big.view(np.uint32))


def test_reduce_noalloc_edge_sizes_and_limit():
    from uavmvs_b200 import api
    rng = np.random.default_rng(22)
    out3 = api.Array.create(3)
    for n in (1, 255, 2048, 2049, 65_537):                        # one element, chunk boundary +- 1, 33 partials
        v = rng.normal(0.0, 1.0, n).astype(np.float32)
        got, ref = _pair(api, v, out3)
        assert got[1] == ref[1] == v.min() and got[2] == ref[2] == v.max()
        exact = float(v.astype(np.float64).sum())
        gate = 1e-6 * float(np.abs(v).astype(np.float64).sum()) + 1e-12     # signed terms: relative to the sum of magnitudes
        assert abs(float(got[0]) - exact) <= gate and abs(float(got[0]) - float(ref[0])) <= 2 * gate
    # beyond the maximum: rejected with a clear error, nothing launched, out3 untouched
    before = out3.download()[:3].copy()
    too_big = api.Array.create(api.REDUCE_NOALLOC_MAX_N + 1)
    with pytest.raises(Exception) as e:
        api.reduce_noalloc(too_big, out3)
    assert "UAVMVS_REDUCE_NOALLOC_MAX_N" in str(e.value)
    assert np.array_equal(out3.download()[:3].view(np.uint32), before.view(np.uint32))
    # an output array that is too short is refused as well
    with pytest.raises(Exception):
        api.reduce_noalloc(api.Array.create(8), api.Array.create(2))
    # the largest accepted size (64 MB of input, all 8192 partials)
    full = api.Array.create(api.REDUCE_NOALLOC_MAX_N)
    ones = np.full(api.REDUCE_NOALLOC_MAX_N, 0.5, np.float32)
    ones[-1] = 7.0
    full.upload(ones)
    api.reduce_noalloc(full, out3)
    got = out3.download()[:3]
    assert got[0] == np.float32(0.5 * (api.REDUCE_NOALLOC_MAX_N - 1) + 7.0) and got[1] == 0.5 and got[2] == 7.0
