"""cavept_sort_pairs_static -- the radix sort with its scratch in a static __device__ array (no cudaMalloc in the call;
include/cavept.h, csrc/cavept_static.cu) -- against the existing entry point cavept_sort_pairs, through the C ABI.

Collected last on purpose (the file name sorts after every other test file): a
failure here must not hide the other tests.  Passes alone in a fresh process and after the rest of the suite.
Bit-exact, as every integer test here is: a stable sort has one answer."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def api():
    import torch

    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import caveengine_b200 as cb

    cb.load_library()  # raises if libcavept.so is missing: no fallback
    return cb


def _pairs(n, bits, seed):
    rng = np.random.default_rng(seed)
    keys = rng.integers(0, 1 << bits, size=n, dtype=np.uint64).astype(np.uint32)
    if n > 1000:
        keys[: n // 2] &= np.uint32(0xFF)  # many duplicates: stability is part of the answer
    return keys, np.arange(n, dtype=np.int32)


def _sort_static(pt, keys, vals, bits):
    """Device arrays come from torch (plumbing: the entry point itself must not allocate); the sort runs on the
    context's stream, so both sides are fenced explicitly."""
    import torch

    dk = torch.from_numpy(keys.view(np.int32).copy()).cuda()
    dv = torch.from_numpy(vals.copy()).cuda()
    torch.cuda.synchronize()
    pt.sort_pairs_static(dk.data_ptr(), dv.data_ptr(), keys.size, bits)
    pt.sync()
    return dk.cpu().numpy().view(np.uint32), dv.cpu().numpy()


def test_static_scratch_sort_equals_the_allocating_sort_on_calls_in_a_row(api):
    """Calls in a row share one scratch array that keeps the previous call's contents: a large call, a small one over
    its leftovers, the maximum size, and an odd size again -- each equal to cavept_sort_pairs bit for bit."""
    with api.PathTracer(16, 16, 1, device=0) as pt:
        for n, bits in [(300_007, 30), (4097, 32), (api.PathTracer.STATIC_SORT_MAX_PAIRS, 30), (5, 8), (100_003, 17)]:
            keys, vals = _pairs(n, bits, seed=n)
            want_k, want_v = keys.copy(), vals.copy()
            pt.sort_pairs(want_k, want_v, bits)  # the existing entry point (allocates its scratch per call)
            got_k, got_v = _sort_static(pt, keys, vals, bits)
            assert np.array_equal(got_k, want_k), f"keys differ at n={n}"
            assert np.array_equal(got_v, want_v), f"values differ at n={n}"
            order = np.argsort(keys, kind="stable")
            assert np.array_equal(got_v, order.astype(np.int32))  # and both are the stable sort
        # the same input twice in a row: identical output
        keys, vals = _pairs(65_536, 32, seed=7)
        a = _sort_static(pt, keys, vals, 32)
        b = _sort_static(pt, keys, vals, 32)
        assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])


def test_static_scratch_sort_rejects_sizes_beyond_its_maximum(api):
    import torch

    n = api.PathTracer.STATIC_SORT_MAX_PAIRS + 1
    dk = torch.zeros(n, dtype=torch.int32, device="cuda")
    dv = torch.zeros(n, dtype=torch.int32, device="cuda")
    with api.PathTracer(16, 16, 1, device=0) as pt:
        with pytest.raises(api.CaveptError) as e:
            pt.sort_pairs_static(dk.data_ptr(), dv.data_ptr(), n, 30)
        assert e.value.code == 5  # CAVEPT_ERR_CAPACITY
        assert "CAVEPT_STATIC_SORT_MAX_PAIRS" in str(e.value)
        with pytest.raises(api.CaveptError) as e:
            pt.sort_pairs_static(0, dv.data_ptr(), 16, 30)
        assert e.value.code == 1  # CAVEPT_ERR_INVALID_ARG
        # zero and one pair: nothing to do, no error
        pt.sort_pairs_static(dk.data_ptr(), dv.data_ptr(), 0, 30)
        pt.sort_pairs_static(dk.data_ptr(), dv.data_ptr(), 1, 30)
        pt.sync()
