"""The allocation-free top-k entry point (onilda_topk_device_noalloc) against the ordinary one.

It exists for callers that must not allocate device memory when they rank a batch - a scorer
inside a Spark executor whose GPU memory already belongs to a pooling allocator, or a streaming
scorer that cannot afford the device-wide stall of cudaMalloc/cudaFree - so its scratch is a
fixed array inside the library that every call on a GPU shares.  What is checked here: two
calls in a row, on different data, each give exactly what onilda_topk gives (rows and order are
bit-exact in this repository's top-k tests, so here too); the second call must not see
anything the first one left in the shared scratch.  The file name sorts last so that a run
that stops at the first failure has already run every other test.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_two_calls_in_a_row_match_the_allocating_entry_point(oni):
    import torch
    rng = np.random.default_rng(11)
    n = 300_000
    pool = rng.random(1500) * 1e-3                       # few distinct values: heavy ties, broken by event index
    first = pool[rng.integers(0, 1500, n)]
    second = rng.random(n // 3) ** 4                     # other size, other distribution, other threshold
    sc = oni.Scorer(np.ones((1, 2)) / 2, np.ones((1, 2)) / 2)
    for scores, thr, k in [(first, 1.0, 1000), (second, 0.25, 4096), (first, float(pool.min()), 50)]:
        dev = torch.tensor(scores, dtype=torch.float64, device="cuda")
        torch.cuda.synchronize()
        idx, val = oni.topk_device_noalloc(dev.data_ptr(), len(scores), thr, k)
        ridx, rval = sc.topk(scores, thr, k)
        assert np.array_equal(idx, ridx) and np.array_equal(val, rval), (thr, k)
        assert len(idx) > 0
    sc.close()


def test_limits_are_refused_with_a_clear_message(oni):
    import torch
    dev = torch.zeros(10, dtype=torch.float64, device="cuda")
    for k in (oni.NOALLOC_MAX_K + 1, -1):
        with pytest.raises(oni.OnildaError) as e:
            oni.topk_device_noalloc(dev.data_ptr(), 10, 1.0, k)
        assert str(oni.NOALLOC_MAX_K) in str(e.value)
    idx, val = oni.topk_device_noalloc(dev.data_ptr(), 10, 1.0, 0)
    assert len(idx) == 0
    idx, val = oni.topk_device_noalloc(dev.data_ptr(), 10, -1.0, 5)      # nothing under the threshold
    assert len(idx) == 0


def test_second_device_has_its_own_scratch(oni):
    """Every GPU that loads the library holds its own copy of the scratch array; the entry point looks the
    address up once per device, with that device current."""
    import torch
    if oni.device_count() < 2:
        pytest.skip("needs two GPUs")
    rng = np.random.default_rng(12)
    scores = rng.random(200_000) ** 3
    sc = oni.Scorer(np.ones((1, 2)) / 2, np.ones((1, 2)) / 2, device=0)
    ridx, rval = sc.topk(scores, 0.5, 777)
    sc.close()
    for dev_id in (1, 0, 1):
        with torch.cuda.device(dev_id):
            dev = torch.tensor(scores, dtype=torch.float64, device=f"cuda:{dev_id}")
            torch.cuda.synchronize()
            idx, val = oni.topk_device_noalloc(dev.data_ptr(), len(scores), 0.5, 777)
        assert np.array_equal(idx, ridx) and np.array_equal(val, rval), dev_id
