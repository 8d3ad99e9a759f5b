"""The allocation-free entry point (fmlk_local_partials_noalloc) against the ordinary one.

Some hosts of this library cannot let it call cudaMalloc: a task manager that has given the whole
GPU memory budget to its own pool, or an operator thread that cannot afford the device-wide
synchronisation cudaMalloc/cudaFree imply.  For them the library keeps a fixed scratch array of its
own (about 10.8 MB per GPU) and offers one entry point that uses only that array plus the caller's
input and output buffers.  This test calls it through the C ABI several times in a row -- each call
with a different shape, so anything the previous call left in the shared array would show -- and
checks every result against the CPU oracle (the same checker the other parity tests use) and
against fmlk_local_partials on a context: counts equal, sums to 1e-12 relative.  The tolerance is
the one the other tests use; it covers the order of the additions, which differs from call to call
because most shapes add in L2 (bulk reductions) in arrival order.  One shape (d = 128, k = 60) is
chosen so that the d-blocked kernel, not only the main one, runs on the fixed scratch.

Collected last (file name) so that a failure here cannot hide any other test under -x."""
import numpy as np
import pytest

from flink_ml_b200 import native
from oracle import oracle as O

pytestmark = pytest.mark.gpu

SUM_RTOL = 1e-12


def _reference(x, cent):
    with native.Context(0) as ctx:
        ctx.append_points(x)
        ctx.set_centroids(cent)
        return ctx.local_partials()


def _noalloc(torch, x, cent):
    dev = torch.device("cuda", 0)
    xd = torch.from_numpy(x).to(dev)
    cd = torch.from_numpy(cent).to(dev)
    k, d = cent.shape
    sums = torch.full((k, d), float("nan"), dtype=torch.float64, device=dev)
    counts = torch.full((k,), float("nan"), dtype=torch.float64, device=dev)
    torch.cuda.synchronize()
    free_before, _ = torch.cuda.mem_get_info(0)
    stream = torch.cuda.current_stream()
    native.local_partials_noalloc(0, stream.cuda_stream, xd.data_ptr(), x.shape[0], d, cd.data_ptr(), k,
                                  sums.data_ptr(), counts.data_ptr())
    stream.synchronize()
    free_after, _ = torch.cuda.mem_get_info(0)
    return sums.cpu().numpy(), counts.cpu().numpy(), free_before, free_after


def test_noalloc_matches_context_path_twice():
    import torch
    rng = np.random.default_rng(3)
    # warm-up call: the first call on a device does the one-time set-up
    _noalloc(torch, rng.random((64, 8)), rng.random((4, 8)))
    # config-2 shape, a smaller ragged one, a long-row one that stays on the main kernel (k <= 32),
    # and one that takes the d-blocked kernel (d >= 128 with more than 32 centroids; 60*128 + 60 <= 8192)
    shapes = [(30000, 64, 100), (4099, 7, 5), (9000, 256, 31), (9000, 128, 60)]
    with native.Context(0) as probe:
        # (a context runs this shape on the ring variant of the d-blocked kernel; the allocation-free entry point,
        # which may not allocate the blocked copy of the model, on assign_wide_kernel itself)
        desc = probe.describe_launch(60, 128, True)
        assert "ring=1" in desc or "wide=1" in desc
        assert "wide=0" in probe.describe_launch(100, 64, True)
    for n, d, k in shapes:
        x = rng.random((n, d))
        cent = x[rng.choice(n, k, replace=False)].copy()
        cent[0] = cent[1]  # an exact tie: rows equal to these centroids take the exact replay
        o_sums, o_counts = O.train_local(x, cent)
        want_sums, want_counts = _reference(x, cent)
        sums, counts, free_before, free_after = _noalloc(torch, x, cent)
        assert np.array_equal(counts, o_counts.astype(np.float64))  # against the oracle
        assert np.allclose(sums, o_sums, rtol=SUM_RTOL, atol=0)
        assert np.array_equal(counts, want_counts.astype(np.float64))  # and against the context path
        assert np.allclose(sums, want_sums, rtol=SUM_RTOL, atol=0)
        assert free_after == free_before  # nothing was allocated or freed on the device
    # an empty batch gives zeros, not the previous call's numbers
    sums, counts, _, _ = _noalloc(torch, np.empty((0, 7)), rng.random((5, 7)))
    assert not sums.any() and not counts.any()


def test_noalloc_rejects_models_beyond_the_limit():
    import torch
    dev = torch.device("cuda", 0)
    x = torch.zeros((16, 64), dtype=torch.float64, device=dev)
    cent = torch.zeros((200, 64), dtype=torch.float64, device=dev)  # 200*64 + 200 > 8192
    out_s = torch.zeros((200, 64), dtype=torch.float64, device=dev)
    out_c = torch.zeros((200,), dtype=torch.float64, device=dev)
    with pytest.raises(native.IllegalArgumentError, match="too large for the allocation-free entry point"):
        native.local_partials_noalloc(0, None, x.data_ptr(), 16, 64, cent.data_ptr(), 200, out_s.data_ptr(),
                                      out_c.data_ptr())
