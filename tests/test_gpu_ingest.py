"""Serialized-row ingest through the C ABI (fmlk_append_points_serialized, fmlk_decode_rows_dev)
against the oracle's decoder: bit-exact for every double, including NaN payloads, signed zeros,
subnormals and infinities; malformed input is refused and leaves the cache as it was."""
import numpy as np
import pytest

from flink_ml_b200 import linalg as L
from flink_ml_b200 import native
from oracle import oracle as O
from test_ingest_format import _special_rows

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    c = native.Context(0)
    yield c
    c.close()


@pytest.mark.parametrize("fmt", [0, 1, 2])
@pytest.mark.parametrize("n,d", [(1, 1), (5, 7), (1000, 64), (100003, 10), (4099, 129), (70000, 100), (21, 7001)])
def test_decode_is_bit_exact(ctx, fmt, n, d):
    x = _special_rows(n, d, 7 * fmt + d)
    enc = O.encode_rows(x, fmt)
    ctx.clear_points()
    assert ctx.append_points_serialized(enc, fmt, d) == n
    assert ctx.num_points() == n
    ids = np.unique(np.concatenate([np.arange(min(n, 50)), np.arange(max(n - 50, 0), n),
                                    np.random.default_rng(1).integers(0, n, 500)]))
    got = ctx.get_points(ids)
    want = O.decode_rows(enc, fmt, d)[ids]
    assert np.array_equal(got.view(np.uint64), want.view(np.uint64))


def test_several_pieces_and_appends(ctx):
    """More than one 32 MiB staging piece, then a second append onto the same cache."""
    n, d = 600_000, 8  # 600k * 77 B = 46 MB of format-2 records
    x = O.generate_dense(2, 0, 1, n, d)
    enc = O.encode_rows(x, 2)
    ctx.clear_points()
    assert ctx.append_points_serialized(enc, 2, d) == n
    assert ctx.append_points_serialized(O.encode_rows(x[:1000], 0), 0, d) == 1000
    assert ctx.num_points() == n + 1000
    ids = np.array([0, 1, 435_000, 435_848, 435_849, 435_850, n - 1, n, n + 999])
    assert np.array_equal(ctx.get_points(ids), np.concatenate([x, x[:1000]])[ids])
    ctx.clear_points()
    assert ctx.append_points_serialized(b"", 2, d) == 0 and ctx.num_points() == 0


def test_fit_on_serialized_rows_equals_fit_on_raw_rows(ctx):
    n, d, k = 50_000, 64, 100
    x = O.generate_dense(2, 0, 1, n, d)
    cent = x[:k].copy()
    ctx.clear_points()
    ctx.append_points(x)
    ctx.set_centroids(cent)
    c0, w0 = ctx.fit(5)
    ctx.clear_points()
    ctx.append_points_serialized(L.serialize_rows(x, L.FMT_VECTOR_WITH_NORM), L.FMT_VECTOR_WITH_NORM, d)
    ctx.set_centroids(cent)
    c1, w1 = ctx.fit(5)
    # same rows in the cache either way; the default accumulation (L2 atomics) leaves only the last bits free
    assert np.allclose(c0, c1, rtol=1e-12, atol=0) and np.array_equal(w0, w1)
    oc, ow = O.fit([x], cent, 5)
    assert np.allclose(c1, oc, rtol=1e-9, atol=0) and np.array_equal(w1, ow)
    ctx.clear_points()


def test_malformed_input_is_refused_and_the_cache_is_unchanged(ctx):
    d = 6
    x = O.generate_dense(2, 0, 1, 100, d)
    ctx.clear_points()
    ctx.append_points(x)
    for fmt in (0, 1, 2):
        enc = O.encode_rows(x, fmt)
        rec = L.record_bytes(fmt, d)
        with pytest.raises(native.IllegalArgumentError, match="Vector size mismatched."):
            ctx.append_points_serialized(enc[:-1], fmt, d)
        bad = enc.copy()
        bad[57 * rec + (0 if fmt == 0 else 1) + 3] = d + 1  # record 57 claims another length
        with pytest.raises(native.IllegalArgumentError, match="Vector size mismatched."):
            ctx.append_points_serialized(bad, fmt, d)
        with pytest.raises(native.IllegalArgumentError, match="Vector size mismatched."):
            ctx.append_points_serialized(O.encode_rows(x[:, :5], fmt), fmt, 5)  # other dimension than the cache
        assert ctx.num_points() == 100
    sparse = O.encode_rows(x, 2)
    sparse[3 * L.record_bytes(2, d)] = 1
    with pytest.raises(native.IllegalArgumentError, match="tag"):
        ctx.append_points_serialized(sparse, 2, d)
    with pytest.raises(native.IllegalArgumentError):
        ctx.append_points_serialized(O.encode_rows(x, 0), 9, d)
    assert ctx.num_points() == 100 and np.array_equal(ctx.get_points(np.arange(100)), x)
    ctx.clear_points()


def test_decode_rows_dev_primitive():
    import torch
    n, d = 20_001, 33
    x = _special_rows(n, d, 5)
    for fmt in (0, 1, 2):
        enc = O.encode_rows(x, fmt)
        rec = torch.from_numpy(enc).cuda()
        out = torch.empty((n, d), dtype=torch.float64, device="cuda:0")
        bad = torch.zeros(1, dtype=torch.int32, device="cuda:0")
        native.decode_rows_dev(0, torch.cuda.current_stream().cuda_stream, rec.data_ptr(), n, d, fmt, out.data_ptr(),
                               bad.data_ptr())
        torch.cuda.synchronize()
        assert int(bad.item()) == 0
        assert np.array_equal(out.cpu().numpy().view(np.uint64), x.view(np.uint64))
        # a base that is only 8-byte aligned takes the direct (unstaged) kernel
        shifted = torch.empty(enc.size + 8, dtype=torch.uint8, device="cuda:0")
        shifted[8:] = rec
        out.zero_()
        native.decode_rows_dev(0, torch.cuda.current_stream().cuda_stream, shifted.data_ptr() + 8, n, d, fmt,
                               out.data_ptr(), bad.data_ptr())
        torch.cuda.synchronize()
        assert int(bad.item()) == 0
        assert np.array_equal(out.cpu().numpy().view(np.uint64), x.view(np.uint64))
