"""The ingest side of the hot path: the reference serializers' byte format.
The reference holds no byte-level test for these serializers, so the known answers below are built
from the format its source defines (DenseVectorSerializer.java:78-93, VectorSerializer.java:80-88,
VectorWithNormSerializer.java:74-78, Bits.java:54-70; DataOutputView.writeInt/writeDouble are
big-endian by the Java specification)."""
import numpy as np
import pytest

from flink_ml_b200 import linalg as L
from oracle import oracle as O

SPECIALS = np.array([0.0, -0.0, 1.0, -2.5, np.inf, -np.inf, 5e-324, -2.2250738585072014e-308, 1.7976931348623157e308,
                     np.pi]).view(np.uint64)


def _special_rows(n, d, seed):
    rng = np.random.default_rng(seed)
    bits = rng.integers(0, 2 ** 63, size=(n, d), dtype=np.uint64) | (rng.integers(0, 2, size=(n, d), dtype=np.uint64) << np.uint64(63))
    pick = rng.integers(0, 4, size=(n, d)) == 0
    bits[pick] = SPECIALS[rng.integers(0, SPECIALS.size, size=int(pick.sum()))]
    return bits.view(np.float64)  # includes NaNs with arbitrary payloads


def test_known_answer_bytes():
    x = np.array([[1.0, -2.5]])
    assert O.encode_rows(x, O.FMT_DENSE_VECTOR).tobytes().hex() == "00000002" "3ff0000000000000" "c004000000000000"
    assert O.encode_rows(x, O.FMT_VECTOR).tobytes().hex() == "00" "00000002" "3ff0000000000000" "c004000000000000"
    want = "00" "00000002" "3ff0000000000000" "c004000000000000" + np.array([O.dnrm2(x[0])]).astype(">f8").tobytes().hex()
    assert O.encode_rows(x, O.FMT_VECTOR_WITH_NORM).tobytes().hex() == want
    assert abs(O.decode_rows(O.encode_rows(x, 2), 2, 2, want_norms=True)[1][0] - np.sqrt(7.25)) < 1e-15


@pytest.mark.parametrize("fmt", [0, 1, 2])
@pytest.mark.parametrize("n,d", [(0, 3), (1, 1), (5, 7), (300, 64), (17, 129)])
def test_oracle_round_trip_is_bit_exact_and_matches_the_host_mirror(fmt, n, d):
    x = _special_rows(n, d, 100 * fmt + d)
    enc = O.encode_rows(x, fmt)
    assert enc.size == n * L.record_bytes(fmt, d)
    back = O.decode_rows(enc, fmt, d)
    assert np.array_equal(back.view(np.uint64), x.view(np.uint64))
    finite = np.nan_to_num(x, nan=1.0, posinf=2.0, neginf=-2.0)
    norms = np.array([O.dnrm2(r) for r in finite]) if fmt == 2 else None
    assert np.array_equal(L.serialize_rows(finite, fmt, norms=norms), O.encode_rows(finite, fmt))
    assert np.array_equal(L.deserialize_rows(enc, fmt, d).view(np.uint64), x.view(np.uint64))


def test_malformed_records_are_rejected():
    x = np.arange(12.0).reshape(3, 4)
    for fmt in (0, 1, 2):
        enc = O.encode_rows(x, fmt)
        with pytest.raises(ValueError, match="Vector size mismatched."):
            O.decode_rows(enc[:-1], fmt, 4)
        with pytest.raises(ValueError, match="Vector size mismatched."):
            O.decode_rows(enc, fmt, 3 if (enc.size % L.record_bytes(fmt, 3)) else 5)
        bad = enc.copy()
        bad[L.record_bytes(fmt, 4) + (0 if fmt == 0 else 1) + 3] = 5  # second record claims length 5
        with pytest.raises(ValueError, match="Vector size mismatched."):
            O.decode_rows(bad, fmt, 4)
        with pytest.raises(ValueError):
            L.deserialize_rows(bad, fmt, 4)
    sparse = O.encode_rows(x, 1)
    sparse[0] = 1
    with pytest.raises(ValueError, match="tag"):
        O.decode_rows(sparse, 1, 4)
