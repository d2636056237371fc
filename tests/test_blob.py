"""DNNw blob writer/parser and the block-sparse 8x4 encoding (host logic)."""
from collections import OrderedDict

import numpy as np
import pytest

from lpcnet_b200 import blob


def test_round_trip():
    A = OrderedDict(a=np.arange(5, dtype=np.float32), b=np.arange(70, dtype=np.int8), c=np.arange(3, dtype=np.int32))
    data = blob.write_blob(A)
    assert len(data) % 64 == 0
    B = blob.parse_blob(data)
    assert list(B) == ["a", "b", "c"]
    for k in A:
        assert np.array_equal(A[k], B[k]) and A[k].dtype == B[k].dtype


def test_rejects_empty_and_truncated():
    with pytest.raises(ValueError):
        blob.write_blob(OrderedDict(a=np.zeros(0, np.float32)))
    data = blob.write_blob(OrderedDict(a=np.arange(5, dtype=np.float32)))
    with pytest.raises(ValueError):
        blob.parse_blob(data[:-1])
    with pytest.raises(ValueError):
        blob.parse_blob(data[:40])


def test_block_sparse_encoding_matches_dense_product():
    rng = np.random.default_rng(0)
    n_in, n_out = 32, 24
    q = rng.integers(-50, 51, size=(n_in, n_out)).astype(np.int32)
    mask = rng.random((n_in // 4, n_out // 8)) < 0.5
    q = q * np.repeat(np.repeat(mask, 4, 0), 8, 1)
    w, idx = blob.block_sparse_encode(q, mask)
    x = rng.integers(0, 256, n_in)
    # decode the stream the way the kernels walk it
    out = np.zeros(n_out, np.int64)
    p = 0
    wp = 0
    for g in range(n_out // 8):
        nb = idx[p]; p += 1
        for _ in range(nb):
            pos = idx[p]; p += 1
            blk = w[wp:wp + 32].reshape(8, 4).astype(np.int64); wp += 32
            out[8 * g:8 * g + 8] += blk @ x[pos:pos + 4]
    assert p == len(idx) and wp == len(w)
    assert np.array_equal(out, q.T.astype(np.int64) @ x)


def test_default_model_shapes_and_saturation():
    A = blob.random_model()
    assert A["sparse_gru_a_recurrent_weights_idx"].size == 144 + A["sparse_gru_a_recurrent_weights"].size // 32
    assert A["gru_b_weights"].size == 32 * 576
    assert blob.saturation_free(A["sparse_gru_a_recurrent_weights"])
    assert blob.saturation_free(A["gru_b_weights"]) and blob.saturation_free(A["gru_b_recurrent_weights"])
    assert not blob.saturation_free(np.array([100, 100, 0, 0], np.int8))


def test_stream_seed_and_features_are_deterministic():
    assert blob.stream_seed(0) == b"LPCNet" and blob.stream_seed(12) == b"LPCNet12"
    assert np.array_equal(blob.synthetic_features(5, 3), blob.synthetic_features(5, 3))
    f = blob.synthetic_features(50, 1)
    assert f.shape == (50, 36) and (f[:, 20:] == 0).all()
