"""Parity of the CUDA 1.6 kb/s encoder (lpcnet_feature_batch_encode / _compute_features, lpcnet_encode,
lpcnet_compute_features; through the C-ABI) against the golden vectors of the unmodified reference
(tests/golden/encode.npz) and the CPU oracle.  Packets are bytes and the bar is bit-exact; the unquantised features
are float32 compared bit for bit."""
import os

import numpy as np
import pytest

from lpcnet_b200 import blob, engine
from oracle import orcpy

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def eng(engine_lib):
    assert engine.device_count() > 0, "no CUDA device: the gpu tests cannot run (there is no CPU fallback)"
    return engine


@pytest.fixture(scope="module")
def cb():
    return blob.synthetic_codebooks()


def bits_equal(a, b):
    return np.array_equal(np.asarray(a).view(np.uint32), np.asarray(b).view(np.uint32))


def test_golden_vectors_of_the_reference(eng, cb):
    g = np.load(os.path.join(HERE, "golden", "encode.npz"))
    fb = eng.FeatureBatch(len(g["pcm"]))
    fb.set_codebooks(cb)
    pk = fb.encode(g["pcm"])
    assert fb.launch_count > 0
    fb.close()
    bad = np.argwhere((pk != g["packets"]).any(axis=2))
    assert bad.size == 0, ("first differing (stream, packet)", bad[0], pk[tuple(bad[0])], g["packets"][tuple(bad[0])])
    fb = eng.FeatureBatch(len(g["pcm"]))
    ft = fb.compute_features(g["pcm"])
    fb.close()
    assert bits_equal(ft, g["features"])


def test_call_splitting_and_single_stream_api(eng, cb):
    g = np.load(os.path.join(HERE, "golden", "encode.npz"))
    x, want = g["pcm"][1], g["packets"][1]
    eng.set_codebooks(cb)
    try:
        enc = eng.Encoder()
        got = np.stack([enc.encode(x[k * 640:(k + 1) * 640]) for k in range(6)])
        enc.close()
        assert np.array_equal(got, want[:6])
        enc = eng.Encoder()
        f = np.concatenate([enc.compute_features(x[k * 640:(k + 1) * 640]) for k in range(4)])
        enc.close()
        assert bits_equal(f, g["features"][1][:16])
    finally:
        eng.set_codebooks(None)
    fb = eng.FeatureBatch(1)
    fb.set_codebooks(cb)
    parts = [fb.encode(x[None, a * 640:b * 640]) for a, b in ((0, 5), (5, 6), (6, 24))]
    fb.close()
    assert np.array_equal(np.concatenate(parts, axis=1)[0], want)


def test_many_streams_match_the_oracle_and_are_batch_invariant(eng, cb):
    B, P = 300, 12                      # 300 streams: chunks of 32 frames -> 8 packets + 4 packets
    pcm = np.stack([(blob.synthetic_pcm(4 * P, 500 + s).astype(np.int32) * (1 + s % 3) // 2).astype(np.int16) for s in range(B)])
    pcm[7, 640 * 3:640 * 8] = 0
    fb = eng.FeatureBatch(B)
    fb.set_codebooks(cb)
    pk, qf = fb.encode(pcm, want_features=True)
    fb.close()
    lib = orcpy.OrcLib()
    for s in (0, 7, 150, 299):
        wp, wf = orcpy.OrcEncoder(lib).encode(pcm[s], cb)
        assert np.array_equal(pk[s], wp), s
        assert bits_equal(qf[s][:, :20], wf.reshape(-1, 36)[:, :20]), s
    pick = [3, 7, 211]
    small = eng.FeatureBatch(len(pick))
    small.set_codebooks(cb)
    again = small.encode(pcm[pick])
    small.close()
    assert np.array_equal(again, pk[pick])


def test_encode_then_decode_closes_the_wire_format_at_full_size(eng, cb, model_blob):
    """4096 streams x 5 packets: lpcnet_batch_decode's decode_packet of the encoder's packets returns the encoder's
    quantised rows bit for bit (a size-independent property; streams are voiced/unvoiced signals, not silence)."""
    B, P = 4096, 5
    base = np.stack([blob.synthetic_pcm(4 * P, 600 + s) for s in range(64)])
    pcm = np.stack([np.roll(base[s % 64], 37 * (s // 64)) for s in range(B)])
    fb = eng.FeatureBatch(B)
    fb.set_codebooks(cb)
    pk, qf = fb.encode(pcm, want_features=True)
    fb.close()
    bt = eng.Batch(model_blob, B)
    bt.set_codebooks(cb)
    _, feats = bt.decode(pk, want_features=True)
    bt.close()
    feats = np.asarray(feats).reshape(B, 4 * P, 36)
    assert bits_equal(feats[:, :, :20], qf[:, :, :20])
    assert len(np.unique(pk.reshape(-1, 8), axis=0)) > B       # the packets are not degenerate


def test_bad_arguments_are_errors(eng, cb):
    fb = eng.FeatureBatch(2)
    with pytest.raises(engine.LPCNetError):
        fb.encode(np.zeros((2, 640), np.int16))                 # no codebooks yet
    with pytest.raises(engine.LPCNetError):
        fb.set_codebooks(cb[:-1])
    fb.close()
