"""The reference's -DEND2END build (src/lpcnet.c:57-78, 108): the frame's LPC are rc2lpc of the first 16 outputs of
feature_dense2 -- no cepstrum, no FEATURES_DELAY line.  tests/golden/end2end.npz holds the outputs of the unmodified
reference compiled that way (oracle/_ref/libref_d2_e2e.so, tests/golden/make_golden.py end2end).  CPU: the oracle's
restatement against it.  GPU: lpcnet_batch_set_end2end against it and against the oracle, through the C-ABI."""
import hashlib
import os

import numpy as np
import pytest

from lpcnet_b200 import blob, engine
from oracle import orcpy

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "end2end.npz")


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


@pytest.fixture(scope="module")
def g(model_blob):
    d = np.load(GOLDEN)
    assert bytes(d["blob_sha256"]) == hashlib.sha256(model_blob).digest()
    return d


@pytest.fixture(scope="module")
def orc_e2e(orclib, model_blob):
    return orcpy.OrcModel(orclib, model_blob, 2, 1.0, None, end2end=True)


def test_oracle_end2end_against_reference(orc_e2e, g):
    st = orcpy.OrcStream(orc_e2e)
    f = blob.synthetic_features(6, 0)
    for k in range(6):
        _, gb, lpc = st.frame_network(f[k])
        assert np.array_equal(bits(lpc), bits(g["fn_lpc"][k])), k
        assert np.array_equal(bits(gb), bits(g["fn_gb"][k]))
    assert np.abs(g["fn_lpc"][0]).max() > 0          # no warm-up zeros: there is no delay line in this build
    for s in g["streams"]:
        pcm, tr, _ = orcpy.OrcStream(orc_e2e, blob.stream_seed(int(s))).run(blob.synthetic_features(10, int(s)), trace=True)
        assert np.array_equal(tr, g[f"trace_{s}"])
        assert np.array_equal(pcm, g[f"pcm_{s}"])


@pytest.mark.gpu
def test_engine_end2end_against_reference(engine_lib, model_blob, g):
    assert engine.device_count() > 0, "no CUDA device: the gpu tests cannot run (there is no CPU fallback)"
    bt = engine.Batch(model_blob, 1)
    bt.set_end2end(True)
    _, gb, lpc = bt.frame_network(blob.synthetic_features(6, 0)[None])
    assert np.array_equal(bits(lpc[0]), bits(g["fn_lpc"]))
    assert np.array_equal(bits(gb[0]), bits(g["fn_gb"]))
    bt.close()
    ids = [int(s) for s in g["streams"]]
    for tile in (8, 16):
        bt = engine.Batch(model_blob, len(ids))
        bt.set_end2end(True)
        bt.set_tile(tile)
        for k, s in enumerate(ids):
            bt.set_seed(k, blob.stream_seed(s))
        feats = np.stack([blob.synthetic_features(10, s) for s in ids])
        pcm, tr, _ = bt.synthesize(feats, trace=True)
        for k, s in enumerate(ids):
            bad = np.nonzero(tr[k] != g[f"trace_{s}"])[0]
            assert bad.size == 0, f"tile {tile} stream {s}: first divergence at step {bad[0]}"
            assert np.array_equal(pcm[k], g[f"pcm_{s}"])
        # the production (no-trace) kernel over two calls, and the switch back
        bt.reset()
        for k, s in enumerate(ids):
            bt.set_seed(k, blob.stream_seed(s))
        a, _, _ = bt.synthesize(feats[:, :4])
        b, _, _ = bt.synthesize(feats[:, 4:])
        assert np.array_equal(np.concatenate([a, b], axis=1), pcm)
        bt.close()
    # process-wide default for the drop-in states
    L = engine.load_library()
    L.lpcnet_b200_set_end2end(1)
    try:
        syn = engine.Synth(model_blob)
        f = blob.synthetic_features(10, ids[0])
        # (lpcnet_create's kiss99 seed is the reference's: compare with the golden's default-seeded frame network only)
        out = np.concatenate([syn.synthesize(f[k]) for k in range(3)])
        assert out.size == 480
    finally:
        L.lpcnet_b200_set_end2end(0)
    bt = engine.Batch(model_blob, 1)
    _, _, lpc = bt.frame_network(blob.synthetic_features(6, 0)[None])
    assert np.all(lpc[0][:2] == 0)                   # FEATURES_DELAY line again
    bt.close()
