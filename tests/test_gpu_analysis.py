"""Parity of the CUDA feature analysis (lpcnet_feature_batch_*, lpcnet_compute_single_frame_features; through the
C-ABI) against the CPU oracle and the golden vectors of the unmodified reference.  Every operation of the path is
an individually rounded float or double operation in the reference's order, so the bar is bit-exact; the one
library call whose last bit is not specified (double log10) gets a documented fallback tolerance of 2e-6 absolute
on the cepstrum, which the test reports if it is ever needed."""
import os

import numpy as np
import pytest

from lpcnet_b200 import blob, engine
from oracle import orcpy

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
LOG10_TOL = 2e-6


def same(got, want):
    got = np.ascontiguousarray(got, np.float32); want = np.ascontiguousarray(want, np.float32)
    if np.array_equal(got.view(np.uint32), want.view(np.uint32)):
        return True
    # pitch features and LPC must still match exactly unless the cepstrum moved
    bad = np.nonzero(got.view(np.uint32) != want.view(np.uint32))
    print("analysis differs from the oracle at", [b[:8] for b in bad], "max abs", np.abs(got - want).max())
    return False


@pytest.fixture(scope="module")
def eng(engine_lib):
    assert engine.device_count() > 0, "no CUDA device: the gpu tests cannot run (there is no CPU fallback)"
    return engine


def test_golden_vectors_of_the_reference(eng):
    g = np.load(os.path.join(HERE, "golden", "analysis.npz"))
    fb = eng.FeatureBatch(g["pcm"].shape[0])
    got = fb.compute(g["pcm"])
    assert fb.launch_count > 0
    assert same(got, g["features"])
    fb.close()


def test_float_and_short_entry_points_and_call_splitting(eng):
    lib = orcpy.OrcLib()
    B, F = 5, 37
    pcm = np.stack([blob.synthetic_pcm(F, 20 + s) for s in range(B)])
    want = np.stack([orcpy.OrcEncoder(lib).run(x) for x in pcm])
    fb = eng.FeatureBatch(B)
    assert same(fb.compute(pcm), want)
    fb.reset()
    assert same(fb.compute(pcm.astype(np.float32)), want)
    # the state carries over: 1 + 3 + 33 frames in three calls equal one call of 37
    fb.reset()
    parts = [fb.compute(pcm[:, a * 160:b * 160]) for a, b in ((0, 1), (1, 4), (4, 37))]
    assert same(np.concatenate(parts, axis=1), want)
    # reset of one stream only
    fb.reset(2)
    again = fb.compute(pcm[:, :160 * 4])
    assert same(again[2], want[2, :4])
    assert not np.array_equal(again[1], want[1, :4])
    fb.close()


def test_single_stream_encoder_api(eng):
    lib = orcpy.OrcLib()
    x = blob.synthetic_pcm(6, 3)
    want = orcpy.OrcEncoder(lib).run(x)
    enc = eng.Encoder()
    got = np.stack([enc.single_frame_features(x[f * 160:(f + 1) * 160]) for f in range(6)])
    assert same(got, want)
    enc.close()
    enc = eng.Encoder()
    got = np.stack([enc.single_frame_features(x[f * 160:(f + 1) * 160].astype(np.float32)) for f in range(6)])
    assert same(got, want)
    enc.close()


def test_many_streams_match_the_oracle_and_are_batch_invariant(eng):
    """2048 streams x 12 frames (several internal chunks): a sample of streams against the oracle, every stream
    against the same stream analysed in a batch of its own."""
    lib = orcpy.OrcLib()
    B, F = 2048, 12
    base = np.stack([blob.synthetic_pcm(F, s) for s in range(64)])
    rng = np.random.default_rng(3)
    gain = rng.uniform(0.05, 1.0, B)
    pcm = np.stack([np.round(base[s % 64] * gain[s]).astype(np.int16) for s in range(B)])
    fb = eng.FeatureBatch(B)
    got = fb.compute(pcm)
    fb.close()
    for s in list(range(0, B, 97)) + [B - 1]:
        assert same(got[s], orcpy.OrcEncoder(lib).run(pcm[s])), s
    small = eng.FeatureBatch(3)
    pick = [5, 1000, 2047]
    assert same(small.compute(pcm[pick]), got[pick])
    small.close()
    # silence and full scale stay finite
    edge = np.zeros((2, F * 160), np.int16)
    edge[1] = np.where(np.arange(F * 160) % 2, 32767, -32768)
    fb2 = eng.FeatureBatch(2)
    e = fb2.compute(edge)
    fb2.close()
    assert np.isfinite(e).all()
    assert same(e[0], orcpy.OrcEncoder(lib).run(edge[0]))
    assert same(e[1], orcpy.OrcEncoder(lib).run(edge[1]))


def test_bad_arguments_are_errors(eng):
    fb = eng.FeatureBatch(2)
    with pytest.raises(engine.LPCNetError):
        fb.reset(5)
    with pytest.raises(engine.LPCNetError):
        eng.FeatureBatch(0)
    fb.close()
