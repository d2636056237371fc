"""Feature-analysis oracle (oracle/orc_analysis.c) against the unmodified reference compiled in this container
(oracle/_ref/libref_d2.so: lpcnet_compute_single_frame_features[_float], src/lpcnet_enc.c:913-933) and against
the committed golden vectors generated from it (tests/golden/analysis.npz, tests/golden/make_golden.py)."""
import os

import numpy as np
import pytest

from lpcnet_b200 import blob
from oracle import orcpy, refpy

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = os.path.join(HERE, "golden", "analysis.npz")

pytestmark = pytest.mark.skipif(not orcpy.available(), reason="oracle not built")


def _signals():
    sig = [blob.synthetic_pcm(40, s) for s in range(4)]
    sig.append(np.zeros(40 * 160, np.int16))                                       # silence
    sig.append(np.random.default_rng(9).integers(-32768, 32767, 40 * 160).astype(np.int16))   # full-scale noise
    x = blob.synthetic_pcm(40, 7).copy()
    x[1600:3200] = 0                                                               # a gap: the energy floor path
    sig.append(x)
    return sig


def test_oracle_matches_golden_vectors():
    g = np.load(GOLD)
    lib = orcpy.OrcLib()
    for i, x in enumerate(g["pcm"]):
        got = orcpy.OrcEncoder(lib).run(x)
        assert np.array_equal(got.view(np.uint32), g["features"][i].view(np.uint32)), i


def test_golden_inputs_are_the_documented_signals():
    g = np.load(GOLD)
    for i, x in enumerate(_signals()):
        assert np.array_equal(g["pcm"][i], x)


@pytest.mark.skipif(not refpy.available(2), reason="reference not compiled here")
def test_oracle_matches_reference_bit_for_bit():
    lib, ref = orcpy.OrcLib(), refpy.RefLib(2)
    assert np.array_equal(orcpy.half_window(lib).view(np.uint32), refpy.ref_half_window(ref).view(np.uint32))
    for x in _signals():
        want = refpy.RefEncoder(ref).run(x)
        got = orcpy.OrcEncoder(lib).run(x)
        assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
        # the float entry point sees the same values
        wantf = refpy.RefEncoder(ref).run(x.astype(np.float32))
        assert np.array_equal(wantf.view(np.uint32), want.view(np.uint32))


def test_features_have_the_documented_layout():
    lib = orcpy.OrcLib()
    f = orcpy.OrcEncoder(lib).run(blob.synthetic_pcm(30, 1))
    period = f[:, 18] * 100 + 200          # .01*(2*period - 200)
    assert ((period >= 66) & (period <= 510)).all()
    assert ((f[:, 19] >= -0.5 - 1e-6) & (f[:, 19] <= 0.5 + 1e-6)).all()
    assert np.isfinite(f).all()
