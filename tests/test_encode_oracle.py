"""1.6 kb/s encoder oracle (oracle/orc_encode.c: lpcnet_encode, lpcnet_compute_features) against the unmodified
reference compiled in this container (oracle/_ref/libref_d2.so, src/lpcnet_enc.c:882-911, with the synthetic codebooks
of oracle/gen_codebooks.py linked in) and against the committed golden vectors generated from it
(tests/golden/encode.npz, tests/golden/make_golden.py).  Packets are bytes: the bar is bit-exact."""
import os

import numpy as np
import pytest

from lpcnet_b200 import blob
from oracle import orcpy, refpy

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = os.path.join(HERE, "golden", "encode.npz")

pytestmark = pytest.mark.skipif(not orcpy.available(), reason="oracle not built")


def _signals():
    n = 24 * 4
    sig = [blob.synthetic_pcm(n, 300 + s) for s in range(4)]
    x = blob.synthetic_pcm(n, 304).copy()
    x[640 * 6:640 * 14] = 0
    sig.append(x)
    sig.append(np.clip(blob.synthetic_pcm(n, 305).astype(np.int32) * 6, -32768, 32767).astype(np.int16))
    sig.append(np.random.default_rng(19).integers(-3000, 3000, n * 160).astype(np.int16))
    return sig


def test_golden_inputs_are_the_documented_signals():
    g = np.load(GOLD)
    for i, x in enumerate(_signals()):
        assert np.array_equal(g["pcm"][i], x)


def test_oracle_matches_golden_vectors():
    g = np.load(GOLD)
    lib, cb = orcpy.OrcLib(), blob.synthetic_codebooks()
    for i, x in enumerate(g["pcm"]):
        pk, _ = orcpy.OrcEncoder(lib).encode(x, cb)
        assert np.array_equal(pk, g["packets"][i]), i
        ft = orcpy.OrcEncoder(lib).compute_features(x)
        assert np.array_equal(ft.view(np.uint32), g["features"][i].view(np.uint32)), i


@pytest.mark.skipif(not refpy.available(2), reason="reference not compiled here")
def test_oracle_matches_reference_bit_for_bit():
    lib, ref, cb = orcpy.OrcLib(), refpy.RefLib(2), blob.synthetic_codebooks()
    for s in range(3):
        x = blob.synthetic_pcm(4 * 30, 400 + s)
        if s == 2:
            x = (x // 50).astype(np.int16)                 # near-silence
        assert np.array_equal(orcpy.OrcEncoder(lib).encode(x, cb)[0], refpy.RefEncoder(ref).encode(x))
        assert np.array_equal(orcpy.OrcEncoder(lib).compute_features(x).view(np.uint32),
                              refpy.RefEncoder(ref).compute_features(x).view(np.uint32))


def test_decoding_the_packets_rebuilds_the_encoders_quantised_rows():
    """decode_packet(lpcnet_encode(x)) returns the encoder's quantised cepstra, pitch and correlation (entries
    [0..20) of the four rows) bit for bit -- the wire format closes (voiced or unvoiced, not digital silence)."""
    lib, cb = orcpy.OrcLib(), blob.synthetic_codebooks()
    g = np.load(GOLD)
    for i in (0, 1, 5, 6):
        pk, ft = orcpy.OrcEncoder(lib).encode(g["pcm"][i], cb)
        vq = np.zeros(18, np.float32)
        for k in range(len(pk)):
            d = np.asarray(lib.decode_packet(vq, pk[k].tobytes(), cb)).reshape(4, 36)
            assert np.array_equal(d[:, :20].view(np.uint32), ft[k][:, :20].view(np.uint32)), (i, k)


def test_packet_fields_are_in_range():
    g = np.load(GOLD)
    bits = np.unpackbits(g["packets"].reshape(-1, 8), axis=1)
    field = lambda a, n: bits[:, a:a + n].dot(1 << np.arange(n - 1, -1, -1))
    assert (field(7, 6) <= 63).all() and (field(48, 13) < 8192).all()
    assert len(np.unique(field(18, 10))) > 20              # the first-stage index actually varies
