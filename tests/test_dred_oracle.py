"""DRED oracle (oracle/orc_dred.c: DRED_rdovae_decode_all) against the golden vectors of the unmodified reference
decoder (tests/golden/dred.npz, generated from oracle/_ref/libref_d2.so: src/dred_rdovae_dec.c compiled with stand-ins
for its generated data files) and, where the reference is compiled, against it live.  float32, compared bit for bit."""
import hashlib
import os

import numpy as np
import pytest

from lpcnet_b200 import blob
from oracle import orcpy, refpy

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = os.path.join(HERE, "golden", "dred.npz")
pytestmark = pytest.mark.skipif(not orcpy.available(), reason="oracle not built")


@pytest.fixture(scope="module")
def dred_blob():
    A = blob.random_model()
    A.update(blob.random_dred_decoder_model())
    return blob.write_blob(A)


def test_oracle_matches_golden_vectors(dred_blob):
    g = np.load(GOLD)
    assert hashlib.sha256(dred_blob).digest() == g["blob_sha256"].tobytes()
    m = orcpy.OrcModel(orcpy.OrcLib(), dred_blob, 2, 1.0, None)
    for k in range(3):
        got = orcpy.dred_decode_all(m, g[f"state_{k}"], g[f"latents_{k}"])
        assert np.array_equal(got.view(np.uint32), g[f"features_{k}"].view(np.uint32)), k
        assert np.isfinite(got).all() and np.abs(got).max() > 0.1


@pytest.mark.skipif(not refpy.available(2), reason="reference not compiled here")
def test_oracle_matches_reference_bit_for_bit(dred_blob):
    R = refpy.RefLib(2)
    m = orcpy.OrcModel(orcpy.OrcLib(), dred_blob, 2, 1.0, None)
    rng = np.random.default_rng(77)
    for n in (2, 13):
        st = (rng.standard_normal(24) * 2).astype(np.float32)       # large values: the tanh / sigmoid clamps
        z = (rng.standard_normal((n, 80)) * 3).astype(np.float32)
        want = refpy.ref_dred_decode_all(R, dred_blob, st, z)
        assert np.array_equal(orcpy.dred_decode_all(m, st, z).view(np.uint32), want.view(np.uint32))


def test_model_without_decoder_is_refused():
    m = orcpy.OrcModel(orcpy.OrcLib(), blob.write_blob(blob.random_model()), 2, 1.0, None)
    with pytest.raises(RuntimeError):
        orcpy.dred_decode_all(m, np.zeros(24, np.float32), np.zeros((1, 80), np.float32))
