"""Parity of the CUDA RDOVAE decoder of DRED (lpcnet_dred_decode_all / _batch through the C-ABI) against the golden
vectors of the unmodified reference decoder (tests/golden/dred.npz) and the CPU oracle; float32 compared bit for bit.
The decoded vectors then drive the causal PLC through lpcnet_plc_fec_add, as DRED's do in the reference."""
import hashlib
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
def dred_blob():
    A = blob.random_model()
    A.update(blob.random_dred_decoder_model())
    A.update(blob.random_plc_model())
    return blob.write_blob(A)


def bits_equal(a, b):
    return np.array_equal(np.asarray(a).view(np.uint32), np.asarray(b).view(np.uint32))


def test_golden_vectors_of_the_reference(eng):
    g = np.load(os.path.join(HERE, "golden", "dred.npz"))
    A = blob.random_model()
    A.update(blob.random_dred_decoder_model())
    b = blob.write_blob(A)
    assert hashlib.sha256(b).digest() == g["blob_sha256"].tobytes()
    dec = eng.DredDecoder(b)
    for k in range(3):
        got = dec.decode_all(g[f"state_{k}"], g[f"latents_{k}"])
        assert bits_equal(got, g[f"features_{k}"]), k
    assert dec.launch_count == 3
    dec.close()


def test_batch_matches_the_oracle_and_single_calls(eng, dred_blob):
    B, n = 700, 12
    rng = np.random.default_rng(5)
    states = (rng.standard_normal((B, 24)) * 1.5).astype(np.float32)
    latents = (rng.standard_normal((B, n, 80)) * 2).astype(np.float32)
    dec = eng.DredDecoder(dred_blob)
    got = dec.decode_batch(states, latents)
    m = orcpy.OrcModel(orcpy.OrcLib(), dred_blob, 2, 1.0, None)
    for s in (0, 1, 333, 699):
        assert bits_equal(got[s], orcpy.dred_decode_all(m, states[s], latents[s])), s
        assert bits_equal(got[s], dec.decode_all(states[s], latents[s])), s
    dec.close()


def test_decoded_vectors_feed_the_causal_plc(eng, dred_blob):
    """DRED -> lpcnet_plc_fec_add -> causal concealment, against the oracle chain fed the same way"""
    F, s = 20, 1
    rng = np.random.default_rng(9)
    dec = eng.DredDecoder(dred_blob)
    feats = dec.decode_all((rng.standard_normal(24)).astype(np.float32), (rng.standard_normal((3, 80))).astype(np.float32))
    dec.close()
    vecs = feats.reshape(-1, 20)[::-1]                 # 12 vectors, oldest first
    fec = {4: list(vecs[:6]), 9: [None] + list(vecs[6:])}
    pcm = blob.synthetic_pcm(F, 77)
    lost = np.zeros(F, bool)
    lost[5:8] = True
    lost[10:13] = True
    plc = eng.PLC(dred_blob, 0)
    out = []
    for f in range(F):
        for v in fec.get(f, []):
            plc.fec_add(v)
        out.append(plc.conceal() if lost[f] else plc.update(pcm[f * 160:(f + 1) * 160]))
    plc.close()
    m = orcpy.OrcModel(orcpy.OrcLib(), dred_blob, 2, 1.0, None)
    want = orcpy.OrcPLC(m, 0, b"LPCNet").run(pcm, lost, fec)
    assert np.array_equal(np.concatenate(out), want)


def test_missing_arrays_are_an_error(eng, model_blob):
    with pytest.raises(engine.LPCNetError):
        eng.DredDecoder(model_blob)
