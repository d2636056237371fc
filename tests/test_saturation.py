"""Models whose int8 weights use the whole range [-127, 127] (blob.saturating_models), as trained models may: the
reference's vpmaddubsw adds adjacent input pairs in a SATURATING s16 (src/vec_avx.h:814-818).  The loader used to refuse
such blobs; now the pairs that can saturate leave the tensor-core / dp4a operands and take an exact s16 pass
(csrc/engine.h DevSat).  Golden vectors: tests/golden/sat.npz, produced by the unmodified reference
(tests/golden/make_golden.py sat); without the saturation the very first active step differs."""
import hashlib
import os

import numpy as np
import pytest

from lpcnet_b200 import blob, engine
from oracle import orcpy

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "sat.npz")


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


@pytest.fixture(scope="module")
def sat_blob():
    return blob.saturating_models()


@pytest.fixture(scope="module")
def g(sat_blob):
    d = np.load(GOLDEN)
    assert bytes(d["blob_sha256"]) == hashlib.sha256(sat_blob).digest()
    return d


def test_model_has_saturating_pairs(sat_blob):
    A = blob.parse_blob(sat_blob)
    for name in ("sparse_gru_a_recurrent_weights", "gru_b_weights", "gru_b_recurrent_weights", "plc_gru1_weights",
                 "plc_gru2_recurrent_weights", "dec_dense2_weights", "dec_dense6_recurrent_weights"):
        assert not blob.saturation_free(A[name]), name


def test_oracle_matches_reference_on_saturating_model(orclib, sat_blob, g):
    om = orcpy.OrcModel(orclib, sat_blob, 2, 1.0, None)
    for s in (0, 5, 9):
        pcm, tr, _ = orcpy.OrcStream(om, blob.stream_seed(s)).run(blob.synthetic_features(10, 400 + s), trace=True)
        assert np.array_equal(tr, g[f"trace_{s}"]) and np.array_equal(pcm, g[f"pcm_{s}"]), s
    om0 = orcpy.OrcModel(orclib, sat_blob, 0, 1.0, None)
    for s in range(2):
        out = orcpy.OrcPLC(om0, 5, blob.stream_seed(s)).run(g[f"plc_in_{s}"], g[f"plc_lost_{s}"])
        assert np.array_equal(out, g[f"plc_out_{s}"]), s
    assert np.array_equal(bits(orcpy.dred_decode_all(om, g["dred_state"], g["dred_latents"])), bits(g["dred_features"]))


@pytest.mark.gpu
@pytest.mark.parametrize("tile", [8, 16, 32])
def test_gpu_synthesis_on_saturating_model(engine_lib, sat_blob, g, tile):
    """Ragged batch on every tile size: the trace / logit kernel (every step's record) and the production kernel (PCM)."""
    ids = [0, 5, 9, 5, 0, 9, 9, 0, 5, 0, 5]
    feats = np.stack([blob.synthetic_features(10, 400 + s) for s in ids])
    for trace in (True, False):
        bt = engine.Batch(sat_blob, len(ids))
        assert bt.saturating_pairs > 1000 and not bt.fixed_slots
        bt.set_tile(tile)
        for k, s in enumerate(ids):
            bt.set_seed(k, blob.stream_seed(s))
        pcm, tr, _ = bt.synthesize(feats, trace=trace)
        for k, s in enumerate(ids):
            if trace:
                bad = np.nonzero(tr[k] != g[f"trace_{s}"])[0]
                assert bad.size == 0, f"slot {k}: first divergence at step {bad[0]}: {tr[k][bad[0]]} vs {g[f'trace_{s}'][bad[0]]}"
            assert np.array_equal(pcm[k], g[f"pcm_{s}"]), (k, trace)
        bt.close()


@pytest.mark.gpu
def test_gpu_large_batch_on_saturating_model(engine_lib, sat_blob, g):
    """2400 streams: whole waves of 16-stream tiles plus the remainder as a sub-range launch of 8-stream tiles."""
    B = 16 * 148 + 32
    ids = np.array([0, 5, 9])[np.arange(B) % 3]
    base = {s: blob.synthetic_features(10, 400 + s) for s in (0, 5, 9)}
    feats = np.stack([base[s] for s in ids])
    bt = engine.Batch(sat_blob, B)
    for k, s in enumerate(ids):
        bt.set_seed(k, blob.stream_seed(int(s)))
    pcm, _, _ = bt.synthesize(feats)
    for s in (0, 5, 9):
        assert (pcm[ids == s] == g[f"pcm_{s}"]).all(), s
    bt.close()


@pytest.mark.gpu
def test_gpu_plc_and_dred_on_saturating_model(engine_lib, sat_blob, g):
    pb = engine.PLCBatch(sat_blob, 2, engine.PLC_NONCAUSAL | engine.PLC_DC_FILTER)
    for s in range(2):
        pb.set_seed(s, blob.stream_seed(s))
    out = pb.process(np.stack([g["plc_in_0"], g["plc_in_1"]]), np.stack([g["plc_lost_0"], g["plc_lost_1"]]))
    for s in range(2):
        bad = np.nonzero(out[s] != g[f"plc_out_{s}"])[0]
        assert bad.size == 0, f"PLC stream {s}: first divergence at sample {bad[0]}"
    pb.close()
    dec = engine.DredDecoder(sat_blob)
    assert np.array_equal(bits(dec.decode_all(g["dred_state"], g["dred_latents"])), bits(g["dred_features"]))
    dec.close()
