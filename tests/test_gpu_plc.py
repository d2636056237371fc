"""Parity of the CUDA packet-loss concealment (lpcnet_plc_batch_*, lpcnet_plc_update / _conceal; through the C-ABI)
against the golden vectors of the unmodified reference (tests/golden/plc.npz, FEATURES_DELAY 0 build) and the CPU
oracle.  The output is PCM (int16): the bar is bit-exact, streams under different loss patterns side by side."""
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
def plc_blob():
    A = blob.random_model()
    A.update(blob.random_plc_model())
    return blob.write_blob(A)


@pytest.fixture(scope="module")
def plc_model(plc_blob):
    return orcpy.OrcModel(orcpy.OrcLib(), plc_blob, 0, 1.0, None)


def first_diff(got, want):
    d = np.nonzero(got != want)
    return None if d[0].size == 0 else tuple(int(v[0]) for v in d)


@pytest.mark.parametrize("opt", [1, 5])
def test_golden_vectors_of_the_reference(eng, plc_blob, opt):
    g = np.load(os.path.join(HERE, "golden", "plc.npz"))
    assert hashlib.sha256(plc_blob).digest() == g["blob_sha256"].tobytes()
    pcm = np.stack([g[f"in_{opt}_{s}"] for s in range(3)])
    lost = np.stack([g[f"lost_{opt}_{s}"] for s in range(3)])
    want = np.stack([g[f"out_{opt}_{s}"] for s in range(3)])
    pb = eng.PLCBatch(plc_blob, 3, opt)
    for s in range(3):
        pb.set_seed(s, blob.stream_seed(s))
    got = pb.process(pcm, lost)
    assert pb.launch_count > 0
    pb.close()
    assert first_diff(got, want) is None, ("first difference (stream, sample)", first_diff(got, want))


def test_call_splitting_and_single_stream_api(eng, plc_blob):
    g = np.load(os.path.join(HERE, "golden", "plc.npz"))
    x, lost, want = g["in_1_0"], g["lost_1_0"], g["out_1_0"]
    # frame by frame through the reference's own entry points (stream 0 = the default "LPCNet" seed)
    plc = eng.PLC(plc_blob, 1)
    out = []
    for f in range(len(lost)):
        out.append(plc.conceal() if lost[f] else plc.update(x[f * 160:(f + 1) * 160]))
    plc.close()
    assert first_diff(np.concatenate(out), want) is None
    # the same in three batched calls of 7 + 1 + 32 frames
    pb = eng.PLCBatch(plc_blob, 1, 1)
    parts = [pb.process(x[None, a * 160:b * 160], lost[None, a:b]) for a, b in ((0, 7), (7, 8), (8, 40))]
    pb.close()
    assert first_diff(np.concatenate(parts, axis=1)[0], want) is None


def test_reset_and_fec_calls(eng, plc_blob):
    """lpcnet_plc_reset returns a used state to its initial one; FEC features are accepted and -- as in the
    reference's non-causal mode, which never reads the queue -- leave the output unchanged."""
    g = np.load(os.path.join(HERE, "golden", "plc.npz"))
    x, lost, want = g["in_1_0"], g["lost_1_0"], g["out_1_0"]
    plc = eng.PLC(plc_blob, 1)
    for f in range(9):
        plc.conceal() if lost[f] else plc.update(x[f * 160:(f + 1) * 160])
    plc.reset()
    out = []
    for f in range(len(lost)):
        plc.fec_add(np.full(20, 0.25, np.float32) if f % 3 else None)
        out.append(plc.conceal() if lost[f] else plc.update(x[f * 160:(f + 1) * 160]))
        if f % 5 == 4:
            plc.fec_clear()
    plc.close()
    assert first_diff(np.concatenate(out), want) is None


def test_many_streams_against_the_oracle(eng, plc_blob, plc_model):
    """96 streams x 30 frames with the DC filter, every stream under its own Gilbert loss pattern (long bursts for
    some): a sample of streams against the oracle, and batch invariance for the rest."""
    B, F, opt = 96, 30, 5
    pcm = np.stack([np.clip(blob.synthetic_pcm(F, 100 + s).astype(np.int32) + 300 * (s % 5), -32768, 32767).astype(np.int16)
                    for s in range(B)])
    lost = np.stack([blob.synthetic_loss_pattern(F, s, 0.6 + 0.3 * (s % 3 == 0), 0.12) for s in range(B)])
    assert lost.sum() > B
    pb = eng.PLCBatch(plc_blob, B, opt)
    for s in range(B):
        pb.set_seed(s, blob.stream_seed(s))
    got = pb.process(pcm, lost)
    pb.close()
    for s in (0, 1, 17, 48, 95):
        want = orcpy.OrcPLC(plc_model, opt, blob.stream_seed(s)).run(pcm[s], lost[s])
        assert first_diff(got[s], want) is None, (s, first_diff(got[s], want))
    pick = [3, 50, 77]
    small = eng.PLCBatch(plc_blob, len(pick), opt)
    for i, s in enumerate(pick):
        small.set_seed(i, blob.stream_seed(s))
    again = small.process(pcm[pick], lost[pick])
    small.close()
    assert first_diff(again, got[pick]) is None


def test_full_size_config4_compaction_is_invisible(eng, plc_blob, plc_model):
    """BASELINE config 4 size: 4096 streams, noncausal_dc, Gilbert losses.  The gathered (compact) sample segments
    and the in-place ones give the same PCM for every stream, and sampled streams match the oracle."""
    B, F, opt = 4096, 14, 5
    base = [blob.synthetic_pcm(F, 100 + s) for s in range(32)]
    pcm = np.stack([base[s % 32] for s in range(B)])
    lost = np.stack([blob.synthetic_loss_pattern(F, 7 + s, 0.6, 0.1) for s in range(B)])
    assert 0.05 < lost.mean() < 0.4

    def run():
        pb = eng.PLCBatch(plc_blob, B, opt)
        for s in range(0, B, 7):
            pb.set_seed(s, blob.stream_seed(s))
        out = pb.process(pcm, lost)
        pb.close()
        return out

    got = run()
    os.environ["LPCNET_B200_NO_COMPACT"] = "1"
    try:
        plain = run()
    finally:
        del os.environ["LPCNET_B200_NO_COMPACT"]
    assert first_diff(got, plain) is None
    for s in (0, 7, 2051, 4095):
        seed = blob.stream_seed(s) if s % 7 == 0 else b"LPCNet"
        want = orcpy.OrcPLC(plc_model, opt, seed).run(pcm[s], lost[s])
        assert first_diff(got[s], want) is None, (s, first_diff(got[s], want))


def test_side_stream_is_invisible_and_deterministic(eng, plc_blob, plc_model):
    """The first-good-frame chain runs on a side CUDA stream beside the main class (plc.cu).  Streams of the two
    chains share tiles, scratch rows and the analysis state arrays, so a missing event or an unmasked write would show
    as a difference from the sequential order (LPCNET_B200_PLC_ONE_STREAM=1) or between repeated runs.  Loss patterns
    with many recoveries per frame (bursts of 1-2 frames at a 35 % loss rate) keep both chains busy in every frame;
    every stream is compared between the modes, sampled ones with the oracle."""
    B, F, opt = 2500, 30, 5
    base = [blob.synthetic_pcm(F, 300 + s) for s in range(32)]
    pcm = np.stack([base[s % 32] for s in range(B)])
    lost = np.stack([blob.synthetic_loss_pattern(F, 900 + s, 0.45, 0.7) for s in range(B)])
    assert 0.2 < lost.mean() < 0.6

    def run():
        pb = eng.PLCBatch(plc_blob, B, opt)
        for s in range(0, B, 5):
            pb.set_seed(s, blob.stream_seed(s))
        out = pb.process(pcm, lost)
        pb.close()
        return out

    runs = [run() for _ in range(3)]
    os.environ["LPCNET_B200_PLC_ONE_STREAM"] = "1"
    try:
        seq = run()
    finally:
        del os.environ["LPCNET_B200_PLC_ONE_STREAM"]
    for k, r in enumerate(runs):
        assert first_diff(r, seq) is None, (k, first_diff(r, seq))
    for s in (0, 5, 1249, 2499):
        seed = blob.stream_seed(s) if s % 5 == 0 else b"LPCNet"
        want = orcpy.OrcPLC(plc_model, opt, seed).run(pcm[s], lost[s])
        assert first_diff(runs[0][s], want) is None, (s, first_diff(runs[0][s], want))


def run_with_fec(pb, pcm, lost, fec):
    """frame by frame, handing each stream's FEC vectors to the queue before the frame they precede"""
    B, F = lost.shape
    out = []
    for f in range(F):
        for s in range(B):
            for v in fec[s].get(f, []):
                pb.fec_add(s, v)
        out.append(pb.process(pcm[:, f * 160:(f + 1) * 160], lost[:, f:f + 1]))
    return np.concatenate(out, axis=1)


@pytest.mark.parametrize("opt", [0, 4, 2, 6])
def test_causal_and_codec_modes_golden_vectors(eng, plc_blob, opt):
    """lpcnet_plc_update_causal / _conceal_causal with the FEC queue (FEATURES_DELAY 2 build of the reference)"""
    g = np.load(os.path.join(HERE, "golden", "plc_causal.npz"))
    assert hashlib.sha256(plc_blob).digest() == g["blob_sha256"].tobytes()
    pcm = np.stack([g[f"in_{opt}_{s}"] for s in range(3)])
    lost = np.stack([g[f"lost_{opt}_{s}"] for s in range(3)])
    want = np.stack([g[f"out_{opt}_{s}"] for s in range(3)])
    pb = eng.PLCBatch(plc_blob, 3, opt)
    for s in range(3):
        pb.set_seed(s, blob.stream_seed(s))
    got = run_with_fec(pb, pcm, lost, [blob.synthetic_fec(lost.shape[1], s) for s in range(3)])
    pb.close()
    assert first_diff(got, want) is None, ("first difference (stream, sample)", first_diff(got, want))


def test_causal_mode_many_streams_in_one_call(eng, plc_blob):
    """64 streams x 24 frames in a single batched call (no FEC), the causal mode with the DC filter: sampled streams
    against the oracle, batch invariance for the rest; and the single-stream entry points with FEC."""
    B, F, opt = 64, 24, 4
    model = orcpy.OrcModel(orcpy.OrcLib(), plc_blob, 2, 1.0, None)
    pcm = np.stack([np.clip(blob.synthetic_pcm(F, 200 + s).astype(np.int32) + 250 * (s % 4), -32768, 32767).astype(np.int16)
                    for s in range(B)])
    lost = np.stack([blob.synthetic_loss_pattern(F, s, 0.55 + 0.3 * (s % 3 == 0), 0.15) for s in range(B)])
    pb = eng.PLCBatch(plc_blob, B, opt)
    for s in range(B):
        pb.set_seed(s, blob.stream_seed(s))
    got = pb.process(pcm, lost)
    pb.close()
    for s in (0, 3, 31, 63):
        want = orcpy.OrcPLC(model, opt, blob.stream_seed(s)).run(pcm[s], lost[s])
        assert first_diff(got[s], want) is None, (s, first_diff(got[s], want))
    pick = [5, 40]
    small = eng.PLCBatch(plc_blob, len(pick), opt)
    for i, s in enumerate(pick):
        small.set_seed(i, blob.stream_seed(s))
    assert first_diff(small.process(pcm[pick], lost[pick]), got[pick]) is None
    small.close()
    # a long call builds its control rows in pieces: forced to 5 frames per piece here
    os.environ["LPCNET_B200_PLC_CTL_FRAMES"] = "5"
    try:
        pb = eng.PLCBatch(plc_blob, B, opt)
        for s in range(B):
            pb.set_seed(s, blob.stream_seed(s))
        assert first_diff(pb.process(pcm, lost), got) is None
        pb.close()
    finally:
        del os.environ["LPCNET_B200_PLC_CTL_FRAMES"]
    # lpcnet_plc_create(LPCNET_PLC_CODEC) + fec_add / update / conceal, one stream
    s, fec = 2, blob.synthetic_fec(F, 2)
    plc = eng.PLC(plc_blob, 2)
    out = []
    for f in range(F):
        for v in fec.get(f, []):
            plc.fec_add(v)
        out.append(plc.conceal() if lost[s, f] else plc.update(pcm[s, f * 160:(f + 1) * 160]))
    plc.close()
    want = orcpy.OrcPLC(model, 2, b"LPCNet").run(pcm[s], lost[s], fec)
    assert first_diff(np.concatenate(out), want) is None


def test_unsupported_modes_and_models_are_errors(eng, plc_blob, model_blob):
    with pytest.raises(engine.LPCNetError):
        eng.PLCBatch(plc_blob, 2, 3)            # 3 is not a mode (lpcnet_plc.c:64-76)
    with pytest.raises(engine.LPCNetError):
        eng.PLCBatch(model_blob, 2, 1)          # no plc_* arrays in the blob
