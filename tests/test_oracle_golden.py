"""The CPU oracle against the golden vectors produced by the unmodified reference
(tests/golden/make_golden.py), and -- when oracle/_ref is present -- against the reference live."""
import hashlib

import numpy as np
import pytest

from lpcnet_b200 import blob
from oracle import orcpy, refpy


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def test_rcp_table_model(orclib, golden_prims):
    g = golden_prims
    assert np.array_equal(orclib.default_rcp_table(), g["rcp_table"])
    orclib.set_rcp_table(None)
    y = orclib.map1("orc_rcp", g["rcp_x"]).astype(np.float32)
    assert np.array_equal(bits(y), bits(g["rcp_y"]))


def test_activations(orclib, golden_prims):
    g = golden_prims
    orclib.set_rcp_table(None)
    assert np.array_equal(bits(orclib.map1("orc_tanh_approx", g["act_x"])), bits(g["tanh"]))
    assert np.array_equal(bits(orclib.map1("orc_sigmoid_approx", g["act_x"])), bits(g["sigmoid"]))


def test_mulaw(orclib, golden_prims):
    g = golden_prims
    got = np.array([orclib.lib.orc_lin2ulaw(orcpy.C.c_float(float(v))) for v in g["lin_x"]], np.int32)
    assert np.array_equal(got, g["lin2ulaw"])
    u = orclib.map1("orc_ulaw2lin", np.arange(256, dtype=np.float32))
    assert np.array_equal(bits(u), bits(g["ulaw2lin"]))


@pytest.mark.parametrize("seed", [b"LPCNet", b"LPCNet17", b"x", b"", b"abcdefgh"])
def test_kiss99(orclib, golden_prims, seed):
    g = golden_prims
    st = orclib.kiss99_srand(seed)
    states = g["kiss_state_" + seed.decode()]
    assert np.array_equal(st, states[0])
    for k in range(16):
        assert orclib.kiss99_rand(st) == int(g["kiss_draws_" + seed.decode()][k])
        assert np.array_equal(st, states[k + 1])


def test_int8_kernels(orclib, golden_prims):
    g = golden_prims
    assert np.array_equal(orclib.ps_to_epi8(g["quant_x"]), g["quant_u8"])
    assert np.array_equal(bits(orclib.sparse_sgemv(g["sp_out0"], g["sp_w"], g["sp_idx"], g["sp_x"])), bits(g["sp_out"]))
    assert np.array_equal(bits(orclib.sgemv8x4(g["sp_out0"], g["de_w"], g["de_x"])), bits(g["de_out"]))


def test_float_matvec(orclib, golden_prims):
    g = golden_prims
    assert np.array_equal(bits(orclib.sgemv(g["f_out0"], g["f_w"], 40, g["f_x"])), bits(g["f_out"]))


def test_lpc_from_cepstrum(orclib, golden_prims):
    g = golden_prims
    got = np.stack([orclib.lpc_from_cepstrum(c) for c in g["ceps"]])
    assert np.array_equal(bits(got), bits(g["lpc"]))


def test_blob_is_the_fixture_blob(model_blob, golden_synth):
    assert np.array_equal(np.frombuffer(hashlib.sha256(model_blob).digest(), np.uint8), golden_synth["blob_sha256"])


@pytest.mark.parametrize("s", [0, 3, 11])
def test_synthesis_free_running(orc_model, golden_synth, s):
    g = golden_synth
    st = orcpy.OrcStream(orc_model, blob.stream_seed(s))
    pcm, tr, lg = st.run(g[f"feat_{s}"], trace=True, logits=True)
    assert np.array_equal(pcm, g[f"pcm_{s}"])
    assert np.array_equal(tr, g[f"trace_{s}"])
    assert np.array_equal(bits(lg[480:576]), bits(g[f"logits_{s}"]))
    snap = st.snapshot()
    ref = g[f"state_{s}"][0]
    for name in snap.dtype.names:
        assert np.array_equal(np.asarray(snap[name]), np.asarray(ref[name])), name


@pytest.mark.parametrize("s", [0, 3, 11])
def test_synthesis_teacher_forced(orc_model, golden_synth, s):
    g = golden_synth
    st = orcpy.OrcStream(orc_model, blob.stream_seed(s))
    pcm, tr, lg = st.run(g[f"feat_{s}"], forced=g[f"forced_{s}"], trace=True, logits=True)
    assert np.array_equal(pcm, g[f"fpcm_{s}"])
    assert np.array_equal(tr, g[f"ftrace_{s}"])
    assert np.array_equal(bits(lg[480:576]), bits(g[f"flogits_{s}"]))


def test_public_api_sequence(orc_model, golden_synth):
    g = golden_synth
    st = orcpy.OrcStream(orc_model)
    f = blob.synthetic_features(6, 0)
    outs = [st.synthesize(f[k], 160 if k < 5 else 80) for k in range(6)]
    assert np.array_equal(np.concatenate(outs), g["api_pcm"])


def test_frame_network(orc_model, golden_synth):
    g = golden_synth
    st = orcpy.OrcStream(orc_model)
    f = blob.synthetic_features(6, 0)
    for k in range(4):
        ga, gb, lpc = st.frame_network(f[k])
        assert np.array_equal(bits(ga), bits(g["fn_ga"][k]))
        assert np.array_equal(bits(gb), bits(g["fn_gb"][k]))
        assert np.array_equal(bits(lpc), bits(g["fn_lpc"][k]))


def test_decode_packet(orclib, golden_decode, codebooks):
    g = golden_decode
    vq = np.zeros(18, np.float32)
    for p, want in zip(g["packets"], g["features"]):
        got = orclib.decode_packet(vq, p.tobytes(), codebooks)
        assert np.array_equal(bits(got), bits(want))


@pytest.mark.skipif(not refpy.available(2), reason="oracle/_ref not built (needs /root/reference)")
def test_live_against_reference(orclib, model_blob):
    """Fresh seeds/features not in the fixtures: oracle == compiled reference, bit for bit."""
    R = refpy.RefLib(2)
    tab = R.rcp_table().view(np.uint32)
    om = orcpy.OrcModel(orclib, model_blob, 2, 1.0, tab)
    for s in (5, 23):
        f = blob.synthetic_features(8, s)
        a = refpy.RefStream(R, model_blob, blob.stream_seed(s)).run(f, trace=True, logits=True)
        b = orcpy.OrcStream(om, blob.stream_seed(s)).run(f, trace=True, logits=True)
        assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1]) and np.array_equal(bits(a[2]), bits(b[2]))


def test_oracle_against_reference_over_10s(orc_model, model_blob):
    """bench.py's workload over its full length (1000 frames): the oracle's PCM hashes to what the unmodified
    reference produced (tests/golden/bench_pcm_hashes.json, written by make_golden.py bench)."""
    import json
    import os
    g = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "bench_pcm_hashes.json")))
    assert g["blob_sha256"] == hashlib.sha256(model_blob).hexdigest()
    for s in (7, 1016):
        pcm = orcpy.OrcStream(orc_model, blob.stream_seed(s)).run(blob.synthetic_features(g["frames"], s % 64))[0]
        n = g["block_samples"]
        got = [hashlib.sha256(pcm[k * n:(k + 1) * n].astype("<i2").tobytes()).hexdigest() for k in range(len(pcm) // n)]
        assert got == g["streams"][str(s)], s


@pytest.mark.skipif(not refpy.available(2), reason="oracle/_ref not built (needs /root/reference)")
def test_out_of_range_pitch_feature_against_reference(orclib, model_blob):
    """features[18] huge / inf / NaN: (int)floor(...) is the host's cvttsd2si in the reference and in the oracle
    (src/lpcnet.c:93-94): the same conditioning, bit for bit."""
    import ctypes as C
    R = refpy.RefLib(2)
    om = orcpy.OrcModel(orclib, model_blob, 2, 1.0, None)
    for bad in (1e9, np.inf, -np.inf, np.nan, 4.2e7):
        f = blob.synthetic_features(4, 60)
        f[:, 18] = bad
        rs, st = refpy.RefStream(R, model_blob), orcpy.OrcStream(om)
        for k in range(4):
            ga = np.zeros(1152, np.float32); gb = np.zeros(48, np.float32); lp = np.zeros(16, np.float32)
            R.lib.ref_frame_network(C.c_void_p(rs.h), refpy._p(np.ascontiguousarray(f[k])), refpy._p(ga), refpy._p(gb), refpy._p(lp))
            oga, ogb, olpc = st.frame_network(f[k])
            assert np.array_equal(bits(ga), bits(oga)) and np.array_equal(bits(gb), bits(ogb)), (bad, k)
