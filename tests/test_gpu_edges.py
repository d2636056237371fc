"""Edge cases of the C-ABI on the GPU: empty calls, bad arguments, extreme inputs (silence, full-scale square waves,
all-ones / all-zero packets, a stream that loses every frame beside one that loses none, a loss on the very first
frame).  Where there is output it is compared bit for bit with the CPU oracle."""
import hashlib
import os

import numpy as np
import pytest

from lpcnet_b200 import blob, engine
from oracle import orcpy

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


@pytest.fixture(scope="module")
def eng(engine_lib):
    assert engine.device_count() > 0, "no CUDA device: the gpu tests cannot run (there is no CPU fallback)"
    return engine


@pytest.fixture(scope="module")
def plc_blob():
    A = blob.random_model()
    A.update(blob.random_plc_model())
    return blob.write_blob(A)


def extreme_pcm(n_frames):
    """silence | full-scale square wave | one click | alternating +-32768/32767 | a ramp through the int16 range"""
    T = n_frames * 160
    t = np.arange(T)
    rows = [np.zeros(T, np.int16),
            np.where((t // 40) % 2 == 0, 32767, -32768).astype(np.int16),
            np.zeros(T, np.int16),
            np.where(t % 2 == 0, 32767, -32768).astype(np.int16),
            np.linspace(-32768, 32767, T).astype(np.int16)]
    rows[2][T // 2] = 32767
    return np.stack(rows)


def test_empty_calls_change_nothing(eng, model_blob, orc_model, codebooks, orclib, plc_blob):
    B, F = 3, 3
    feats = np.stack([blob.synthetic_features(F, s) for s in range(B)])
    seeds = [blob.stream_seed(s) for s in range(B)]
    bt = eng.Batch(model_blob, B)
    bt.set_codebooks(codebooks)
    for s in range(B):
        bt.set_seed(s, seeds[s])
    n0 = bt.launch_count
    pcm, _, _ = bt.synthesize(feats[:, :0])
    assert pcm.shape == (B, 0)
    pcm, ft = bt.decode(np.zeros((B, 0, 8), np.uint8), want_features=True)
    assert pcm.shape == (B, 0) and ft.shape == (B, 0, 36)
    assert bt.launch_count == n0                     # nothing ran
    pcm, _, _ = bt.synthesize(feats)
    for s in range(B):
        assert np.array_equal(pcm[s], orcpy.OrcStream(orc_model, seeds[s]).run(feats[s])[0])
    bt.close()

    x = np.stack([blob.synthetic_pcm(8, 40 + s) for s in range(B)])
    fb = eng.FeatureBatch(B)
    fb.set_codebooks(codebooks)
    n0 = fb.launch_count
    assert fb.compute(x[:, :0]).shape == (B, 0, 36)
    pk = fb.encode(x[:, :0])
    assert pk.shape == (B, 0, 8)
    assert fb.launch_count == n0
    pk = fb.encode(x)
    for s in range(B):
        want, _ = orcpy.OrcEncoder(orclib).encode(x[s], codebooks)
        assert np.array_equal(pk[s], want)
    fb.close()

    pb = eng.PLCBatch(plc_blob, B, 5)
    n0 = pb.launch_count
    out = pb.process(x[:, :0], np.zeros((B, 0), np.uint8))
    assert out.shape == (B, 0) and pb.launch_count == n0
    pb.close()


def test_bad_arguments_are_errors_with_a_message(eng, model_blob):
    L = eng.load_library()
    bt = eng.Batch(model_blob, 2)
    f = np.zeros((2, 1, 36), np.float32)
    pcm = np.zeros((2, 160), np.int16)
    P = eng._p
    assert L.lpcnet_batch_synthesize(bt.h, P(f), 36, 1, None, None, None, None) != 0          # no output buffer
    assert L.lpcnet_b200_last_error()
    assert L.lpcnet_batch_synthesize(bt.h, None, 36, 1, P(pcm), None, None, None) != 0         # no features
    assert L.lpcnet_batch_synthesize(bt.h, P(f), 19, 1, P(pcm), None, None, None) != 0         # stride < NB_FEATURES
    assert L.lpcnet_batch_synthesize(None, P(f), 36, 1, P(pcm), None, None, None) != 0         # no batch
    with pytest.raises(eng.LPCNetError):
        bt.set_seed(2, b"x")                                                                    # stream out of range
    with pytest.raises(eng.LPCNetError):
        bt.set_tile(12)
    with pytest.raises(eng.LPCNetError):
        bt.get_state(-1)
    with pytest.raises(eng.LPCNetError):
        bt.decode(np.zeros((2, 1, 8), np.uint8))                                                # no codebooks yet
    with pytest.raises(eng.LPCNetError):
        bt.set_codebooks(np.zeros(100, np.float32))                                             # wrong size
    # the batch is still usable after the refused calls
    got, _, _ = bt.synthesize(np.zeros((2, 2, 36), np.float32))
    assert got.shape == (2, 320)
    bt.close()
    with pytest.raises(eng.LPCNetError):
        eng.Batch(model_blob, 0)
    with pytest.raises(eng.LPCNetError):
        eng.Batch(model_blob[:1000], 1)                                                         # truncated blob
    with pytest.raises(eng.LPCNetError):
        eng.FeatureBatch(0)


def test_extreme_packets_decode_like_the_oracle(eng, model_blob, codebooks, orc_model, orclib):
    """All-zero, all-ones and alternating packets: every index field at its smallest / largest value (pitch index
    63, interpolation id 7, the forbidden-interpolation remap)."""
    pats = [bytes(8), b"\xff" * 8, b"\xaa" * 8, b"\x55" * 8, bytes(range(0, 256, 32))]
    packets = np.stack([np.frombuffer(p * 1, np.uint8).reshape(1, 8).repeat(3, 0) for p in pats])
    packets[4, 1] = 255 - packets[4, 1]
    bt = eng.Batch(model_blob, len(pats))
    bt.set_codebooks(codebooks)
    pcm, feats = bt.decode(packets, want_features=True)
    bt.close()
    for s in range(len(pats)):
        vq_mem = np.zeros(18, np.float32)
        want = np.stack([orclib.decode_packet(vq_mem, packets[s, k].tobytes(), codebooks) for k in range(3)])
        assert np.array_equal(bits(feats[s]), bits(want.reshape(12, 36))), s
        assert np.array_equal(pcm[s], orcpy.OrcStream(orc_model).run(feats[s])[0]), s


def test_extreme_signals_through_analysis_and_encoder(eng, codebooks, orclib):
    x = extreme_pcm(12)
    B = x.shape[0]
    fb = eng.FeatureBatch(B)
    fb.set_codebooks(codebooks)
    got = fb.compute(x)
    fb.reset()
    pk, qf = fb.encode(x, want_features=True)
    fb.close()
    for s in range(B):
        want = orcpy.OrcEncoder(orclib).run(x[s])
        assert np.array_equal(bits(got[s]), bits(want)), ("features", s)
        wpk, wqf = orcpy.OrcEncoder(orclib).encode(x[s], codebooks)
        assert np.array_equal(pk[s], wpk), ("packets", s)
        # the LPC entries [20, 36) of the encoder's quantised rows never leave lpcnet_encode
        assert np.array_equal(bits(qf[s][:, :20]), bits(wqf.reshape(-1, 36)[:, :20])), ("quantised features", s)


@pytest.mark.parametrize("opt", [5, 0, 4, 2])
def test_plc_all_lost_none_lost_and_first_frame_lost(eng, plc_blob, opt):
    """tests/golden/plc_edges.npz (the unmodified reference): stream 0 loses every frame, 1 none, 2 the very first
    frame, 3 everything but the first, 4 alternates from the first; silence, full-scale square waves, a click, a ramp.
    In the causal and codec modes the early losses meet a synthesiser that is still inside its FEATURES_DELAY warm-up
    (src/lpcnet.c:239-243: zeros out, state untouched)."""
    g = np.load(os.path.join(HERE, "golden", "plc_edges.npz"))
    assert hashlib.sha256(plc_blob).digest() == g["blob_sha256"].tobytes()
    x, lost, want = g["in"], g["lost"], g[f"out_{opt}"]
    B = x.shape[0]
    pb = eng.PLCBatch(plc_blob, B, opt)
    for s in range(B):
        pb.set_seed(s, blob.stream_seed(s))
    got = pb.process(x, lost)
    pb.close()
    for s in range(B):
        d = np.nonzero(got[s] != want[s])[0]
        assert d.size == 0, ("stream, first differing sample", s, int(d[0]))
    # frame by frame through lpcnet_plc_update / lpcnet_plc_conceal, the stream that loses its first frame
    s = 2
    plc = eng.PLC(plc_blob, opt)
    out = [plc.conceal() if lost[s, f] else plc.update(x[s, f * 160:(f + 1) * 160]) for f in range(lost.shape[1])]
    plc.close()
    # the single-stream state is seeded "LPCNet" like the reference's; the fixture used stream seeds, so compare with
    # the oracle on the default seed
    model = orcpy.OrcModel(orcpy.OrcLib(), plc_blob, 0 if opt & 1 else 2, 1.0, None)
    ref = orcpy.OrcPLC(model, opt, b"LPCNet").run(x[s], lost[s])
    assert np.array_equal(np.concatenate(out), ref)


def test_silence_and_full_scale_features_synthesize_like_the_oracle(eng, model_blob, orc_model):
    """Feature rows of zeros, of large magnitude (saturating the pitch index clamp) and with NaN-free extremes."""
    F = 4
    base = blob.synthetic_features(F, 0)
    rows = [np.zeros_like(base), base * 8.0, -base * 8.0, base.copy()]
    rows[3][:, 18] = 5.0          # pitch feature far above the range: the index clamps (src/lpcnet.c:109-111)
    feats = np.stack(rows)
    bt = eng.Batch(model_blob, len(rows))
    pcm, tr, _ = bt.synthesize(feats, trace=True)
    bt.close()
    for s in range(len(rows)):
        ref = orcpy.OrcStream(orc_model).run(feats[s], trace=True)
        assert np.array_equal(pcm[s], ref[0]), s
        assert np.array_equal(tr[s], ref[1]), s


def overflow_fec(F):
    """more vectors than PLC_MAX_FEC (100) in one go, steady arrivals afterwards, a few "skip one" markers"""
    rng = np.random.default_rng(77)
    fec = {f: [(rng.standard_normal(20) * 0.5).astype(np.float32) for _ in range(3)] for f in range(F)}
    fec[0] = [(rng.standard_normal(20) * 0.5).astype(np.float32) for _ in range(120)]
    fec[5] = [None] * 4 + fec[5]
    return fec


@pytest.mark.parametrize("opt", [0, 2])
def test_fec_queue_overflow_and_compaction(eng, plc_blob, opt):
    """lpcnet_plc_fec_add past PLC_MAX_FEC: the reference drops vectors while nothing has been consumed, then moves
    the kept tail to the front (lpcnet_plc.c:111-127)."""
    F = 40
    x = blob.synthetic_pcm(F, 3)
    lost = blob.synthetic_loss_pattern(F, 1, 0.6, 0.2)
    assert lost.sum() > 5
    fec = overflow_fec(F)
    model = orcpy.OrcModel(orcpy.OrcLib(), plc_blob, 2, 1.0, None)
    want = orcpy.OrcPLC(model, opt, b"LPCNet").run(x, lost, fec)
    plc = eng.PLC(plc_blob, opt)
    out = []
    for f in range(F):
        for v in fec.get(f, []):
            plc.fec_add(v)
        out.append(plc.conceal() if lost[f] else plc.update(x[f * 160:(f + 1) * 160]))
    plc.close()
    d = np.nonzero(np.concatenate(out) != want)[0]
    assert d.size == 0, int(d[0])
    # the same through the batch entry points, the stream beside one without FEC
    pb = eng.PLCBatch(plc_blob, 2, opt)
    pb.set_seed(0, b"LPCNet")
    pb.set_seed(1, b"LPCNet")
    got = []
    for f in range(F):
        for v in fec.get(f, []):
            pb.fec_add(0, v)
        got.append(pb.process(np.stack([x, x])[:, f * 160:(f + 1) * 160], np.stack([lost, lost])[:, f:f + 1]))
    pb.close()
    got = np.concatenate(got, axis=1)
    assert np.array_equal(got[0], want)
    assert np.array_equal(got[1], orcpy.OrcPLC(model, opt, b"LPCNet").run(x, lost))


@pytest.mark.parametrize("opt", [4, 6, 5])
def test_plc_calls_of_any_length_agree(eng, plc_blob, opt):
    """one call of 30 frames == calls of 1, 7, 2 and 20 frames (state carried across calls), all modes"""
    B, F = 6, 30
    x = np.stack([blob.synthetic_pcm(F, 60 + s) for s in range(B)])
    lost = np.stack([blob.synthetic_loss_pattern(F, 20 + s, 0.6, 0.2) for s in range(B)])
    lost[1, :3] = 1
    outs = []
    for cuts in ((0, 30), (0, 1, 8, 10, 30)):
        pb = eng.PLCBatch(plc_blob, B, opt)
        for s in range(B):
            pb.set_seed(s, blob.stream_seed(s))
        outs.append(np.concatenate([pb.process(x[:, a * 160:b * 160], lost[:, a:b]) for a, b in zip(cuts[:-1], cuts[1:])], axis=1))
        pb.close()
    assert np.array_equal(outs[0], outs[1])
    model = orcpy.OrcModel(orcpy.OrcLib(), plc_blob, 0 if opt & 1 else 2, 1.0, None)
    for s in (1, 4):
        assert np.array_equal(outs[0][s], orcpy.OrcPLC(model, opt, blob.stream_seed(s)).run(x[s], lost[s])), s


def test_reset_of_one_stream_and_two_batches_side_by_side(eng, model_blob, orc_model, codebooks):
    """lpcnet_batch_reset(stream) in the middle of a run restarts that stream only; a second batch (other size, other
    tile, with codebooks) used in between leaves the first untouched (no process-wide scratch)."""
    B, F = 5, 6
    feats = np.stack([blob.synthetic_features(F, s) for s in range(B)])
    seeds = [blob.stream_seed(s) for s in range(B)]
    b1 = eng.Batch(model_blob, B)
    b2 = eng.Batch(model_blob, 19)
    b2.set_tile(16)
    b2.set_codebooks(codebooks)
    for s in range(B):
        b1.set_seed(s, seeds[s])
    first, _, _ = b1.synthesize(feats[:, :3])
    other, _, _ = b2.synthesize(np.stack([blob.synthetic_features(4, 30 + s) for s in range(19)]))
    b2.decode(np.zeros((19, 1, 8), np.uint8))
    b1.reset(2)
    b1.set_seed(2, seeds[2])
    second, _, _ = b1.synthesize(feats[:, 3:])
    for s in range(B):
        st = orcpy.OrcStream(orc_model, seeds[s])
        want = st.run(feats[s])[0]
        if s == 2:
            want = np.concatenate([want[:480], orcpy.OrcStream(orc_model, seeds[s]).run(feats[s, 3:])[0]])
        assert np.array_equal(np.concatenate([first[s], second[s]]), want), s
    assert np.array_equal(other[7], orcpy.OrcStream(orc_model).run(blob.synthetic_features(4, 37))[0])
    b1.close()
    b2.close()


def test_dred_decoder_edges(eng):
    """no latents, one latent, the reference's maximum of 26 per packet (DRED_NUM_REDUNDANCY_FRAMES / 2), saturating
    and zero inputs"""
    A = blob.random_model()
    A.update(blob.random_dred_decoder_model())
    b = blob.write_blob(A)
    m = orcpy.OrcModel(orcpy.OrcLib(), b, 2, 1.0, None)
    dec = eng.DredDecoder(b)
    n0 = dec.launch_count
    assert dec.decode_all(np.zeros(24, np.float32), np.zeros((0, 80), np.float32)).shape == (0, 80)
    assert dec.launch_count == n0
    rng = np.random.default_rng(2)
    cases = [(np.zeros(24, np.float32), np.zeros((1, 80), np.float32)),
             (np.full(24, 50.0, np.float32), np.full((26, 80), -50.0, np.float32)),
             ((rng.standard_normal(24) * 1e4).astype(np.float32), (rng.standard_normal((26, 80)) * 1e4).astype(np.float32)),
             ((rng.standard_normal(24) * 1e-30).astype(np.float32), (rng.standard_normal((3, 80)) * 1e-30).astype(np.float32))]
    for k, (st, z) in enumerate(cases):
        got = dec.decode_all(st, z)
        want = orcpy.dred_decode_all(m, st, z)
        assert np.array_equal(bits(got), bits(want)), k
    dec.close()
    with pytest.raises(eng.LPCNetError):
        eng.DredDecoder(blob.write_blob(blob.random_model()))        # no dred_rdovae_dec_* arrays


def test_states_on_different_threads_do_not_interfere(eng, model_blob, plc_blob, codebooks, orc_model):
    """SURVEY 8(b) threading: the reference is re-entrant per state.  Four host threads, each with its own handles
    (batches of different sizes and tiles, a single-stream LPCNetState, an encoder batch, a PLC batch), call the
    library at the same time (ctypes drops the GIL in the calls); every thread's output equals the one it produces
    alone."""
    import threading

    F = 5
    def work_synth(n, tile, reps, out):
        feats = np.stack([blob.synthetic_features(F, (3 * n + s) % 16) for s in range(n)])
        bt = eng.Batch(model_blob, n)
        if tile:
            bt.set_tile(tile)
        for _ in range(reps):
            bt.reset()
            for s in range(n):
                bt.set_seed(s, blob.stream_seed(s))
            out.append(bt.synthesize(feats)[0].copy())
        bt.close()

    def work_single(reps, out):
        f = blob.synthetic_features(F, 2)
        syn = eng.Synth(model_blob)
        for _ in range(reps):
            syn.reset()
            out.append(np.concatenate([syn.synthesize(f[k]) for k in range(F)]))
        syn.close()

    def work_encode(reps, out):
        x = np.stack([blob.synthetic_pcm(8, 70 + s) for s in range(40)])
        fb = eng.FeatureBatch(40)
        fb.set_codebooks(codebooks)
        for _ in range(reps):
            fb.reset()
            out.append(fb.encode(x).copy())
        fb.close()

    def work_plc(reps, out):
        x = np.stack([blob.synthetic_pcm(10, 80 + s) for s in range(12)])
        lost = np.stack([blob.synthetic_loss_pattern(10, 30 + s, 0.6, 0.25) for s in range(12)])
        for _ in range(reps):
            pb = eng.PLCBatch(plc_blob, 12, 4)
            for s in range(12):
                pb.set_seed(s, blob.stream_seed(s))
            out.append(pb.process(x, lost).copy())
            pb.close()

    jobs = [(work_synth, (19, 16, 4)), (work_synth, (150, 0, 3)), (work_single, (3,)), (work_encode, (6,)), (work_plc, (3,))]
    alone = []
    for fn, args in jobs:                       # each job on its own first
        out = []
        fn(*args[:-1], 1, out) if len(args) > 1 else fn(1, out)
        alone.append(out[0])
    results = [[] for _ in jobs]
    errors = []

    def guarded(fn, args, out):
        try:
            fn(*args, out)
        except Exception as e:                  # surfaced below: an exception in a thread must fail the test
            errors.append(repr(e))

    threads = [threading.Thread(target=guarded, args=(fn, args, results[k])) for k, (fn, args) in enumerate(jobs)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not errors, errors
    for k, (fn, args) in enumerate(jobs):
        assert len(results[k]) == args[-1]
        for r in results[k]:
            assert np.array_equal(r, alone[k]), (k, fn.__name__)
    # and the single-stream thread is what the oracle gives
    assert np.array_equal(alone[2], orcpy.OrcStream(orc_model).run(blob.synthetic_features(F, 2))[0])


def test_pinned_destination_takes_chunks_as_they_finish(eng, model_blob):
    """lpcnet_batch_synthesize into pinned host memory copies every 32-frame chunk out on its own stream beside the
    next chunk's kernel; into pageable memory it copies once at the end.  Same samples either way."""
    import torch
    B, F = 21, 70                                  # three chunks, a ragged last tile
    feats = np.stack([blob.synthetic_features(F, s % 16) for s in range(B)])
    outs = []
    for pinned in (False, True):
        bt = eng.Batch(model_blob, B)
        for s in range(B):
            bt.set_seed(s, blob.stream_seed(s))
        buf = None
        if pinned:
            hold = torch.empty((B, F * 160), dtype=torch.int16, pin_memory=True)
            hold.fill_(-1)
            buf = hold.numpy()
        pcm, _, _ = bt.synthesize(feats, pcm_out=buf)
        outs.append(np.array(pcm))
        # a second call continues the streams: the copies of the first are complete and ordered
        pcm2, _, _ = bt.synthesize(feats[:, :3], pcm_out=buf[:, :480].copy() if pinned else None)
        outs.append(np.array(pcm2))
        bt.close()
    assert np.array_equal(outs[0], outs[2]) and np.array_equal(outs[1], outs[3])
    assert np.abs(outs[0]).max() > 0


def test_decode_into_pinned_memory(eng, model_blob, codebooks):
    import torch
    rng = np.random.default_rng(12)
    packets = rng.integers(0, 256, size=(9, 20, 8), dtype=np.uint8)      # 80 frames: three chunks
    outs = []
    for pinned in (False, True):
        bt = eng.Batch(model_blob, 9)
        bt.set_codebooks(codebooks)
        hold = torch.empty((9, 20 * 640), dtype=torch.int16, pin_memory=True) if pinned else None
        pcm, _ = bt.decode(packets, pcm_out=hold.numpy() if pinned else None)
        outs.append(np.array(pcm))
        bt.close()
    assert np.array_equal(outs[0], outs[1]) and np.abs(outs[0]).max() > 0


@pytest.mark.parametrize("opt", [5, 4])
def test_plc_in_place_ignores_the_content_of_lost_frames(eng, plc_blob, opt):
    """lpcnet_plc_batch_process works in place like lpcnet_plc_update / _conceal; what the buffer holds in a lost
    frame is never read (the reference's lpcnet_plc_conceal only writes)."""
    B, F = 4, 12
    x = np.stack([blob.synthetic_pcm(F, 90 + s) for s in range(B)])
    lost = np.stack([blob.synthetic_loss_pattern(F, 40 + s, 0.6, 0.25) for s in range(B)])
    assert lost.sum() > 0
    outs = []
    for junk in (False, True):
        pb = eng.PLCBatch(plc_blob, B, opt)
        for s in range(B):
            pb.set_seed(s, blob.stream_seed(s))
        if junk:
            buf = x.copy()
            buf.reshape(B, F, 160)[lost != 0] = 12345
            outs.append(pb.process_inplace(buf, lost).copy())
        else:
            outs.append(pb.process(x, lost))
        pb.close()
    assert np.array_equal(outs[0], outs[1])


def test_out_of_range_pitch_feature_follows_x86(model_blob, orc_model):
    """features[18] huge: the reference's (int)floor(...) is x86's cvttsd2si (INT_MIN for everything out of range), so
    the pitch index clamps to 33 -- not to 255 as CUDA's saturating conversion would give (src/lpcnet.c:93-94).
    Conditioning compared bit for bit with the oracle (which runs the host's own conversion).  Finite values only: an
    infinite or NaN feature also enters the convolution as a value, and the reference's conditioning is then NaN
    throughout (its SSE min/max let NaN through) -- noise either way, outside the contract."""
    F = 6
    feats = np.stack([blob.synthetic_features(F, 60 + s) for s in range(5)])
    feats[0, :, 18] = 1e9
    feats[1, :, 18] = 3e38
    feats[2, :, 18] = -1e9
    feats[3, 2:, 18] = -3e38
    feats[4, :, 18] = 4.2e7            # 50*x + 100 just below 2^31: in range, clamps to 255
    bt = engine.Batch(model_blob, 5)
    ga, gb, lpc = bt.frame_network(feats)
    for s in range(5):
        st = orcpy.OrcStream(orc_model)
        for f in range(F):
            oga, ogb, olpc = st.frame_network(feats[s, f])
            assert np.array_equal(ga[s, f].view(np.uint32), oga.view(np.uint32)), (s, f)
            assert np.array_equal(gb[s, f].view(np.uint32), ogb.view(np.uint32)), (s, f)
    bt.close()


def test_single_stream_decoder_takes_process_wide_codebooks(model_blob, golden_decode, codebooks, orc_model):
    """lpcnet_decoder_create + lpcnet_decoder_load_model of a blob WITHOUT ceps_codebook* arrays: the tables of
    lpcnet_b200_set_codebooks are installed (an opaque LPCNetDecState cannot reach lpcnet_batch_set_codebooks)."""
    engine.set_codebooks(codebooks)
    try:
        dec = engine.Decoder(model_blob)
        pk = golden_decode["packets"][:2]
        got = np.concatenate([dec.decode(p.tobytes()) for p in pk])
        feats = golden_decode["features"][:2].reshape(8, 36)
        assert np.array_equal(got, orcpy.OrcStream(orc_model).run(feats)[0])
        dec.close()
    finally:
        engine.set_codebooks(None)
    dec = engine.Decoder(model_blob)              # no tables anywhere: an error, not a crash
    with pytest.raises(engine.LPCNetError):
        dec.decode(golden_decode["packets"][0].tobytes())
    dec.close()
