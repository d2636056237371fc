"""BASELINE configs[1] at its stated size and length: 1024 streams x 10 s (1000 frames) through the PRODUCTION
instantiation of the sample kernel (no trace / logits => the variant bench.py times), compared sample for sample
with the CPU oracle and with hashes of the unmodified reference's output (tests/golden/bench_pcm_hashes.json),
free-running and teacher-forced, with a first-divergence report per stream (north_star's wording).

Loop under test: /root/reference/src/lpcnet.c:235-271 (lpcnet_synthesize_tail_impl); frame_count saturates at
1000 (lpcnet.c:119), which the 1005-frame run crosses."""
import hashlib
import json
import os
from concurrent.futures import ThreadPoolExecutor

import numpy as np
import pytest

from lpcnet_b200 import blob, engine
from oracle import orcpy

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))
PDF_TOL = 1e-4
F10S = 1000


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def tree_to_pdf(logits):
    """training_tf2/lpcnet.py:75-85: pdf of the 8-level binary tree from its 255 node logits."""
    p = 1.0 / (1.0 + np.exp(-logits.astype(np.float64)))
    pdf = np.ones((logits.shape[0], 1))
    for lvl in range(8):
        nodes = p[:, (1 << lvl):(2 << lvl)]
        pdf = np.stack([pdf * (1 - nodes), pdf * nodes], axis=-1).reshape(logits.shape[0], -1)
    return pdf


def oracle_pool(orc_model, jobs):
    """jobs: [(features, seed, forced | None)] -> PCM rows, streams spread over the host cores (ctypes drops the GIL)."""
    def one(j):
        f, seed, forced = j
        return orcpy.OrcStream(orc_model, seed).run(f, forced=forced)[0]
    with ThreadPoolExecutor(max(1, min(32, os.cpu_count() or 1))) as ex:
        return list(ex.map(one, jobs))


def first_divergence(a, b):
    bad = np.nonzero(np.asarray(a) != np.asarray(b))[0]
    return None if bad.size == 0 else int(bad[0])


def report(title, rows):
    """rows: [(stream, first divergent sample | None, n_samples)]"""
    print(f"\n{title}")
    for s, d, n in rows:
        print(f"  stream {s:5d}: " + (f"no divergence in {n} samples" if d is None else f"FIRST DIVERGENCE at sample {d} (frame {d // 160}, step {d % 160})"))
    out = os.path.join(os.path.dirname(HERE), "gpurun_out")
    if os.path.isdir(out):
        with open(os.path.join(out, "first_divergence.jsonl"), "a") as f:
            f.write(json.dumps({"run": title, "streams": [{"stream": s, "first_divergence": d, "samples": n} for s, d, n in rows]}) + "\n")


@pytest.fixture(scope="module")
def eng(engine_lib):
    assert engine.device_count() > 0, "no CUDA device: the gpu tests cannot run (there is no CPU fallback)"
    return engine


# the streams compared: every slot of the first and last 8-stream tile, and one slot of 16 tiles in between
CHECK = sorted(set(list(range(8)) + list(range(1016, 1024)) + [8 * (61 * k % 128) + (k % 8) for k in range(1, 17)]))


def bench_batch(eng, model_blob, B, n_frames):
    """bench.py's workload: stream s = feature track s % 64, seed "LPCNet<s>"."""
    tracks = np.stack([blob.synthetic_features(n_frames, k) for k in range(64)])
    feats = tracks[np.arange(B) % 64]
    bt = eng.Batch(model_blob, B)
    for s in range(B):
        bt.set_seed(s, blob.stream_seed(s))
    return bt, feats


def test_config2_production_kernel_10s(eng, model_blob, orc_model):
    """1024 streams x 1000 frames, production kernel (8-stream tiles, arithmetic rcpps), free-running and then
    teacher-forced on a perturbed copy of the free-running output."""
    B = 1024
    assert len(CHECK) == 32
    bt, feats = bench_batch(eng, model_blob, B, F10S)
    assert bt.tile == 8 and bt.rcp_arith, "not the instantiation bench.py times"
    pcm, tr, lg = bt.synthesize(feats)
    assert tr is None and lg is None
    # (1) against the unmodified reference: SHA-256 of every second of 8 fixed streams
    g = json.load(open(os.path.join(HERE, "golden", "bench_pcm_hashes.json")))
    assert g["blob_sha256"] == hashlib.sha256(model_blob).hexdigest()
    for s, hs in g["streams"].items():
        got = [hashlib.sha256(pcm[int(s), k * 16000:(k + 1) * 16000].astype("<i2").tobytes()).hexdigest() for k in range(10)]
        bad = [k for k in range(10) if got[k] != hs[k]]
        assert not bad, f"stream {s}: first second that differs from the reference: {bad[0]}"
    # (2) against the oracle, sample for sample, 32 streams over all tile slots
    ref = oracle_pool(orc_model, [(feats[s], blob.stream_seed(s), None) for s in CHECK])
    rows = [(s, first_divergence(pcm[s], ref[i]), pcm.shape[1]) for i, s in enumerate(CHECK)]
    report("config 2, free-running, 1024 streams x 10 s, production kernel (8 streams per CTA)", rows)
    assert all(d is None for _, d, _ in rows)
    assert (pcm[:, :320] == 0).all() and (pcm[:, 320:] != 0).any()
    # (3) teacher-forced on a perturbed copy (so that forcing changes the path)
    forced = (pcm.astype(np.int32) * 3 // 4).astype(np.int16)
    bt.reset()
    for s in range(B):
        bt.set_seed(s, blob.stream_seed(s))
    fpcm, _, _ = bt.synthesize(feats, forced=forced)
    ref = oracle_pool(orc_model, [(feats[s], blob.stream_seed(s), forced[s]) for s in CHECK])
    rows = [(s, first_divergence(fpcm[s], ref[i]), fpcm.shape[1]) for i, s in enumerate(CHECK)]
    report("config 2, teacher-forced, 1024 streams x 10 s, production kernel", rows)
    assert all(d is None for _, d, _ in rows)
    # the state after 10 s (GRU states, signal history, kiss99) equals the oracle's too
    st = orcpy.OrcStream(orc_model, blob.stream_seed(CHECK[-1]))
    st.run(feats[CHECK[-1]], forced=forced[CHECK[-1]])
    snap, osnap = bt.get_state(CHECK[-1]), st.snapshot()
    for name in snap.dtype.names:
        assert np.array_equal(np.asarray(snap[name]), np.asarray(osnap[name])), name
    bt.close()


@pytest.mark.parametrize("tile,n_frames", [(16, 1005), (32, 1000)])
def test_large_tiles_10s(eng, model_blob, orc_model, tile, n_frames):
    """64 streams x 10 s on 16- and 32-stream tiles (production kernel); the 16-stream run is 1005 frames long, past
    the frame counter's saturation at 1000 (lpcnet.c:119)."""
    B = 64
    bt, feats = bench_batch(eng, model_blob, B, n_frames)
    bt.set_tile(tile)
    pcm, _, _ = bt.synthesize(feats)
    check = list(range(0, 64, 4)) + [63]
    ref = oracle_pool(orc_model, [(feats[s], blob.stream_seed(s), None) for s in check])
    rows = [(s, first_divergence(pcm[s], ref[i]), pcm.shape[1]) for i, s in enumerate(check)]
    report(f"64 streams x {n_frames} frames, free-running, {tile} streams per CTA", rows)
    assert all(d is None for _, d, _ in rows)
    assert int(bt.get_state(0)["frame_count"]) == 1000
    bt.close()


def test_debug_kernel_10s_pdf(eng, model_blob, orc_model):
    """2 streams x 10 s through the trace / logits instantiation: every step's trace record bit-exact and the
    256-way pdf within 1e-4 of the oracle's (it is 0: the node logits are bit-identical)."""
    B = 2
    bt, feats = bench_batch(eng, model_blob, B, F10S)
    pcm, tr, lg = bt.synthesize(feats, trace=True, logits=True)
    for s in range(B):
        rp, rt, rl = orcpy.OrcStream(orc_model, blob.stream_seed(s)).run(feats[s], trace=True, logits=True)
        d = first_divergence(tr[s], rt)
        assert d is None, f"stream {s}: first divergence at step {d}: {tr[s][d]} vs {rt[d]}"
        assert np.array_equal(pcm[s], rp)
        worst = 0.0
        for k in range(320, pcm.shape[1], 16000):
            worst = max(worst, float(np.abs(tree_to_pdf(lg[s][k:k + 16000]) - tree_to_pdf(rl[k:k + 16000])).max()))
        assert worst <= PDF_TOL, worst
        assert np.array_equal(bits(lg[s]), bits(rl))
    bt.close()
