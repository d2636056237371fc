"""Generates the golden fixtures in tests/golden/ by running the UNMODIFIED reference
(oracle/_ref/libref_d2.so = /root/reference/src/*.c compiled in place by oracle/Makefile with
-O2 -mavx2 -mfma -ffp-contract=off, plus oracle/ref_harness.c).  Run in the build container
(needs /root/reference):   python tests/golden/make_golden.py
The fixtures travel to the GPU box; the reference does not."""
import hashlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from lpcnet_b200 import blob  # noqa: E402
from oracle import refpy  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))


def prims(R):
    rng = np.random.default_rng(7)
    x = np.concatenate([rng.standard_normal(3000) * 3, rng.uniform(-12, 12, 1000),
                        [0, -0.0, 1e-8, -1e-8, 40, -40, 1e4, -1e4]]).astype(np.float32)
    d = {"act_x": x, "tanh": R.vec("ref_vec_tanh", x), "sigmoid": R.vec("ref_vec_sigmoid", x)}
    lx = np.concatenate([rng.standard_normal(2000) * 3000, rng.uniform(-40000, 40000, 1000),
                         [0, -0.0, 1, -1, 32767, -32768, 1e6, -1e6]]).astype(np.float32)
    d["lin_x"] = lx
    d["lin2ulaw"] = np.array([R.lib.ref_lin2ulaw(refpy.C.c_float(float(v))) for v in lx], np.int32)
    d["ulaw2lin"] = np.array([R.lib.ref_ulaw2lin(refpy.C.c_float(float(u))) for u in range(256)], np.float32)
    d["rcp_table"] = R.rcp_table().view(np.uint32)
    # rcp at random mantissas/exponents: shows the estimate depends only on the top 11 bits
    rx = (rng.uniform(1, 2, 500) * 2.0 ** rng.integers(-3, 40, 500)).astype(np.float32)
    d["rcp_x"] = rx
    d["rcp_y"] = np.array([R.lib.ref_rcp_ss(refpy.C.c_float(float(v))) for v in rx], np.float32)
    for seed in (b"LPCNet", b"LPCNet17", b"x", b"", b"abcdefgh"):
        st = R.kiss99_srand(seed)
        seq = [st.copy()]
        draws = []
        for _ in range(16):
            draws.append(R.kiss99_rand(st))
            seq.append(st.copy())
        d["kiss_state_" + seed.decode()] = np.stack(seq)
        d["kiss_draws_" + seed.decode()] = np.array(draws, np.uint32)
    q = rng.uniform(-1.3, 1.3, 512).astype(np.float32)
    d["quant_x"] = q
    d["quant_u8"] = R.ps_to_epi8(q)
    # int8 kernels on a small block-sparse matrix
    n_in, n_out = 64, 48
    qw = rng.integers(-60, 61, size=(n_in, n_out)).astype(np.int32)
    mask = rng.random((n_in // 4, n_out // 8)) < 0.4
    qw = qw * np.repeat(np.repeat(mask, 4, 0), 8, 1)
    wq, idx = blob.block_sparse_encode(qw, mask)
    xin = rng.uniform(-1, 1, n_in).astype(np.float32)
    out0 = (rng.standard_normal(n_out) * 0.5).astype(np.float32)
    d["sp_w"], d["sp_idx"], d["sp_x"], d["sp_out0"] = wq, idx, xin, out0
    d["sp_out"] = R.sparse_sgemv(out0, wq, idx, xin)
    qd = rng.integers(-60, 61, size=(16, 48)).astype(np.int32)
    wd = blob.dotp_layout(qd)
    xd = rng.uniform(-1, 1, 16).astype(np.float32)
    d["de_w"], d["de_x"] = wd, xd
    d["de_out"] = R.sgemv8x4(out0, wd, xd)
    wf = (rng.standard_normal((40, 32)) * 0.2).astype(np.float32)  # [cols][rows]
    xf = rng.standard_normal(40).astype(np.float32)
    of = rng.standard_normal(32).astype(np.float32)
    d["f_w"], d["f_x"], d["f_out0"] = wf, xf, of
    d["f_out"] = R.sgemv16(of, wf, 40, xf)
    ceps = (rng.standard_normal((32, 18)) * 0.5).astype(np.float32)
    ceps[:, 0] += 2
    d["ceps"] = ceps
    d["lpc"] = np.stack([R.lpc_from_cepstrum(c) for c in ceps])
    np.savez_compressed(os.path.join(OUT, "prims.npz"), **d)


def synth(R):
    A = blob.random_model()
    b = blob.write_blob(A)
    d = {"blob_sha256": np.frombuffer(hashlib.sha256(b).digest(), np.uint8)}
    F = 12
    streams = [0, 3, 11]
    d["streams"] = np.array(streams)
    for s in streams:
        f = blob.synthetic_features(F, s)
        d[f"feat_{s}"] = f
        rs = refpy.RefStream(R, b, blob.stream_seed(s))
        pcm, tr, lg = rs.run(f, trace=True, logits=True)
        d[f"pcm_{s}"], d[f"trace_{s}"] = pcm, tr
        d[f"logits_{s}"] = lg[3 * 160:3 * 160 + 96].copy()   # steps 480..575
        d[f"state_{s}"] = np.array([rs.snapshot()])
        # teacher forced on a perturbed copy of its own output (so forcing actually changes the path)
        forced = (pcm.astype(np.int32) // 2).astype(np.int16)
        rs2 = refpy.RefStream(R, b, blob.stream_seed(s))
        pcm2, tr2, lg2 = rs2.run(f, forced=forced, trace=True, logits=True)
        d[f"forced_{s}"], d[f"fpcm_{s}"], d[f"ftrace_{s}"] = forced, pcm2, tr2
        d[f"flogits_{s}"] = lg2[3 * 160:3 * 160 + 96].copy()
    # the public API on stream 0, with a short last call (N=80)
    rs = refpy.RefStream(R, b)
    f = blob.synthetic_features(6, 0)
    outs = [rs.synthesize(f[k], 160 if k < 5 else 80) for k in range(6)]
    d["api_pcm"] = np.concatenate(outs)
    # frame network alone
    rs = refpy.RefStream(R, b)
    ga = np.zeros((4, 1152), np.float32); gb = np.zeros((4, 48), np.float32); lp = np.zeros((4, 16), np.float32)
    for k in range(4):
        R.lib.ref_frame_network(refpy.C.c_void_p(rs.h), refpy._p(np.ascontiguousarray(f[k])), refpy._p(ga[k]), refpy._p(gb[k]), refpy._p(lp[k]))
    d["fn_ga"], d["fn_gb"], d["fn_lpc"] = ga, gb, lp
    d["logit_table"] = rs.logit_table()
    np.savez_compressed(os.path.join(OUT, "synth.npz"), **d)


def decode(R):
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    rng = np.random.default_rng(5)
    pk = rng.integers(0, 256, size=(24, 8)).astype(np.uint8)
    pk[5, 1] &= 0xF8   # modulation field (bits 13..15) = 0 -> unvoiced branch
    vq = np.zeros(18, np.float32)
    feats = []
    for p in pk:
        feats.append(R.decode_packet(vq, p.tobytes()))
    np.savez_compressed(os.path.join(OUT, "decode.npz"), packets=pk, features=np.stack(feats))


def analysis_signals():
    """The inputs of analysis.npz (tests/test_analysis_oracle.py rebuilds and compares them)."""
    sig = [blob.synthetic_pcm(40, s) for s in range(4)]
    sig.append(np.zeros(40 * 160, np.int16))                                       # silence
    sig.append(np.random.default_rng(9).integers(-32768, 32767, 40 * 160).astype(np.int16))   # full-scale noise
    x = blob.synthetic_pcm(40, 7).copy()
    x[1600:3200] = 0                                                               # a gap: the energy floor path
    sig.append(x)
    return sig


def analysis(R):
    """lpcnet_compute_single_frame_features of the reference on 7 signals x 40 frames."""
    sig = analysis_signals()
    feats = [refpy.RefEncoder(R).run(x) for x in sig]
    np.savez_compressed(os.path.join(OUT, "analysis.npz"), pcm=np.stack(sig), features=np.stack(feats))


def plc_causal_fec(F, s):
    """FEC features handed to lpcnet_plc_fec_add before frame f (None = skip one); none for stream 0."""
    if s == 0:
        return {}
    rng = np.random.default_rng(50 + s)
    return {f: [None if rng.random() < 0.2 else (rng.standard_normal(20) * 0.5).astype(np.float32)
                for _ in range(int(rng.integers(0, 3)))] for f in range(F)}


def plc_causal(R):
    """lpcnet_plc_update / lpcnet_plc_conceal of the reference's FEATURES_DELAY 2 build (libref_d2.so) in the causal
    and codec modes, without and with the DC filter: 3 streams x 40 frames under Gilbert loss, streams 1 and 2 with FEC
    features in the queue."""
    b = plc_model_blob()
    d = {"blob_sha256": np.frombuffer(hashlib.sha256(b).digest(), np.uint8)}
    F = 40
    for opt in (0, 4, 2, 6):
        for s in range(3):
            x = blob.synthetic_pcm(F, s)
            if opt & 4:
                x = np.clip(x.astype(np.int32) + 700, -32768, 32767).astype(np.int16)
            lost = blob.synthetic_loss_pattern(F, s, 0.6 + 0.25 * (s == 2), 0.15)
            d[f"in_{opt}_{s}"], d[f"lost_{opt}_{s}"] = x, lost
            d[f"out_{opt}_{s}"] = refpy.RefPLC(R, b, opt, blob.stream_seed(s)).run(x, lost, plc_causal_fec(F, s))
    np.savez_compressed(os.path.join(OUT, "plc_causal.npz"), **d)


def dred_inputs():
    """The inputs of dred.npz: 3 packets (initial state + latent vectors) of 1, 9 and 26 latents."""
    rng = np.random.default_rng(31)
    return [((rng.standard_normal(24) * 0.8).astype(np.float32), (rng.standard_normal((n, 80)) * (0.5 + 0.5 * k)).astype(np.float32))
            for k, n in enumerate((1, 9, 26))]


def dred_model_blob():
    A = blob.random_model()
    A.update(blob.random_dred_decoder_model())
    return blob.write_blob(A)


def dred(R):
    """The reference's RDOVAE decoder (dred_rdovae_dec_init_states + dred_rdovae_decode_qframe, src/dred_rdovae_dec.c)
    on the random-init decoder of blob.random_dred_decoder_model()."""
    b = dred_model_blob()
    d = {"blob_sha256": np.frombuffer(hashlib.sha256(b).digest(), np.uint8)}
    for k, (st, z) in enumerate(dred_inputs()):
        d[f"state_{k}"], d[f"latents_{k}"] = st, z
        d[f"features_{k}"] = refpy.ref_dred_decode_all(R, b, st, z)
    np.savez_compressed(os.path.join(OUT, "dred.npz"), **d)


def encode_signals():
    """The inputs of encode.npz: 7 signals x 24 packets (tests/test_encode_oracle.py rebuilds and compares them)."""
    n = 24 * 4
    sig = [blob.synthetic_pcm(n, 300 + s) for s in range(4)]
    x = blob.synthetic_pcm(n, 304).copy()
    x[640 * 6:640 * 14] = 0                                                        # digital silence: NaN pitch contour
    sig.append(x)
    sig.append(np.clip(blob.synthetic_pcm(n, 305).astype(np.int32) * 6, -32768, 32767).astype(np.int16))   # clipping
    sig.append(np.random.default_rng(19).integers(-3000, 3000, n * 160).astype(np.int16))                # unvoiced
    return sig


def encode(R):
    """lpcnet_encode and lpcnet_compute_features of the reference (lpcnet_enc.c:882-911) with the synthetic
    codebooks of oracle/gen_codebooks.py linked in."""
    sig = encode_signals()
    pk = [refpy.RefEncoder(R).encode(x) for x in sig]
    ft = [refpy.RefEncoder(R).compute_features(x) for x in sig]
    np.savez_compressed(os.path.join(OUT, "encode.npz"), pcm=np.stack(sig), packets=np.stack(pk), features=np.stack(ft))


def plc_model_blob():
    A = blob.random_model()
    A.update(blob.random_plc_model())
    return blob.write_blob(A)


def plc_edges(R):
    """Losses where the concealment has the least to stand on, through the reference (FEATURES_DELAY 0 build for the
    non-causal mode, 2 for the causal and codec modes): stream 0 loses every frame, 1 none, 2 the very first frame,
    3 everything but the first, 4 every other frame starting with the first; inputs are silence, full-scale square
    waves, a click and a ramp (the synthesiser is still in its FEATURES_DELAY warm-up when the early losses arrive,
    src/lpcnet.c:239-243)."""
    b = plc_model_blob()
    F = 14
    T = F * 160
    t = np.arange(T)
    x = np.stack([blob.synthetic_pcm(F, 9),
                  np.where((t // 40) % 2 == 0, 32767, -32768).astype(np.int16),
                  np.zeros(T, np.int16),
                  np.where(t % 2 == 0, 32767, -32768).astype(np.int16),
                  np.linspace(-32768, 32767, T).astype(np.int16)])
    x[2, T // 2] = 32767
    lost = np.zeros((5, F), np.uint8)
    lost[0, :] = 1
    lost[2, 0] = 1
    lost[3, 1:] = 1
    lost[4, ::2] = 1
    d = {"blob_sha256": np.frombuffer(hashlib.sha256(b).digest(), np.uint8), "in": x, "lost": lost}
    R0 = refpy.RefLib(0)
    for opt in (5, 0, 4, 2):
        lib = R0 if opt & 1 else R
        d[f"out_{opt}"] = np.stack([refpy.RefPLC(lib, b, opt, blob.stream_seed(s)).run(x[s], lost[s]) for s in range(5)])
    np.savez_compressed(os.path.join(OUT, "plc_edges.npz"), **d)


def plc(R):
    """lpcnet_plc_update / lpcnet_plc_conceal of the reference's FEATURES_DELAY 0 build (libref_d0.so), non-causal
    mode without and with the DC filter, 3 streams x 40 frames under Gilbert loss; plus burg_cepstral_analysis and
    the predictor network on its own."""
    R0 = refpy.RefLib(0)
    b = plc_model_blob()
    d = {"blob_sha256": np.frombuffer(hashlib.sha256(b).digest(), np.uint8)}
    F = 40
    for opt in (1, 5):
        for s in range(3):
            x = blob.synthetic_pcm(F, s)
            if opt == 5:
                x = np.clip(x.astype(np.int32) + 700, -32768, 32767).astype(np.int16)   # a DC offset to remove
            lost = blob.synthetic_loss_pattern(F, s)
            d[f"in_{opt}_{s}"], d[f"lost_{opt}_{s}"] = x, lost
            d[f"out_{opt}_{s}"] = refpy.RefPLC(R0, b, opt, blob.stream_seed(s)).run(x, lost)
    x = blob.synthetic_pcm(12, 5).astype(np.float32)
    d["burg_in"] = x
    d["burg_out"] = np.stack([refpy.ref_burg_cepstral_analysis(R0, x[f * 160:(f + 1) * 160]) for f in range(12)])
    rp = refpy.RefPLC(R0, b, 1)
    rng = np.random.default_rng(11)
    g1 = np.zeros(256, np.float32); g2 = np.zeros(256, np.float32)
    pin = (rng.standard_normal((16, 57)) * 0.7).astype(np.float32)
    pin[::4] = 0
    d["pred_in"] = pin
    d["pred_out"] = np.stack([rp.pred(g1, g2, v) for v in pin])
    d["pred_gru1"], d["pred_gru2"] = g1.copy(), g2.copy()
    np.savez_compressed(os.path.join(OUT, "plc.npz"), **d)


BENCH_STEPS = 26
BENCH_STREAMS = [0, 7, 8, 63, 64, 511, 1016, 1023]   # first / last tile of bench.py's 1024 streams, first / last slot


def bench_hashes(R):
    """bench.py's workload (BASELINE configs[1]: stream s = feature track s % 64, seed "LPCNet<s>", 1000 frames)
    through the UNMODIFIED reference for a few fixed streams: SHA-256 of every second of PCM (little-endian int16),
    so that bench.py and tests/test_gpu_long.py can check the timed kernel's output without the reference at hand
    and name the first second that differs ("streams": the first call on fresh states); "steps": SHA-256 of the whole
    PCM of call k = 0..BENCH_STEPS-1 on the same states, i.e. of every warm-up and timed step of a bench run."""
    import json
    from concurrent.futures import ThreadPoolExecutor
    b = blob.write_blob(blob.random_model())
    F = 1000

    def one(s):
        rs = refpy.RefStream(R, b, blob.stream_seed(s))
        f = blob.synthetic_features(F, s % 64)
        first, steps = None, []
        for k in range(BENCH_STEPS):          # bench.py calls the engine again and again on the same states
            pcm = rs.run(f)[0]
            if k == 0:
                first = [hashlib.sha256(pcm[j * 16000:(j + 1) * 16000].astype("<i2").tobytes()).hexdigest() for j in range(F // 100)]
            steps.append(hashlib.sha256(pcm.astype("<i2").tobytes()).hexdigest())
        return first, steps

    with ThreadPoolExecutor(8) as ex:
        hs = list(ex.map(one, BENCH_STREAMS))
    d = {"blob_sha256": hashlib.sha256(b).hexdigest(), "frames": F, "block_samples": 16000,
         "source": "oracle/_ref/libref_d2.so (the unmodified reference, -O2 -mavx2 -mfma -ffp-contract=off)",
         "streams": {str(s): h[0] for s, h in zip(BENCH_STREAMS, hs)},
         "steps": {str(s): h[1] for s, h in zip(BENCH_STREAMS, hs)}}
    json.dump(d, open(os.path.join(OUT, "bench_pcm_hashes.json"), "w"), indent=1)


def sat(R):
    """Models whose int8 weights use the whole range (blob.saturating_models): vpmaddubsw's s16 pair sums saturate.
    Synthesis (3 streams x 10 frames, traces), non-causal PLC with the DC filter (FEATURES_DELAY 0 build, 30 frames) and
    the RDOVAE decoder through the unmodified reference."""
    b = blob.saturating_models()
    d = {"blob_sha256": np.frombuffer(hashlib.sha256(b).digest(), np.uint8)}
    for s in (0, 5, 9):
        f = blob.synthetic_features(10, 400 + s)
        pcm, tr, _ = refpy.RefStream(R, b, blob.stream_seed(s)).run(f, trace=True)
        d[f"pcm_{s}"], d[f"trace_{s}"] = pcm, tr
    R0 = refpy.RefLib(0)
    for s in range(2):
        x = blob.synthetic_pcm(30, 40 + s)
        lost = blob.synthetic_loss_pattern(30, 40 + s)
        d[f"plc_in_{s}"], d[f"plc_lost_{s}"] = x, lost
        d[f"plc_out_{s}"] = refpy.RefPLC(R0, b, 5, blob.stream_seed(s)).run(x, lost)
    st, z = dred_inputs()[1]
    d["dred_state"], d["dred_latents"] = st, z
    d["dred_features"] = refpy.ref_dred_decode_all(R, b, st, z)
    np.savez_compressed(os.path.join(OUT, "sat.npz"), **d)


def demo_blob():
    """weights_blob.bin for the demo tests: the model compiled into oracle/_ref/lpcnet_demo_parity_d* (random LPCNet +
    PLC predictor) plus the VQ codebooks the reference links in (ceps_codebooks.c), as arrays under their C names."""
    A = blob.random_model()
    A.update(blob.random_plc_model())
    cb = blob.synthetic_codebooks()
    A["ceps_codebook1"], A["ceps_codebook2"], A["ceps_codebook3"] = cb[:1024 * 17], cb[1024 * 17:2 * 1024 * 17], cb[2 * 1024 * 17:3 * 1024 * 17]
    A["ceps_codebook_diff4"] = cb[3 * 1024 * 17:]
    return blob.write_blob(A)


def demo_inputs():
    """{mode: (argv after the binary, {file name: bytes})} for every mode of src/lpcnet_demo.c"""
    feats = blob.synthetic_features(20, 3)
    pk = np.random.default_rng(21).integers(0, 256, size=(6, 8), dtype=np.uint8)
    pcm = blob.synthetic_pcm(40, 11)
    lost = blob.synthetic_loss_pattern(40, 11)
    loss_txt = ("\n".join(str(int(x)) for x in lost[::2]) + "\n").encode()
    nolpc = blob.synthetic_features(12, 8)
    return {
        "synthesis": (["-synthesis", "in", "out"], {"in": feats.tobytes()}),
        "decode": (["-decode", "in", "out"], {"in": pk.tobytes()}),
        "features": (["-features", "in", "out"], {"in": pcm.tobytes()}),
        "encode": (["-encode", "in", "out"], {"in": pcm.tobytes()}),
        "addlpc": (["-addlpc", "in", "out"], {"in": nolpc.tobytes()}),
        "plc_noncausal_dc": (["-plc_file", "noncausal_dc", "loss", "in", "out"], {"in": pcm.tobytes(), "loss": loss_txt}),
        "plc_causal_dc": (["-plc_file", "causal_dc", "loss", "in", "out"], {"in": pcm.tobytes(), "loss": loss_txt}),
    }


def run_demo(binary, argv, files, extra=None):
    import subprocess
    import tempfile
    with tempfile.TemporaryDirectory() as d:
        for name, data in {**files, **(extra or {})}.items():
            with open(os.path.join(d, name), "wb") as f:
                f.write(data)
        r = subprocess.run([binary] + argv, cwd=d, capture_output=True)
        if r.returncode != 0:
            raise RuntimeError(f"{binary} {argv}: rc {r.returncode}: {r.stderr.decode()[-400:]}")
        return open(os.path.join(d, "out"), "rb").read()


def demo(R):
    """Every mode of the reference's own command-line program (src/lpcnet_demo.c, unmodified, pinned-semantics flags,
    compiled-in model: oracle/_ref/lpcnet_demo_parity_d2; _d0, the FEATURES_DELAY 0 build, for the non-causal PLC)."""
    ref = os.path.join(ROOT, "oracle", "_ref")
    d = {"blob_sha256": np.frombuffer(hashlib.sha256(demo_blob()).digest(), np.uint8)}
    for mode, (argv, files) in demo_inputs().items():
        binary = os.path.join(ref, "lpcnet_demo_parity_d0" if mode == "plc_noncausal_dc" else "lpcnet_demo_parity_d2")
        d[mode] = np.frombuffer(run_demo(binary, argv, files), np.uint8)
    np.savez_compressed(os.path.join(OUT, "demo.npz"), **d)


def end2end(_R):
    """The reference built -DEND2END (oracle/_ref/libref_d2_e2e.so; src/lpcnet.c:57-78, 108): the frame network's LPC of
    six frames, and the PCM and per-step trace of three streams over 10 frames."""
    R = refpy.RefLib("2_e2e")
    b = blob.write_blob(blob.random_model())
    d = {"blob_sha256": np.frombuffer(hashlib.sha256(b).digest(), np.uint8)}
    streams = [0, 5, 9]
    d["streams"] = np.array(streams)
    for s in streams:
        f = blob.synthetic_features(10, s)
        rs = refpy.RefStream(R, b, blob.stream_seed(s))
        pcm, tr, _ = rs.run(f, trace=True, logits=False)
        d[f"pcm_{s}"], d[f"trace_{s}"] = pcm, tr
    rs = refpy.RefStream(R, b)
    f = blob.synthetic_features(6, 0)
    ga = np.zeros((6, 1152), np.float32); gb = np.zeros((6, 48), np.float32); lp = np.zeros((6, 16), np.float32)
    for k in range(6):
        R.lib.ref_frame_network(refpy.C.c_void_p(rs.h), refpy._p(np.ascontiguousarray(f[k])), refpy._p(ga[k]), refpy._p(gb[k]), refpy._p(lp[k]))
    d["fn_lpc"], d["fn_gb"] = lp, gb
    np.savez_compressed(os.path.join(OUT, "end2end.npz"), **d)


if __name__ == "__main__":
    R = refpy.RefLib(2)
    which = sys.argv[1:] or ["prims", "synth", "decode", "analysis", "plc", "encode", "plc_causal", "dred", "plc_edges", "bench", "sat", "demo", "end2end"]
    for name in which:
        {"prims": prims, "synth": synth, "decode": decode, "analysis": analysis, "plc": plc, "encode": encode, "plc_causal": plc_causal, "dred": dred, "plc_edges": plc_edges, "bench": bench_hashes, "sat": sat, "demo": demo, "end2end": end2end}[name](R)
    for f in sorted(os.listdir(OUT)):
        print(f, os.path.getsize(os.path.join(OUT, f)))
