"""The "DNNw" weight-blob format and random-init LPCNet models written in it.

Blob layout (reference: src/write_lpcnet_weights.c:47-67, src/nnet.h:41-61,
src/parse_lpcnet_weights.c:36-77): a sequence of records, each a 64-byte head
``{char head[4]="DNNw"; int32 version; int32 type; int32 size; int32 block_size;
char name[44]}`` followed by ``size`` payload bytes zero-padded to
``block_size = ceil(size/64)*64``.  Arrays are looked up by name and exact byte
size (parse_lpcnet_weights.c:79-88).

The array list and per-array layouts follow training_tf2/dump_lpcnet.py
(printSparseVector 85-121, dump_sparse_gru 131-169, dump_grub 172-218,
dump_mdense_layer 257-282, main 440-470).  Models cannot be downloaded offline, so
`random_model` draws random-init weights in exactly that layout.
"""
from __future__ import annotations

import struct
from collections import OrderedDict

import numpy as np

WEIGHT_TYPE_FLOAT = 0
WEIGHT_TYPE_INT = 1
WEIGHT_TYPE_QWEIGHT = 2
WEIGHT_BLOCK_SIZE = 64

_TYPE_OF_DTYPE = {
    np.dtype(np.float32): WEIGHT_TYPE_FLOAT,
    np.dtype(np.int32): WEIGHT_TYPE_INT,
    np.dtype(np.int8): WEIGHT_TYPE_QWEIGHT,
}


def write_blob(arrays: "OrderedDict[str, np.ndarray]") -> bytes:
    """Serialise named arrays (float32 / int32 / int8) as a DNNw blob."""
    out = bytearray()
    for name, arr in arrays.items():
        arr = np.ascontiguousarray(arr)
        wtype = _TYPE_OF_DTYPE[arr.dtype]
        payload = arr.tobytes()
        size = len(payload)
        if size == 0:
            # parse_weights aborts on a zero-size record (parse_lpcnet_weights.c:62-73)
            raise ValueError(f"array {name} is empty; the reference parser rejects it")
        block = (size + WEIGHT_BLOCK_SIZE - 1) // WEIGHT_BLOCK_SIZE * WEIGHT_BLOCK_SIZE
        bname = name.encode()
        if len(bname) > 43:
            raise ValueError(f"array name too long: {name}")
        out += struct.pack("<4siiii44s", b"DNNw", 0, wtype, size, block, bname)
        out += payload
        out += b"\0" * (block - size)
    return bytes(out)


def parse_blob(data: bytes) -> "OrderedDict[str, np.ndarray]":
    """Inverse of `write_blob`, with parse_record's validation (36-51)."""
    arrays: "OrderedDict[str, np.ndarray]" = OrderedDict()
    off = 0
    n = len(data)
    while n - off > 0:
        if n - off < WEIGHT_BLOCK_SIZE:
            raise ValueError("truncated record head")
        _head, _ver, wtype, size, block, bname = struct.unpack_from("<4siiii44s", data, off)
        if block < size or block > n - off - WEIGHT_BLOCK_SIZE or bname[-1] != 0 or size <= 0:
            raise ValueError("bad record")
        name = bname.split(b"\0", 1)[0].decode()
        dtype = {0: np.float32, 1: np.int32, 2: np.int8}.get(wtype, np.uint8)
        arrays[name] = np.frombuffer(data, dtype=dtype, count=size // np.dtype(dtype).itemsize,
                                     offset=off + WEIGHT_BLOCK_SIZE).copy()
        off += WEIGHT_BLOCK_SIZE + block
    return arrays


def block_sparse_encode(q: np.ndarray, mask: np.ndarray):
    """Encode q[in][out] (int8 values) into the 8x4 block-sparse stream.

    `mask[in/4][out/8]` says which blocks are present.  Emission order follows
    printSparseVector (dump_lpcnet.py:98-118): for each 8-output group, a count, then
    per kept 4-input group its position 4*j and the 32 weights ordered
    [output r][input c] (the DOT_PROD transpose, line 110).
    """
    n_in, n_out = q.shape
    w = []
    idx = []
    for i in range(n_out // 8):
        js = np.nonzero(mask[:, i])[0]
        idx.append(len(js))
        for j in js:
            idx.append(4 * int(j))
            blk = q[4 * j:4 * j + 4, 8 * i:8 * i + 8]  # [c][r]
            w.append(blk.T.reshape(-1))               # [r][c]
    wq = np.concatenate(w).astype(np.int8) if w else np.zeros((0,), np.int8)
    return wq, np.asarray(idx, dtype=np.int32)


def dotp_layout(q: np.ndarray) -> np.ndarray:
    """Dense int8 'dotp' layout [out/8][in/4][8][4] (dump_lpcnet.py:56-60)."""
    n_in, n_out = q.shape
    v = q.reshape(n_in // 4, 4, n_out // 8, 8).transpose(2, 0, 3, 1)
    return np.ascontiguousarray(v).reshape(-1).astype(np.int8)


def _rand_mask(rng, n_in4, n_out8, density):
    total = n_in4 * n_out8
    k = int(round(density * total))
    m = np.zeros(total, dtype=bool)
    m[rng.choice(total, size=k, replace=False)] = True
    return m.reshape(n_in4, n_out8)


def random_model(seed: int = 1234, N: int = 384, NB: int = 16, C: int = 128,
                 density=(0.05, 0.05, 0.2), qmax=(40, 40, 20), qmax_b_in: int = 8,
                 qmax_b_rec: int = 40, gru_b_density: float = 1.0,
                 sigma: float = 0.1) -> "OrderedDict[str, np.ndarray]":
    """Random-init LPCNet model in the reference's array layout.

    int8 blocks are uniform in [-qmax, qmax] so that |q[2k]|+|q[2k+1]| <= 127 and the
    AVX2 `vpmaddubsw` s16 intermediate never saturates (training_tf2/lpcnet.py:287-309).
    """
    rng = np.random.default_rng(seed)
    f32 = np.float32
    A: "OrderedDict[str, np.ndarray]" = OrderedDict()

    def normal(shape, s=sigma):
        return (rng.standard_normal(shape) * s).astype(f32)

    # --- order mirrors dump_lpcnet.py main (440-470) ---
    for nm in ("sig", "pred", "exc"):
        A[f"gru_a_embed_{nm}_weights"] = normal((256, 3 * N))
    A["gru_a_dense_feature_weights"] = normal((C, 3 * N))
    A["gru_a_dense_feature_bias"] = normal((3 * N,))
    A["gru_b_dense_feature_weights"] = normal((C, 3 * NB))
    A["gru_b_dense_feature_bias"] = np.zeros((3 * NB,), f32)

    # GRU-B: input weights (block-sparse stream, dense by default), recurrent (dotp)
    qb = rng.integers(-qmax_b_in, qmax_b_in + 1, size=(N, 3 * NB)).astype(np.int32)
    mask_b = _rand_mask(rng, N // 4, 3 * NB // 8, gru_b_density)
    qb = qb * np.repeat(np.repeat(mask_b, 4, axis=0), 8, axis=1)
    wq, idx = block_sparse_encode(qb, mask_b)
    A["gru_b_weights"] = wq
    A["gru_b_weights_idx"] = idx
    qb2 = rng.integers(-qmax_b_rec, qmax_b_rec + 1, size=(NB, 3 * NB)).astype(np.int32)
    A["gru_b_recurrent_weights"] = dotp_layout(qb2)
    bias_b = normal((2, 3 * NB))
    A["gru_b_bias"] = bias_b.reshape(-1)
    sub = bias_b.astype(np.float64).copy()
    sub[0] -= qb.sum(axis=0) * (1.0 / 128.0)
    sub[1] -= qb2.sum(axis=0) * (1.0 / 128.0)
    A["gru_b_subias"] = sub.astype(f32).reshape(-1)

    A["feature_conv1_weights"] = normal((3, 84, C))
    A["feature_conv1_bias"] = normal((C,))
    A["feature_conv2_weights"] = normal((3, C, C))
    A["feature_conv2_bias"] = normal((C,))
    A["embed_pitch_weights"] = normal((256, 64))
    A["embed_sig_weights"] = normal((256, 128))
    A["feature_dense1_weights"] = normal((C, C))
    A["feature_dense1_bias"] = normal((C,))
    A["feature_dense2_weights"] = normal((C, C))
    A["feature_dense2_bias"] = normal((C,))
    A["dual_fc_weights"] = normal((256, 2, NB), 0.5)
    A["dual_fc_bias"] = normal((2, 256))
    A["dual_fc_factor"] = (2.0 + 0.2 * rng.standard_normal((2, 256))).astype(f32)

    # GRU-A recurrent: diag + block-sparse int8 per gate
    A["sparse_gru_a_recurrent_weights_diag"] = normal((3 * N,), 0.3)
    qa = np.zeros((N, 3 * N), np.int32)
    mask_a = np.zeros((N // 4, 3 * N // 8), bool)
    for g in range(3):
        m = _rand_mask(rng, N // 4, N // 8, density[g])
        mask_a[:, g * (N // 8):(g + 1) * (N // 8)] = m
        qg = rng.integers(-qmax[g], qmax[g] + 1, size=(N, N)).astype(np.int32)
        qa[:, g * N:(g + 1) * N] = qg * np.repeat(np.repeat(m, 4, axis=0), 8, axis=1)
    wq, idx = block_sparse_encode(qa, mask_a)
    A["sparse_gru_a_recurrent_weights"] = wq
    A["sparse_gru_a_recurrent_weights_idx"] = idx
    bias_a = normal((2, 3 * N))
    A["sparse_gru_a_bias"] = bias_a.reshape(-1)
    sub = bias_a.astype(np.float64).copy()
    sub[1] -= qa.sum(axis=0) * (1.0 / 128.0)
    A["sparse_gru_a_subias"] = sub.astype(f32).reshape(-1)

    for k in list(A.keys()):
        A[k] = np.ascontiguousarray(A[k]).reshape(-1)
    return A


def random_plc_model(seed: int = 4321, cond: int = 128, units: int = 256, sigma: float = 0.1,
                     qmax_in: int = 12, qmax_rec: int = 8) -> "OrderedDict[str, np.ndarray]":
    """Random-init PLC predictor in the reference's array layout (training_tf2/dump_plc.py:160-208, 331-368;
    src/lpcnet_plc.c:135-145): dense 57 -> cond (tanh), GRU cond -> units, GRU units -> units (both int8, the
    input matrix in the block-sparse stream format with every block present, the recurrent one in the dense
    8x4 layout), dense units -> 20 (linear).  Append to a model dict before write_blob: lpcnet_plc_load_model
    reads both models from one blob (src/lpcnet_plc.c:88-98)."""
    rng = np.random.default_rng(seed)
    f32 = np.float32
    A: "OrderedDict[str, np.ndarray]" = OrderedDict()

    def normal(shape, s=sigma):
        return (rng.standard_normal(shape) * s).astype(f32)

    A["plc_dense1_weights"] = normal((57, cond))
    A["plc_dense1_bias"] = normal((cond,))
    for name, nin in (("plc_gru1", cond), ("plc_gru2", units)):
        q = rng.integers(-qmax_in, qmax_in + 1, size=(nin, 3 * units)).astype(np.int32)
        mask = np.ones((nin // 4, 3 * units // 8), bool)
        wq, idx = block_sparse_encode(q, mask)
        A[f"{name}_weights"] = wq
        A[f"{name}_weights_idx"] = idx
        q2 = rng.integers(-qmax_rec, qmax_rec + 1, size=(units, 3 * units)).astype(np.int32)
        A[f"{name}_recurrent_weights"] = dotp_layout(q2)
        bias = normal((2, 3 * units))
        A[f"{name}_bias"] = bias.reshape(-1)
        sub = bias.astype(np.float64).copy()
        sub[0] -= q.sum(axis=0) * (1.0 / 128.0)
        sub[1] -= q2.sum(axis=0) * (1.0 / 128.0)
        A[f"{name}_subias"] = sub.astype(f32).reshape(-1)
    A["plc_out_weights"] = normal((units, 20), 0.05)
    A["plc_out_bias"] = np.concatenate([[1.0], np.zeros(17), [0.3, 0.0]]).astype(f32) + normal((20,), 0.02)
    for k in list(A.keys()):
        A[k] = np.ascontiguousarray(A[k]).reshape(-1)
    return A


def saturation_free(wq: np.ndarray) -> bool:
    """True if no adjacent input pair of any block row can saturate vpmaddubsw's s16:
    255*(|q0|+|q1|) <= 32767, i.e. |q0|+|q1| <= 128."""
    w = wq.astype(np.int32).reshape(-1, 2)
    return bool((np.abs(w).sum(axis=1) * 255 <= 32767).all())


def synthetic_features(n_frames: int, stream: int = 0, total: int = 36) -> np.ndarray:
    """Synthetic feature file content for one stream: [n_frames][36]."""
    rng = np.random.default_rng(1000 + stream)
    f = np.zeros((n_frames, total), np.float32)
    f[:, 0] = 2.0 + 0.5 * rng.standard_normal(n_frames)
    f[:, 1:18] = 0.5 * rng.standard_normal((n_frames, 17))
    f[:, 18] = rng.uniform(-1.0, 2.0, n_frames)
    f[:, 19] = rng.uniform(-0.5, 0.5, n_frames)
    return f


def stream_seed(stream: int) -> bytes:
    """Per-stream kiss99 seed string; stream 0 is the reference default "LPCNet"
    (lpcnet.c:176-181)."""
    return b"LPCNet" if stream == 0 else b"LPCNet" + str(stream).encode()


def synthetic_pcm(n_frames: int, stream: int = 0) -> np.ndarray:
    """Synthetic 16 kHz input for the analysis side (band-limited noise + a sinusoid with a
    moving fundamental, slow amplitude envelope): int16 [n_frames*160]."""
    rng = np.random.default_rng(5000 + stream)
    n = n_frames * 160
    t = np.arange(n)
    f0 = 80 + 200 * rng.random()
    ph = 2 * np.pi * f0 / 16000 * t + 3 * np.sin(2 * np.pi * (1 + 2 * rng.random()) / 16000 * t)
    x = 4000 * np.sin(ph) + 1500 * np.sin(2.01 * ph) + 700 * np.sin(3.02 * ph)
    noise = rng.standard_normal(n + 8)
    x += 600 * np.convolve(noise, np.ones(8) / 8, mode="valid")[:n] + 100 * noise[:n]
    env = 0.55 + 0.45 * np.sin(2 * np.pi * 1.3 / 16000 * t + 6 * rng.random())
    return np.clip(np.round(x * env), -32768, 32767).astype(np.int16)


def synthetic_loss_pattern(n_frames: int, stream: int = 0, p_stay: float = 0.6, p_enter: float = 0.1) -> np.ndarray:
    """Two-state Gilbert loss pattern per 10 ms frame, seed 7 + stream; the first three
    frames are always received.  bool [n_frames]."""
    rng = np.random.default_rng(7 + stream)
    lost = np.zeros(n_frames, bool)
    state = False
    for f in range(n_frames):
        state = bool(rng.random() < (p_stay if state else p_enter))
        lost[f] = state
    lost[:3] = False
    return lost


def synthetic_codebooks(seed: int = 99) -> np.ndarray:
    """Synthetic VQ codebooks for the 1.6 kb/s codec (the real ceps_codebooks.c ships in the model tarball, which is
    not available offline): ceps_codebook1/2/3 [1024][17] and ceps_codebook_diff4 [4096][18] (lpcnet_private.h:109-112),
    N(0, 0.3^2), concatenated -- the same values oracle/gen_codebooks.py compiles into the reference build."""
    rng = np.random.default_rng(seed)
    cb = [rng.standard_normal(1024 * 17) * 0.3 for _ in range(3)]
    cb.append(rng.standard_normal(4096 * 18) * 0.3)
    return np.concatenate([c.astype(np.float32) for c in cb])


def synthetic_fec(n_frames: int, stream: int) -> dict:
    """Synthetic FEC feature vectors for the causal PLC tests: {frame: [20 floats | None, ...]} handed to
    lpcnet_plc_fec_add before that frame (None = "skip one"); none for stream 0.  Seed 50 + stream."""
    if stream == 0:
        return {}
    rng = np.random.default_rng(50 + stream)
    return {f: [None if rng.random() < 0.2 else (rng.standard_normal(20) * 0.5).astype(np.float32)
                for _ in range(int(rng.integers(0, 3)))] for f in range(n_frames)}


def random_dred_decoder_model(seed: int = 8765, cond: int = 256, latent: int = 80, state: int = 24, sigma: float = 0.1,
                              qmax_in: int = 10, qmax_rec: int = 8) -> "OrderedDict[str, np.ndarray]":
    """Random-init RDOVAE decoder (DRED) in the reference's array layout (torch/rdovae/export_rdovae_weights.py:178-222
    over CoreDecoder, torch/rdovae/rdovae/rdovae.py:342-413; consumed by src/dred_rdovae_dec.c): state1..3 state -> cond
    (tanh), dec_dense1 latent -> cond, three int8 GRUs cond -> cond (dec_dense2/4/6) with dense layers between them
    (dec_dense3/5/7/8), dec_final 8*cond -> 4*20 (linear).  Append to a model dict before write_blob: the reference's
    write_lpcnet_weights.c puts every model into one blob too."""
    rng = np.random.default_rng(seed)
    f32 = np.float32
    A: "OrderedDict[str, np.ndarray]" = OrderedDict()

    def normal(shape, s=sigma):
        return (rng.standard_normal(shape) * s).astype(f32)

    for k in (1, 2, 3):
        A[f"state{k}_weights"] = normal((state, cond), 0.3)
        A[f"state{k}_bias"] = normal((cond,))
    A["dec_dense1_weights"] = normal((latent, cond))
    A["dec_dense1_bias"] = normal((cond,))
    for name in ("dec_dense3", "dec_dense5", "dec_dense7", "dec_dense8"):
        A[f"{name}_weights"] = normal((cond, cond), 0.08)
        A[f"{name}_bias"] = normal((cond,))
    A["dec_final_weights"] = normal((8 * cond, 80), 0.03)
    A["dec_final_bias"] = normal((80,), 0.2)
    for name in ("dec_dense2", "dec_dense4", "dec_dense6"):
        q = rng.integers(-qmax_in, qmax_in + 1, size=(cond, 3 * cond)).astype(np.int32)
        mask = rng.random((cond // 4, 3 * cond // 8)) < 0.5            # input_sparse: half of the 8x4 blocks present
        q = q * np.repeat(np.repeat(mask, 4, axis=0), 8, axis=1)
        wq, idx = block_sparse_encode(q, mask)
        A[f"{name}_weights"] = wq
        A[f"{name}_weights_idx"] = idx
        q2 = rng.integers(-qmax_rec, qmax_rec + 1, size=(cond, 3 * cond)).astype(np.int32)
        A[f"{name}_recurrent_weights"] = dotp_layout(q2)
        bias = normal((2, 3 * cond))
        A[f"{name}_bias"] = bias.reshape(-1)
        sub = bias.astype(np.float64).copy()
        sub[0] -= q.sum(axis=0) * (1.0 / 128.0)
        sub[1] -= q2.sum(axis=0) * (1.0 / 128.0)
        A[f"{name}_subias"] = sub.astype(f32).reshape(-1)
    for k in list(A.keys()):
        A[k] = np.ascontiguousarray(A[k]).reshape(-1)
    return A


def saturating_models(qmax: int = 127) -> bytes:
    """One blob (LPCNet + PLC predictor + RDOVAE decoder) whose int8 matrices use the whole range [-qmax, qmax], like
    trained models may (the training-time WeightClip bounds adjacent OUTPUT weights; the pairs vpmaddubsw adds are
    adjacent INPUT weights): pairs with |q0| + |q1| > 128 saturate its s16 sum whenever both activations are large
    (src/vec_avx.h:814-818).  The parity tests on it pin that behaviour (engine.h DevSat)."""
    A = random_model(seed=77, qmax=(qmax, qmax, qmax), qmax_b_in=qmax, qmax_b_rec=qmax)
    A.update(random_plc_model(seed=78, qmax_in=qmax, qmax_rec=qmax))
    A.update(random_dred_decoder_model(seed=79, qmax_in=qmax, qmax_rec=qmax))
    return write_blob(A)
