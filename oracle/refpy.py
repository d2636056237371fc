"""ctypes binding to oracle/_ref/libref_d{0,2}.so (the unmodified reference + ref_harness.c).

TEST INFRASTRUCTURE ONLY.  Importable only from tests/, tools that generate tests/golden/
fixtures, __graft_entry__.smoke() and bench.py's cpu_baseline leg.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))

TRACE_DTYPE = np.dtype([("exc", "<i4"), ("u_sig", "<i4"), ("u_pred", "<i4"), ("rng_z", "<u4"),
                        ("rng_w", "<u4"), ("xa_hash", "<u4"), ("xb_hash", "<u4"), ("pcm", "<i4")])

SNAP_DTYPE = np.dtype([("conv1_state", "<f4", 168), ("conv2_state", "<f4", 256),
                       ("gru_a_state", "<f4", 384), ("gru_b_state", "<f4", 16),
                       ("last_sig", "<f4", 16), ("lpc", "<f4", 16), ("old_lpc", "<f4", (2, 16)),
                       ("gru_a_condition", "<f4", 1152), ("gru_b_condition", "<f4", 48),
                       ("deemph_mem", "<f4"), ("last_exc", "<i4"), ("frame_count", "<i4"),
                       ("rng", "<u4", 4)])


def _p(a, t=C.c_void_p):
    return a.ctypes.data_as(t)


def available(delay: int = 2) -> bool:
    return os.path.exists(os.path.join(HERE, "_ref", f"libref_d{delay}.so"))


class RefLib:
    def __init__(self, delay=2):
        # delay: 2 or 0 (FEATURES_DELAY of the build), or "2_e2e": FEATURES_DELAY 2 with -DEND2END
        self.lib = C.CDLL(os.path.join(HERE, "_ref", f"libref_d{delay}.so"))
        L = self.lib
        L.ref_new.restype = C.c_void_p
        L.ref_new.argtypes = [C.c_char_p, C.c_int]
        L.ref_free.argtypes = [C.c_void_p]
        L.ref_seed.argtypes = [C.c_void_p, C.c_char_p, C.c_int]
        L.ref_reset.argtypes = [C.c_void_p]
        L.ref_synthesize.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int]
        L.ref_public_synthesize.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
        L.ref_get.argtypes = [C.c_void_p, C.c_void_p]
        L.ref_set.argtypes = [C.c_void_p, C.c_void_p]
        L.ref_run.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                              C.c_void_p, C.c_void_p]
        L.ref_all_node_logits.argtypes = [C.c_void_p, C.c_void_p]
        L.ref_tanh_approx.restype = C.c_float
        L.ref_tanh_approx.argtypes = [C.c_float]
        L.ref_sigmoid_approx.restype = C.c_float
        L.ref_sigmoid_approx.argtypes = [C.c_float]
        L.ref_lin2ulaw.argtypes = [C.c_float]
        L.ref_ulaw2lin.restype = C.c_float
        L.ref_ulaw2lin.argtypes = [C.c_float]
        L.ref_rcp_ss.restype = C.c_float
        L.ref_rcp_ss.argtypes = [C.c_float]
        L.ref_kiss99_rand.restype = C.c_uint32
        assert L.ref_snapshot_size() == SNAP_DTYPE.itemsize, (L.ref_snapshot_size(), SNAP_DTYPE.itemsize)
        assert L.ref_features_delay() == int(str(delay).split("_")[0])

    # -- kernel-level helpers --
    def vec(self, name, x):
        x = np.ascontiguousarray(x, np.float32)
        y = np.empty_like(x)
        getattr(self.lib, name)(_p(y), _p(x), C.c_int(x.size))
        return y

    def ps_to_epi8(self, x):
        x = np.ascontiguousarray(x, np.float32)
        y = np.zeros(x.size + 32, np.uint8)
        self.lib.ref_ps_to_epi8(_p(y), _p(x), C.c_int(x.size))
        return y[:x.size]

    def sparse_sgemv(self, out, w, idx, x):
        out = np.ascontiguousarray(out, np.float32).copy()
        self.lib.ref_sparse_sgemv_accum8x4(_p(out), _p(np.ascontiguousarray(w, np.int8)), C.c_int(out.size),
                                           C.c_int(x.size), _p(np.ascontiguousarray(idx, np.int32)),
                                           _p(np.ascontiguousarray(x, np.float32)))
        return out

    def sgemv8x4(self, out, w, x):
        out = np.ascontiguousarray(out, np.float32).copy()
        self.lib.ref_sgemv_accum8x4(_p(out), _p(np.ascontiguousarray(w, np.int8)), C.c_int(out.size),
                                    C.c_int(x.size), _p(np.ascontiguousarray(x, np.float32)))
        return out

    def sgemv16(self, out, w, cols, x):
        out = np.ascontiguousarray(out, np.float32).copy()
        self.lib.ref_sgemv_accum16(_p(out), _p(np.ascontiguousarray(w, np.float32)), C.c_int(out.size),
                                   C.c_int(cols), C.c_int(out.size), _p(np.ascontiguousarray(x, np.float32)))
        return out

    def kiss99_srand(self, seed: bytes):
        st = np.zeros(4, np.uint32)
        self.lib.ref_kiss99_srand(_p(st), seed, C.c_int(len(seed)))
        return st

    def kiss99_rand(self, st):
        return int(self.lib.ref_kiss99_rand(_p(st)))

    def lpc_from_cepstrum(self, ceps):
        ceps = np.ascontiguousarray(ceps, np.float32)
        lpc = np.zeros(16, np.float32)
        self.lib.ref_lpc_from_cepstrum(_p(lpc), _p(ceps))
        return lpc

    def rcp_table(self):
        """rcpps of the 2048 bucket representatives 1 + i/2048 (this CPU's estimate table)."""
        t = np.zeros(2048, np.float32)
        for i in range(2048):
            t[i] = self.lib.ref_rcp_ss(C.c_float(1.0 + i / 2048.0))
        return t

    def decode_packet(self, vq_mem, buf: bytes):
        feats = np.zeros((4, 36), np.float32)
        self.lib.ref_decode_packet(_p(feats), _p(vq_mem), buf)
        return feats


class RefStream:
    """One reference LPCNetState with a model loaded from `blob`."""

    def __init__(self, lib: RefLib, blob: bytes, seed: bytes | None = None):
        self.L = lib
        self.h = lib.lib.ref_new(blob, len(blob))
        if not self.h:
            raise RuntimeError("reference lpcnet_load_model rejected the blob")
        if seed is not None:
            lib.lib.ref_seed(self.h, seed, len(seed))

    def __del__(self):
        if getattr(self, "h", None):
            self.L.lib.ref_free(self.h)
            self.h = None

    def run(self, feats, forced=None, trace=False, logits=False):
        feats = np.ascontiguousarray(feats, np.float32)
        n_frames, stride = feats.shape
        pcm = np.zeros(n_frames * 160, np.int16)
        tr = np.zeros(n_frames * 160, TRACE_DTYPE) if trace else None
        lg = np.zeros((n_frames * 160, 256), np.float32) if logits else None
        if forced is not None:
            forced = np.ascontiguousarray(forced, np.int16)
            assert forced.size == n_frames * 160
        self.L.lib.ref_run(self.h, _p(feats), stride, n_frames, _p(pcm),
                           _p(forced) if forced is not None else None,
                           _p(tr) if trace else None, _p(lg) if logits else None)
        return pcm, tr, lg

    def synthesize(self, features, n=160):
        f = np.ascontiguousarray(features, np.float32)
        out = np.zeros(n, np.int16)
        self.L.lib.ref_public_synthesize(self.h, _p(f), _p(out), n)
        return out

    def snapshot(self):
        s = np.zeros(1, SNAP_DTYPE)
        self.L.lib.ref_get(self.h, _p(s))
        return s[0]

    def restore(self, snap):
        s = np.zeros(1, SNAP_DTYPE)
        s[0] = snap
        self.L.lib.ref_set(self.h, _p(s))

    def logit_table(self):
        t = np.zeros(256, np.float32)
        self.L.lib.ref_sampling_logit_table(self.h, _p(t))
        return t


class RefEncoder:
    """The reference's own LPCNetEncState (lpcnet_encoder_create, lpcnet_compute_single_frame_features[_float])."""

    def __init__(self, lib: RefLib):
        self.L = lib
        L = lib.lib
        L.lpcnet_encoder_create.restype = C.c_void_p
        L.lpcnet_encoder_destroy.argtypes = [C.c_void_p]
        L.lpcnet_compute_single_frame_features.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.lpcnet_compute_single_frame_features_float.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        self.h = L.lpcnet_encoder_create()

    def __del__(self):
        if getattr(self, "h", None):
            self.L.lib.lpcnet_encoder_destroy(self.h)
            self.h = None

    def run(self, pcm):
        pcm = np.asarray(pcm)
        n = pcm.size // 160
        out = np.zeros((n, 36), np.float32)
        for f in range(n):
            if pcm.dtype == np.int16:
                x = np.ascontiguousarray(pcm[f * 160:(f + 1) * 160])
                self.L.lib.lpcnet_compute_single_frame_features(self.h, _p(x), _p(out[f]))
            else:
                x = np.ascontiguousarray(pcm[f * 160:(f + 1) * 160], np.float32)
                self.L.lib.lpcnet_compute_single_frame_features_float(self.h, _p(x), _p(out[f]))
        return out


    def encode(self, pcm):
        """the reference's lpcnet_encode (lpcnet_enc.c:882-893) over [n_packets*640] int16 -> packets [n][8]"""
        x = np.ascontiguousarray(pcm, np.int16)
        n = x.size // 640
        pk = np.zeros((n, 8), np.uint8)
        self.L.lib.lpcnet_encode.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        for k in range(n):
            self.L.lib.lpcnet_encode(self.h, _p(x[k * 640:(k + 1) * 640]), _p(pk[k]))
        return pk

    def compute_features(self, pcm):
        """the reference's lpcnet_compute_features (lpcnet_enc.c:895-911) -> [n_packets*4][36]"""
        x = np.ascontiguousarray(pcm, np.int16)
        n = x.size // 640
        ft = np.zeros((n, 4, 36), np.float32)
        self.L.lib.lpcnet_compute_features.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        for k in range(n):
            self.L.lib.lpcnet_compute_features(self.h, _p(x[k * 640:(k + 1) * 640]), _p(ft[k]))
        return ft.reshape(n * 4, 36)


def ref_half_window(lib: RefLib):
    return np.ctypeslib.as_array((C.c_float * 160).in_dll(lib.lib, "half_window")).copy()


def ref_burg_cepstral_analysis(lib: RefLib, x):
    x = np.ascontiguousarray(x, np.float32)
    out = np.zeros(36, np.float32)
    lib.lib.ref_burg_cepstral_analysis(_p(out), _p(x))
    return out


class RefPLC:
    """The reference's LPCNetPLCState (lpcnet_plc_create / _load_model / _update / _conceal); use RefLib(0)
    for the non-causal mode (it needs a FEATURES_DELAY 0 build, src/lpcnet_plc.c:357-361)."""

    def __init__(self, lib: RefLib, blob: bytes, options: int = 1, seed: bytes | None = None):
        self.L = lib
        L = lib.lib
        L.ref_plc_new.restype = C.c_void_p
        L.ref_plc_new.argtypes = [C.c_char_p, C.c_int, C.c_int]
        L.ref_plc_free.argtypes = [C.c_void_p]
        L.ref_plc_seed.argtypes = [C.c_void_p, C.c_char_p, C.c_int]
        L.ref_plc_update.argtypes = [C.c_void_p, C.c_void_p]
        L.ref_plc_conceal.argtypes = [C.c_void_p, C.c_void_p]
        L.ref_plc_pred.argtypes = [C.c_void_p] * 5
        self.h = L.ref_plc_new(blob, len(blob), options)
        if not self.h:
            raise RuntimeError("reference rejected the PLC blob / options")
        if seed is not None:
            L.ref_plc_seed(self.h, seed, len(seed))

    def __del__(self):
        if getattr(self, "h", None):
            self.L.lib.ref_plc_free(self.h)
            self.h = None

    def run(self, pcm, lost, fec=None):
        """fec: {frame: [features20 | None, ...]} handed to lpcnet_plc_fec_add before that frame is processed"""
        out = np.ascontiguousarray(pcm, np.int16).copy()
        self.L.lib.ref_plc_fec_add.argtypes = [C.c_void_p, C.c_void_p]
        for f, l in enumerate(lost):
            for v in (fec or {}).get(f, []):
                self.L.lib.ref_plc_fec_add(self.h, None if v is None else _p(np.ascontiguousarray(v, np.float32)))
            fr = out[f * 160:(f + 1) * 160]
            if l:
                fr[:] = 0
                self.L.lib.ref_plc_conceal(self.h, _p(fr))
            else:
                self.L.lib.ref_plc_update(self.h, _p(fr))
        return out

    def pred(self, gru1, gru2, x):
        out = np.zeros(20, np.float32)
        self.L.lib.ref_plc_pred(self.h, _p(gru1), _p(gru2), _p(out), _p(np.ascontiguousarray(x, np.float32)))
        return out


def ref_dred_decode_all(lib: RefLib, blob: bytes, state, latents):
    """the reference's dred_rdovae_dec_init_states + dred_rdovae_decode_qframe (src/dred_rdovae_dec.c) over a model
    parsed from `blob`: state [24], latents [n][80] -> features [n][80]"""
    z = np.ascontiguousarray(latents, np.float32)
    st = np.ascontiguousarray(state, np.float32)
    out = np.zeros((z.shape[0], 80), np.float32)
    lib.lib.ref_dred_decode_all.argtypes = [C.c_char_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
    if lib.lib.ref_dred_decode_all(blob, len(blob), _p(out), _p(st), _p(z), z.shape[0]) != 0:
        raise RuntimeError("reference rejected the RDOVAE decoder arrays")
    return out
