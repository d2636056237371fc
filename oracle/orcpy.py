"""ctypes binding to oracle/liblpcnet_oracle.so (our CPU restatement of the hot path).

TEST INFRASTRUCTURE ONLY: importable from tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs, never from lpcnet_b200/.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from .refpy import SNAP_DTYPE, TRACE_DTYPE, _p

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "liblpcnet_oracle.so")


def available() -> bool:
    return os.path.exists(LIB_PATH)


class OrcLib:
    def __init__(self):
        self.lib = L = C.CDLL(LIB_PATH)
        L.orc_model_load.restype = C.c_void_p
        L.orc_model_load.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_float, C.c_void_p]
        L.orc_model_free.argtypes = [C.c_void_p]
        L.orc_model_set_end2end.argtypes = [C.c_void_p, C.c_int]
        L.orc_model_nblocks_a.argtypes = [C.c_void_p]
        L.orc_model_nblocks_b.argtypes = [C.c_void_p]
        L.orc_state_new.restype = C.c_void_p
        L.orc_state_new.argtypes = [C.c_void_p]
        L.orc_state_free.argtypes = [C.c_void_p]
        L.orc_state_reset.argtypes = [C.c_void_p]
        L.orc_seed.argtypes = [C.c_void_p, C.c_char_p, C.c_int]
        L.orc_get.argtypes = [C.c_void_p, C.c_void_p]
        L.orc_set.argtypes = [C.c_void_p, C.c_void_p]
        L.orc_synthesize.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int]
        L.orc_run.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p,
                              C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_frame_network.argtypes = [C.c_void_p] * 6
        for n in ("orc_rcp", "orc_tanh_approx", "orc_sigmoid_approx", "orc_ulaw2lin"):
            getattr(L, n).restype = C.c_float
            getattr(L, n).argtypes = [C.c_float]
        L.orc_lin2ulaw.argtypes = [C.c_float]
        L.orc_kiss99_rand.restype = C.c_uint32
        L.orc_set_rcp_table.argtypes = [C.c_void_p]

    def default_rcp_table(self):
        t = np.zeros(2048, np.uint32)
        self.lib.orc_default_rcp_table(_p(t))
        return t

    def set_rcp_table(self, tab):
        self.lib.orc_set_rcp_table(_p(np.ascontiguousarray(tab, np.uint32)) if tab is not None else None)

    def map1(self, name, x):
        f = getattr(self.lib, name)
        return np.array([f(C.c_float(float(v))) for v in np.asarray(x, np.float32)])

    def ps_to_epi8(self, x):
        x = np.ascontiguousarray(x, np.float32)
        y = np.zeros(x.size, np.uint8)
        self.lib.orc_ps_to_epi8(_p(y), _p(x), C.c_int(x.size))
        return y

    def sparse_sgemv(self, out, w, idx, x):
        out = np.ascontiguousarray(out, np.float32).copy()
        self.lib.orc_sparse_sgemv_accum8x4(_p(out), _p(np.ascontiguousarray(w, np.int8)), C.c_int(out.size),
                                           C.c_int(x.size), _p(np.ascontiguousarray(idx, np.int32)),
                                           _p(np.ascontiguousarray(x, np.float32)))
        return out

    def sgemv8x4(self, out, w, x):
        out = np.ascontiguousarray(out, np.float32).copy()
        self.lib.orc_sgemv_accum8x4(_p(out), _p(np.ascontiguousarray(w, np.int8)), C.c_int(out.size),
                                    C.c_int(x.size), _p(np.ascontiguousarray(x, np.float32)))
        return out

    def sgemv(self, out, w, cols, x):
        out = np.ascontiguousarray(out, np.float32).copy()
        self.lib.orc_sgemv_accum(_p(out), _p(np.ascontiguousarray(w, np.float32)), C.c_int(out.size),
                                 C.c_int(cols), C.c_int(out.size), _p(np.ascontiguousarray(x, np.float32)))
        return out

    def kiss99_srand(self, seed: bytes):
        st = np.zeros(4, np.uint32)
        self.lib.orc_kiss99_srand(_p(st), seed, C.c_int(len(seed)))
        return st

    def kiss99_rand(self, st):
        return int(self.lib.orc_kiss99_rand(_p(st)))

    def lpc_from_cepstrum(self, ceps):
        ceps = np.ascontiguousarray(ceps, np.float32)
        lpc = np.zeros(16, np.float32)
        self.lib.orc_lpc_from_cepstrum(_p(lpc), _p(ceps))
        return lpc

    def fft_tables(self):
        tw = np.zeros((320, 2), np.float32)
        br = np.zeros(320, np.int16)
        dct = np.zeros(18 * 18, np.float32)
        self.lib.orc_fft_tables(_p(tw), _p(br), _p(dct))
        return tw, br, dct

    def decode_packet(self, vq_mem, buf: bytes, codebooks):
        feats = np.zeros((4, 36), np.float32)
        self.lib.orc_decode_packet(_p(feats), _p(vq_mem), buf, _p(np.ascontiguousarray(codebooks, np.float32)))
        return feats


class OrcModel:
    def __init__(self, lib: OrcLib, blob: bytes, features_delay: int = 2, lpc_gamma: float = 1.0, rcp_table=None,
                 end2end: bool = False):
        self.L = lib
        tab = np.ascontiguousarray(rcp_table, np.uint32) if rcp_table is not None else None
        self.h = lib.lib.orc_model_load(blob, len(blob), features_delay, C.c_float(lpc_gamma),
                                        _p(tab) if tab is not None else None)
        if not self.h:
            raise RuntimeError("oracle rejected the blob")
        self.features_delay = features_delay
        if end2end:
            lib.lib.orc_model_set_end2end(self.h, 1)

    def __del__(self):
        if getattr(self, "h", None):
            self.L.lib.orc_model_free(self.h)
            self.h = None

    @property
    def nblocks_a(self):
        return self.L.lib.orc_model_nblocks_a(self.h)

    @property
    def nblocks_b(self):
        return self.L.lib.orc_model_nblocks_b(self.h)


class OrcStream:
    def __init__(self, model: OrcModel, seed: bytes | None = None):
        self.m = model
        self.L = model.L
        self.h = self.L.lib.orc_state_new(model.h)
        if seed is not None:
            self.L.lib.orc_seed(self.h, seed, len(seed))

    def __del__(self):
        if getattr(self, "h", None):
            self.L.lib.orc_state_free(self.h)
            self.h = None

    def run(self, feats, forced=None, trace=False, logits=False):
        feats = np.ascontiguousarray(feats, np.float32)
        n_frames, stride = feats.shape
        pcm = np.zeros(n_frames * 160, np.int16)
        tr = np.zeros(n_frames * 160, TRACE_DTYPE) if trace else None
        lg = np.zeros((n_frames * 160, 256), np.float32) if logits else None
        if forced is not None:
            forced = np.ascontiguousarray(forced, np.int16)
        self.L.lib.orc_run(self.m.h, self.h, _p(feats), stride, n_frames, _p(pcm),
                           _p(forced) if forced is not None else None,
                           _p(tr) if trace else None, _p(lg) if logits else None)
        return pcm, tr, lg

    def synthesize(self, features, n=160, preload=0, out=None):
        f = np.ascontiguousarray(features, np.float32)
        if out is None:
            out = np.zeros(n, np.int16)
        self.L.lib.orc_synthesize(self.m.h, self.h, _p(f), _p(out), n, preload)
        return out

    def frame_network(self, features):
        f = np.ascontiguousarray(features, np.float32)
        ga = np.zeros(1152, np.float32)
        gb = np.zeros(48, np.float32)
        lpc = np.zeros(16, np.float32)
        self.L.lib.orc_frame_network(self.m.h, self.h, _p(f), _p(ga), _p(gb), _p(lpc))
        return ga, gb, lpc

    def snapshot(self):
        s = np.zeros(1, SNAP_DTYPE)
        self.L.lib.orc_get(self.h, _p(s))
        return s[0]

    def restore(self, snap):
        s = np.zeros(1, SNAP_DTYPE)
        s[0] = snap
        self.L.lib.orc_set(self.h, _p(s))


class OrcEncoder:
    """Feature analysis oracle (orc_analysis.c): lpcnet_compute_single_frame_features of one stream."""

    def __init__(self, lib: OrcLib):
        self.L = lib
        lib.lib.orc_enc_new.restype = C.c_void_p
        lib.lib.orc_enc_free.argtypes = [C.c_void_p]
        lib.lib.orc_enc_single_frame.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        self.h = lib.lib.orc_enc_new()

    def __del__(self):
        if getattr(self, "h", None):
            self.L.lib.orc_enc_free(self.h)
            self.h = None

    def run(self, pcm):
        """pcm: [n_frames*160] (int16 or float) -> [n_frames][36]"""
        x = np.ascontiguousarray(pcm, np.float32)
        n = x.size // 160
        out = np.zeros((n, 36), np.float32)
        for f in range(n):
            self.L.lib.orc_enc_single_frame(self.h, _p(x[f * 160:(f + 1) * 160]), _p(out[f]))
        return out


    def encode(self, pcm, codebooks):
        """lpcnet_encode over [n_packets*640] samples -> (packets [n][8] uint8, quantised features [n][4][36])"""
        x = np.ascontiguousarray(pcm, np.float32)
        cb = np.ascontiguousarray(codebooks, np.float32)
        n = x.size // 640
        pk = np.zeros((n, 8), np.uint8)
        ft = np.zeros((n, 4, 36), np.float32)
        self.L.lib.orc_enc_superframe.argtypes = [C.c_void_p] * 4 + [C.c_int, C.c_void_p]
        for k in range(n):
            self.L.lib.orc_enc_superframe(self.h, _p(x[k * 640:(k + 1) * 640]), _p(pk[k]), _p(ft[k]), 1, _p(cb))
        return pk, ft

    def compute_features(self, pcm):
        """lpcnet_compute_features over [n_packets*640] samples -> [n*4][36]"""
        x = np.ascontiguousarray(pcm, np.float32)
        n = x.size // 640
        ft = np.zeros((n, 4, 36), np.float32)
        self.L.lib.orc_enc_superframe.argtypes = [C.c_void_p] * 4 + [C.c_int, C.c_void_p]
        for k in range(n):
            self.L.lib.orc_enc_superframe(self.h, _p(x[k * 640:(k + 1) * 640]), None, _p(ft[k]), 0, None)
        return ft.reshape(n * 4, 36)


def half_window(lib: OrcLib):
    lib.lib.orc_half_window.restype = C.POINTER(C.c_float)
    return np.ctypeslib.as_array(lib.lib.orc_half_window(), (160,)).copy()


def burg_cepstral_analysis(lib: OrcLib, x):
    x = np.ascontiguousarray(x, np.float32)
    out = np.zeros(36, np.float32)
    lib.lib.orc_burg_cepstral_analysis(_p(out), _p(x))
    return out


def plc_pred(model: OrcModel, gru1, gru2, x):
    """compute_plc_pred: updates gru1/gru2 (float32[256]) in place, returns the 20 predicted features."""
    out = np.zeros(20, np.float32)
    model.L.lib.orc_plc_pred.argtypes = [C.c_void_p] * 5
    model.L.lib.orc_plc_pred(model.h, _p(gru1), _p(gru2), _p(out), _p(np.ascontiguousarray(x, np.float32)))
    return out


class OrcPLC:
    """Packet-loss concealment of one stream (orc_plc.c): options 1 | 4 need a model loaded with features_delay=0, the
    causal (0) and codec (2) modes take any delay up to 2."""

    def __init__(self, model: OrcModel, options: int = 1, seed: bytes | None = None):
        self.m = model
        L = model.L.lib
        L.orc_plc_new.restype = C.c_void_p
        L.orc_plc_new.argtypes = [C.c_void_p, C.c_int]
        L.orc_plc_free.argtypes = [C.c_void_p]
        L.orc_plc_seed.argtypes = [C.c_void_p, C.c_char_p, C.c_int]
        L.orc_plc_update.argtypes = [C.c_void_p, C.c_void_p]
        L.orc_plc_conceal.argtypes = [C.c_void_p, C.c_void_p]
        self.h = L.orc_plc_new(model.h, options)
        if not self.h:
            raise RuntimeError("orc_plc_new: needs plc_* arrays, a mode in {0, 1, 2}, and features_delay 0 for the non-causal mode")
        if seed is not None:
            L.orc_plc_seed(self.h, seed, len(seed))

    def __del__(self):
        if getattr(self, "h", None):
            self.m.L.lib.orc_plc_free(self.h)
            self.h = None

    def peek(self):
        L = self.m.L.lib
        L.orc_plc_peek.argtypes = [C.c_void_p] * 6
        f = np.zeros(20, np.float32); g1 = np.zeros(256, np.float32); g2 = np.zeros(256, np.float32)
        lc = np.zeros(1, np.int32); dc = np.zeros(2, np.float64)
        L.orc_plc_peek(self.h, _p(f), _p(g1), _p(g2), _p(lc), _p(dc))
        return f, g1, g2, dc

    def run(self, pcm, lost, fec=None):
        """pcm int16 [n_frames*160], lost bool [n_frames] -> concealed output int16 [n_frames*160]
        (the non-causal mode delays the signal by 80 samples)."""
        out = np.ascontiguousarray(pcm, np.int16).copy()
        L = self.m.L.lib
        L.orc_plc_fec_add.argtypes = [C.c_void_p, C.c_void_p]
        for f, l in enumerate(lost):
            for v in (fec or {}).get(f, []):
                L.orc_plc_fec_add(self.h, None if v is None else _p(np.ascontiguousarray(v, np.float32)))
            fr = out[f * 160:(f + 1) * 160]
            if l:
                fr[:] = 0
                L.orc_plc_conceal(self.h, _p(fr))
            else:
                L.orc_plc_update(self.h, _p(fr))
        return out


def dred_decode_all(model: "OrcModel", state, latents):
    """DRED_rdovae_decode_all (orc_dred.c): state [24], latents [n][80] -> features [n][80]"""
    z = np.ascontiguousarray(latents, np.float32)
    st = np.ascontiguousarray(state, np.float32)
    out = np.zeros((z.shape[0], 80), np.float32)
    L = model.L.lib
    L.orc_dred_decode_all.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
    if L.orc_dred_decode_all(model.h, _p(out), _p(st), _p(z), z.shape[0]) != 0:
        raise RuntimeError("orc_dred_decode_all: the model has no RDOVAE decoder arrays")
    return out
