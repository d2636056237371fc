"""ctypes host mirror of include/lpcnet_b200.h (the C-ABI of liblpcnet_b200.so).

`Batch` mirrors the batched multi-stream API; `Synth` / `Decoder` mirror the reference's
single-stream `lpcnet_create / lpcnet_load_model / lpcnet_synthesize / lpcnet_decode` calls
(include/lpcnet.h:160-188, 67-96) with the same argument meaning and error behaviour.

There is no CPU fallback: if the library is missing, or no CUDA device is present, the calls
raise.  Nothing here imports the test oracle.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "liblpcnet_b200.so")

FRAME_SIZE = 160
NB_FEATURES = 20
NB_TOTAL_FEATURES = 36

TRACE_DTYPE = np.dtype([("exc", "<i4"), ("u_sig", "<i4"), ("u_pred", "<i4"), ("rng_z", "<u4"),
                        ("rng_w", "<u4"), ("xa_hash", "<u4"), ("xb_hash", "<u4"), ("pcm", "<i4")])

STATE_DTYPE = np.dtype([("conv1_state", "<f4", 168), ("conv2_state", "<f4", 256),
                        ("gru_a_state", "<f4", 384), ("gru_b_state", "<f4", 16),
                        ("last_sig", "<f4", 16), ("lpc", "<f4", 16), ("old_lpc", "<f4", (2, 16)),
                        ("gru_a_condition", "<f4", 1152), ("gru_b_condition", "<f4", 48),
                        ("deemph_mem", "<f4"), ("last_exc", "<i4"), ("frame_count", "<i4"),
                        ("rng", "<u4", 4)])

# every symbol include/lpcnet_b200.h declares (checked by tests/test_abi.py)
EXPORTED_SYMBOLS = [
    "lpcnet_get_size", "lpcnet_init", "lpcnet_create", "lpcnet_destroy", "lpcnet_reset",
    "lpcnet_load_model", "lpcnet_synthesize", "lpcnet_decoder_get_size", "lpcnet_decoder_init",
    "lpcnet_decoder_create", "lpcnet_decoder_destroy", "lpcnet_decoder_load_model", "lpcnet_decode",
    "lpcnet_batch_create", "lpcnet_batch_destroy", "lpcnet_batch_n_streams", "lpcnet_batch_reset",
    "lpcnet_batch_set_seed", "lpcnet_batch_set_cuda_stream", "lpcnet_batch_set_tile",
    "lpcnet_batch_synthesize", "lpcnet_batch_synthesize_device", "lpcnet_batch_frame_network",
    "lpcnet_batch_decode", "lpcnet_batch_set_codebooks", "lpcnet_batch_get_state",
    "lpcnet_batch_set_state", "lpcnet_batch_sync", "lpcnet_batch_nblocks", "lpcnet_batch_launch_count",
    "lpcnet_batch_set_profiling", "lpcnet_batch_read_profile", "lpcnet_batch_profile_launches", "lpcnet_b200_measure_l2_gather",
    "lpcnet_b200_set_rcp_table", "lpcnet_b200_default_rcp_table", "lpcnet_b200_set_model_constants",
    "lpcnet_b200_set_end2end", "lpcnet_batch_set_end2end",
    "lpcnet_b200_last_error", "lpcnet_b200_device_count", "lpcnet_b200_lpc_from_cepstrum_noalloc",
    "lpcnet_batch_create_multi", "lpcnet_multi_destroy", "lpcnet_multi_n_streams", "lpcnet_multi_n_devices",
    "lpcnet_multi_shard", "lpcnet_multi_set_seed", "lpcnet_multi_reset", "lpcnet_multi_set_codebooks",
    "lpcnet_multi_synthesize", "lpcnet_multi_decode",
    "lpcnet_encoder_get_size", "lpcnet_encoder_init", "lpcnet_encoder_create", "lpcnet_encoder_destroy",
    "lpcnet_compute_single_frame_features", "lpcnet_compute_single_frame_features_float",
    "lpcnet_feature_batch_create", "lpcnet_feature_batch_destroy", "lpcnet_feature_batch_n_streams",
    "lpcnet_feature_batch_launch_count", "lpcnet_feature_batch_reset", "lpcnet_feature_batch_compute",
    "lpcnet_feature_batch_compute_float", "lpcnet_feature_batch_compute_device",
    "lpcnet_b200_set_codebooks", "lpcnet_encode", "lpcnet_compute_features", "lpcnet_feature_batch_set_codebooks",
    "lpcnet_feature_batch_encode", "lpcnet_feature_batch_compute_features",
    "lpcnet_plc_get_size", "lpcnet_plc_init", "lpcnet_plc_create", "lpcnet_plc_destroy", "lpcnet_plc_load_model",
    "lpcnet_plc_update", "lpcnet_plc_conceal", "lpcnet_plc_reset", "lpcnet_plc_fec_add", "lpcnet_plc_fec_clear",
    "lpcnet_plc_batch_reset", "lpcnet_plc_batch_fec_add", "lpcnet_plc_batch_fec_clear",
    "lpcnet_dred_decoder_create", "lpcnet_dred_decoder_destroy", "lpcnet_dred_decoder_launch_count",
    "lpcnet_dred_decode_all", "lpcnet_dred_decode_batch", "lpcnet_plc_batch_create", "lpcnet_plc_batch_destroy",
    "lpcnet_plc_batch_n_streams", "lpcnet_plc_batch_set_seed", "lpcnet_plc_batch_launch_count",
    "lpcnet_plc_batch_process", "lpcnet_plc_batch_peek", "lpcnet_plc_batch_peek_frame",
]

_lib = None


PACKET_SAMPLES = 640         # LPCNET_PACKET_SAMPLES (include/lpcnet.h)


class LPCNetError(RuntimeError):
    pass


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def load_library() -> C.CDLL:
    """dlopen liblpcnet_b200.so (built in-tree by lpcnet_b200/build.py).  Fails loudly."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise LPCNetError(f"{LIB_PATH} is missing: run `python -m lpcnet_b200.build` "
                          "(the engine has no CPU fallback)")
    L = C.CDLL(LIB_PATH)
    vp, ci, cf = C.c_void_p, C.c_int, C.c_float
    L.lpcnet_b200_last_error.restype = C.c_char_p
    L.lpcnet_batch_create.restype = vp
    L.lpcnet_batch_create.argtypes = [vp, ci, ci, ci, ci, cf]
    L.lpcnet_batch_destroy.argtypes = [vp]
    L.lpcnet_batch_n_streams.argtypes = [vp]
    L.lpcnet_batch_reset.argtypes = [vp, ci]
    L.lpcnet_batch_set_seed.argtypes = [vp, ci, C.c_char_p, ci]
    L.lpcnet_batch_set_cuda_stream.argtypes = [vp, vp]
    L.lpcnet_batch_set_tile.argtypes = [vp, ci]
    L.lpcnet_batch_synthesize.argtypes = [vp, vp, ci, ci, vp, vp, vp, vp]
    L.lpcnet_batch_synthesize_device.argtypes = [vp, vp, ci, ci, ci, vp, vp, vp, vp]
    L.lpcnet_batch_frame_network.argtypes = [vp, vp, ci, ci, vp, vp, vp]
    L.lpcnet_batch_decode.argtypes = [vp, vp, ci, vp, vp]
    L.lpcnet_batch_set_codebooks.argtypes = [vp, vp, ci]
    L.lpcnet_batch_get_state.argtypes = [vp, ci, vp]
    L.lpcnet_batch_set_state.argtypes = [vp, ci, vp]
    L.lpcnet_batch_sync.argtypes = [vp]
    L.lpcnet_batch_nblocks.argtypes = [vp, ci]
    L.lpcnet_batch_launch_count.restype = C.c_longlong
    L.lpcnet_batch_launch_count.argtypes = [vp]
    L.lpcnet_batch_set_profiling.argtypes = [vp, ci]
    L.lpcnet_batch_read_profile.argtypes = [vp, vp, vp, vp, vp]
    L.lpcnet_batch_profile_launches.argtypes = [vp, vp, vp, ci]
    L.lpcnet_b200_measure_l2_gather.argtypes = [ci, vp]
    L.lpcnet_b200_lpc_from_cepstrum_noalloc.argtypes = [ci, vp, ci, vp, ci, vp]
    L.lpcnet_batch_create_multi.restype = vp
    L.lpcnet_batch_create_multi.argtypes = [vp, ci, ci, C.c_uint, ci, cf]
    L.lpcnet_multi_destroy.argtypes = [vp]
    L.lpcnet_multi_n_streams.argtypes = [vp]
    L.lpcnet_multi_n_devices.argtypes = [vp]
    L.lpcnet_multi_shard.argtypes = [vp, ci, vp, vp, vp]
    L.lpcnet_multi_set_seed.argtypes = [vp, ci, C.c_char_p, ci]
    L.lpcnet_multi_reset.argtypes = [vp, ci]
    L.lpcnet_multi_set_codebooks.argtypes = [vp, vp, ci]
    L.lpcnet_multi_synthesize.argtypes = [vp, vp, ci, ci, vp]
    L.lpcnet_multi_decode.argtypes = [vp, vp, ci, vp]
    L.lpcnet_b200_set_rcp_table.argtypes = [vp]
    L.lpcnet_b200_default_rcp_table.argtypes = [vp]
    L.lpcnet_b200_set_model_constants.argtypes = [ci, cf]
    L.lpcnet_b200_set_end2end.argtypes = [ci]
    L.lpcnet_batch_set_end2end.argtypes = [vp, ci]
    L.lpcnet_create.restype = vp
    L.lpcnet_destroy.argtypes = [vp]
    L.lpcnet_reset.argtypes = [vp]
    L.lpcnet_load_model.argtypes = [vp, vp, ci]
    L.lpcnet_synthesize.argtypes = [vp, vp, vp, ci]
    L.lpcnet_synthesize.restype = None
    L.lpcnet_decoder_create.restype = vp
    L.lpcnet_decoder_destroy.argtypes = [vp]
    L.lpcnet_decoder_load_model.argtypes = [vp, vp, ci]
    L.lpcnet_decode.argtypes = [vp, vp, vp]
    L.lpcnet_encoder_create.restype = vp
    L.lpcnet_encoder_destroy.argtypes = [vp]
    L.lpcnet_compute_single_frame_features.argtypes = [vp, vp, vp]
    L.lpcnet_compute_single_frame_features_float.argtypes = [vp, vp, vp]
    L.lpcnet_feature_batch_create.restype = vp
    L.lpcnet_feature_batch_create.argtypes = [ci, ci]
    L.lpcnet_feature_batch_destroy.argtypes = [vp]
    L.lpcnet_feature_batch_n_streams.argtypes = [vp]
    L.lpcnet_feature_batch_launch_count.restype = C.c_longlong
    L.lpcnet_feature_batch_launch_count.argtypes = [vp]
    L.lpcnet_feature_batch_reset.argtypes = [vp, ci]
    L.lpcnet_feature_batch_compute.argtypes = [vp, vp, ci, vp]
    L.lpcnet_feature_batch_compute_float.argtypes = [vp, vp, ci, vp]
    L.lpcnet_feature_batch_compute_device.argtypes = [vp, vp, vp, ci, vp, vp]
    L.lpcnet_b200_set_codebooks.argtypes = [vp, ci]
    L.lpcnet_encode.argtypes = [vp, vp, vp]
    L.lpcnet_compute_features.argtypes = [vp, vp, vp]
    L.lpcnet_feature_batch_set_codebooks.argtypes = [vp, vp, ci]
    L.lpcnet_feature_batch_encode.argtypes = [vp, vp, ci, vp, vp]
    L.lpcnet_feature_batch_compute_features.argtypes = [vp, vp, ci, vp]
    L.lpcnet_plc_create.restype = vp
    L.lpcnet_plc_create.argtypes = [ci]
    L.lpcnet_plc_destroy.argtypes = [vp]
    L.lpcnet_plc_load_model.argtypes = [vp, vp, ci]
    L.lpcnet_plc_update.argtypes = [vp, vp]
    L.lpcnet_plc_reset.argtypes = [vp]
    L.lpcnet_plc_reset.restype = None
    L.lpcnet_plc_fec_add.argtypes = [vp, vp]
    L.lpcnet_plc_fec_add.restype = None
    L.lpcnet_plc_fec_clear.argtypes = [vp]
    L.lpcnet_plc_fec_clear.restype = None
    L.lpcnet_plc_batch_reset.argtypes = [vp]
    L.lpcnet_plc_batch_fec_add.argtypes = [vp, ci, vp]
    L.lpcnet_plc_batch_fec_clear.argtypes = [vp, ci]
    L.lpcnet_dred_decoder_create.restype = vp
    L.lpcnet_dred_decoder_create.argtypes = [vp, ci, ci]
    L.lpcnet_dred_decoder_destroy.argtypes = [vp]
    L.lpcnet_dred_decoder_launch_count.restype = C.c_longlong
    L.lpcnet_dred_decoder_launch_count.argtypes = [vp]
    L.lpcnet_dred_decode_all.argtypes = [vp, vp, vp, vp, ci]
    L.lpcnet_dred_decode_batch.argtypes = [vp, ci, vp, vp, vp, ci]
    L.lpcnet_plc_conceal.argtypes = [vp, vp]
    L.lpcnet_plc_batch_create.restype = vp
    L.lpcnet_plc_batch_create.argtypes = [vp, ci, ci, ci, ci, cf]
    L.lpcnet_plc_batch_destroy.argtypes = [vp]
    L.lpcnet_plc_batch_n_streams.argtypes = [vp]
    L.lpcnet_plc_batch_set_seed.argtypes = [vp, ci, C.c_char_p, ci]
    L.lpcnet_plc_batch_launch_count.restype = C.c_longlong
    L.lpcnet_plc_batch_launch_count.argtypes = [vp]
    L.lpcnet_plc_batch_process.argtypes = [vp, vp, vp, ci]
    L.lpcnet_plc_batch_peek.argtypes = [vp, ci, vp, vp, vp, vp]
    L.lpcnet_plc_batch_peek_frame.argtypes = [vp, ci, vp, vp]
    _lib = L
    return L


def last_error() -> str:
    return load_library().lpcnet_b200_last_error().decode()


def device_count() -> int:
    return load_library().lpcnet_b200_device_count()


def default_rcp_table() -> np.ndarray:
    t = np.zeros(2048, np.uint32)
    load_library().lpcnet_b200_default_rcp_table(_p(t))
    return t


def set_rcp_table(table) -> None:
    L = load_library()
    if table is None:
        L.lpcnet_b200_set_rcp_table(None)
    else:
        t = np.ascontiguousarray(table, np.uint32)
        assert t.size == 2048
        L.lpcnet_b200_set_rcp_table(_p(t))


def measure_l2_gather(device: int = 0) -> float:
    v = C.c_double(0)
    if load_library().lpcnet_b200_measure_l2_gather(device, C.byref(v)) != 0:
        raise LPCNetError(f"lpcnet_b200_measure_l2_gather: {last_error()}")
    return v.value


NOALLOC_MAX_ROWS = 4096      # LPCNET_B200_NOALLOC_MAX_ROWS


def lpc_from_cepstrum_noalloc(d_cepstrum: int, stride: int, d_lpc: int, n_rows: int, cuda_stream: int = 0,
                              device: int = -1) -> None:
    """lpcnet_b200_lpc_from_cepstrum_noalloc: device pointers in and out, no allocation inside the call."""
    rc = load_library().lpcnet_b200_lpc_from_cepstrum_noalloc(device, C.c_void_p(d_cepstrum), stride, C.c_void_p(d_lpc),
                                                              n_rows, C.c_void_p(cuda_stream) if cuda_stream else None)
    if rc != 0:
        raise LPCNetError(f"lpcnet_b200_lpc_from_cepstrum_noalloc: {last_error()}")


def _check(rc: int, what: str) -> None:
    if rc != 0:
        raise LPCNetError(f"{what}: {last_error()}")


class Batch:
    """n_streams independent LPCNet synthesis streams advanced in lock step on one GPU."""

    def __init__(self, blob: bytes, n_streams: int, device: int = -1, features_delay: int = 2,
                 lpc_gamma: float = 1.0):
        self.L = load_library()
        self._blob = np.frombuffer(blob, np.uint8)
        self.h = self.L.lpcnet_batch_create(_p(self._blob), len(blob), n_streams, device, features_delay,
                                            C.c_float(lpc_gamma))
        if not self.h:
            raise LPCNetError(f"lpcnet_batch_create: {last_error()}")
        self.n_streams = n_streams
        self.features_delay = features_delay

    def close(self):
        if getattr(self, "h", None):
            self.L.lpcnet_batch_destroy(self.h)
            self.h = None

    def __del__(self):
        self.close()

    def reset(self, stream: int = -1):
        _check(self.L.lpcnet_batch_reset(self.h, stream), "lpcnet_batch_reset")

    def set_seed(self, stream: int, seed: bytes):
        _check(self.L.lpcnet_batch_set_seed(self.h, stream, seed, len(seed)), "lpcnet_batch_set_seed")

    def set_end2end(self, on: bool = True):
        """the reference's -DEND2END build: LPC from rc2lpc(condition[0..16)) (src/lpcnet.c:57-78, 108)"""
        _check(self.L.lpcnet_batch_set_end2end(self.h, 1 if on else 0), "lpcnet_batch_set_end2end")

    def set_tile(self, streams_per_cta: int):
        _check(self.L.lpcnet_batch_set_tile(self.h, streams_per_cta), "lpcnet_batch_set_tile")

    def set_cuda_stream(self, handle: int):
        _check(self.L.lpcnet_batch_set_cuda_stream(self.h, C.c_void_p(handle)), "lpcnet_batch_set_cuda_stream")

    def sync(self):
        _check(self.L.lpcnet_batch_sync(self.h), "lpcnet_batch_sync")

    @property
    def nblocks_a(self):
        return self.L.lpcnet_batch_nblocks(self.h, 0)

    @property
    def nblocks_b(self):
        return self.L.lpcnet_batch_nblocks(self.h, 1)

    @property
    def chunks_a(self):
        return self.L.lpcnet_batch_nblocks(self.h, 4)

    @property
    def tile(self):
        """streams per CTA the sample kernel runs with (set_tile, or the automatic choice)"""
        return self.L.lpcnet_batch_nblocks(self.h, 5)

    @property
    def rcp_arith(self):
        """the fast kernel variants run: rcpps computed (verified on this GPU) and conversion-free int32 seeds (the
        model's sums proven inside +-2^22)"""
        return bool(self.L.lpcnet_batch_nblocks(self.h, 2)) and bool(self.L.lpcnet_batch_nblocks(self.h, 7))

    @property
    def fixed_slots(self):
        return bool(self.L.lpcnet_batch_nblocks(self.h, 3))

    @property
    def saturating_pairs(self):
        """int8 input pairs of the sample network that can saturate vpmaddubsw's s16 (0 for clipped models)"""
        return self.L.lpcnet_batch_nblocks(self.h, 6)

    @property
    def launch_count(self):
        return int(self.L.lpcnet_batch_launch_count(self.h))

    def set_profiling(self, on: bool):
        _check(self.L.lpcnet_batch_set_profiling(self.h, int(on)), "lpcnet_batch_set_profiling")

    def profile_launches(self):
        """(ms, steps) arrays of the sample-kernel launches summed by the last read_profile()."""
        n = self.L.lpcnet_batch_profile_launches(self.h, None, None, 0)
        ms = np.zeros(max(n, 1), np.float32)
        steps = np.zeros(max(n, 1), np.int32)
        self.L.lpcnet_batch_profile_launches(self.h, _p(ms), _p(steps), n)
        return ms[:n], steps[:n]

    def read_profile(self):
        """(sample_kernel_ms, n_sample_launches, frame_network_ms, n_frame_groups) since the last read."""
        sm, fm = C.c_double(0), C.c_double(0)
        sn, fn = C.c_int(0), C.c_int(0)
        _check(self.L.lpcnet_batch_read_profile(self.h, C.byref(sm), C.byref(sn), C.byref(fm), C.byref(fn)),
               "lpcnet_batch_read_profile")
        return sm.value, sn.value, fm.value, fn.value

    def synthesize(self, features, forced=None, trace=False, logits=False, pcm_out=None):
        """features: [n_streams][n_frames][>=20] float32 (host).  Returns (pcm, trace, logits)."""
        f = np.ascontiguousarray(features, np.float32)
        assert f.ndim == 3 and f.shape[0] == self.n_streams, f.shape
        n_frames, stride = f.shape[1], f.shape[2]
        T = n_frames * FRAME_SIZE
        pcm = pcm_out if pcm_out is not None else np.zeros((self.n_streams, T), np.int16)
        fo = None
        if forced is not None:
            fo = np.ascontiguousarray(forced, np.int16)
            assert fo.shape == (self.n_streams, T)
        tr = np.zeros((self.n_streams, T), TRACE_DTYPE) if trace else None
        lg = np.zeros((self.n_streams, T, 256), np.float32) if logits else None
        _check(self.L.lpcnet_batch_synthesize(self.h, _p(f), stride, n_frames, _p(pcm), _p(fo), _p(tr), _p(lg)),
               "lpcnet_batch_synthesize")
        return pcm, tr, lg

    def synthesize_device(self, d_features: int, stride: int, n_frames: int, d_pcm: int,
                          samples_per_frame: int = FRAME_SIZE, d_forced: int = 0):
        """Device-pointer entry point (asynchronous on the batch's CUDA stream)."""
        _check(self.L.lpcnet_batch_synthesize_device(self.h, C.c_void_p(d_features), stride, n_frames,
                                                     samples_per_frame, C.c_void_p(d_pcm),
                                                     C.c_void_p(d_forced) if d_forced else None, None, None),
               "lpcnet_batch_synthesize_device")

    def frame_network(self, features):
        f = np.ascontiguousarray(features, np.float32)
        n_frames, stride = f.shape[1], f.shape[2]
        ga = np.zeros((self.n_streams, n_frames, 1152), np.float32)
        gb = np.zeros((self.n_streams, n_frames, 48), np.float32)
        lpc = np.zeros((self.n_streams, n_frames, 16), np.float32)
        _check(self.L.lpcnet_batch_frame_network(self.h, _p(f), stride, n_frames, _p(ga), _p(gb), _p(lpc)),
               "lpcnet_batch_frame_network")
        return ga, gb, lpc

    def set_codebooks(self, codebooks):
        cb = np.ascontiguousarray(codebooks, np.float32).reshape(-1)
        _check(self.L.lpcnet_batch_set_codebooks(self.h, _p(cb), cb.size), "lpcnet_batch_set_codebooks")

    def decode(self, packets, want_features=False, pcm_out=None):
        """packets: [n_streams][n_packets][8] uint8.  Returns (pcm [n][n_packets*640], features).  pcm_out: a
        caller's buffer of that shape (pinned memory receives the chunks while the next ones are computed)."""
        pk = np.ascontiguousarray(packets, np.uint8)
        assert pk.ndim == 3 and pk.shape[0] == self.n_streams and pk.shape[2] == 8
        n_packets = pk.shape[1]
        pcm = pcm_out if pcm_out is not None else np.zeros((self.n_streams, n_packets * 640), np.int16)
        assert pcm.shape == (self.n_streams, n_packets * 640) and pcm.dtype == np.int16 and pcm.flags.c_contiguous
        feats = np.zeros((self.n_streams, n_packets * 4, 36), np.float32) if want_features else None
        _check(self.L.lpcnet_batch_decode(self.h, _p(pk), n_packets, _p(pcm), _p(feats)), "lpcnet_batch_decode")
        return pcm, feats

    def get_state(self, stream: int):
        s = np.zeros(1, STATE_DTYPE)
        _check(self.L.lpcnet_batch_get_state(self.h, stream, _p(s)), "lpcnet_batch_get_state")
        return s[0]

    def set_state(self, stream: int, state):
        s = np.zeros(1, STATE_DTYPE)
        s[0] = state
        _check(self.L.lpcnet_batch_set_state(self.h, stream, _p(s)), "lpcnet_batch_set_state")


class MultiBatch:
    """n_streams sharded over the GPUs of one box behind one handle (lpcnet_batch_create_multi): contiguous shards,
    one feeder thread per GPU, host buffers in and out."""

    def __init__(self, blob: bytes, n_streams: int, device_mask: int = 0, features_delay: int = 2, lpc_gamma: float = 1.0):
        self.L = load_library()
        self._blob = np.frombuffer(blob, np.uint8)
        self.h = self.L.lpcnet_batch_create_multi(_p(self._blob), len(blob), n_streams, device_mask, features_delay,
                                                  C.c_float(lpc_gamma))
        if not self.h:
            raise LPCNetError(f"lpcnet_batch_create_multi: {last_error()}")
        self.n_streams = n_streams

    def close(self):
        if getattr(self, "h", None):
            self.L.lpcnet_multi_destroy(self.h)
            self.h = None

    def __del__(self):
        self.close()

    @property
    def shards(self):
        """[(device, first_stream, n_streams)]"""
        out = []
        for k in range(self.L.lpcnet_multi_n_devices(self.h)):
            d, f, n = C.c_int(0), C.c_int(0), C.c_int(0)
            _check(self.L.lpcnet_multi_shard(self.h, k, C.byref(d), C.byref(f), C.byref(n)), "lpcnet_multi_shard")
            out.append((d.value, f.value, n.value))
        return out

    def set_seed(self, stream: int, seed: bytes):
        _check(self.L.lpcnet_multi_set_seed(self.h, stream, seed, len(seed)), "lpcnet_multi_set_seed")

    def reset(self, stream: int = -1):
        _check(self.L.lpcnet_multi_reset(self.h, stream), "lpcnet_multi_reset")

    def set_codebooks(self, codebooks):
        cb = np.ascontiguousarray(codebooks, np.float32).reshape(-1)
        _check(self.L.lpcnet_multi_set_codebooks(self.h, _p(cb), cb.size), "lpcnet_multi_set_codebooks")

    def synthesize(self, features, pcm_out=None):
        f = np.ascontiguousarray(features, np.float32)
        assert f.ndim == 3 and f.shape[0] == self.n_streams, f.shape
        pcm = pcm_out if pcm_out is not None else np.zeros((self.n_streams, f.shape[1] * FRAME_SIZE), np.int16)
        _check(self.L.lpcnet_multi_synthesize(self.h, _p(f), f.shape[2], f.shape[1], _p(pcm)), "lpcnet_multi_synthesize")
        return pcm

    def decode(self, packets, pcm_out=None):
        pk = np.ascontiguousarray(packets, np.uint8)
        assert pk.ndim == 3 and pk.shape[0] == self.n_streams and pk.shape[2] == 8
        pcm = pcm_out if pcm_out is not None else np.zeros((self.n_streams, pk.shape[1] * 640), np.int16)
        _check(self.L.lpcnet_multi_decode(self.h, _p(pk), pk.shape[1], _p(pcm)), "lpcnet_multi_decode")
        return pcm


class Synth:
    """The reference's single-stream API: lpcnet_create + lpcnet_load_model + lpcnet_synthesize."""

    def __init__(self, blob: bytes):
        self.L = load_library()
        self.h = self.L.lpcnet_create()
        self._blob = np.frombuffer(blob, np.uint8)
        if self.L.lpcnet_load_model(self.h, _p(self._blob), len(blob)) != 0:
            err = last_error()
            self.L.lpcnet_destroy(self.h)
            self.h = None
            raise LPCNetError(f"lpcnet_load_model: {err}")

    def close(self):
        if getattr(self, "h", None):
            self.L.lpcnet_destroy(self.h)
            self.h = None

    def __del__(self):
        self.close()

    def reset(self):
        self.L.lpcnet_reset(self.h)

    def synthesize(self, features, n: int = FRAME_SIZE):
        f = np.ascontiguousarray(features, np.float32)
        assert f.size >= NB_FEATURES
        out = np.zeros(n, np.int16)
        last_error()      # (a void entry point: its failure shows as a new error text, so drop any stale one first)
        self.L.lpcnet_synthesize(self.h, _p(f), _p(out), n)
        err = last_error()
        if err:
            raise LPCNetError(f"lpcnet_synthesize: {err}")
        return out


class Decoder:
    """lpcnet_decoder_create + lpcnet_decode (include/lpcnet.h:67-96)."""

    def __init__(self, blob: bytes):
        self.L = load_library()
        self.h = self.L.lpcnet_decoder_create()
        self._blob = np.frombuffer(blob, np.uint8)
        if self.L.lpcnet_decoder_load_model(self.h, _p(self._blob), len(blob)) != 0:
            err = last_error()
            self.L.lpcnet_decoder_destroy(self.h)
            self.h = None
            raise LPCNetError(f"lpcnet_decoder_load_model: {err}")

    def close(self):
        if getattr(self, "h", None):
            self.L.lpcnet_decoder_destroy(self.h)
            self.h = None

    def __del__(self):
        self.close()

    def decode(self, packet: bytes):
        assert len(packet) == 8
        pcm = np.zeros(640, np.int16)
        buf = np.frombuffer(packet, np.uint8)
        _check(self.L.lpcnet_decode(self.h, _p(buf), _p(pcm)), "lpcnet_decode")
        return pcm


class FeatureBatch:
    """Batched per-frame feature analysis (lpcnet_compute_single_frame_features for n_streams states at once)."""

    def __init__(self, n_streams: int, device: int = 0):
        self.L = load_library()
        self.n_streams = n_streams
        self.h = self.L.lpcnet_feature_batch_create(n_streams, device)
        if not self.h:
            raise LPCNetError(f"lpcnet_feature_batch_create: {last_error()}")

    def close(self):
        if getattr(self, "h", None):
            self.L.lpcnet_feature_batch_destroy(self.h)
            self.h = None

    def __del__(self):
        self.close()

    def reset(self, stream: int = -1):
        _check(self.L.lpcnet_feature_batch_reset(self.h, stream), "lpcnet_feature_batch_reset")

    @property
    def launch_count(self):
        return int(self.L.lpcnet_feature_batch_launch_count(self.h))

    def compute(self, pcm):
        """pcm: [n_streams][n_frames*160] int16 or float32 -> features [n_streams][n_frames][36] float32."""
        pcm = np.asarray(pcm)
        assert pcm.ndim == 2 and pcm.shape[0] == self.n_streams and pcm.shape[1] % FRAME_SIZE == 0
        n_frames = pcm.shape[1] // FRAME_SIZE
        out = np.zeros((self.n_streams, n_frames, NB_TOTAL_FEATURES), np.float32)
        if pcm.dtype == np.int16:
            x = np.ascontiguousarray(pcm)
            _check(self.L.lpcnet_feature_batch_compute(self.h, _p(x), n_frames, _p(out)), "lpcnet_feature_batch_compute")
        else:
            x = np.ascontiguousarray(pcm, np.float32)
            _check(self.L.lpcnet_feature_batch_compute_float(self.h, _p(x), n_frames, _p(out)),
                   "lpcnet_feature_batch_compute_float")
        return out

    def set_codebooks(self, codebooks):
        cb = np.ascontiguousarray(codebooks, np.float32).reshape(-1)
        _check(self.L.lpcnet_feature_batch_set_codebooks(self.h, _p(cb), cb.size), "lpcnet_feature_batch_set_codebooks")

    def encode(self, pcm, want_features=False):
        """lpcnet_encode for every stream: pcm [n_streams][n_packets*640] int16 -> packets [n_streams][n_packets][8]
        uint8 (and the encoder's quantised feature rows [n_streams][n_packets*4][36] if asked)."""
        x = np.ascontiguousarray(pcm, np.int16)
        assert x.ndim == 2 and x.shape[0] == self.n_streams and x.shape[1] % PACKET_SAMPLES == 0
        n = x.shape[1] // PACKET_SAMPLES
        pk = np.zeros((self.n_streams, n, 8), np.uint8)
        ft = np.zeros((self.n_streams, n * 4, NB_TOTAL_FEATURES), np.float32) if want_features else None
        _check(self.L.lpcnet_feature_batch_encode(self.h, _p(x), n, _p(pk), _p(ft) if want_features else None),
               "lpcnet_feature_batch_encode")
        return (pk, ft) if want_features else pk

    def compute_features(self, pcm):
        """lpcnet_compute_features for every stream: pcm [n_streams][n_packets*640] int16 -> [n_streams][n_packets*4][36]."""
        x = np.ascontiguousarray(pcm, np.int16)
        assert x.ndim == 2 and x.shape[0] == self.n_streams and x.shape[1] % PACKET_SAMPLES == 0
        n = x.shape[1] // PACKET_SAMPLES
        ft = np.zeros((self.n_streams, n * 4, NB_TOTAL_FEATURES), np.float32)
        _check(self.L.lpcnet_feature_batch_compute_features(self.h, _p(x), n, _p(ft)), "lpcnet_feature_batch_compute_features")
        return ft

    def compute_device(self, d_pcm16: int, d_pcmf: int, n_frames: int, d_features: int, cuda_stream: int = 0):
        _check(self.L.lpcnet_feature_batch_compute_device(self.h, d_pcm16 or None, d_pcmf or None, n_frames,
                                                          d_features, cuda_stream or None),
               "lpcnet_feature_batch_compute_device")


class Encoder:
    """lpcnet_encoder_create + lpcnet_compute_single_frame_features[_float] (include/lpcnet.h:98-158)."""

    def __init__(self):
        self.L = load_library()
        self.h = self.L.lpcnet_encoder_create()
        if not self.h:
            raise LPCNetError("lpcnet_encoder_create failed")

    def close(self):
        if getattr(self, "h", None):
            self.L.lpcnet_encoder_destroy(self.h)
            self.h = None

    def __del__(self):
        self.close()

    def single_frame_features(self, pcm):
        pcm = np.asarray(pcm)
        assert pcm.size == FRAME_SIZE
        out = np.zeros(NB_TOTAL_FEATURES, np.float32)
        if pcm.dtype == np.int16:
            x = np.ascontiguousarray(pcm)
            _check(self.L.lpcnet_compute_single_frame_features(self.h, _p(x), _p(out)),
                   "lpcnet_compute_single_frame_features")
        else:
            x = np.ascontiguousarray(pcm, np.float32)
            _check(self.L.lpcnet_compute_single_frame_features_float(self.h, _p(x), _p(out)),
                   "lpcnet_compute_single_frame_features_float")
        return out


    def encode(self, pcm):
        """lpcnet_encode (include/lpcnet.h:113-120): 640 int16 samples -> 8 bytes"""
        x = np.ascontiguousarray(pcm, np.int16)
        assert x.size == PACKET_SAMPLES
        buf = np.zeros(8, np.uint8)
        _check(self.L.lpcnet_encode(self.h, _p(x), _p(buf)), "lpcnet_encode")
        return buf

    def compute_features(self, pcm):
        """lpcnet_compute_features (include/lpcnet.h:122-131): 640 int16 samples -> [4][36]"""
        x = np.ascontiguousarray(pcm, np.int16)
        assert x.size == PACKET_SAMPLES
        out = np.zeros((4, NB_TOTAL_FEATURES), np.float32)
        _check(self.L.lpcnet_compute_features(self.h, _p(x), _p(out)), "lpcnet_compute_features")
        return out


def set_codebooks(codebooks) -> None:
    """Process-wide VQ tables for encoder states created afterwards (lpcnet_b200_set_codebooks); None clears them."""
    L = load_library()
    if codebooks is None:
        _check(L.lpcnet_b200_set_codebooks(None, 0), "lpcnet_b200_set_codebooks")
        return
    cb = np.ascontiguousarray(codebooks, np.float32).reshape(-1)
    _check(L.lpcnet_b200_set_codebooks(_p(cb), cb.size), "lpcnet_b200_set_codebooks")


PLC_CAUSAL = 0
PLC_NONCAUSAL = 1
PLC_CODEC = 2
PLC_DC_FILTER = 4


class PLCBatch:
    """Batched packet-loss concealment (non-causal mode): n_streams PLC states, each with its own loss pattern."""

    def __init__(self, blob: bytes, n_streams: int, options: int = PLC_NONCAUSAL, device: int = 0,
                 lpc_gamma: float = 1.0):
        self.L = load_library()
        self.n_streams = n_streams
        self._blob = np.frombuffer(blob, np.uint8)
        self.h = self.L.lpcnet_plc_batch_create(_p(self._blob), len(blob), n_streams, device, options, lpc_gamma)
        if not self.h:
            raise LPCNetError(f"lpcnet_plc_batch_create: {last_error()}")

    def close(self):
        if getattr(self, "h", None):
            self.L.lpcnet_plc_batch_destroy(self.h)
            self.h = None

    def __del__(self):
        self.close()

    def set_seed(self, stream: int, seed: bytes):
        _check(self.L.lpcnet_plc_batch_set_seed(self.h, stream, seed, len(seed)), "lpcnet_plc_batch_set_seed")

    @property
    def launch_count(self):
        return int(self.L.lpcnet_plc_batch_launch_count(self.h))

    def fec_add(self, stream: int, features):
        """lpcnet_plc_fec_add for one stream (features: 20 floats, or None = skip one)."""
        f = None if features is None else np.ascontiguousarray(features, np.float32)
        _check(self.L.lpcnet_plc_batch_fec_add(self.h, stream, None if f is None else _p(f)), "lpcnet_plc_batch_fec_add")

    def fec_clear(self, stream: int):
        _check(self.L.lpcnet_plc_batch_fec_clear(self.h, stream), "lpcnet_plc_batch_fec_clear")

    def reset(self):
        """lpcnet_plc_reset for every stream (seeds return to the default "LPCNet")."""
        _check(self.L.lpcnet_plc_batch_reset(self.h), "lpcnet_plc_batch_reset")

    def peek(self, stream: int):
        """(features[20], gru1[256], gru2[256], [dc_mem, syn_dc]) of one stream."""
        f = np.zeros(20, np.float32); g1 = np.zeros(256, np.float32); g2 = np.zeros(256, np.float32)
        dc = np.zeros(2, np.float64)
        _check(self.L.lpcnet_plc_batch_peek(self.h, stream, _p(f), _p(g1), _p(g2), _p(dc)), "lpcnet_plc_batch_peek")
        return f, g1, g2, dc

    def peek_frame(self, stream: int):
        burg = np.zeros(36, np.float32); enc = np.zeros(36, np.float32)
        _check(self.L.lpcnet_plc_batch_peek_frame(self.h, stream, _p(burg), _p(enc)), "lpcnet_plc_batch_peek_frame")
        return burg, enc

    def process(self, pcm, lost):
        """pcm int16 [n_streams][n_frames*160], lost [n_streams][n_frames] -> concealed int16, same shape."""
        out = np.ascontiguousarray(pcm, np.int16).copy()
        lost = np.ascontiguousarray(lost, np.uint8)
        assert out.shape[0] == self.n_streams and lost.shape[0] == self.n_streams
        n_frames = lost.shape[1]
        assert out.shape[1] == n_frames * FRAME_SIZE
        out.reshape(self.n_streams, n_frames, FRAME_SIZE)[lost != 0] = 0     # a lost frame carries no input
        _check(self.L.lpcnet_plc_batch_process(self.h, _p(lost), _p(out), n_frames), "lpcnet_plc_batch_process")
        return out

    def process_inplace(self, pcm, lost):
        """lpcnet_plc_batch_process as the C-ABI has it: pcm int16 [n_streams][n_frames*160] is read and overwritten
        (lost frames are not read).  A pinned buffer avoids the staging copies of pageable memory."""
        lost = np.ascontiguousarray(lost, np.uint8)
        assert pcm.dtype == np.int16 and pcm.flags.c_contiguous and pcm.shape == (self.n_streams, lost.shape[1] * FRAME_SIZE)
        _check(self.L.lpcnet_plc_batch_process(self.h, _p(lost), _p(pcm), lost.shape[1]), "lpcnet_plc_batch_process")
        return pcm


class PLC:
    """lpcnet_plc_create + lpcnet_plc_load_model + lpcnet_plc_update / lpcnet_plc_conceal (include/lpcnet.h:190-215)."""

    def __init__(self, blob: bytes, options: int = PLC_NONCAUSAL):
        self.L = load_library()
        self.h = self.L.lpcnet_plc_create(options)
        self._blob = np.frombuffer(blob, np.uint8)
        if self.L.lpcnet_plc_load_model(self.h, _p(self._blob), len(blob)) != 0:
            err = last_error()
            self.L.lpcnet_plc_destroy(self.h)
            self.h = None
            raise LPCNetError(f"lpcnet_plc_load_model: {err}")

    def close(self):
        if getattr(self, "h", None):
            self.L.lpcnet_plc_destroy(self.h)
            self.h = None

    def __del__(self):
        self.close()

    def reset(self):
        self.L.lpcnet_plc_reset(self.h)

    def fec_add(self, features):
        """lpcnet_plc_fec_add: accepted; the non-causal mode never reads the queue (as in the reference)."""
        f = None if features is None else np.ascontiguousarray(features, np.float32)
        self.L.lpcnet_plc_fec_add(self.h, None if f is None else _p(f))

    def fec_clear(self):
        self.L.lpcnet_plc_fec_clear(self.h)

    def update(self, pcm):
        x = np.ascontiguousarray(pcm, np.int16).copy()
        _check(self.L.lpcnet_plc_update(self.h, _p(x)), "lpcnet_plc_update")
        return x

    def conceal(self):
        x = np.zeros(FRAME_SIZE, np.int16)
        _check(self.L.lpcnet_plc_conceal(self.h, _p(x)), "lpcnet_plc_conceal")
        return x


class DredDecoder:
    """RDOVAE decoder of DRED (DRED_rdovae_decode_all): latent vectors -> FEC feature vectors."""

    def __init__(self, blob: bytes, device: int = 0):
        self.L = load_library()
        self._blob = np.frombuffer(blob, np.uint8)
        self.h = self.L.lpcnet_dred_decoder_create(_p(self._blob), len(blob), device)
        if not self.h:
            raise LPCNetError(f"lpcnet_dred_decoder_create: {last_error()}")

    def close(self):
        if getattr(self, "h", None):
            self.L.lpcnet_dred_decoder_destroy(self.h)
            self.h = None

    def __del__(self):
        self.close()

    @property
    def launch_count(self):
        return int(self.L.lpcnet_dred_decoder_launch_count(self.h))

    def decode_all(self, state, latents):
        """state [24], latents [n][80] -> features [n][80] (four 20-float vectors per latent, reverse order)"""
        z = np.ascontiguousarray(latents, np.float32)
        st = np.ascontiguousarray(state, np.float32)
        assert st.size == 24 and z.ndim == 2 and z.shape[1] == 80
        out = np.zeros((z.shape[0], 80), np.float32)
        _check(self.L.lpcnet_dred_decode_all(self.h, _p(out), _p(st), _p(z), z.shape[0]), "lpcnet_dred_decode_all")
        return out

    def decode_batch(self, states, latents):
        """states [B][24], latents [B][n][80] -> features [B][n][80]"""
        z = np.ascontiguousarray(latents, np.float32)
        st = np.ascontiguousarray(states, np.float32)
        assert z.ndim == 3 and z.shape[2] == 80 and st.shape == (z.shape[0], 24)
        out = np.zeros(z.shape[:2] + (80,), np.float32)
        _check(self.L.lpcnet_dred_decode_batch(self.h, z.shape[0], _p(out), _p(st), _p(z), z.shape[1]), "lpcnet_dred_decode_batch")
        return out
