"""The C-ABI library builds, loads and exports every symbol include/lpcnet_b200.h declares.
No compute is issued here (this file runs without a GPU)."""
import ctypes
import os
import re

import numpy as np

from lpcnet_b200 import engine

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "lpcnet_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return re.findall(r"LPCNET_B200_EXPORT[^;(]*?\b(lpcnet_\w+)\s*\(", text)


def test_header_and_python_mirror_agree():
    syms = declared_symbols()
    assert len(syms) >= 30
    assert sorted(syms) == sorted(engine.EXPORTED_SYMBOLS)


def test_library_exports_every_declared_symbol(engine_lib):
    for name in declared_symbols():
        assert hasattr(engine_lib, name), name


def test_struct_sizes_match_header(engine_lib):
    # LPCNetB200StreamState / LPCNetB200Trace layouts mirrored as numpy dtypes
    assert engine.STATE_DTYPE.itemsize == (168 + 256 + 384 + 16 + 16 + 16 + 32 + 1152 + 48 + 1) * 4 + 8 + 16
    assert engine.TRACE_DTYPE.itemsize == 32
    assert engine_lib.lpcnet_get_size() > 0 and engine_lib.lpcnet_decoder_get_size() >= engine_lib.lpcnet_get_size()


def test_rcp_table_export_matches_oracle_model(engine_lib, orclib):
    assert np.array_equal(engine.default_rcp_table(), orclib.default_rcp_table())


def test_fails_loudly_without_gpu_or_on_bad_blob(engine_lib, model_blob):
    if engine.device_count() == 0:
        # no CPU fallback: creation must fail with a message, not silently compute elsewhere
        try:
            engine.Batch(model_blob, 4)
            raise AssertionError("Batch() succeeded without a GPU")
        except engine.LPCNetError as e:
            assert "no CUDA device" in str(e)
        st = engine_lib.lpcnet_create()
        arr = np.frombuffer(model_blob, np.uint8)
        assert engine_lib.lpcnet_load_model(st, arr.ctypes.data_as(ctypes.c_void_p), len(model_blob)) == -1
        out = np.zeros(160, np.int16)
        f = np.zeros(36, np.float32)
        engine_lib.lpcnet_synthesize(st, f.ctypes.data_as(ctypes.c_void_p), out.ctypes.data_as(ctypes.c_void_p), 160)
        assert b"no model loaded" in engine_lib.lpcnet_b200_last_error()
        engine_lib.lpcnet_destroy(st)
    else:
        try:
            engine.Batch(model_blob[:1000], 4)
            raise AssertionError("truncated blob accepted")
        except engine.LPCNetError:
            pass
