"""The C-ABI library builds, loads and exports every symbol include/lpcnet_b200.h declares.
No compute is issued here (this file runs without a GPU)."""
import ctypes
import os
import re

import numpy as np
import pytest

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


def test_links_as_liblpcnet(engine_lib, tmp_path):
    """The reference installs `liblpcnet` + `lpcnet.pc` (Makefile.am:9, lpcnet.pc.in); the build leaves the same names
    beside the library so that a program's `-llpcnet` resolves to it.  A C program that uses only the reference's API
    (declared here as include/lpcnet.h:160-215 declares it) links that way and runs -- no compute: get_size only, and
    create + destroy, which need no GPU."""
    import shutil
    import subprocess
    here = os.path.dirname(engine.LIB_PATH)
    assert os.path.islink(os.path.join(here, "liblpcnet.so")) and os.path.exists(os.path.join(here, "lpcnet.pc"))
    if shutil.which("gcc") is None:
        pytest.skip("no gcc")
    src = tmp_path / "user.c"
    src.write_text(
        "#include <stdio.h>\n"
        "typedef struct LPCNetState LPCNetState;\n"
        "int lpcnet_get_size(void); LPCNetState *lpcnet_create(void); void lpcnet_destroy(LPCNetState *st);\n"
        "int lpcnet_plc_get_size(void); int lpcnet_encoder_get_size(void); int lpcnet_decoder_get_size(void);\n"
        "int main(void) { LPCNetState *st = lpcnet_create(); printf(\"%d %d %d %d %d\\n\", lpcnet_get_size(),\n"
        "  lpcnet_decoder_get_size(), lpcnet_encoder_get_size(), lpcnet_plc_get_size(), st != NULL); lpcnet_destroy(st); return 0; }\n")
    exe = tmp_path / "user"
    libs = open(os.path.join(here, "lpcnet.pc")).read()
    assert "-llpcnet" in libs
    subprocess.run(["gcc", "-o", str(exe), str(src), f"-L{here}", "-llpcnet", f"-Wl,-rpath,{here}"], check=True)
    out = subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout.split()
    assert all(int(v) > 0 for v in out), out
