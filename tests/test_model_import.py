"""SURVEY.md 8 f4, second half: real models reach the engine through generated C (`nnet_data.c`, `plc_data.c` of a
model tarball) -> tools/write_weights_shim.c -> weights_blob.bin.  The tarball is not in /root/reference, so the
generated files are produced here by tools/gen_nnet_data.py (the exporters' format) from the random-init model; the
shim's blob must equal lpcnet_b200.blob.write_blob byte for byte, and so must the blob written by the reference's own
exporter src/write_lpcnet_weights.c compiled unmodified over the same files."""
import os
import subprocess

import pytest

from lpcnet_b200 import blob

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"
SHIM = os.path.join(ROOT, "oracle", "_ref", "write_weights_shim")
MODEL = os.path.join(ROOT, "oracle", "_ref", "model_gen")


def expected_blob():
    A = blob.random_model()
    A.update(blob.random_plc_model())
    return blob.write_blob(A)


@pytest.mark.skipif(not os.path.exists(SHIM), reason="oracle/_ref/write_weights_shim not built (make -C oracle ref)")
def test_shim_writes_the_blob(tmp_path):
    out = tmp_path / "weights_blob.bin"
    subprocess.run([SHIM, str(out)], check=True, capture_output=True)
    assert out.read_bytes() == expected_blob()


@pytest.mark.skipif(not (os.path.isdir(REF + "/src") and os.path.exists(os.path.join(MODEL, "nnet_data.c"))),
                    reason="needs /root/reference and the generated model files")
def test_reference_exporter_agrees(tmp_path):
    """src/write_lpcnet_weights.c, unmodified, over the generated files (+ empty tables for the two DRED files it also
    includes): the same bytes."""
    for name, table in (("dred_rdovae_enc_data.c", "rdovae_enc_arrays"), ("dred_rdovae_dec_data.c", "rdovae_dec_arrays")):
        (tmp_path / name).write_text(f'#include "nnet.h"\nconst WeightArray {table}[] = {{{{NULL, 0, 0, NULL}}}};\n')
    exe = tmp_path / "dump_weights_blob"
    subprocess.run(["gcc", "-O1", "-w", "-DDOT_PROD", "-DDUMP_BINARY_WEIGHTS", "-mavx2", "-mfma", f"-I{tmp_path}", f"-I{MODEL}", f"-I{ROOT}/oracle/gen",
                    f"-I{REF}/src", f"-I{REF}/include", f"{REF}/src/write_lpcnet_weights.c", "-o", str(exe)], check=True)
    subprocess.run([str(exe)], cwd=tmp_path, check=True)
    assert (tmp_path / "weights_blob.bin").read_bytes() == expected_blob()


def test_generated_file_shape(tmp_path):
    """gen_nnet_data.py writes what dump_lpcnet.py:56-82, 478-493 writes: guarded arrays, the name table, the init."""
    import sys
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import gen_nnet_data
    import numpy as np
    arrays = {"a_weights": np.array([0.1, -2.5e-7, 3.0], np.float32), "a_idx": np.array([1, 2], np.int32),
              "a_q": np.array([-128, 127], np.int8)}
    gen_nnet_data.write_data_file(str(tmp_path / "x.c"), arrays, "lpcnet_arrays", "nnet_data.h", "int init_x(void) { return 0; }\n")
    src = (tmp_path / "x.c").read_text()
    assert "static const float a_weights[3] = {\n   0.1, -2.5e-07, 3.0\n};" in src
    assert "#define WEIGHTS_a_q_TYPE WEIGHT_TYPE_qweight" in src and "static const qweight a_q[2]" in src
    assert '  {"a_idx", WEIGHTS_a_idx_TYPE, sizeof(a_idx), a_idx},' in src and "{NULL, 0, 0, NULL}" in src
