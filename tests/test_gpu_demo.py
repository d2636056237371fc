"""The drop-in proof: the reference's UNMODIFIED command-line program (src/lpcnet_demo.c, compiled with its own
-DUSE_WEIGHTS_FILE by oracle/Makefile into oracle/_ref/lpcnet_demo_b200) linked against liblpcnet_b200.so instead of
the reference library, run in every one of its modes, and its output files compared byte for byte with those of the
reference demo itself (tests/golden/demo.npz, from oracle/_ref/lpcnet_demo_parity_d*: same sources, pinned flags).

-decode never loads a model in the demo (src/lpcnet_demo.c:176-189: the reference has one compiled in); the library
takes ./weights_blob.bin as that compiled-in model (csrc/api.cu ensure_default_model)."""
import hashlib
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import make_golden  # noqa: E402

pytestmark = pytest.mark.gpu

DEMO = os.path.join(ROOT, "oracle", "_ref", "lpcnet_demo_b200")


@pytest.fixture(scope="module")
def golden():
    return np.load(os.path.join(ROOT, "tests", "golden", "demo.npz"))


@pytest.fixture(scope="module")
def weights(golden, engine_lib):
    b = make_golden.demo_blob()
    assert bytes(golden["blob_sha256"]) == hashlib.sha256(b).digest()
    return b


@pytest.mark.skipif(not os.path.exists(DEMO), reason="oracle/_ref/lpcnet_demo_b200 not built (make -C oracle ref needs /root/reference)")
@pytest.mark.parametrize("mode", ["synthesis", "decode", "features", "encode", "addlpc", "plc_noncausal_dc", "plc_causal_dc"])
def test_unmodified_reference_demo_on_this_library(golden, weights, mode):
    argv, files = make_golden.demo_inputs()[mode]
    out = np.frombuffer(make_golden.run_demo(DEMO, argv, files, {"weights_blob.bin": weights}), np.uint8)
    want = golden[mode]
    assert out.size == want.size, (out.size, want.size)
    if mode in ("synthesis", "decode") or mode.startswith("plc"):
        a, b = out.view(np.int16), want.view(np.int16)
        bad = np.nonzero(a != b)[0]
        assert bad.size == 0, f"{mode}: first divergence at sample {bad[0]}: {a[bad[0]]} vs {b[bad[0]]}"
    else:
        bad = np.nonzero(out != want)[0]
        assert bad.size == 0, f"{mode}: first differing byte {bad[0]}"
