import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def orclib():
    """The CPU oracle (oracle/liblpcnet_oracle.so), built on demand with gcc."""
    from oracle import orcpy
    if not orcpy.available():
        subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "oracle"), "oracle"], check=True)
    return orcpy.OrcLib()


@pytest.fixture(scope="session")
def model_blob():
    from lpcnet_b200 import blob
    return blob.write_blob(blob.random_model())


@pytest.fixture(scope="session")
def orc_model(orclib, model_blob):
    from oracle import orcpy
    return orcpy.OrcModel(orclib, model_blob, 2, 1.0, None)


@pytest.fixture(scope="session")
def golden_prims():
    return np.load(os.path.join(GOLDEN, "prims.npz"))


@pytest.fixture(scope="session")
def golden_synth():
    return np.load(os.path.join(GOLDEN, "synth.npz"))


@pytest.fixture(scope="session")
def golden_decode():
    return np.load(os.path.join(GOLDEN, "decode.npz"))


@pytest.fixture(scope="session")
def codebooks():
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import gen_codebooks
    return np.concatenate(gen_codebooks.codebooks())


@pytest.fixture(scope="session")
def engine_lib():
    """liblpcnet_b200.so, built in-tree with nvcc if stale."""
    from lpcnet_b200 import build, engine
    build.build()
    return engine.load_library()
