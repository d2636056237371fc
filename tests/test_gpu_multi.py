"""lpcnet_batch_create_multi (include/lpcnet_b200.h): one handle over the GPUs of a box, contiguous stream shards, one
feeder thread per GPU.  Streams never interact (src/lpcnet.c:146-167 touches one LPCNetState), so the sharded result
must be the single-GPU result bit for bit, whatever the number of GPUs -- and that is the oracle's (test_gpu_parity)."""
import ctypes as C

import numpy as np
import pytest

from lpcnet_b200 import blob, engine, sharding

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def eng(engine_lib):
    assert engine.device_count() > 0, "no CUDA device: the gpu tests cannot run (there is no CPU fallback)"
    return engine


def seeds(n):
    return [bytes([s & 255, (s >> 8) & 255, 7]) for s in range(n)]


def single(eng, model_blob, feats):
    B = feats.shape[0]
    bt = eng.Batch(model_blob, B, device=0)
    for s, sd in enumerate(seeds(B)):
        bt.set_seed(s, sd)
    pcm, _, _ = bt.synthesize(feats)
    bt.close()
    return pcm


@pytest.mark.parametrize("mask", [1, 0])
def test_multi_equals_single(eng, model_blob, mask):
    B, F = 21, 6
    feats = np.stack([blob.synthetic_features(F, 900 + s) for s in range(B)])
    want = single(eng, model_blob, feats)
    mb = eng.MultiBatch(model_blob, B, device_mask=mask)
    shards = mb.shards
    world = 1 if mask == 1 else min(eng.device_count(), 32, B)
    assert len(shards) == world
    for k, (dev, first, n) in enumerate(shards):                    # the same split as the process-per-GPU path
        assert (first, first + n) == sharding.shard_range(B, k, world)
    assert sum(n for _, _, n in shards) == B
    for s, sd in enumerate(seeds(B)):
        mb.set_seed(s, sd)
    got = mb.synthesize(feats)
    assert np.array_equal(got, want)
    # a second call continues every stream's state; reset + reseed reproduces the first
    again = mb.synthesize(feats)
    assert not np.array_equal(again, got)
    mb.reset(-1)
    for s, sd in enumerate(seeds(B)):
        mb.set_seed(s, sd)
    assert np.array_equal(mb.synthesize(feats), want)
    mb.close()


def test_multi_decode_equals_single(eng, model_blob):
    B, P = 9, 2
    rng = np.random.default_rng(5)
    packets = rng.integers(0, 256, (B, P, 8), dtype=np.uint8)
    cb = blob.synthetic_codebooks()
    bt = eng.Batch(model_blob, B, device=0)
    bt.set_codebooks(cb)
    for s, sd in enumerate(seeds(B)):
        bt.set_seed(s, sd)
    want, _ = bt.decode(packets)
    bt.close()
    assert np.abs(want).max() > 0
    mb = eng.MultiBatch(model_blob, B, device_mask=0)
    mb.set_codebooks(cb)
    for s, sd in enumerate(seeds(B)):
        mb.set_seed(s, sd)
    assert np.array_equal(mb.decode(packets), want)
    mb.close()


def test_multi_errors(eng, model_blob):
    L = eng.load_library()
    b = np.frombuffer(model_blob, np.uint8)
    p = b.ctypes.data_as(C.c_void_p)
    n = eng.device_count()
    if n < 32:
        assert not L.lpcnet_batch_create_multi(p, len(model_blob), 4, 1 << n, 2, C.c_float(1.0))
        assert "device_mask" in eng.last_error()
    assert not L.lpcnet_batch_create_multi(p, len(model_blob), 0, 0, 2, C.c_float(1.0))
    assert not L.lpcnet_batch_create_multi(p, 100, 4, 1, 2, C.c_float(1.0))      # truncated blob: the GPU is named
    assert "GPU 0" in eng.last_error()
    mb = eng.MultiBatch(model_blob, 3, device_mask=1)
    assert L.lpcnet_multi_set_seed(mb.h, 3, b"x", 1) == -1
    assert L.lpcnet_multi_shard(mb.h, 1, None, None, None) == -1
    assert "no such shard" in eng.last_error()
    mb.close()
