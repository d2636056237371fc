"""lpcnet_b200_lpc_from_cepstrum_noalloc (include/lpcnet_b200.h): the entry point that allocates nothing (static device
scratch of LPCNET_B200_NOALLOC_MAX_ROWS rows) -- no counterpart in the
reference.  Checked through the C-ABI against the existing entry point that computes the same thing,
lpcnet_batch_frame_network (with FEATURES_DELAY 0 and gamma 1 its LPC output is lpc_from_cepstrum of each frame), bit
for bit like the other parity tests, on two calls in a row (the second on other data, so that what the first left in
the shared scratch would show).  Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

from lpcnet_b200 import blob, engine

pytestmark = pytest.mark.gpu


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


@pytest.fixture(scope="module")
def eng(engine_lib):
    assert engine.device_count() > 0, "no CUDA device: the gpu tests cannot run (there is no CPU fallback)"
    return engine


def test_noalloc_matches_frame_network_twice(eng, model_blob):
    import torch      # device buffers of the CALLER (the entry point itself allocates nothing)
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    B, F = 37, 9                       # 333 rows: a ragged last block
    n = B * F
    bt = eng.Batch(model_blob, B, device=0, features_delay=0, lpc_gamma=1.0)
    d_lpc = torch.empty((n, 16), dtype=torch.float32, device=dev)
    d_in = torch.empty((n, 36), dtype=torch.float32, device=dev)
    eng.lpc_from_cepstrum_noalloc(0, 18, 0, 0, device=0)          # set-up only (n_rows = 0)
    torch.cuda.synchronize()
    for call in range(2):
        feats = np.stack([blob.synthetic_features(F, 200 + 50 * call + s) for s in range(B)])
        if call == 1:
            feats[:, :, 0] += 1.5      # louder frames: other spectra than the first call's
        _, _, want = bt.frame_network(feats)                        # existing entry point, [B][F][16]
        d_in.copy_(torch.from_numpy(feats.reshape(n, 36)))
        d_lpc.fill_(float("nan"))
        torch.cuda.synchronize()
        free0 = torch.cuda.mem_get_info(dev)[0]
        eng.lpc_from_cepstrum_noalloc(d_in.data_ptr(), 36, d_lpc.data_ptr(), n, device=0)
        torch.cuda.synchronize()
        # (from the second call on: the first launch of a kernel may make the driver load its code, lazily)
        assert call == 0 or torch.cuda.mem_get_info(dev)[0] == free0, "device memory in use changed across the call"
        got = d_lpc.cpu().numpy().reshape(B, F, 16)
        assert np.isfinite(got).all()
        assert np.array_equal(bits(got), bits(want)), f"call {call}: differs from lpcnet_batch_frame_network"
    assert np.abs(want).max() > 0.1
    bt.close()


def test_noalloc_full_size_and_limit(eng, model_blob):
    import torch
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    n = eng.NOALLOC_MAX_ROWS
    rng = np.random.default_rng(3)
    ceps = (rng.standard_normal((n, 18)) * 0.5).astype(np.float32)
    ceps[:, 0] += 2
    d_in = torch.from_numpy(ceps).to(dev)
    d_lpc = torch.zeros((n, 16), dtype=torch.float32, device=dev)
    stream = torch.cuda.Stream(device=dev)
    stream.wait_stream(torch.cuda.current_stream(dev))
    eng.lpc_from_cepstrum_noalloc(d_in.data_ptr(), 18, d_lpc.data_ptr(), n, cuda_stream=stream.cuda_stream, device=0)
    stream.synchronize()
    got = d_lpc.cpu().numpy()
    # the same rows through the existing entry point, 64 streams x 64 frames
    bt = eng.Batch(model_blob, 64, device=0, features_delay=0, lpc_gamma=1.0)
    feats = np.zeros((64, 64, 20), np.float32)
    feats[:, :, :18] = ceps.reshape(64, 64, 18)
    _, _, want = bt.frame_network(feats)
    bt.close()
    assert np.array_equal(bits(got), bits(want.reshape(n, 16)))
    # one row more than the static scratch holds: a clear error, nothing launched
    L = eng.load_library()
    rc = L.lpcnet_b200_lpc_from_cepstrum_noalloc(0, C.c_void_p(d_in.data_ptr()), 18, C.c_void_p(d_lpc.data_ptr()), n + 1, None)
    assert rc == -1
    msg = eng.last_error()
    assert "LPCNET_B200_NOALLOC_MAX_ROWS" in msg and str(n) in msg
    assert L.lpcnet_b200_lpc_from_cepstrum_noalloc(0, C.c_void_p(d_in.data_ptr()), 17, C.c_void_p(d_lpc.data_ptr()), 4, None) == -1
    assert "stride" in eng.last_error()
