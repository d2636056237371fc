"""Where the host-buffer call loses time against the device-pointer call (bench.py e2e vs value): wall time of a
1024-stream x 1000-frame call against the sum of its sample-kernel launches, for a few variants of the buffers."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np
import torch
from lpcnet_b200 import blob, engine

B, F = 1024, 1000
bl = blob.write_blob(blob.random_model())
base = np.stack([blob.synthetic_features(F, s) for s in range(64)])
hf = torch.empty((B, F, 36), dtype=torch.float32, pin_memory=True); hf.numpy()[...] = base[np.arange(B) % 64]
hp = torch.empty((B, F * 160), dtype=torch.int16, pin_memory=True)
df = hf.cuda(); dp = torch.empty((B, F * 160), dtype=torch.int16, device="cuda")
bt = engine.Batch(bl, B)
bt.set_profiling(True)
def host(): bt.synthesize(hf.numpy(), pcm_out=hp.numpy())
def dev():
    bt.synthesize_device(df.data_ptr(), 36, F, dp.data_ptr()); bt.L.lpcnet_batch_sync(bt.h)
for name, fn in (("device pointers", dev), ("host pinned", host), ("device pointers", dev), ("host pinned", host)):
    fn(); bt.read_profile()
    t0 = time.perf_counter(); fn(); dt = (time.perf_counter() - t0) * 1e3
    sm, sn, fm, fn_ = bt.read_profile()
    print(f"{name:16s} wall {dt:8.2f} ms | sample kernels {sm:8.2f} ms in {sn} launches | frame network {fm:7.2f} ms | wall - sample {dt - sm:6.2f} ms")
