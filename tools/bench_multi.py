"""Throughput of the in-library multi-GPU handle (lpcnet_batch_create_multi, include/lpcnet_b200.h): one process, one
feeder thread per GPU, HOST (pinned) buffers in and out, copies inside the timed region -- the same measurement as
bench.py's `e2e`, for an integrator that cannot run one process per GPU.  Weak scaling: --streams per GPU.  Wall clock
around the synchronous C call (it returns when every GPU's PCM is in the caller's buffer).  One JSON line.

    python tools/bench_multi.py [--streams 1024] [--frames 32] [--calls 20] [--mask 0]
"""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

from lpcnet_b200 import blob, engine


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--streams", type=int, default=1024, help="streams per GPU")
    ap.add_argument("--frames", type=int, default=32)
    ap.add_argument("--calls", type=int, default=20)
    ap.add_argument("--mask", type=int, default=0)
    a = ap.parse_args()
    import torch
    n_dev = bin(a.mask).count("1") if a.mask else engine.device_count()
    B = a.streams * n_dev
    bl = blob.write_blob(blob.random_model())
    mb = engine.MultiBatch(bl, B, device_mask=a.mask)
    for s in range(B):
        mb.set_seed(s, blob.stream_seed(s))
    base = np.stack([blob.synthetic_features(a.frames, 100 + s) for s in range(64)])
    tf = torch.empty((B, a.frames, base.shape[2]), dtype=torch.float32, pin_memory=True)
    feats = tf.numpy()
    feats[...] = base[np.arange(B) % 64]
    tp = torch.empty((B, a.frames * 160), dtype=torch.int16, pin_memory=True)
    pcm = tp.numpy()
    for _ in range(3):
        mb.synthesize(feats, pcm_out=pcm)
    times = []
    for _ in range(a.calls):
        t0 = time.perf_counter()
        mb.synthesize(feats, pcm_out=pcm)
        times.append(time.perf_counter() - t0)
    med = float(np.median(times))
    n = B * a.frames * 160
    print(json.dumps({"workload": "lpcnet_multi_synthesize, host pinned buffers", "n_gpus": len(mb.shards), "shards": mb.shards,
                      "streams": B, "frames_per_call": a.frames, "calls": a.calls, "samples_per_s": n / med,
                      "samples_per_s_best": n / min(times), "ms_per_call": med * 1e3,
                      "h2d_bytes_per_call": int(feats.nbytes), "d2h_bytes_per_call": int(pcm.nbytes)}))
    mb.close()


if __name__ == "__main__":
    main()
