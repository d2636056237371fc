"""Throughput of the rows either side of the synthesis path, through the C-ABI with HOST buffers (host<->device copies
inside the timed region): packet-loss concealment (BASELINE config 4: 4096 streams, noncausal_dc, Gilbert losses),
feature analysis (lpcnet_compute_single_frame_features batched), packet decode + synthesis (config 3) and -- when the
library exports it -- the 1.6 kb/s encoder.  One JSON line per workload.  Not the bench contract (bench.py is); the
numbers land in DESIGN.md and profiles/.

    python tools/bench_aux.py [plc|features|decode|encode ...] [--streams N] [--frames F]
"""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

from lpcnet_b200 import blob, engine


def pinned(a):
    """a copy of `a` in page-locked host memory (the C-ABI takes plain host pointers; DMA from pageable memory is staged)"""
    import torch
    t = torch.empty(a.shape, dtype=torch.from_numpy(a[:0].copy()).dtype, pin_memory=True)
    out = t.numpy()
    out[...] = a
    out_holder.append(t)
    return out


out_holder = []


def timed(fn, reps):
    fn()                                    # warm-up (allocations, first-launch overheads)
    best = []
    for _ in range(reps):
        t0 = time.perf_counter()
        fn()
        best.append(time.perf_counter() - t0)
    return float(np.median(best)), float(min(best))


def bench_plc(B, F, reps, options=5):
    A = blob.random_model()
    A.update(blob.random_plc_model())
    bl = blob.write_blob(A)
    base = [blob.synthetic_pcm(F, 100 + s) for s in range(64)]
    pcm = np.stack([base[s % 64] for s in range(B)])
    lost = np.stack([blob.synthetic_loss_pattern(F, 7 + s, 0.6, 0.1) for s in range(B)])
    pb = engine.PLCBatch(bl, B, options)
    for s in range(B):
        pb.set_seed(s, blob.stream_seed(s))
    n0 = pb.launch_count
    buf = pinned(pcm)                      # in place, as the C-ABI works; the concealed output of one call is the next one's input
    med, mn = timed(lambda: pb.process_inplace(buf, lost), reps)
    launches = (pb.launch_count - n0) // (reps + 1)
    pb.close()
    return {"workload": {5: "plc noncausal_dc", 4: "plc causal_dc", 6: "plc codec_dc"}.get(options, f"plc options {options}"), "pinned_inplace": True, "streams": B, "frames": F, "loss_rate": float(lost.mean()),
            "samples_per_s": B * F * 160 / med, "x_realtime": B * F * 160 / med / 16000, "ms_per_frame": med / F * 1e3,
            "best_s": mn, "median_s": med, "launches_per_call": int(launches)}


def bench_features(B, F, reps):
    base = [blob.synthetic_pcm(F, 100 + s) for s in range(64)]
    pcm = pinned(np.stack([base[s % 64] for s in range(B)]))
    fb = engine.FeatureBatch(B)
    n0 = fb.launch_count
    med, mn = timed(lambda: fb.compute(pcm), reps)
    launches = (fb.launch_count - n0) // (reps + 1)
    fb.close()
    return {"workload": "feature analysis", "pinned_input": True, "streams": B, "frames": F, "samples_per_s": B * F * 160 / med,
            "x_realtime": B * F * 160 / med / 16000, "frames_per_s": B * F / med, "best_s": mn, "median_s": med,
            "launches_per_call": int(launches)}


def bench_decode(B, P, reps):
    bl = blob.write_blob(blob.random_model())
    rng = np.random.default_rng(5)
    packets = rng.integers(0, 256, size=(B, P, 8), dtype=np.uint8)
    bt = engine.Batch(bl, B)
    bt.set_codebooks(blob.synthetic_codebooks())
    import torch
    hold = torch.empty((B, P * 640), dtype=torch.int16, pin_memory=True)     # pinned PCM destination
    out = hold.numpy()
    med, mn = timed(lambda: bt.decode(packets, pcm_out=out), reps)
    bt.close()
    return {"workload": "decode 1.6 kb/s + synthesis", "streams": B, "packets": P, "pinned_output": True,
            "samples_per_s": B * P * 640 / med,
            "x_realtime": B * P * 640 / med / 16000, "best_s": mn, "median_s": med}


def bench_encode(B, P, reps):
    base = [blob.synthetic_pcm(P * 4, 100 + s) for s in range(64)]
    pcm = pinned(np.stack([base[s % 64] for s in range(B)]))
    eb = engine.FeatureBatch(B)
    eb.set_codebooks(blob.synthetic_codebooks())
    n0 = eb.launch_count
    med, mn = timed(lambda: eb.encode(pcm), reps)
    launches = (eb.launch_count - n0) // (reps + 1)
    eb.close()
    return {"workload": "encode 1.6 kb/s", "pinned_input": True, "streams": B, "packets": P, "samples_per_s": B * P * 640 / med,
            "x_realtime": B * P * 640 / med / 16000, "packets_per_s": B * P / med, "best_s": mn, "median_s": med,
            "launches_per_call": int(launches)}


def bench_dred(B, n, reps):
    A = blob.random_model()
    A.update(blob.random_dred_decoder_model())
    bl = blob.write_blob(A)
    rng = np.random.default_rng(3)
    states = rng.standard_normal((B, 24)).astype(np.float32)
    latents = rng.standard_normal((B, n, 80)).astype(np.float32)
    dec = engine.DredDecoder(bl)
    med, mn = timed(lambda: dec.decode_batch(states, latents), reps)
    dec.close()
    return {"workload": "dred rdovae decode", "packets": B, "latents_per_packet": n, "feature_vectors_per_s": B * n * 4 / med,
            "x_realtime": B * n * 4 / med / 100, "best_s": mn, "median_s": med}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("what", nargs="*", default=["plc", "features", "encode"])
    ap.add_argument("--streams", type=int, default=4096)
    ap.add_argument("--frames", type=int, default=100)
    ap.add_argument("--reps", type=int, default=3)
    a = ap.parse_args()
    if engine.device_count() == 0:
        raise SystemExit("no CUDA device (the engine has no CPU fallback)")
    for w in a.what:
        if w == "plc":
            r = bench_plc(a.streams, a.frames, a.reps)
        elif w == "plc_causal":
            r = bench_plc(a.streams, a.frames, a.reps, 4)
        elif w == "features":
            r = bench_features(a.streams, a.frames, a.reps)
        elif w == "decode":
            r = bench_decode(a.streams, max(1, a.frames // 4), a.reps)
        elif w == "dred":
            r = bench_dred(a.streams, 26, a.reps)
        elif w == "encode":
            r = bench_encode(a.streams, max(1, a.frames // 4), a.reps)
        else:
            raise SystemExit(f"unknown workload {w}")
        print(json.dumps(r), flush=True)


if __name__ == "__main__":
    main()
