"""Real-time streaming operation (north_star: ">= 10^5 concurrent real-time 16 kHz streams on one 8xB200 box" means
12 500 LIVE streams per GPU): one `lpcnet_batch_synthesize(n_frames = 1)` call per 10 ms frame for all streams of the
GPU -- the reference's own calling convention, one frame per call (src/lpcnet.c:279-281, include/lpcnet.h:188) -- with
HOST buffers (pinned features in, pinned PCM out, copies inside the timed call).  A stream count is sustainable when
the p99 wall time of a call stays under 10 ms.

    python tools/bench_stream.py [--streams 12500] [--calls 1000] [--sweep 12500,15000,17500,20000]
One JSON line per stream count; bench.py --mode stream wraps this and adds the clock record.
"""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

from lpcnet_b200 import blob, engine


def run(B, calls, device=0, warm=20, graph=True):
    import torch
    model_blob = blob.write_blob(blob.random_model())
    bt = engine.Batch(model_blob, B, device=device)
    for s in range(0, B, 61):
        bt.set_seed(s, blob.stream_seed(s))
    if hasattr(bt, "set_stream_mode"):
        bt.set_stream_mode(graph)
    tracks = np.stack([blob.synthetic_features(calls + warm, k) for k in range(64)])     # [64][F][36]
    hold_f = torch.empty((B, 36), dtype=torch.float32, pin_memory=True)
    hold_p = torch.empty((B, 160), dtype=torch.int16, pin_memory=True)
    hf, hp = hold_f.numpy(), hold_p.numpy()
    idx = np.arange(B) % 64
    l0 = bt.launch_count
    ts = []
    nz = 0
    for f in range(calls + warm):
        hf[...] = tracks[idx, f]                      # the frame's features arrive (not timed: the caller's work)
        t0 = time.perf_counter()
        bt.synthesize(hf[:, None, :], pcm_out=hp)     # H2D + frame network + 160 sample steps + D2H, synchronous
        dt = time.perf_counter() - t0
        if f >= warm:
            ts.append(dt)
            nz += int(np.count_nonzero(hp[:: max(1, B // 64)]))
    ts = np.sort(np.array(ts)) * 1e3
    out = {"workload": "real-time streaming: one lpcnet_batch_synthesize(n_frames=1) call per 10 ms frame, host buffers",
           "streams": B, "calls": calls, "tile_streams_per_cta": bt.tile,
           "ms_per_call_p50": float(ts[len(ts) // 2]), "ms_per_call_p99": float(ts[min(len(ts) - 1, int(0.99 * len(ts)))]),
           "ms_per_call_max": float(ts[-1]), "ms_per_call_mean": float(ts.mean()),
           "real_time": bool(ts[min(len(ts) - 1, int(0.99 * len(ts)))] < 10.0),
           "samples_per_s": B * 160 / (ts.mean() * 1e-3), "launches_per_call": (bt.launch_count - l0) / (calls + warm),
           "h2d_bytes_per_call": int(hf.nbytes), "d2h_bytes_per_call": int(hp.nbytes), "nonzero_checked": nz > 0}
    bt.close()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--streams", type=int, default=12500)
    ap.add_argument("--calls", type=int, default=1000)
    ap.add_argument("--sweep", default="")
    ap.add_argument("--device", type=int, default=0)
    ap.add_argument("--gpus", type=int, default=1, help="run the same stream count on N GPUs at once, one process each")
    a = ap.parse_args()
    if engine.device_count() == 0:
        raise SystemExit("no CUDA device (the engine has no CPU fallback)")
    if a.gpus > 1:
        # the box's real-time capacity: every GPU holds --streams live streams at the same time (processes are
        # independent, as the streams are); the figure that counts is the worst GPU's p99
        import subprocess
        procs = [subprocess.Popen([sys.executable, os.path.abspath(__file__), "--streams", str(a.streams), "--calls", str(a.calls),
                                   "--device", str(k)], stdout=subprocess.PIPE, text=True) for k in range(a.gpus)]
        per = [json.loads(p.communicate()[0].strip().splitlines()[-1]) for p in procs]
        print(json.dumps({"workload": per[0]["workload"], "n_gpus": a.gpus, "streams_per_gpu": a.streams,
                          "streams_total": a.streams * a.gpus, "calls": a.calls,
                          "ms_per_call_p50_worst_gpu": max(x["ms_per_call_p50"] for x in per),
                          "ms_per_call_p99_worst_gpu": max(x["ms_per_call_p99"] for x in per),
                          "ms_per_call_max_worst_gpu": max(x["ms_per_call_max"] for x in per),
                          "real_time": all(x["real_time"] for x in per),
                          "per_gpu_p99": [x["ms_per_call_p99"] for x in per]}), flush=True)
        return
    for B in ([int(x) for x in a.sweep.split(",")] if a.sweep else [a.streams]):
        print(json.dumps(run(B, a.calls, a.device)), flush=True)


if __name__ == "__main__":
    main()
