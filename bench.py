#!/usr/bin/env python
"""bench.py -- BASELINE.json's metric on BASELINE.json's config.

A "step" is one pass of the hot path over one batch: `--streams` (default 1024) independent
16 kHz streams x `--seconds` (default 10 s = 1000 frames) of synthetic features through the
batched synthesis engine (frame network + per-sample network), random-init weights in the
reference's blob layout.  That is BASELINE.json configs[1] ("batched -synthesis, 1024 streams
x 10 s on 1 B200").  Under torchrun (N>1) every rank runs the same number of streams on its own
GPU (weak scaling, no data-path collective: streams are independent).

  value : aggregate samples/s, inputs resident in HBM (device-pointer entry point)
  e2e   : the same through the host-buffer C-ABI call (pinned host features in, PCM out,
          copies inside the timed region)
  roofline / cpu_baseline / clocks / gpu_launches: see DESIGN.md "Measurement".

`--impl reference` times the reference's own CPU implementation (oracle/_ref/lpcnet_demo_*,
the unmodified sources built with the README's -Ofast flags) on all host cores.
"""
from __future__ import annotations

import argparse
import hashlib
import json
import os
import shutil
import statistics
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

FRAME = 160
METRIC = "aggregate 16 kHz samples/sec"
UNIT = "samples/s"


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"hbm_gbs": d["hbm_gbs"], "bf16_tflops": d["bf16_tflops"],
                "bf16_tflops_sustained": d.get("bf16_tflops_sustained", d["bf16_tflops"]), "src": "measured"}
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "src": "fallback"}


from tools.cpu_baseline import CpuReference, cpu_model  # noqa: E402  (the reference's lpcnet_demo on the host cores)


def measure_int8_peak(dev):
    """Dense int8 tensor throughput of this GPU (SURVEY.md 8d asks for a measurement, not the datasheet): an
    8192^3 s8 x s8 -> s32 GEMM through cuBLASLt (torch._int_mm), best of 10, CUDA events.  None if unavailable."""
    import torch
    try:
        n = 8192
        a = torch.randint(-127, 127, (n, n), dtype=torch.int8, device=dev)
        b = torch.randint(-127, 127, (n, n), dtype=torch.int8, device=dev)
        for _ in range(3):
            torch._int_mm(a, b)
        best = 1e9
        for _ in range(10):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            torch._int_mm(a, b)
            e1.record()
            e1.synchronize()
            best = min(best, e0.elapsed_time(e1))
        return 2.0 * n ** 3 / (best * 1e-3) / 1e12
    except Exception:
        return None


def golden_hashes():
    """SHA-256 of the unmodified reference's PCM for fixed streams of this workload, per call on the same states
    (tests/golden/bench_pcm_hashes.json, written by tests/golden/make_golden.py bench from oracle/_ref)."""
    try:
        return json.load(open(os.path.join(ROOT, "tests", "golden", "bench_pcm_hashes.json")))
    except OSError:
        return None


# ------------------------------------------------------------------------------------------
class ClockSampler:
    FIELDS = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
              "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
              "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.idx = gpu_index
        self.path = tempfile.mktemp(prefix="clocks_", suffix=".csv")
        self.proc = None

    def start(self):
        if shutil.which("nvidia-smi") is None:
            return
        self.f = open(self.path, "w")
        self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits",
                                      "-lms", "200", "-i", str(self.idx)], stdout=self.f, stderr=subprocess.DEVNULL)

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        if self.proc is None:
            return out
        self.proc.terminate()
        try:
            self.proc.wait(5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        self.f.close()
        sm, mx, reasons, power = [], [], set(), []
        for line in open(self.path):
            c = [x.strip() for x in line.split(",")]
            if len(c) < 9:
                continue
            try:
                sm.append(float(c[1])); mx.append(float(c[2])); power.append(float(c[3]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), c[5:9]):
                if v == "Active":
                    reasons.add(name)
        os.unlink(self.path)
        if sm:
            out = {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "reasons": sorted(reasons),
                   "power_w_max": max(power), "samples": len(sm)}
        return out


def run_reference(args):
    """--impl reference: the reference's CPU implementation on the host cores (rank 0 only)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from lpcnet_b200 import blob
    b = blob.write_blob(blob.random_model())
    n_frames = int(round(args.seconds * 100))
    ref = CpuReference("synthesis", n_frames, b)
    try:
        for _ in range(args.warmup):
            ref.run_once()
        t0 = time.perf_counter()
        runs = [ref.run_once() for _ in range(args.steps)]
        dt = time.perf_counter() - t0
        value = statistics.mean(r[0] for r in runs)
        per_core = [v for r in runs for v in r[1]]
        cores = len(ref.cores) if ref.binary else 1
        what = os.path.basename(ref.binary) if ref.binary else "oracle port (liblpcnet_oracle.so)"
        line = {
            "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8*s8->s32 + f32",
            "data": "synthetic", "x_realtime": value / 16000.0,
            "config": {"workload": f"lpcnet_demo -synthesis, {cores} streams x {args.seconds:g} s per step "
                                   f"(bounded sample of batched -synthesis, {args.streams} streams x {args.seconds:g} s)",
                       "model": "random-init, N=384 NB=16 density (0.05,0.05,0.2), reference blob layout",
                       "cpu": cpu_model()},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": ref.kind,
                             "per_core_min": min(per_core), "per_core_median": statistics.median(per_core),
                             "sample": f"{args.steps} x ({cores} concurrent processes, one per host core, each lpcnet_demo "
                                       f"-synthesis of one {args.seconds:g} s stream) with {what}", "cpu": cpu_model()},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        }
        print(json.dumps(line), flush=True)
    finally:
        ref.close()


def step_latency_percentiles(launch_ms, launch_steps):
    """Per-sample step latency of one resident stream tile, per sample-kernel launch (launch duration / steps in
    the launch; the two warm-up frames of a call run no steps and are excluded): [median, p99] in microseconds."""
    lat = [1e3 * float(m) / int(n) for m, n in zip(launch_ms, launch_steps) if int(n) >= 160 * 4]
    if not lat:
        return None
    lat.sort()
    return [lat[len(lat) // 2], lat[min(len(lat) - 1, int(0.99 * len(lat)))]]


def run_aux(args, dev, local_rank, model_blob):
    """BASELINE configs[2], [3] and a configs[4] point, measured in this process with the same clock sampling and each
    beside the reference's own CPU run of that mode (tools/cpu_baseline.py).  Short by design (a few seconds each):
    the headline's step count is untouched.  Host-buffer C-ABI calls (copies inside the timed region) throughout."""
    import torch
    from lpcnet_b200 import blob, engine
    out = []

    def pinned(shape, dtype):
        t = torch.empty(shape, dtype=dtype, pin_memory=True)
        return t, t.numpy()

    def timed(fn, reps):
        fn()
        torch.cuda.synchronize()
        ts = []
        clocks = ClockSampler(local_rank)
        clocks.start()
        for _ in range(reps):
            t0 = time.perf_counter()
            fn()                                   # every call ends with a stream synchronise inside the library
            ts.append(time.perf_counter() - t0)
        return statistics.median(ts), clocks.stop()

    def cpu(mode, n_frames, blob_bytes=None):
        if args.no_cpu_baseline:
            return None
        ref = CpuReference(mode, n_frames, blob_bytes)
        try:
            return ref.measure(1) if ref.binary else None
        finally:
            ref.close()

    # ---- configs[2]: batched -decode, 8192 streams (1.6 kb/s packets -> lpcnet_dec + synthesis) ----
    B, P = 8192, 25
    bt = engine.Batch(model_blob, B, device=local_rank)
    bt.set_codebooks(blob.synthetic_codebooks())
    packets = np.random.default_rng(5).integers(0, 256, size=(B, P, 8), dtype=np.uint8)
    hold, pcm = pinned((B, P * 640), torch.int16)
    l0 = bt.launch_count
    med, clk = timed(lambda: bt.decode(packets, pcm_out=pcm), 3)
    v = B * P * 640 / med
    out.append({"config": "configs[2]: batched -decode from synthetic 1.6 kb/s bitstreams, 8192 streams x 1 s",
                "metric": METRIC, "unit": UNIT, "value": v, "value_is_e2e": True, "x_realtime": v / 16000.0,
                "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": int(packets.nbytes), "d2h_bytes_per_step": int(pcm.nbytes)},
                "ms_per_call": med * 1e3, "gpu_launches": (bt.launch_count - l0) // 4, "clocks": clk,
                "cpu_baseline": cpu("decode", 1000)})
    bt.close()
    del hold

    # ---- configs[3]: PLC noncausal_dc, 4096 streams, Gilbert losses ----
    B, F = 4096, 100
    A = blob.random_model()
    A.update(blob.random_plc_model())
    pb = engine.PLCBatch(blob.write_blob(A), B, engine.PLC_NONCAUSAL | engine.PLC_DC_FILTER, device=local_rank)
    for s in range(B):
        pb.set_seed(s, blob.stream_seed(s))
    base = [blob.synthetic_pcm(F, 100 + s) for s in range(64)]
    lost = np.stack([blob.synthetic_loss_pattern(F, 7 + s, 0.6, 0.1) for s in range(B)])
    hold, buf = pinned((B, F * FRAME), torch.int16)
    for s in range(B):
        buf[s] = base[s % 64]
    l0 = pb.launch_count
    med, clk = timed(lambda: pb.process_inplace(buf, lost), 2)      # in place, as the C-ABI works
    v = B * F * FRAME / med
    out.append({"config": "configs[3]: PLC noncausal_dc batched with synthetic loss patterns, 4096 streams x 1 s",
                "metric": METRIC, "unit": UNIT, "value": v, "value_is_e2e": True, "x_realtime": v / 16000.0,
                "ms_per_10ms_frame": med / F * 1e3, "loss_rate": float(lost.mean()),
                "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": int(buf.nbytes + lost.nbytes), "d2h_bytes_per_step": int(buf.nbytes)},
                "ms_per_call": med * 1e3, "gpu_launches": (pb.launch_count - l0) // 3, "clocks": clk,
                "cpu_baseline": cpu("plc", 1000)})
    pb.close()
    del hold

    # ---- configs[4] point: 65536 streams on this GPU, free-running synthesis ----
    B, F = 65536, 100
    tracks = np.stack([blob.synthetic_features(F, k) for k in range(64)])
    hold_f, hf = pinned((B, F, 36), torch.float32)
    hf[...] = tracks[np.arange(B) % 64]
    hold_p, hp = pinned((B, F * FRAME), torch.int16)
    bt = engine.Batch(model_blob, B, device=local_rank)
    for s in range(0, B, 997):
        bt.set_seed(s, blob.stream_seed(s))
    d_feat = hold_f.to(dev)
    d_pcm = torch.zeros((B, F * FRAME), dtype=torch.int16, device=dev)
    l0 = bt.launch_count

    def dev_step():
        bt.synthesize_device(d_feat.data_ptr(), 36, F, d_pcm.data_ptr())
        bt.sync()
    med_d, clk = timed(dev_step, 2)
    launches = (bt.launch_count - l0) // 3
    med_h, _ = timed(lambda: bt.synthesize(hf, pcm_out=hp), 1)
    v, e = B * F * FRAME / med_d, B * F * FRAME / med_h
    out.append({"config": "configs[4] point: throughput sweep, 65536 streams x 1 s on one B200",
                "metric": METRIC, "unit": UNIT, "value": v, "x_realtime": v / 16000.0, "tile_streams_per_cta": bt.tile,
                "e2e": {"value": e, "unit": UNIT, "h2d_bytes_per_step": int(hf.nbytes), "d2h_bytes_per_step": int(hp.nbytes)},
                "ms_per_call": med_d * 1e3, "gpu_launches": launches, "clocks": clk,
                "cpu_baseline": "the headline's cpu_baseline (same mode: lpcnet_demo -synthesis per core)"})
    bt.close()

    # ---- north_star's target as stated: CONCURRENT REAL-TIME streams -- 12 500 per GPU (10^5 per 8-GPU box), one
    # lpcnet_batch_synthesize(n_frames = 1) call per 10 ms frame (the reference's own calling convention,
    # src/lpcnet.c:279-281), host buffers; sustainable when the p99 wall time of a call stays under 10 ms ----
    from tools.bench_stream import run as stream_run
    clocks = ClockSampler(local_rank)
    clocks.start()
    sr = stream_run(12500, 200, device=local_rank, warm=10)
    sr["clocks"] = clocks.stop()
    sr["config"] = "real-time streaming: 12500 live streams on one B200, one 10 ms frame per call"
    sr["metric"], sr["unit"], sr["value"] = "p99 wall time of one 10 ms frame of all streams", "ms", sr["ms_per_call_p99"]
    sr["higher_is_better"] = False
    sr["cpu_baseline"] = "one stream per host core in real time by definition when the headline's per_core_min >= 16000 samples/s"
    out.append(sr)
    return out


def run_gpu(args):
    import torch
    import torch.distributed as dist
    from lpcnet_b200 import blob, engine, sharding

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available() or engine.device_count() == 0:
        raise SystemExit("bench.py: no CUDA device; the engine has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    B = args.streams                       # per GPU (weak scaling)
    n_frames = int(round(args.seconds * 100))
    T = n_frames * FRAME
    stride = 36                            # the feature-file frame: 36 floats, 20 used (lpcnet_demo.c:213-215)
    model_blob = blob.write_blob(blob.random_model())
    lo, _ = sharding.shard_range(B * world, rank, world)

    # synthetic inputs: 64 distinct feature tracks cycled over the streams, per-stream kiss99 seeds
    tracks = np.stack([blob.synthetic_features(n_frames, k) for k in range(64)])
    h_feat = torch.empty((B, n_frames, stride), dtype=torch.float32, pin_memory=True)
    hf = h_feat.numpy()
    for s in range(B):
        hf[s] = tracks[(lo + s) % 64]
    h_pcm = torch.empty((B, T), dtype=torch.int16, pin_memory=True)
    d_feat = h_feat.to(dev)
    d_pcm = torch.zeros((B, T), dtype=torch.int16, device=dev)

    bt = engine.Batch(model_blob, B, device=local_rank)
    if args.tile:
        bt.set_tile(args.tile)
    for s in range(B):
        bt.set_seed(s, blob.stream_seed(lo + s))
    # a dedicated (non-default) stream: the library and the timing events share it
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)
    assert stream.cuda_stream != 0
    bt.set_cuda_stream(stream.cuda_stream)

    def step_device():
        bt.synthesize_device(d_feat.data_ptr(), stride, n_frames, d_pcm.data_ptr())

    def step_host():
        bt.synthesize(hf, pcm_out=h_pcm.numpy())

    # ---- parity of the timed run itself: SHA-256 of fixed streams' PCM against the unmodified reference's ----
    gold = golden_hashes() if (rank == 0 and n_frames == 1000 and not args.tile) else None
    parity = {"streams": 0, "ok": None, "checked_steps": [], "note": "no golden hashes for this workload"}
    if gold and gold["blob_sha256"] == hashlib.sha256(model_blob).hexdigest():
        gstreams = [int(k) for k in gold["steps"] if int(k) < B]
        parity = {"streams": len(gstreams), "ok": True, "checked_steps": [], "mismatches": [],
                  "against": "tests/golden/bench_pcm_hashes.json: " + gold["source"]}

    def check_parity(call_index, src):
        """call_index: how many engine calls these stream states have seen before this one"""
        if parity["ok"] is None or call_index >= len(next(iter(gold["steps"].values()))):
            return
        rows = src[gstreams].cpu().numpy() if hasattr(src, "cpu") else src[gstreams]
        for k, s_ in enumerate(gstreams):
            if hashlib.sha256(rows[k].astype("<i2").tobytes()).hexdigest() != gold["steps"][str(s_)][call_index]:
                parity["ok"] = False
                parity["mismatches"].append({"stream": s_, "call": call_index})
        parity["checked_steps"].append(call_index)

    # ---- warm-up ----
    for k in range(args.warmup):
        step_device()
        if k == 0:
            barrier()
            check_parity(0, d_pcm)
    barrier()

    # ---- timed: HBM-resident ----
    clocks = ClockSampler(local_rank)
    if rank == 0:
        clocks.start()
    bt.set_profiling(True)
    bt.read_profile()
    l0 = bt.launch_count
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record(stream)
    for _ in range(args.steps):
        step_device()
    e1.record(stream)
    barrier()
    ms = e0.elapsed_time(e1)
    launches = bt.launch_count - l0
    sample_ms, n_sample, frame_ms, n_frame = bt.read_profile()
    launch_ms, launch_steps = bt.profile_launches()
    bt.set_profiling(False)
    clk = clocks.stop() if rank == 0 else None
    nonzero = int((d_pcm[:, 320:] != 0).sum().item())
    if nonzero == 0:
        raise SystemExit("bench.py: output is all zeros -- the kernel did not run")
    check_parity(args.warmup + args.steps - 1, d_pcm)     # the output of the LAST TIMED step

    # ---- timed: end to end through the host-buffer C-ABI call ----
    step_host()
    barrier()
    check_parity(args.warmup + args.steps, h_pcm.numpy())  # the host-buffer path, first call
    k_e2e = max(1, min(args.steps, 2))
    f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    f0.record(stream)
    for _ in range(k_e2e):
        step_host()
    f1.record(stream)
    barrier()
    ms_e2e = f0.elapsed_time(f1)

    t = torch.tensor([ms, ms_e2e], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, ms_e2e = float(t[0]), float(t[1])
    samples_per_step = B * world * T
    value = samples_per_step * args.steps / (ms * 1e-3)
    e2e_value = samples_per_step * k_e2e / (ms_e2e * 1e-3)

    if rank == 0:
        pk = peaks()
        # ---- roofline of the dominant kernel (sample_kernel) ----
        # algorithmic int8 ops per stream-sample: 2*(32*n_blk_A + 32*n_blk_B + 3*NB*NB)  
        active_frames = n_frames - bt.features_delay
        ops_per_sample = 2 * (32 * bt.nblocks_a + 32 * bt.nblocks_b + 3 * 16 * 16)
        kernel_s = sample_ms * 1e-3 / max(1, n_sample)
        samples_per_launch = B * active_frames * FRAME * args.steps / max(1, n_sample)
        int8_tops = ops_per_sample * samples_per_launch / kernel_s / 1e12
        int8_meas = measure_int8_peak(dev)                   # cuBLASLt s8 x s8 -> s32, 8192^3, on this GPU
        int8_peak = int8_meas if int8_meas else 2.0 * pk["bf16_tflops_sustained"]
        # on-chip bytes the algorithm must move per stream-sample: 3 embedding rows + the conditioning row
        gather_bytes = 4 * 1152 * 4
        gather_gbs = gather_bytes * samples_per_launch / kernel_s / 1e9
        try:
            l2_peak = engine.measure_l2_gather(local_rank)
        except Exception:
            l2_peak = None
        hbm_bytes_per_sample = 2.0 + (stride * 4 + (1152 + 48 + 16) * 4 * 2) / FRAME   # pcm + features + conditioning w/r
        # DRAM bytes of one sample_kernel launch from the committed ncu list of this workload (profiles/traffic.json:
        # caches left alone, because the conditioning rows are meant to be found in L2; the cold-cache figure of the
        # `--set full` capture is in the same file); null when the workload differs from it.
        traffic = None
        tile_used = bt.tile
        n_sms = torch.cuda.get_device_properties(dev).multi_processor_count
        try:
            tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))["sample_kernel"]
            frames_per_launch = active_frames * args.steps / max(1, n_sample)
            if tj["streams"] == B and tj["tile_streams_per_cta"] == tile_used and abs(tj["frames_per_launch"] - frames_per_launch) < 2:
                traffic = tj["dram_bytes_read"] + tj["dram_bytes_write"]
        except Exception:
            traffic = None
        # L1/shared-memory data pipe (128 B/clk/SM): bytes the algorithm moves through it per stream-sample --
        # the four gathered input rows, the A fragments (4 B per block per stream), the B fragments (256 B per
        # chunk, shared by the streams of a CTA) and the 30 dual-FC nodes the sampler evaluates
        chunks = bt.chunks_a + bt.nblocks_b / 8.0 + 6
        smem_bytes = gather_bytes + 32.0 * chunks + 256.0 * chunks / tile_used + 30 * 128
        sm_clock = (clk or {}).get("sm_mhz") or 1965.0
        smem_peak = 128.0 * n_sms * sm_clock * 1e6 / 1e9
        smem_gbs = smem_bytes * samples_per_launch / kernel_s / 1e9
        roofline = {"bound": "tensor", "kernel": "sample_kernel", "achieved": int8_tops, "peak": int8_peak,
                    "unit": "TFLOP/s", "frac": int8_tops / int8_peak, "traffic": traffic,
                    "peak_source": ("torch._int_mm 8192^3 s8*s8->s32 measured in this run" if int8_meas
                                    else f"2 x bf16_tflops_sustained ({pk['src']})"),
                    "ops_per_stream_sample": ops_per_sample, "avg_launch_ms": kernel_s * 1e3,
                    "launches": n_sample, "share_of_step": sample_ms / ms,
                    "note": "int8 u8*s8->s32 tensor ops; the kernel is bound by on-chip gather bandwidth and issue "
                            "slots, not by tensor throughput (DESIGN.md)"}
        roof_l2 = {"bound": "l2_gather", "achieved": gather_gbs, "peak": l2_peak, "unit": "GB/s",
                   "frac": (gather_gbs / l2_peak) if l2_peak else None,
                   "bytes_per_stream_sample": gather_bytes,
                   "peak_source": "lpcnet_b200_measure_l2_gather (random 4608 B rows from a 3.5 MB L2-resident table)"}
        roof_smem = {"bound": "l1_shared_data_pipe", "achieved": smem_gbs, "peak": smem_peak, "unit": "GB/s",
                     "frac": smem_gbs / smem_peak, "bytes_per_stream_sample": smem_bytes,
                     "peak_source": f"128 B/clk/SM x {n_sms} SMs x {sm_clock:.0f} MHz (SM clock sampled during the run)"}
        # SURVEY.md 8(d)'s algorithmic on-chip bytes per stream-sample (each datum once: 3 embedding rows, one 4-byte
        # activation word per block, state read+write, conditioning amortised over the frame) against the same pipe:
        # a roofline fraction, where roofline_smem above counts what THIS implementation moves (pipe utilisation)
        algo_bytes = 3 * 1152 * 4 + 4 * (bt.nblocks_a + bt.nblocks_b + 24) + 2 * (384 + 16) * 4 + (1152 + 48 + 16) * 4 / FRAME
        roof_onchip = {"bound": "on_chip_bytes_algorithmic", "achieved": algo_bytes * samples_per_launch / kernel_s / 1e9,
                       "peak": smem_peak, "unit": "GB/s", "frac": algo_bytes * samples_per_launch / kernel_s / 1e9 / smem_peak,
                       "bytes_per_stream_sample": algo_bytes}
        # issue slots: what this kernel is bound by.  Warp-instructions per stream-sample from the
        # committed ncu capture of this workload (profiles/traffic.json), executed per second in this run, against one
        # instruction per scheduler and clock on the SMs of the device.
        roof_issue = None
        try:
            if traffic is not None and tj.get("warp_instructions_per_launch"):
                wi = tj["warp_instructions_per_launch"] / (tj["streams"] * tj["frames_per_launch"] * FRAME)
                ach = wi * samples_per_launch / kernel_s
                peak_i = 4.0 * n_sms * sm_clock * 1e6
                roof_issue = {"bound": "issue_slots", "achieved": ach / 1e9, "peak": peak_i / 1e9, "unit": "G warp-inst/s",
                              "frac": ach / peak_i, "warp_instructions_per_stream_sample": wi,
                              "frac_of_sms_in_use": ach / (peak_i * min(1.0, -(-B // tile_used) / n_sms)),
                              "peak_source": f"4 schedulers x {n_sms} SMs x {sm_clock:.0f} MHz"}
        except Exception:
            roof_issue = None
        roof_hbm = {"bound": "hbm", "achieved": hbm_bytes_per_sample * value / world / 1e9, "peak": pk["hbm_gbs"],
                    "unit": "GB/s", "frac": hbm_bytes_per_sample * value / world / 1e9 / pk["hbm_gbs"]}
        # ---- CPU baseline: bounded sample of the same workload on the host cores ----
        cpu = None
        if not args.no_cpu_baseline and world == 1:     # the reported baseline is an N=1 figure
            ref = CpuReference("synthesis", n_frames, model_blob)
            try:
                cpu = ref.measure(3)
            finally:
                ref.close()
        aux = None
        if world == 1 and not args.no_aux and B == 1024:
            try:
                aux = run_aux(args, dev, local_rank, model_blob)
            except Exception as e:      # the headline stands on its own; say why the side measurements are missing
                aux = [{"error": f"{type(e).__name__}: {e}"}]
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u8*s8->s32 + f32", "data": "synthetic",
            "x_realtime": value / 16000.0,
            "step_latency_us": kernel_s * 1e6 / (active_frames * FRAME * args.steps / max(1, n_sample)),
            "step_latency_us_p50_p99": step_latency_percentiles(launch_ms, launch_steps),
            "config": {"workload": f"batched -synthesis, {B} streams x {args.seconds:g} s per GPU "
                                   + ("(BASELINE configs[1])" if B == 1024 else "(BASELINE configs[4]: throughput sweep point)"),
                       "streams_per_gpu": B, "seconds": args.seconds, "tile_streams_per_cta": bt.tile,
                       "model": "random-init, N=384 NB=16 density (0.05,0.05,0.2), reference blob layout",
                       "parallelism": f"streams sharded over {world} GPU(s), no collective",
                       "l2": f"inputs larger than L2 (features {h_feat.numel() * 4 / 2**20:.0f} MiB + pcm "
                             f"{h_pcm.numel() * 2 / 2**20:.0f} MiB per step)"},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h_feat.numel() * 4 * world,
                    "d2h_bytes_per_step": h_pcm.numel() * 2 * world, "steps": k_e2e},
            "gpu_launches": launches, "clocks": clk,
            "parity_check": parity,
            "roofline": roofline, "roofline_issue": roof_issue, "roofline_onchip": roof_onchip, "roofline_smem": roof_smem, "roofline_l2": roof_l2,
            "roofline_hbm": roof_hbm,
            # the frame network of chunk k+1 runs on a second stream beside the sample kernel of chunk k: its time
            # (here as a fraction of the step) is concurrent with the sample kernel, not additive
            "frame_network_share_of_step": frame_ms / ms, "frame_network_overlapped": True,
            "cpu_baseline": cpu,
            "aux": aux,
        }
        print(json.dumps(line), flush=True)
    bt.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if parity["ok"] is False:
        raise SystemExit(f"bench.py: the timed run's PCM differs from the reference's: {parity['mismatches']}")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--streams", type=int, default=1024, help="streams per GPU")
    ap.add_argument("--seconds", type=float, default=10.0, help="audio per stream per step")
    ap.add_argument("--tile", type=int, default=0, help="streams per CTA (0 = auto)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-aux", action="store_true", help="skip the configs[2]/[3]/[4] side measurements")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "b200":
        args.warmup = 3
    if args.impl == "reference":
        run_reference(args)
    else:
        run_gpu(args)


if __name__ == "__main__":
    main()
