"""The timed CPU baselines of bench.py: the reference's own `lpcnet_demo` (unmodified sources, compiled under
oracle/_ref by oracle/Makefile with the README's -Ofast flags), one process per host core, all cores at once
(BASELINE.md section 3).  Used only by bench.py's `cpu_baseline` / `--impl reference` legs.

  synthesis : lpcnet_demo -synthesis feat.f32 out.pcm               (src/lpcnet_demo.c:202-219)   configs[0], [1], [4]
  decode    : lpcnet_demo -decode packets.bin out.pcm               (src/lpcnet_demo.c:176-189)   configs[2]
  plc       : lpcnet_demo -plc_file noncausal_dc loss.txt in out    (src/lpcnet_demo.c:220-262)   configs[3]
"""
from __future__ import annotations

import os
import shutil
import statistics
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
REF = os.path.join(ROOT, "oracle", "_ref")
FRAME = 160


def cpu_model() -> str:
    """model name + family/model/stepping (cloud hosts often hide the name) + the vector ISA the demo can use"""
    name, fam, mod, step, flags = "unknown", "?", "?", "?", ""
    try:
        for line in open("/proc/cpuinfo"):
            k, _, v = line.partition(":")
            k, v = k.strip(), v.strip()
            if k == "model name" and name == "unknown":
                name = v
            elif k == "cpu family" and fam == "?":
                fam = v
            elif k == "model" and mod == "?":
                mod = v
            elif k == "stepping" and step == "?":
                step = v
            elif k == "flags" and not flags:
                flags = v
    except OSError:
        pass
    isa = [f for f in ("avx2", "fma", "avx512f", "avx512_vnni", "avx_vnni", "amx_int8") if f in flags.split()]
    return f"{name} (family {fam} model {mod} stepping {step}; {' '.join(isa)})"


class CpuReference:
    """mode: 'synthesis' | 'decode' | 'plc'.  n_frames of 10 ms per process."""

    def __init__(self, mode: str, n_frames: int, blob_bytes: bytes | None = None):
        from lpcnet_b200 import blob
        self.mode, self.n_frames = mode, n_frames
        self.dir = tempfile.mkdtemp(prefix=f"lpcnet_ref_{mode}_")
        self.cores = sorted(os.sched_getaffinity(0))
        self.blob_bytes = blob_bytes
        if blob_bytes is not None:
            with open(os.path.join(self.dir, "weights_blob.bin"), "wb") as f:
                f.write(blob_bytes)       # lpcnet_demo (USE_WEIGHTS_FILE) reads ./weights_blob.bin (lpcnet_demo.c:113)
        for i, _ in enumerate(self.cores):
            if mode == "synthesis":
                blob.synthetic_features(n_frames, i).tofile(os.path.join(self.dir, f"in{i}"))
            elif mode == "decode":
                np.random.default_rng(5 + i).integers(0, 256, size=(n_frames // 4, 8), dtype=np.uint8).tofile(os.path.join(self.dir, f"in{i}"))
            else:
                blob.synthetic_pcm(n_frames, 100 + i).tofile(os.path.join(self.dir, f"in{i}"))
                lost = blob.synthetic_loss_pattern(n_frames, 7 + i, 0.6, 0.1)
                # the demo reads one flag per 20 ms (lpcnet_demo.c:236-239)
                with open(os.path.join(self.dir, f"loss{i}"), "w") as f:
                    f.write("\n".join(str(int(x)) for x in lost[::2]) + "\n")
        self.binary, self.kind = self._pick_binary()

    def _cmd(self, binary, i):
        if self.mode == "synthesis":
            return [binary, "-synthesis", f"in{i}", f"out{i}"]
        if self.mode == "decode":
            return [binary, "-decode", f"in{i}", f"out{i}"]
        return [binary, "-plc_file", "noncausal_dc", f"loss{i}", f"in{i}", f"out{i}"]

    def _pick_binary(self):
        names = {"synthesis": ("lpcnet_demo_fast_native", "lpcnet_demo_fast_v3"), "decode": ("lpcnet_demo_embed_d2_native", "lpcnet_demo_embed_d2"),
                 "plc": ("lpcnet_demo_embed_d0_native", "lpcnet_demo_embed_d0")}[self.mode]
        want = self.n_frames * FRAME * 2 if self.mode != "decode" else (self.n_frames // 4) * 640 * 2
        for name in names:
            p = os.path.join(REF, name)
            if os.path.exists(p):
                # -march=native was resolved on the build container; fall back if this host lacks the ISA
                r = subprocess.run(self._cmd(p, 0), cwd=self.dir, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
                if r.returncode == 0 and os.path.getsize(os.path.join(self.dir, "out0")) == want:
                    return p, "reference"
        return None, "port"

    def run_once(self):
        """All cores at once, one stream each.  Returns (aggregate samples/s, per-core samples/s list)."""
        samples = self.n_frames * FRAME
        if not self.binary:
            if self.mode != "synthesis":
                raise RuntimeError(f"no reference binary for {self.mode} (oracle/_ref did not travel)")
            # the oracle port, single thread (only when oracle/_ref did not travel)
            from lpcnet_b200 import blob
            from oracle import orcpy
            om = orcpy.OrcModel(orcpy.OrcLib(), self.blob_bytes, 2, 1.0, None)
            t0 = time.perf_counter()
            orcpy.OrcStream(om).run(blob.synthetic_features(self.n_frames, 0))
            v = samples / (time.perf_counter() - t0)
            return v, [v]
        taskset = shutil.which("taskset")
        t0 = time.perf_counter()
        procs = []
        for i, core in enumerate(self.cores):
            cmd = self._cmd(self.binary, i)
            if taskset:
                cmd = [taskset, "-c", str(core)] + cmd
            procs.append(subprocess.Popen(cmd, cwd=self.dir, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL))
        done = [None] * len(procs)
        while any(d is None for d in done):
            for k, p in enumerate(procs):
                if done[k] is None and p.poll() is not None:
                    if p.returncode != 0:
                        raise RuntimeError("reference lpcnet_demo failed")
                    done[k] = time.perf_counter() - t0
            time.sleep(0.002)
        dt = time.perf_counter() - t0
        return len(procs) * samples / dt, [samples / d for d in done]

    def measure(self, runs: int, warm: int = 1):
        for _ in range(warm):
            self.run_once()
        agg, per_core = [], []
        for _ in range(runs):
            a, pc = self.run_once()
            agg.append(a)
            per_core += pc
        n = len(self.cores) if self.binary else 1
        what = os.path.basename(self.binary) if self.binary else "oracle port (liblpcnet_oracle.so)"
        arg = {"synthesis": "-synthesis", "decode": "-decode", "plc": "-plc_file noncausal_dc"}[self.mode]
        return {"value": statistics.mean(agg), "unit": "samples/s", "cores": n, "kind": self.kind,
                "per_core_min": min(per_core), "per_core_median": statistics.median(per_core),
                "sample": f"{runs} x ({n} concurrent processes, one per host core, each lpcnet_demo {arg} of one "
                          f"{self.n_frames / 100:.0f} s stream) with {what}",
                "cpu": cpu_model()}

    def close(self):
        shutil.rmtree(self.dir, ignore_errors=True)
