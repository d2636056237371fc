"""Dynamic SASS opcode histogram (and phase split by BAR.SYNC) from an ncu report with source info.
   python tools/sass_hist.py <report.ncu-rep> [top]"""
import csv, subprocess, sys, collections, io
rep = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
lines = out.splitlines()
start = next(i for i, l in enumerate(lines) if l.startswith('"Address"'))
rows = list(csv.DictReader(io.StringIO("\n".join(lines[start:]))))
ops = collections.Counter(); samp = collections.Counter()
total = 0; wf = 0; wfi = 0
seg = []; cur = collections.Counter(); cur_n = 0; cur_s = 0
for r in rows:
    src = r["Source"].strip()
    n = int(r["Instructions Executed"] or 0); s = int(r["# Samples"] or 0)
    toks = src.split()
    op = toks[0] if not toks[0].startswith("@") else toks[1]
    op = op.split(".")[0] if not op.startswith(("LDS", "LDG", "STS", "STG")) else ".".join(op.split(".")[:2])
    ops[op] += n; samp[op] += s; total += n
    wf += int(r.get("L1 Wavefronts Shared") or 0); wfi += int(r.get("L1 Wavefronts Shared Ideal") or 0)
    cur[op] += n; cur_n += n; cur_s += s
    if op == "BAR":
        seg.append((cur_n, cur_s, cur)); cur = collections.Counter(); cur_n = 0; cur_s = 0
seg.append((cur_n, cur_s, cur))
ts = sum(samp.values())
print(f"total warp-instructions {total}, samples {ts}, shared wavefronts {wf} (ideal {wfi})")
for op, n in ops.most_common(top):
    print(f"  {op:12s} {100.0 * n / total:5.1f}% inst  {100.0 * samp[op] / max(ts, 1):5.1f}% samples")
print("segments between BAR.SYNC (static order): share of executed instructions / samples, top opcodes")
for n, s, c in seg:
    if n * 200 < total: continue
    print(f"  {100.0 * n / total:5.1f}% inst {100.0 * s / max(ts, 1):5.1f}% smp : " + ", ".join(f"{o} {100.0 * k / n:.0f}%" for o, k in c.most_common(8)))
