"""Per-kernel totals of an `ncu --metrics gpu__time_duration.sum --csv` launch list.
usage: python tools/launch_summary.py launches.csv [skip_first_n_launches]"""
import collections, csv, sys
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
skip = int(sys.argv[2]) if len(sys.argv) > 2 else 0
hdr = rows[0]
ik, iv, iu = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
tot, cnt = collections.Counter(), collections.Counter()
for r in rows[1 + skip:]:
    v = float(r[iv].replace(",", ""))
    v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3, "nsecond": 1e-3, "usecond": 1.0, "msecond": 1e3, "second": 1e6}.get(r[iu], 1.0)
    name = r[ik].split("(")[0][:60]
    tot[name] += v; cnt[name] += 1
T = sum(tot.values())
print(f"kernel,launches,total_us,share,avg_us   (all: {T:.0f} us, {sum(cnt.values())} launches)")
for k, v in tot.most_common():
    print(f"{k},{cnt[k]},{v:.1f},{v / T:.3f},{v / cnt[k]:.1f}")
