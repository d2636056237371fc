"""Per-kernel DRAM traffic of an `ncu --cache-control none --metrics dram__bytes_read.sum,dram__bytes_write.sum,
gpu__time_duration.sum --csv` launch list (tools/prof_l2.sh), over the second half of the list (the first call of
tools/prof_run.py includes warm-up frames).   python tools/dram_summary.py list.csv <output samples in that half>"""
import collections
import csv
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
samples = float(sys.argv[2])
h = rows[0]
ik, im, iv, iid = h.index("Kernel Name"), h.index("Metric Name"), h.index("Metric Value"), h.index("ID")
per = collections.defaultdict(dict)
for r in rows[1:]:
    per[int(r[iid])][r[im]] = float(r[iv].replace(",", ""))
    per[int(r[iid])]["k"] = r[ik].split("(")[0].replace("lpcb::", "").replace("<unnamed>::", "")
ids = sorted(per)
sel = ids[len(ids) // 2:]
agg = collections.defaultdict(lambda: [0, 0.0, 0.0, 0.0])
for i in sel:
    a = agg[per[i]["k"]]
    a[0] += 1; a[1] += per[i]["dram__bytes_read.sum"]; a[2] += per[i]["dram__bytes_write.sum"]; a[3] += per[i]["gpu__time_duration.sum"]
print(f"kernel,launches,dram_read_MB,dram_write_MB,read_MB_per_launch,write_MB_per_launch,bytes_per_output_sample,time_ms   ({samples:.0f} output samples)")
tr = tw = 0.0
for k, (n, rd, wr, t) in sorted(agg.items(), key=lambda kv: -(kv[1][1] + kv[1][2])):
    tr += rd; tw += wr
    print(f"{k},{n},{rd/1e6:.2f},{wr/1e6:.2f},{rd/n/1e6:.3f},{wr/n/1e6:.3f},{(rd+wr)/samples:.2f},{t/1e6:.2f}")
print(f"all,,{tr/1e6:.2f},{tw/1e6:.2f},,,{(tr+tw)/samples:.2f},")
