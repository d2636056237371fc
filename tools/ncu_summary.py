"""Summarise an .ncu-rep: key raw metrics + per-source-line instruction/stall shares.
usage: python tools/ncu_summary.py rep.ncu-rep [topN]"""
import csv, collections, subprocess, sys, io
rep = sys.argv[1]; topn = int(sys.argv[2]) if len(sys.argv) > 2 else 30
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, data = rows[0], rows[1], rows[2:]
keys = ['gpu__time_duration.sum', 'launch__registers_per_thread', 'launch__grid_size', 'smsp__inst_executed.sum',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active',
        'dram__bytes_read.sum', 'dram__bytes_write.sum', 'l1tex__t_sector_hit_rate.pct', 'lts__t_sector_hit_rate.pct',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum',
        'l1tex__throughput.avg.pct_of_peak_sustained_active', 'lts__throughput.avg.pct_of_peak_sustained_elapsed',
        'l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum', 'l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum',
        'lts__t_bytes.sum', 'sm__cycles_elapsed.max', 'sm__inst_executed_pipe_alu.sum', 'sm__inst_executed_pipe_fma.sum',
        'sm__inst_executed_pipe_lsu.sum', 'sm__inst_executed_pipe_xu.sum', 'sm__inst_executed_pipe_tensor.sum',
        'smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_wait_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_dispatch_stall_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio']
for d in data[:1]:
    print('kernel:', d[hdr.index('Kernel Name')][:70])
    for k in keys:
        if k in hdr:
            i = hdr.index(k); print(f'  {k} = {d[i]} {units[i]}')
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
sec = None; h = None
agg = collections.defaultdict(lambda: [0, 0]); txt = {}
for r in csv.reader(io.StringIO(src)):
    if not r: continue
    if r[0] == "File Path": sec = r[1].split('/')[-1]; continue
    if r[0] == "Line No": h = r; continue
    if h is None: continue
    try: ln = int(r[0])
    except ValueError: continue
    d = dict(zip(h, r))
    try: inst = int(d.get("Instructions Executed") or 0); smp = int(d.get("# Samples") or 0)
    except ValueError: continue
    agg[(sec, ln)][0] += inst; agg[(sec, ln)][1] += smp; txt[(sec, ln)] = r[1].strip()[:100]
tot = sum(v[0] for v in agg.values()) or 1; tots = sum(v[1] for v in agg.values()) or 1
print(f"total warp-inst {tot}  samples {tots}")
for f in sorted(set(k[0] for k in agg)):
    print(f"  file {f}: inst {100*sum(v[0] for k,v in agg.items() if k[0]==f)/tot:.1f}% samples {100*sum(v[1] for k,v in agg.items() if k[0]==f)/tots:.1f}%")
for k, v in sorted(agg.items(), key=lambda kv: -(kv[1][0] / tot + kv[1][1] / tots))[:topn]:
    print(f"{k[0]:13s}:{k[1]:4d} inst {100*v[0]/tot:5.1f}% smp {100*v[1]/tots:5.1f}%  {txt[k]}")
