# Round-2 evidence run (one GPU): launch lists + one `ncu --set full` capture per kernel DESIGN.md quotes.
# Each capture follows a plain run of the same command that exited 0.   bash tools/prof_round2.sh <tag>
set -x
T=${1:-r2_v3}
O=gpurun_out
B="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-aux"
NCU="ncu --clock-control none"
$B > $O/${T}_plain.json 2> $O/${T}_plain.err || exit 1
$NCU --metrics gpu__time_duration.sum -c 800 --csv --log-file $O/${T}_launches.csv $B > $O/ncu_a.log 2>&1
$NCU --set full --import-source on -k regex:sample_kernel -s 4 -c 1 -o $O/${T}_t8 -f $B > $O/ncu_b.log 2>&1
$NCU --set full --import-source on -k regex:cond_kernel -s 4 -c 1 -o $O/${T}_cond -f $B > $O/ncu_c.log 2>&1
P="python tools/bench_aux.py plc --streams 4096 --frames 20 --reps 1"
$P > $O/${T}_plc_plain.json 2>&1 || exit 1
$NCU --metrics gpu__time_duration.sum -c 6000 --csv --log-file $O/${T}_plc_launches.csv $P > $O/ncu_d.log 2>&1
$NCU --set full --import-source on -k regex:plc_pred_kernel -s 40 -c 1 -o $O/${T}_plc_pred -f $P > $O/ncu_e.log 2>&1
E="python tools/bench_aux.py encode --streams 4096 --frames 16 --reps 1"
$E > $O/${T}_enc_plain.json 2>&1 || exit 1
$NCU --metrics gpu__time_duration.sum -c 2000 --csv --log-file $O/${T}_enc_launches.csv $E > $O/ncu_f.log 2>&1
$NCU --set full --import-source on -k regex:pitch_kernel -s 1 -c 1 -o $O/${T}_pitch -f $E > $O/ncu_g.log 2>&1
$NCU --set full --import-source on -k regex:vq_end_kernel -s 1 -c 1 -o $O/${T}_vq_end -f $E > $O/ncu_h.log 2>&1
cat $O/${T}_plain.json $O/${T}_plc_plain.json $O/${T}_enc_plain.json
ls -la $O/*.ncu-rep
