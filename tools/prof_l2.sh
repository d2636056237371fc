# DRAM traffic of the bench workload's kernels as they run (caches left alone, one metric pass: no replay), and one
# `--set full` capture of the sample kernel for the other sections.   bash tools/prof_l2.sh <tag>
T=${1:-r2_v5}
O=gpurun_out
python tools/prof_run.py 1024 8 160 > $O/${T}_prof_plain.log 2>&1 || exit 1
ncu --cache-control none --clock-control none --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum -c 2000 --csv --log-file $O/${T}_dram_per_launch.csv python tools/prof_run.py 1024 8 160 > $O/ncu_l2a.log 2>&1
LPCNET_B200_L2_PERSIST_MB=0 ncu --cache-control none --clock-control none --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum -c 2000 --csv --log-file $O/${T}_dram_per_launch_l2off.csv python tools/prof_run.py 1024 8 160 > $O/ncu_l2b.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:sample_kernel -s 45 -c 1 -o $O/${T}_t8 -f python tools/prof_run.py 1024 8 160 > $O/ncu_l2c.log 2>&1
ls -la $O/${T}*
