set -x
python bench.py > gpurun_out/bench_r1_v4.json 2> gpurun_out/bench_r1_v4.err || exit 1
python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/b_plain.json 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file gpurun_out/r1_v4_launches.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:sample_kernel -s 4 -c 1 -o gpurun_out/r1_v4_t8 -f python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu2.log 2>&1
python tools/prof_run.py 4736 16 4 > gpurun_out/p16.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:sample_kernel -s 1 -c 1 -o gpurun_out/r1_v4_t16 -f python tools/prof_run.py 4736 16 4 > gpurun_out/ncu3.log 2>&1
cat gpurun_out/bench_r1_v4.json
