# the L2-sized chunking against the round-1 chunking (LPCNET_B200_L2_PERSIST_MB=0): throughput (plain runs) and
# DRAM bytes of the sample + cond kernels (ncu, caches left alone)
for mb in default 0; do
  if [ $mb = default ]; then unset LPCNET_B200_L2_PERSIST_MB; else export LPCNET_B200_L2_PERSIST_MB=$mb; fi
  export LPCNET_B200_DEBUG=1
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-aux > gpurun_out/l2_$mb.json 2> gpurun_out/l2_$mb.err
  grep "^L2" gpurun_out/l2_$mb.err | head -1
  python -c "import json; d=json.load(open('gpurun_out/l2_$mb.json')); print('mb=$mb value', d['value'], 'e2e', d['e2e']['value'], d['gpu_launches'], d['parity_check']['ok'])"
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-aux --streams 2048 --seconds 5 > gpurun_out/l2_2048_$mb.json 2> gpurun_out/l2_2048_$mb.err
  python -c "import json; d=json.load(open('gpurun_out/l2_2048_$mb.json')); print('2048 mb=$mb value', d['value'], 'e2e', d['e2e']['value'], d['gpu_launches'])"
  ncu --cache-control none --clock-control none --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum -k regex:"sample_kernel|cond_kernel" -c 400 --csv --log-file gpurun_out/l2_dram_$mb.csv python tools/prof_run.py 1024 8 160 > /dev/null 2>&1
done
