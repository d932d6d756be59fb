#!/bin/bash
# One C5 step with prefix sharing under ncu --set full: the pre-pass (lcp_kernel, group_kernel), the tcgen05
# group kernel and the split kernel, i.e. every launch of the step (DRAM bytes per kernel -> profiles/traffic.json).
CMD="python bench.py --workload c5 --share-prefix on --no-cpu-baseline --no-e2e --no-neighbours --steps 3 --warmup 3 --sustain-s 0 --parity-seqs 0"
$CMD > gpurun_out/plain_c5.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:'paged_decode|lcp_kernel|group_kernel' -s 8 -c 4 -f -o gpurun_out/r02_c5_step $CMD > gpurun_out/ncu_c5.log 2>&1
tail -3 gpurun_out/ncu_c5.log
ncu -i gpurun_out/r02_c5_step.ncu-rep --page raw --csv --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum 2>/dev/null | cut -d, -f5,12- | tail -6
