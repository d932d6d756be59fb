#!/bin/bash
CMD="python bench.py --workload c5 --share-prefix on --no-cpu-baseline --no-e2e --steps 3 --warmup 3 --sustain-s 0 --parity-seqs 0"
$CMD > gpurun_out/plain_c5.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 20 -c 12 --csv --log-file gpurun_out/r02_c5_launches.csv $CMD > gpurun_out/ncu_c5_l.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:group_tc5 -s 3 -c 1 -f -o gpurun_out/r02_c5_group_tc5 $CMD > gpurun_out/ncu_c5.log 2>&1
grep -v "^==" gpurun_out/r02_c5_launches.csv | cut -d, -f5,12- | tail -14
