#!/bin/bash
python scripts/diag_e2e.py c3 > gpurun_out/r02_diag_e2e.jsonl 2> gpurun_out/r02_diag_e2e.err
python scripts/diag_e2e.py c4 >> gpurun_out/r02_diag_e2e.jsonl 2>> gpurun_out/r02_diag_e2e.err
cat gpurun_out/r02_diag_e2e.jsonl
CMD="python bench.py --workload c3 --shard-of 8 --steps 3 --warmup 3 --no-cpu-baseline --no-neighbours --no-e2e --sustain-s 0 --parity-seqs 0"
$CMD --kernel stream > gpurun_out/plain_stream.log 2>&1 && $CMD --kernel grid > gpurun_out/plain_grid.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:paged_decode -s 4 -c 1 -f -o gpurun_out/r02_shard8_stream $CMD --kernel stream > gpurun_out/ncu_stream.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:paged_decode -s 4 -c 1 -f -o gpurun_out/r02_shard8_grid $CMD --kernel grid > gpurun_out/ncu_grid.log 2>&1
tail -3 gpurun_out/ncu_stream.log gpurun_out/ncu_grid.log
