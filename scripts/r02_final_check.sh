#!/bin/bash
# Round-2 closing check on ONE B200: the suite exactly as the driver runs it, the default bench line, and the
# launch list of the default command restricted to this library's kernels.
O=gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -8 > $O/r02_pytest_gpu_full.txt
python bench.py --steps 20 --warmup 5 > $O/r02_bench_c3_driver_args.json 2> $O/r02_bench_c3_driver_args.err
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-neighbours --no-e2e --sustain-s 0 --parity-seqs 0"
$CMD > $O/plain_c3.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'paged_decode|store_kvcache|combine|lcp_kernel|group_kernel|peer_' -c 40 --csv --log-file $O/r02_ncu_launches_c3_own_kernels.csv $CMD > $O/ncu_c3_l.log 2>&1
tail -3 $O/r02_pytest_gpu_full.txt
tail -c 1500 $O/r02_bench_c3_driver_args.json
grep -c paged_decode $O/r02_ncu_launches_c3_own_kernels.csv
