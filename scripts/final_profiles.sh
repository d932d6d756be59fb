# evidence for the tensor-core kernel: ncu launch list + full capture (C3, C4), sustained power trace,
# engine loop, default bench
CMD="python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e"
$CMD > gpurun_out/plain_c3.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 30 --csv --log-file gpurun_out/launches_c3_mma.csv $CMD > /dev/null 2>&1
$CMD > gpurun_out/plain_c3b.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:paged_decode_split -s 3 -c 1 -o gpurun_out/prof_split_c3_mma $CMD > gpurun_out/ncu_c3.log 2>&1
CMD4="python bench.py --workload c4 --steps 3 --warmup 3 --no-cpu-baseline --no-e2e"
$CMD4 > gpurun_out/plain_c4.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:paged_decode_split -s 3 -c 1 -o gpurun_out/prof_split_c4_mma $CMD4 > gpurun_out/ncu_c4.log 2>&1
tail -1 gpurun_out/ncu_c4.log
python scripts/power_trace.py c3 2>&1 | tail -26 > gpurun_out/power_trace_mma.txt; head -14 gpurun_out/power_trace_mma.txt
show() { python -c "
import json,sys
d=json.load(open('$1')); r=d['roofline']
print('$2', 'ms/step %.3f'%d['ms_per_step'], 'tok/s %.0f'%d['value'], 'GB/s %.0f'%r['achieved'], 'frac %.3f'%r['frac'], 'e2e %.0f'%(d['e2e'] or {'value':0})['value'], d['clocks'])" || tail -5 $1; }
timeout 300 python bench.py --workload c2e --steps 100 --warmup 10 > gpurun_out/c2e.log 2>&1; tail -1 gpurun_out/c2e.log > gpurun_out/c2e.json; show gpurun_out/c2e.json c2e-graph
timeout 600 python bench.py 2>&1 | tail -1 > gpurun_out/bench_default.json; show gpurun_out/bench_default.json default-c3-200steps
timeout 600 python bench.py --cuda-cores 2>&1 | tail -1 > gpurun_out/bench_default_ffma.json; show gpurun_out/bench_default_ffma.json default-c3-200steps-cuda-cores
