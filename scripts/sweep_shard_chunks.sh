show() { python -c "
import json,sys
d=json.load(open('$1')); r=d['roofline']
print('$2', 'ms/step %.4f'%d['ms_per_step'], 'tok/s %.0f'%d['value'], 'GB/s %.0f'%r['achieved'], 'frac %.3f'%r['frac'], r.get('kernel_ms'), d['config'].get('chunk_tokens'))" || tail -5 $1; }
for c in 0 1024 2048 4096 8192; do timeout 300 python bench.py --shard-of 8 --chunk $c --steps 200 --warmup 10 --no-cpu-baseline --no-e2e 2>&1 | tail -1 > gpurun_out/s8_$c.json; show gpurun_out/s8_$c.json "c3 shard8 chunk $c"; done
for c in 0 2048 4096 8192; do timeout 300 python bench.py --shard-of 4 --chunk $c --steps 200 --warmup 10 --no-cpu-baseline --no-e2e 2>&1 | tail -1 > gpurun_out/s4_$c.json; show gpurun_out/s4_$c.json "c3 shard4 chunk $c"; done
for c in 0 2048 4096 8192; do timeout 300 python bench.py --shard-of 2 --chunk $c --steps 200 --warmup 10 --no-cpu-baseline --no-e2e 2>&1 | tail -1 > gpurun_out/s2_$c.json; show gpurun_out/s2_$c.json "c3 shard2 chunk $c"; done
