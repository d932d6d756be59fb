timeout 300 python scripts/sweep_small.py 0 2>&1 | tail -15 > gpurun_out/sweep_small_mma.jsonl; cat gpurun_out/sweep_small_mma.jsonl
show() { python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']
print('$1', d['config']['chunk_tokens'], 'ms/step %.3f'%d['ms_per_step'], 'frac %.3f'%r['frac'])"; }
for w in c2 c3 c4 c5; do python bench.py --workload $w --steps 30 --warmup 5 --no-cpu-baseline --no-e2e 2>&1 | tail -1 | show $w; done
for n in 4 8; do python bench.py --shard-of $n --steps 50 --warmup 5 --no-cpu-baseline --no-e2e 2>&1 | tail -1 | show shard$n; done
