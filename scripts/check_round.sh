# full GPU test suite + default-heuristic benches (1 GPU)
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
show() { python -c "
import json,sys
d=json.load(open('$1')); r=d['roofline']
print('$2', 'ms/step %.3f'%d['ms_per_step'], 'tok/s %.0f'%d['value'], 'GB/s %.0f'%r['achieved'], 'frac %.3f'%r['frac'], 'e2e %.0f'%(d['e2e'] or {'value':0})['value'], 'kern_ms', r.get('kernel_ms'), 'chunk', d['config'].get('chunk_tokens'))" || tail -5 $1; }
timeout 300 python bench.py --workload c2e --steps 100 --warmup 10 > gpurun_out/c2e.log 2>&1; tail -1 gpurun_out/c2e.log > gpurun_out/c2e.json; show gpurun_out/c2e.json c2e-graph
timeout 300 python bench.py --workload c2e --steps 100 --warmup 10 --eager > gpurun_out/c2e_eager.log 2>&1; tail -1 gpurun_out/c2e_eager.log > gpurun_out/c2e_eager.json; show gpurun_out/c2e_eager.json c2e-eager
for w in c1 c2 c3 c4 c5; do timeout 300 python bench.py --workload $w --steps 50 --warmup 10 --no-cpu-baseline 2>&1 | tail -1 > gpurun_out/d_$w.json; show gpurun_out/d_$w.json $w; done
for n in 2 4 8; do timeout 300 python bench.py --shard-of $n --steps 100 --warmup 10 --no-cpu-baseline --no-e2e 2>&1 | tail -1 > gpurun_out/sh_$n.json; show gpurun_out/sh_$n.json "c3 shard$n"; done
