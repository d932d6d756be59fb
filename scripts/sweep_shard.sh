# engine loop test + c2e bench + chunk sweep on the per-rank shard of an 8-way head split (1 GPU)
timeout 600 python -m pytest tests/test_gpu_engine.py -m gpu -x -q 2>&1 | tail -3
show() { python -c "
import json,sys
d=json.load(open('$1')); r=d['roofline']
print('$2', 'ms/step %.3f'%d['ms_per_step'], 'tok/s %.0f'%d['value'], 'GB/s %.0f'%r['achieved'], 'frac %.3f'%r['frac'], 'e2e %.0f'%(d['e2e'] or {'value':0})['value'], r.get('kernel_ms'), d['config'].get('chunk_tokens'))" || tail -5 $1; }
timeout 300 python bench.py --workload c2e --steps 100 --warmup 10 > gpurun_out/c2e.log 2>&1; tail -1 gpurun_out/c2e.log > gpurun_out/c2e.json; show gpurun_out/c2e.json c2e-graph
timeout 300 python bench.py --workload c2e --steps 100 --warmup 10 --eager > gpurun_out/c2e_eager.log 2>&1; tail -1 gpurun_out/c2e_eager.log > gpurun_out/c2e_eager.json; show gpurun_out/c2e_eager.json c2e-eager
for c in 0 128 256 512 1024 2048; do timeout 300 python bench.py --shard-of 8 --chunk $c --steps 100 --warmup 10 --no-cpu-baseline --no-e2e 2>&1 | tail -1 > gpurun_out/s8_$c.json; show gpurun_out/s8_$c.json "c3 shard8 chunk $c"; done
for c in 0 512 1024 2048; do timeout 300 python bench.py --shard-of 4 --chunk $c --steps 100 --warmup 10 --no-cpu-baseline --no-e2e 2>&1 | tail -1 > gpurun_out/s4_$c.json; show gpurun_out/s4_$c.json "c3 shard4 chunk $c"; done
