# full GPU suite with the tensor-core kernel as default, then chunk sweep for it
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
run() { timeout 300 python bench.py --workload $1 --chunk $2 $3 --steps 30 --warmup 5 --no-cpu-baseline --no-e2e 2>&1 | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']
print('$1 $3 chunk $2 ->', d['config']['chunk_tokens'], 'ms/step %.3f'%d['ms_per_step'], 'GB/s %.0f'%r['achieved'], 'frac %.3f'%r['frac'])"; }
for c in 0 512 1024 2048 4096; do run c3 $c; done
for c in 0 256 512 1024; do run c2 $c; done
for c in 0 1024 2048 4096; do run c4 $c; done
for c in 0 1024 2048 4096; do run c5 $c; done
for c in 0 256 512 1024 2048; do run c3 $c "--shard-of 8"; done
for c in 0 512 1024 2048; do run c3 $c "--shard-of 4"; done
run c1 0
