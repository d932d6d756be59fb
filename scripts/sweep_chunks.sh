# chunk-size sweep used to set the default split heuristic (minivllm_b200/csrc/abi.cu make_plan)
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -2
run() { timeout 300 python bench.py --workload $1 --chunk $2 --steps 30 --warmup 5 --no-cpu-baseline --no-e2e 2>&1 | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']
print('$1 chunk $2 ->', d['config']['chunk_tokens'], 'ms/step %.3f'%d['ms_per_step'], 'tok/s %.0f'%d['value'], 'kern_ms %.3f'%r['kernel_ms'], 'GB/s %.0f'%r['achieved'], 'frac %.3f'%r['frac'])"; }
for w in c1 c2 c3 c4 c5; do run $w 0; done
run c3 2048; run c4 2048; run c4 1024
