# 8-GPU regression check: multi-rank tests, then C3 / C4 head-sharded with the fused peer-store gather
timeout 300 python -m pytest tests/test_gpu_multi.py -m gpu -x -q 2>&1 | tail -2
show() { python -c "
import json,sys
d=json.load(open('$1')); r=d['roofline']
print('$1', d['n_gpus'], d['config']['workload'], 'ms/step %.4f'%d['ms_per_step'], 'tok/s %.0f'%d['value'], 'kern_ms %.4f'%r['kernel_ms'], 'frac %.3f'%r['frac'], 'e2e %.0f'%(d['e2e'] or {'value':0})['value'], '|', d['config'].get('gather'))" || tail -3 $1; }
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
run() { timeout 300 $TR --nproc-per-node $1 --master-port $((29500+RANDOM%400)) bench.py --gpus $1 --workload $2 --gather $3 --steps $4 --warmup 10 > gpurun_out/n$1_$2_$3.log 2>&1; tail -1 gpurun_out/n$1_$2_$3.log > gpurun_out/n$1_$2_$3.json; show gpurun_out/n$1_$2_$3.json; }
run 8 c3 peer 200
run 8 c4 peer 50
run 4 c3 peer 200
