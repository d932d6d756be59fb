# 2-GPU check: multi-rank tests, then head-sharded bench with the fused peer-store gather vs NCCL
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
show() { python -c "
import json,sys
d=json.load(open('$1')); r=d['roofline']
print('$1', d['n_gpus'], d['config']['workload'], 'ms/step %.3f'%d['ms_per_step'], 'tok/s %.0f'%d['value'], 'kern_ms %.3f'%r['kernel_ms'], 'frac %.3f'%r['frac'], 'e2e %.0f'%(d['e2e'] or {'value':0})['value'], '|', d['config'].get('gather'))" || tail -3 $1; }
for w in c3 c5 c1 c2; do python bench.py --workload $w --steps 50 --warmup 10 --no-cpu-baseline 2>&1 | tail -1 > gpurun_out/b_$w.json; show gpurun_out/b_$w.json; done
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29511 bench.py --gpus 2 --steps 100 --warmup 10 --gather peer > gpurun_out/n2_peer.log 2>&1; tail -1 gpurun_out/n2_peer.log > gpurun_out/n2_peer.json; show gpurun_out/n2_peer.json
timeout 600 $TR --master-port 29512 bench.py --gpus 2 --steps 100 --warmup 10 --gather nccl > gpurun_out/n2_nccl.log 2>&1; tail -1 gpurun_out/n2_nccl.log > gpurun_out/n2_nccl.json; show gpurun_out/n2_nccl.json
