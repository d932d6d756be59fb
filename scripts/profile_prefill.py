"""One launch of the prefill kernel at a long-sequence shape, for ncu."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from minivllm_b200.layers.attention import flash_attention_prefill
dev = torch.device("cuda", 0)
H, KVH, D, nseq, n = 32, 8, 128, 2, int(sys.argv[1]) if len(sys.argv) > 1 else 8192
flags = int(sys.argv[2]) if len(sys.argv) > 2 else 0
T = nseq * n
q = torch.randn(T, H, D, device=dev).to(torch.bfloat16)
k = torch.randn(T, KVH, D, device=dev).to(torch.bfloat16)
v = torch.randn_like(k)
cu = torch.arange(0, T + 1, n, dtype=torch.int32, device=dev)
for _ in range(2):
    flash_attention_prefill(q, k, v, cu, D ** -0.5, H, KVH, D, max_seqlen=n, flags=flags)
torch.cuda.synchronize()
