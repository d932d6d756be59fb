"""Compare the tcgen05 prefill kernel with the mma.sync kernel and a float64 torch evaluation."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from minivllm_b200 import _ext
from minivllm_b200.layers.attention import flash_attention_prefill
dev = torch.device("cuda", 0)


def ref(q, k, v, lens, scale):
    out = torch.zeros_like(q, dtype=torch.float64)
    H, KVH = q.shape[1], k.shape[1]
    s = 0
    for n in lens:
        qs = q[s:s + n].double().permute(1, 0, 2)
        ks = k[s:s + n].double().permute(1, 0, 2).repeat_interleave(H // KVH, 0)
        vs = v[s:s + n].double().permute(1, 0, 2).repeat_interleave(H // KVH, 0)
        sc = (qs @ ks.transpose(1, 2)) * scale
        sc = sc.masked_fill(torch.ones(n, n, dtype=torch.bool, device=q.device).triu(1), float("-inf"))
        out[s:s + n] = (torch.softmax(sc, -1) @ vs).permute(1, 0, 2)
        s += n
    return out


for dtype in (torch.bfloat16, torch.float16):
    for (H, KVH, D, lens) in ((1, 1, 128, [128]), (1, 1, 128, [256]), (2, 1, 64, [128]), (4, 2, 128, [100, 300, 129]),
                              (8, 2, 64, [513, 77, 1000]), (4, 1, 128, [2048, 1, 255, 256, 257])):
        T = sum(lens)
        g = torch.Generator(device="cpu").manual_seed(1)
        q = torch.randn(T, H, D, generator=g).to(dtype).to(dev)
        k = torch.randn(T, KVH, D, generator=g).to(dtype).to(dev)
        v = torch.randn(T, KVH, D, generator=g).to(dtype).to(dev)
        cu = torch.tensor([0] + list(torch.tensor(lens).cumsum(0)), dtype=torch.int32, device=dev)
        want = ref(q, k, v, lens, D ** -0.5)
        a = flash_attention_prefill(q, k, v, cu, D ** -0.5, H, KVH, D, max_seqlen=max(lens), flags=4)
        b = flash_attention_prefill(q, k, v, cu, D ** -0.5, H, KVH, D, max_seqlen=max(lens), flags=_ext.PREFILL_MMA_SYNC)
        c = flash_attention_prefill(q, k, v, cu, D ** -0.5, H, KVH, D, max_seqlen=max(lens), flags=2)
        torch.cuda.synchronize()
        print("   single-tile kernel err", (c.double() - want).abs().max().item(), "pair vs single differing", (a != c).sum().item())
        ea, eb = (a.double() - want).abs().max().item(), (b.double() - want).abs().max().item()
        nd = (a != b).sum().item()
        print(f"{dtype} H{H} KVH{KVH} D{D} lens{lens}: tcgen05 err {ea:.3e}  mma.sync err {eb:.3e}  differing outputs {nd}/{a.numel()}", flush=True)
        if not ea < 2e-2:
            bad = ((a.double() - want).abs() > 2e-2).nonzero()
            print("  first bad (tok, head, d):", bad[:5].tolist(), "count", len(bad))
            print("  got", a[bad[0][0], bad[0][1], :8].float().tolist())
            print("  want", want[bad[0][0], bad[0][1], :8].tolist())
