"""GPU (-m gpu): the mirror installed behind the reference's OWN import paths and driven by the
reference's OWN callers (needs the unmodified reference package under baseline/_ref; skipped otherwise).

What runs is the reference's code: ``Sequence`` objects, ``ModelRunner.prepare_decode``
(engine/model_runner.py:231-256, called unbound on a stand-in runner) building the decode context through
``set_context``, and ``Qwen3Attention.forward`` (models/qwen3.py:80-116: QKV projection, q/k RMSNorm,
rotary embedding, ``self.attention(q, k, v)``, output projection).  ``minivllm_b200.dropin.install()``
swaps ``Attention`` / ``store_kvcache`` / ``paged_attention_decode`` / the context functions underneath
it.  Checked: the attention output against the float64 oracle on the q/k/v the reference's modules
produced; the whole layer against the same layer run on the reference's Triton kernels (cache bit for
bit, output within bf16 tolerance); and this package's fused core (``Qwen3AttentionCore``) against the
reference layer's attention input/output.
"""
import os
import sys
import types

import pytest
import torch

pytestmark = pytest.mark.gpu
_REF = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "baseline", "_ref")


@pytest.fixture(scope="module")
def reference():
    if not os.path.isdir(os.path.join(_REF, "myvllm")):
        pytest.skip("reference package not installed under baseline/_ref")
    import torch.distributed as dist
    sys.path.insert(0, _REF)
    made_group = False
    try:
        if not dist.is_initialized():   # Qwen3Attention reads dist.get_world_size() (qwen3.py:25)
            dist.init_process_group("gloo", store=dist.HashStore(), rank=0, world_size=1)
            made_group = True
        try:
            import myvllm.engine.model_runner as mr
            import myvllm.engine.sequence as sq
            import myvllm.models.qwen3 as q3
        except Exception as exc:
            pytest.skip(f"reference package does not import here: {exc!r}")
        yield types.SimpleNamespace(mr=mr, sq=sq, q3=q3)
    finally:
        from minivllm_b200 import dropin
        dropin.uninstall()
        if made_group:
            dist.destroy_process_group()
        sys.path.remove(_REF)
        for name in [n for n in sys.modules if n == "myvllm" or n.startswith("myvllm.")]:
            del sys.modules[name]


def _layer(ref, dev, hidden, H, KVH, D, max_pos, dtype):
    torch.manual_seed(5)
    layer = ref.q3.Qwen3Attention(hidden, H, D, scale=1.0, num_kv_heads=KVH, base=10000, max_position=max_pos)
    layer = layer.to(dev)
    with torch.no_grad():   # the reference never loads weights (torch.empty parameters, SURVEY A-1)
        for p in layer.parameters():
            p.copy_(torch.randn(p.shape, generator=torch.Generator().manual_seed(p.numel())) * p.shape[-1] ** -0.5)
    return layer.to(dtype)


def _sequences(ref, BS, lens, num_blocks):
    ref.sq.Sequence.block_size = BS
    g = torch.Generator().manual_seed(3)
    perm = torch.randperm(num_blocks, generator=g).tolist()
    seqs = []
    for n in lens:
        s = ref.sq.Sequence(list(range(n)))
        s.block_table = [perm.pop() for _ in range((n + BS - 1) // BS)]
        seqs.append(s)
    return seqs


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_reference_callers_run_on_the_mirror(cuda_device, reference, dtype):
    """fp32 is the reference's own regime (main.py runs the model in the default dtype); with a bf16
    model its rotary embedding hands fp32 q / k to attention beside bf16 v and caches, which both its
    Triton kernels and the mirror accept (operands cast to the cache's dtype)."""
    import torch._dynamo
    from minivllm_b200 import dropin, reset_context, set_context
    from minivllm_b200.layers import attention as ours
    from minivllm_b200.layers.pre_attention import Qwen3AttentionCore
    from oracle import c_oracle
    ref, dev = reference, cuda_device
    hidden, H, KVH, D, BS, NB = 2048, 16, 8, 128, 16, 64     # hidden = H * D (qwen3.py:38); the reference builds Attention with BS 16
    lens = [5, 17, 31, 33, 250, 129, 2, 63]                   # never a multiple of BS (SURVEY A-4)
    B = len(lens)
    old_disable = torch._dynamo.config.disable
    torch._dynamo.config.disable = True                       # the reference's @torch.compile'd norms, eagerly
    try:
        layer = _layer(ref, dev, hidden, H, KVH, D, 512, dtype)
        g = torch.Generator().manual_seed(11)
        kc0 = torch.randn(NB, BS, KVH, D, generator=g).to(dtype).to(dev)
        vc0 = torch.randn(NB, BS, KVH, D, generator=g).to(dtype).to(dev)
        x = torch.randn(B, hidden, generator=g).to(dtype).to(dev)
        positions = torch.tensor([n - 1 for n in lens], device=dev)
        runner = types.SimpleNamespace(block_size=BS)

        def run():
            seqs = _sequences(ref, BS, lens, NB)
            ref.mr.ModelRunner.prepare_decode(runner, seqs)   # the reference's own context producer
            layer.attention.k_cache, layer.attention.v_cache = kc0.clone(), vc0.clone()
            grabbed = {}
            hook = layer.attention.register_forward_hook(
                lambda m, args, out: grabbed.update(q=args[0].clone(), k=args[1].clone(), v=args[2].clone(), o=out.clone()))
            try:
                with torch.no_grad():
                    y = layer(x, positions)
            finally:
                hook.remove()
            return y, grabbed, layer.attention.k_cache, layer.attention.v_cache

        # 1. the reference end to end: its Attention, its Triton kernels
        assert type(layer.attention).__module__ == "myvllm.layers.attention"
        try:
            y_ref, g_ref, kc_ref, vc_ref = run()
        except Exception as exc:
            if dtype == torch.float32:
                raise
            pytest.skip(f"the reference's own kernels do not take its mixed-dtype hand-over here: {exc!r}")
        ctx_ref = ref.mr.get_context()
        assert type(ctx_ref).__module__ == "myvllm.utils.context"

        # 2. the same callers on the mirror
        patched = dropin.install()
        assert "myvllm.models.qwen3.Attention" in patched and "myvllm.engine.model_runner.set_context" in patched
        assert ref.q3.Attention is ours.Attention
        new_att = ref.q3.Attention(layer.attention.num_heads, D, 1.0, layer.attention.num_kv_heads)
        layer.attention = new_att.to(dev)     # what Qwen3Attention.__init__ would have built (qwen3.py:56-61)
        y, got, kc, vc = run()
        ctx = ref.mr.get_context()
        assert type(ctx).__module__ == "minivllm_b200.utils.context" and not ctx.is_prefill
        assert ctx.slot_mapping.dtype == torch.int64 and ctx.block_tables.dtype == torch.int32
    finally:
        torch._dynamo.config.disable = old_disable
        dropin.uninstall()
        reset_context()

    # the modules in front of attention are the reference's in both runs: same q, k, v bit for bit
    assert torch.equal(got["q"], g_ref["q"]) and torch.equal(got["k"], g_ref["k"]) and torch.equal(got["v"], g_ref["v"])
    # KV-cache write: bit-exact against the reference's Triton store kernel
    assert torch.equal(kc, kc_ref) and torch.equal(vc, vc_ref)
    # attention output: the float64 oracle on the cache the step left behind (scale = 1/sqrt(D), attention.py:512)
    q = got["q"].view(B, H, D).to(dtype)      # what the kernel saw (a bf16 model hands over fp32 q)
    tables = ctx.block_tables.cpu()
    want = c_oracle.paged_decode(q.cpu(), kc.cpu(), vc.cpu(), tables, ctx.context_lens.cpu(), 1.0 / D ** 0.5,
                                 H, KVH, D, BS)
    err = (got["o"].view(B, H, D).float().cpu() - want).abs()
    a, r = (1e-4, 0.0) if dtype == torch.float32 else (1e-3, 2.0 ** -8)
    assert float((err - (a + r * want.abs())).max()) <= 0
    # and against the reference's own kernels: attention output and the whole layer (after o_proj);
    # north_star: 1e-4 abs with both sides fp32, 2e-2 for bf16
    tol = 1e-4 if dtype == torch.float32 else 2e-2
    assert got["o"].dtype == g_ref["o"].dtype
    assert (got["o"].float() - g_ref["o"].float()).abs().max() <= tol
    scale_y = max(1.0, float(y_ref.float().abs().max()))
    assert (y.float() - y_ref.float()).abs().max() <= tol * scale_y

    # 3. this package's fused core on the reference layer's projections: same attention output
    core = Qwen3AttentionCore(H, KVH, D, eps=1e-5, base=10000, max_position=512, block_size=BS).to(dev)
    core.attention.k_cache, core.attention.v_cache = kc0.clone(), vc0.clone()
    with torch.no_grad():
        qkv = layer.qkv_projection(x)
    set_context(False, slot_mapping=ctx.slot_mapping, context_lens=ctx.context_lens, block_tables=ctx.block_tables)
    try:
        o_core = core(qkv, positions)
    finally:
        reset_context()
    # (norm + RoPE in fp32 with one rounding here, op-by-op rounding in the reference's modules)
    assert (o_core.view(B, H, D).float().cpu() - want).abs().max() <= (1e-3 if dtype == torch.float32 else 2e-2)
    assert torch.equal(core.attention.v_cache, vc)
