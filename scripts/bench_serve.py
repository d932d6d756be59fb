"""The whole attention-path loop at BASELINE config 2's shape (Qwen3-0.6B: 16Q/8KV/D128, 28 layers,
256-slot pages, bf16): scheduler + prefix-caching block manager + prepare_prefill / prepare_decode +
prefill kernels + CUDA-graph decode, driven to completion on synthetic prompts that share a prefix.
The model around attention is two stubs that cost nothing (preallocated activations, fixed next
token), so the figure is what this path — kernels AND host bookkeeping — sustains.
Prints one JSON line: prefill tokens/s, decode tokens/s (= sequences per decode step / step time, the
reference's definition, llm_engine.py:58,79-83), host share."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from minivllm_b200.engine import AttentionEngine, DecodeRunner, Scheduler

dev = torch.device("cuda", 0)
L, H, KVH, D, BS = 28, 16, 8, 128, 256
NSEQ, PROMPT, SHARED, NEW = 64, 2048, 1024, 128
if "--small" in sys.argv:
    NSEQ, PROMPT, SHARED, NEW = 16, 512, 256, 32
EAGER = "--eager" in sys.argv
max_len = PROMPT + NEW + 1
num_blocks = NSEQ * ((max_len + BS - 1) // BS) + 8
runner = DecodeRunner(L, H, KVH, D, BS, num_blocks, NSEQ, max_len + BS, torch.bfloat16, dev, enforce_eager=EAGER)
if not EAGER:
    runner.capture_cudagraph([NSEQ])
sch = Scheduler(max_num_sequences=NSEQ, max_num_batched_tokens=16384, max_cached_blocks=num_blocks, block_size=BS,
                eos=-1)
g = torch.Generator().manual_seed(0)
pool = torch.randn(L, 16384 + 64, H + 2 * KVH, D, device=dev).to(torch.bfloat16)   # stand-in for the projections


def activations(toks, pos):
    x = pool[:, :len(toks)]
    return x[:, :, :H], x[:, :, H:H + KVH], x[:, :, H + KVH:]


# warm-up on a throw-away engine: first-use costs (kernel attributes, module load, pinned buffers)
_r = DecodeRunner(L, H, KVH, D, BS, 16, 4, 4 * BS, torch.bfloat16, dev, enforce_eager=True)
_s = Scheduler(4, 16384, 16, BS, eos=-1)
_e = AttentionEngine(_r, _s, activations, lambda seqs, out: [1] * len(seqs))
_p = torch.randint(0, 30000, (2 * BS,), generator=g).tolist()
for tail in (5, 300):
    _e.add_prompt(_p + list(range(tail)), max_tokens=3, ignore_eos=True)
    _e.step()
_e.run_to_completion()
torch.cuda.synchronize()
del _r, _s, _e

eng = AttentionEngine(runner, sch, activations, lambda seqs, out: [1] * len(seqs))
shared = torch.randint(0, 30000, (SHARED,), generator=g).tolist()
for i in range(NSEQ):
    eng.add_prompt(shared + torch.randint(0, 30000, (PROMPT - SHARED,), generator=g).tolist(), max_tokens=NEW,
                   ignore_eos=True)
torch.cuda.synchronize()
t_prefill = t_decode = 0.0
n_prefill_steps = n_decode_steps = 0
decode_seqs = 0
t_all0 = time.perf_counter()
while not sch.is_finished():
    t0 = time.perf_counter()
    seqs, is_prefill, out = eng.step()
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    if is_prefill:
        t_prefill += dt; n_prefill_steps += 1
        if "--verbose" in sys.argv:
            print(f"prefill step: {len(seqs)} prompts, {1e3 * dt:.2f} ms", file=sys.stderr)
    else:
        t_decode += dt; n_decode_steps += 1; decode_seqs += len(seqs)
wall = time.perf_counter() - t_all0
print(json.dumps({
    "workload": f"{NSEQ} prompts x {PROMPT} tokens ({SHARED} shared), {NEW} generated each; {L} layers {H}Q/{KVH}KV/D{D} bf16, BS={BS}",
    "decode": "eager" if EAGER else "CUDA graphs",
    "prefill_tokens_computed": eng.prefill_tokens, "prefix_cache_hits_blocks": sch.block_manager.hits,
    "prefill_steps": n_prefill_steps, "prefill_tokens_per_s": eng.prefill_tokens / t_prefill,
    "decode_steps": n_decode_steps, "decode_ms_per_step": 1e3 * t_decode / max(n_decode_steps, 1),
    "decode_tokens_per_s": decode_seqs / t_decode, "preemptions": sch.num_preemptions, "wall_s": wall}))
