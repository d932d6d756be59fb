"""CPU: the host-side mirror of the reference interface and the C-ABI surface (no compute calls).

Reference interface being mirrored: src/myvllm/utils/context.py:5-27 and
src/myvllm/layers/attention.py:67-74, 418-429, 477-484.
"""
import ctypes
import inspect
import os
import re

import numpy as np
import pytest
import torch

import minivllm_b200 as pkg
from minivllm_b200 import _ext
from minivllm_b200.layers import attention as att
from minivllm_b200.utils import context as ctxmod

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "mvl_abi.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(mvl_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    names = _declared_symbols()
    assert "mvl_store_kvcache" in names and "mvl_paged_decode" in names
    raw = ctypes.CDLL(_ext.LIB_PATH)
    for name in names:
        assert hasattr(raw, name), f"{name} declared in mvl_abi.h but not exported"
    assert sorted(_ext.PROTOTYPES) == names, "ctypes binding and header disagree"


def _ctype_of(param, ret=False):
    import ctypes as C
    t = param.strip()
    if t in ("void", ""):
        return None
    if "*" in t:
        return C.c_char_p if (ret and "char" in t) else C.c_void_p
    t = re.sub(r"\b(const|[a-z_][a-z_0-9]*)$", lambda m: m.group(0) if m.group(0) in
               ("int", "float", "unsigned", "size_t", "int64_t", "int32_t", "const") else "", t).replace("const", "").strip()
    return {"int": C.c_int, "float": C.c_float, "unsigned": C.c_uint, "unsigned int": C.c_uint, "size_t": C.c_size_t,
            "int64_t": C.c_int64, "int32_t": C.c_int32}[t]


def test_ctypes_prototypes_match_the_header_argument_by_argument():
    """Every declaration of include/mvl_abi.h, parsed: return type and each parameter's type class (pointer, int,
    int64_t, size_t, float, unsigned) equal the ctypes prototype the Python side calls through — a swapped or
    missing argument in the binding would otherwise pass garbage without any error."""
    text = open(os.path.join(ROOT, "include", "mvl_abi.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    text = re.sub(r"//[^\n]*", "", text)
    decls = re.findall(r"([A-Za-z_][A-Za-z_0-9 \*]*?)\b(mvl_[a-z0-9_]+)\s*\(([^;{}]*)\)\s*;", text)
    assert len(decls) == len(_ext.PROTOTYPES)
    for ret, name, params in decls:
        want_ret = _ctype_of(ret, ret=True)
        want_args = [c for c in (_ctype_of(x) for x in params.split(",")) if c is not None]
        got_ret, got_args = _ext.PROTOTYPES[name]
        assert got_ret == want_ret, f"{name}: return type"
        assert len(got_args) == len(want_args), f"{name}: {len(got_args)} arguments bound, {len(want_args)} declared"
        for i, (g, w) in enumerate(zip(got_args, want_args)):
            assert g == w, f"{name}: argument {i} bound as {g.__name__}, declared as {w.__name__}"


def test_header_is_plain_c_and_the_library_is_callable_from_c(tmp_path):
    """The boundary is a C ABI: the header must compile as C99 (what cgo / JNI / a Rust bindgen would feed on), and
    a C program that only has the header and the .so must be able to call it (argument validation needs no GPU)."""
    import shutil
    import subprocess
    if shutil.which("gcc") is None:
        pytest.skip("no gcc")
    inc = os.path.join(ROOT, "include")
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-pedantic", "-fsyntax-only", "-x", "c",
                    os.path.join(inc, "mvl_abi.h")], check=True, capture_output=True)
    src = tmp_path / "use_abi.c"
    src.write_text(r"""
#include <dlfcn.h>
#include <stdio.h>
#include <string.h>
#include "mvl_abi.h"
typedef int (*version_fn)(void);
typedef const char* (*strerror_fn)(int);
typedef int (*store_fn)(const void*, const void*, void*, void*, const void*, int, int64_t, int, int, int64_t,
                        int64_t, int64_t, int, void*);
int main(int argc, char** argv) {
  void* h = dlopen(argv[1], RTLD_NOW);
  if (!h) { fprintf(stderr, "%s\n", dlerror()); return 2; }
  version_fn version = (version_fn)dlsym(h, "mvl_abi_version");
  strerror_fn err = (strerror_fn)dlsym(h, "mvl_strerror");
  store_fn store = (store_fn)dlsym(h, "mvl_store_kvcache");
  if (!version || !err || !store) return 3;
  if (version() != MVL_ABI_VERSION) return 4;
  if (strcmp(err(MVL_OK), "ok") != 0) return 5;
  if (store(0, 0, 0, 0, 0, 1, 0, 8, 128, 1024, 1024, 64, MVL_DTYPE_BF16, 0) != MVL_OK) return 6;     /* no tokens */
  if (store(0, 0, 0, 0, 0, 1, 4, 8, 128, 1024, 1024, 64, MVL_DTYPE_BF16, 0) != MVL_ERR_NULL_POINTER) return 7;
  if (store(0, 0, 0, 0, 0, 1, 4, 8, 128, 1024, 1024, 64, 9, 0) != MVL_ERR_DTYPE) return 8;
  printf("abi %d\n", version());
  return 0;
}
""")
    exe = tmp_path / "use_abi"
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-I", inc, str(src), "-o", str(exe), "-ldl"], check=True,
                   capture_output=True)
    r = subprocess.run([str(exe), _ext.LIB_PATH], capture_output=True, text=True)
    assert r.returncode == 0, (r.returncode, r.stderr)
    assert r.stdout.strip() == f"abi {_ext.ABI_VERSION}"


def test_abi_version_and_strerror():
    lib = _ext.lib()
    assert lib.mvl_abi_version() == _ext.ABI_VERSION
    assert lib.mvl_strerror(0) == b"ok"
    for code in range(-7, 0):
        assert lib.mvl_strerror(code) not in (b"ok", b"unknown mvl_status")
    assert lib.mvl_strerror(-99) == b"unknown mvl_status"


def test_argument_validation_needs_no_gpu():
    lib = _ext.lib()
    # num_tokens == 0 is a no-op; bad dtype / shape are rejected before any CUDA call
    assert lib.mvl_store_kvcache(None, None, None, None, None, 1, 0, 8, 128, 1024, 1024, 64, 2, None) == 0
    assert lib.mvl_store_kvcache(None, None, None, None, None, 1, 4, 8, 128, 1024, 1024, 64, 9, None) == -2
    assert lib.mvl_store_kvcache(None, None, None, None, None, 1, 4, 8, 128, 1024, 1024, 64, 2, None) == -1
    buf = ctypes.create_string_buffer(64)
    p = ctypes.addressof(buf)
    p += (-p) % 16
    assert lib.mvl_store_kvcache(p, p, p, p, p, 1, 4, 8, 128, 8, 1024, 64, 2, None) == -3  # stride < row
    assert lib.mvl_store_kvcache(p + 2, p, p, p, p, 1, 4, 8, 128, 1024, 1024, 64, 2, None) == -4
    with pytest.raises(ValueError):
        _ext.check(-3)
    with pytest.raises(_ext.MvlError):
        _ext.check(-5)


def test_context_matches_reference_interface():
    assert [f for f in ctxmod.Context.__dataclass_fields__] == [
        "is_prefill", "cu_seqlens_q", "cu_seqlens_k", "max_seqlen_q", "max_seqlen_k", "slot_mapping",
        "context_lens", "block_tables"]
    assert list(inspect.signature(ctxmod.set_context).parameters) == [
        "is_prefill", "cu_seqlens_q", "cu_seqlens_k", "max_seqlen_q", "max_seqlen_k", "slot_mapping",
        "context_lens", "block_tables"]
    pkg.reset_context()
    c0 = pkg.get_context()
    assert c0.is_prefill is False and c0.slot_mapping is None and c0.max_seqlen_k == 0
    slots = torch.tensor([3, -1])
    pkg.set_context(False, slot_mapping=slots, context_lens=torch.tensor([4, 9]))
    c1 = pkg.get_context()
    assert c1 is not c0 and c1.slot_mapping is slots and c1.block_tables is None
    pkg.set_context(True)  # replaces wholesale, does not merge
    assert pkg.get_context().slot_mapping is None and pkg.get_context().is_prefill is True
    pkg.reset_context()
    assert pkg.get_context().is_prefill is False


def test_operator_signatures_match_reference():
    assert list(inspect.signature(att.store_kvcache).parameters) == [
        "key", "value", "k_cache", "v_cache", "slot_mapping", "block_size"]
    assert list(inspect.signature(att.paged_attention_decode).parameters)[:10] == [
        "query", "k_cache", "v_cache", "block_tables", "context_lens", "scale", "num_heads",
        "num_kv_heads", "head_dim", "block_size"]
    sig = inspect.signature(att.Attention.__init__)
    assert list(sig.parameters)[1:] == ["num_heads", "head_dim", "scale", "num_kv_heads", "block_size"]
    assert sig.parameters["scale"].default == 1.0 and sig.parameters["block_size"].default == 16
    layer = att.Attention(16, 128, num_kv_heads=None)
    assert layer.num_kv_heads == 16 and layer.k_cache.numel() == 0 and layer.v_cache.numel() == 0
    assert hasattr(layer, "k_cache") and hasattr(layer, "v_cache")  # the runner's duck-typing hook


def test_no_cpu_fallback():
    q = torch.zeros(2, 4, 64)
    kc = torch.zeros(3, 16, 2, 64)
    bt = torch.zeros(2, 1, dtype=torch.int32)
    cl = torch.ones(2, dtype=torch.int32)
    with pytest.raises(ValueError, match="CUDA"):
        att.paged_attention_decode(q, kc, kc, bt, cl, 1.0, 4, 2, 64, 16)
    with pytest.raises(ValueError, match="CUDA"):
        att.store_kvcache(torch.zeros(2, 2, 64), torch.zeros(2, 2, 64), kc, kc.clone(),
                          torch.zeros(2, dtype=torch.int64), 16)


def test_product_does_not_import_the_oracle():
    pat = re.compile(r"^\s*(from|import)\s+oracle\b|oracle/|paged_oracle|c_oracle", re.M)
    for dirpath, _, files in os.walk(os.path.join(ROOT, "minivllm_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                assert not pat.search(open(os.path.join(dirpath, f)).read()), f"{f} uses the oracle"


def test_block_pool_bookkeeping():
    from minivllm_b200.engine import BlockPool, SeqState
    pool = BlockPool(8, 16)
    s = SeqState()
    pool.allocate(s, 16)
    assert len(s.block_table) == 1 and len(s) == 16
    pool.append_token(s)                 # token 17 starts page 2
    assert len(s.block_table) == 2 and len(s) == 17
    for _ in range(15):
        pool.append_token(s)
    assert len(s.block_table) == 2 and len(s) == 32
    pool.deallocate(s)
    assert len(pool.free) == 8 and len(s) == 0
    with pytest.raises(MemoryError):
        pool.allocate(s, 16 * 9)


def test_prefill_metadata_matches_per_token_loop():
    """Slot of prompt token j is table[j // BS] * BS + j % BS (reference model_runner.py:201-213,
    written there as per-block ranges); ragged prompts incl. exact page multiples and one token."""
    from minivllm_b200.engine import BlockPool, SeqState, prefill_metadata
    BS = 16
    pool = BlockPool(64, BS, seed=3)
    lens = [1, 15, 16, 17, 48, 5]
    seqs = [SeqState() for _ in lens]
    for s, n in zip(seqs, lens):
        pool.allocate(s, n)
    m = prefill_metadata(seqs, BS)
    assert m.cu_seqlens_q.tolist() == np.concatenate([[0], np.cumsum(lens)]).tolist()
    assert m.cu_seqlens_k.tolist() == m.cu_seqlens_q.tolist() and not m.any_cached
    assert (m.max_seqlen_q, m.max_seqlen_k) == (48, 48)
    want = [s.block_table[j // BS] * BS + j % BS for s in seqs for j in range(len(s))]
    assert m.slot_mapping.tolist() == want and len(set(want)) == len(want)
    with pytest.raises(ValueError):
        prefill_metadata([], BS)
    seqs[0].block_table = []
    with pytest.raises(ValueError):
        prefill_metadata(seqs, BS)


def test_prefix_block_manager_shares_and_frees_pages():
    """N2: leading full blocks of a prompt that are already resident are shared (ref-counted), the
    block of the last token never is, the chain hash makes a block's identity depend on its whole
    prefix, pages return when their last user leaves and can be revived by a later prompt."""
    from minivllm_b200.engine import PrefixBlockManager, PrefixSeq, prefill_metadata
    BS = 4
    mgr = PrefixBlockManager(16, BS)
    base = list(range(100, 110))                       # 10 tokens: 2 full blocks + 2
    a = PrefixSeq(base)
    assert mgr.can_allocate(a)
    mgr.allocate(a)
    assert a.num_cached_tokens == 0 and len(a.block_table) == 3 and len(mgr.free_block_ids) == 13
    b = PrefixSeq(base[:8] + [7, 8, 9])                # same first two blocks, different tail
    mgr.allocate(b)
    assert b.num_cached_tokens == 8 and b.block_table[:2] == a.block_table[:2] and b.block_table[2] != a.block_table[2]
    assert [mgr.blocks[i].ref_count for i in a.block_table[:2]] == [2, 2]
    c = PrefixSeq(base[:8])                            # exactly two full blocks: the last one is recomputed
    mgr.allocate(c)
    assert c.num_cached_tokens == 4 and c.block_table[0] == a.block_table[0] and c.block_table[1] != a.block_table[1]
    d = PrefixSeq([1, 2, 3, 4] + base[4:8] + [5])      # second block has the same tokens, different prefix
    mgr.allocate(d)
    assert d.num_cached_tokens == 0 and not set(d.block_table) & set(a.block_table)
    # metadata of a batch with cached prefixes: queries = the uncached tokens, slots at their positions
    m = prefill_metadata([b, d], BS)
    assert m.any_cached and m.cu_seqlens_q.tolist() == [0, 3, 12] and m.cu_seqlens_k.tolist() == [0, 11, 20]
    assert m.slot_mapping.tolist()[:3] == [b.block_table[2] * BS + j for j in range(3)]
    # generated tokens: a page is opened when a token starts a block, the hash is published when one fills
    for t in (50, 51):
        a.append_token(t)
        assert mgr.can_append(a)
        mgr.append(a)
    assert len(a) == 12 and len(a.block_table) == 3 and mgr.blocks[a.block_table[2]].hash is not None
    a.append_token(52)
    mgr.append(a)
    assert len(a.block_table) == 4
    e = PrefixSeq(base + [50, 51, 99])                 # reuses a's three full blocks, incl. the one filled by appends
    mgr.allocate(e)
    assert e.num_cached_tokens == 12 and e.block_table[:3] == a.block_table[:3]
    # release everything: all pages come back, and a cached page on the free list is revived on a hit
    for s in (a, b, c, d, e):
        mgr.deallocate(s)
    assert len(mgr.free_block_ids) == 16 and all(blk.ref_count == 0 for blk in mgr.blocks)
    f = PrefixSeq(base)
    mgr.allocate(f)
    assert f.num_cached_tokens == 8
    with pytest.raises(MemoryError):
        mgr.allocate(PrefixSeq(list(range(1000, 1000 + 4 * 20))))


def test_scheduler_prefill_first_decode_preempt_finish():
    """N1: the reference's scheduling policy (scheduler.py:25-82) over the prefix block manager —
    prefill first under the token budget (counting only uncached tokens), then one token per running
    sequence, preemption of the youngest when the pool runs dry, retirement at max_tokens / EOS —
    with lengths and tables that actually advance (A-3)."""
    from minivllm_b200.engine import EngineSeq, Scheduler
    BS = 4
    sch = Scheduler(max_num_sequences=3, max_num_batched_tokens=16, max_cached_blocks=9, block_size=BS, eos=99)
    a = EngineSeq(list(range(10)), max_tokens=3)
    b = EngineSeq(list(range(8)) + [50, 51], max_tokens=8)      # shares two blocks with a
    c = EngineSeq(list(range(200, 207)), max_tokens=2)
    for s in (a, b, c):
        sch.add_sequence(s)
    seqs, is_prefill = sch.schedule()
    # a: 10 new tokens; b: 2 new (8 cached, published by a's allocation) -> 12 <= 16; c: 7 more would exceed 16
    assert is_prefill and seqs == [a, b] and b.num_cached_tokens == 8 and b.block_table[:2] == a.block_table[:2]
    sch.postprocess(seqs, [1, 2])
    assert len(a) == 11 and len(b) == 11 and a.num_completion_tokens == 1
    seqs, is_prefill = sch.schedule()
    assert is_prefill and seqs == [c]                           # prefill goes first while anything waits
    sch.postprocess(seqs, [3])
    seqs, is_prefill = sch.schedule()                           # decode: a, b, c in admission order
    assert not is_prefill and seqs == [a, b, c]
    assert [len(s.block_table) for s in seqs] == [3, 3, 2]      # 11 tokens -> 3 pages, 8 tokens -> 2 pages
    sch.postprocess(seqs, [4, 5, 99])                           # c hits EOS
    assert c.status == "finished" and c not in sch.running and not c.block_table
    seqs, _ = sch.schedule()
    assert seqs == [a, b] and len(a) == 12
    sch.postprocess(seqs, [6, 7])                               # a reaches max_tokens = 3
    assert a.status == "finished" and sch.block_manager.blocks[b.block_table[0]].ref_count == 1
    # b (grows to 18 tokens, 5 pages) and d (to 32 tokens, 8 pages) cannot both stay in 9 pages: preemption
    d = EngineSeq(list(range(300, 312)), max_tokens=20)
    sch.add_sequence(d)
    seqs, is_prefill = sch.schedule()
    assert is_prefill and seqs == [d]
    sch.postprocess(seqs, [8])
    steps = 0
    while not sch.is_finished() and steps < 200:
        seqs, is_prefill = sch.schedule()
        assert seqs
        sch.postprocess(seqs, [9] * len(seqs))
        steps += 1
    assert sch.is_finished() and sch.num_preemptions >= 1
    assert b.num_completion_tokens == 8 and d.num_completion_tokens == 20
    assert len(sch.block_manager.free_block_ids) == 9


def test_dropin_registers_the_reference_import_paths():
    """Without the reference package on the path, ``dropin.install()`` registers stand-in modules so the
    reference's import lines resolve to the mirror; ``uninstall()`` removes them again."""
    import importlib
    import sys

    from minivllm_b200 import dropin
    from minivllm_b200.layers import attention as att
    from minivllm_b200.utils import context as ctx
    if importlib.util.find_spec("myvllm") is not None:
        pytest.skip("a real myvllm is importable here; the patching path is covered by tests/test_gpu_dropin.py")
    try:
        names = dropin.install()
        assert "myvllm.layers.attention.Attention" in names and "myvllm.utils.set_context" in names
        from myvllm.layers.attention import Attention, paged_attention_decode, store_kvcache
        from myvllm.utils import get_context, reset_context, set_context
        from myvllm.utils.context import Context
        import myvllm.layers
        assert Attention is att.Attention and store_kvcache is att.store_kvcache
        assert paged_attention_decode is att.paged_attention_decode and myvllm.layers.Attention is att.Attention
        assert (get_context, set_context, reset_context, Context) == (ctx.get_context, ctx.set_context,
                                                                      ctx.reset_context, ctx.Context)
        assert dropin.paged_attention_decode_triton is att.paged_attention_decode
        assert dropin.install() == []          # idempotent
    finally:
        dropin.uninstall()
    assert "myvllm" not in sys.modules and "myvllm.layers.attention" not in sys.modules
