"""GPU (-m gpu): KV-head sharding.  The concatenation of per-rank outputs must equal the unsharded
run BIT FOR BIT (same reduction order per head).  With one GPU the ranks run one
after the other on it; with >= 2 GPUs a second test runs real ranks over NCCL and checks
``gather_heads`` on every rank.
"""
import os
import sys

import pytest
import torch

from cases import make_inputs
from minivllm_b200.head_shard import shard_plan
from minivllm_b200.layers.attention import paged_attention_decode, store_kvcache

pytestmark = pytest.mark.gpu


CHUNK = 128  # pinned: the default split size depends on the shard's size, and bit-identity of
             # sharded vs unsharded results needs the same reduction tree on both sides


def _local_decode(inp, sh, dev, chunk=CHUNK):
    q = inp["q"][:, sh.q_heads].contiguous().to(dev)
    kc = inp["k_cache"][:, :, sh.kv_heads].contiguous().to(dev)
    vc = inp["v_cache"][:, :, sh.kv_heads].contiguous().to(dev)
    return paged_attention_decode(q, kc, vc, inp["block_tables"].to(dev), inp["context_lens"].to(dev),
                                  inp["scale"], sh.num_heads, sh.num_kv_heads, inp["D"], inp["BS"],
                                  chunk_tokens=chunk)


@pytest.mark.parametrize("name,world", [("ragged_bf16", 2), ("ragged_bf16", 4), ("ragged_bf16", 8),
                                        ("long_split", 2), ("g8_ragged", 2)])
def test_sharded_equals_unsharded_bitwise(cuda_device, name, world):
    inp = make_inputs(name)
    full = _local_decode(inp, shard_plan(inp["H"], inp["KVH"], 1, 0), cuda_device)
    parts = [_local_decode(inp, shard_plan(inp["H"], inp["KVH"], world, r), cuda_device)
             for r in range(world)]
    assert torch.equal(torch.cat(parts, dim=1), full)
    # with each side's own default split size the results agree to rounding, not bit for bit
    auto = torch.cat([_local_decode(inp, shard_plan(inp["H"], inp["KVH"], world, r), cuda_device, 0)
                      for r in range(world)], dim=1)
    assert (auto.float() - full.float()).abs().max() <= 2e-2


def test_sharded_store_is_a_slice_of_the_full_store(cuda_device):
    inp = make_inputs("ragged_bf16")
    KVH, D, BS = inp["KVH"], inp["D"], inp["BS"]
    g = torch.Generator().manual_seed(9)
    k = torch.randn(8, KVH, D, generator=g).to(torch.bfloat16)
    v = torch.randn(8, KVH, D, generator=g).to(torch.bfloat16)
    slots = torch.tensor([5, -1, 77, 300, 2, 19, 640, 1])
    kc, vc = inp["k_cache"].to(cuda_device), inp["v_cache"].to(cuda_device)
    store_kvcache(k.to(cuda_device), v.to(cuda_device), kc, vc, slots.to(cuda_device), BS)
    for r in range(4):
        sh = shard_plan(inp["H"], KVH, 4, r)
        kc_r = inp["k_cache"][:, :, sh.kv_heads].contiguous().to(cuda_device)
        vc_r = inp["v_cache"][:, :, sh.kv_heads].contiguous().to(cuda_device)
        store_kvcache(k[:, sh.kv_heads].to(cuda_device), v[:, sh.kv_heads].to(cuda_device), kc_r, vc_r,
                      slots.to(cuda_device), BS)
        assert torch.equal(kc_r, kc[:, :, sh.kv_heads]) and torch.equal(vc_r, vc[:, :, sh.kv_heads])


@pytest.mark.parametrize("name,world", [("ragged_bf16", 2), ("long_split", 2), ("ragged_bf16", 8)])
def test_peer_stores_gather_on_one_gpu(cuda_device, name, world):
    """mvl_paged_decode_peers with every "peer" buffer on this GPU: each emulated rank writes its
    heads at its head offset into all destination buffers; all buffers must end up equal to the
    unsharded result (includes a zero-length row, written by the zero-fill path)."""
    import ctypes

    from minivllm_b200 import _ext
    from minivllm_b200.layers import attention as att
    inp = make_inputs(name)
    inp["context_lens"] = inp["context_lens"].clone()
    inp["context_lens"][1] = 0
    full = _local_decode(inp, shard_plan(inp["H"], inp["KVH"], 1, 0), cuda_device)
    B, H, D = inp["q"].shape
    bufs = [torch.full((B, H, D), float("nan"), dtype=inp["q"].dtype, device=cuda_device)
            for _ in range(world)]
    ptrs = (ctypes.c_void_p * world)(*[t.data_ptr() for t in bufs])
    bt, cl = inp["block_tables"].to(cuda_device), inp["context_lens"].to(cuda_device)
    lib = _ext.lib()
    for r in range(world):
        sh = shard_plan(H, inp["KVH"], world, r)
        q = inp["q"][:, sh.q_heads].contiguous().to(cuda_device)
        kc = inp["k_cache"][:, :, sh.kv_heads].contiguous().to(cuda_device)
        vc = inp["v_cache"][:, :, sh.kv_heads].contiguous().to(cuda_device)
        nbytes = lib.mvl_paged_decode_workspace_bytes(B, sh.num_heads, sh.num_kv_heads, D, bt.shape[1],
                                                      inp["BS"], att._dtype_code(q), CHUNK)
        ws = torch.zeros(max(nbytes, 16), dtype=torch.uint8, device=cuda_device)  # counters at zero
        _ext.check(lib.mvl_paged_decode_peers(
            ctypes.cast(ptrs, ctypes.c_void_p), world, H, sh.q_heads.start, q.data_ptr(), kc.data_ptr(),
            vc.data_ptr(), bt.data_ptr(), cl.data_ptr(), int(cl.dtype == torch.int64), B, bt.shape[1],
            bt.stride(0), sh.num_heads, sh.num_kv_heads, D, inp["BS"], kc.shape[0], float(inp["scale"]),
            att._dtype_code(q), ws.data_ptr(), nbytes, CHUNK, _ext.DECODE_FUSED_COMBINE,
            torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()
    for buf in bufs:
        assert torch.equal(buf, full)


@pytest.mark.parametrize("chunk", [0, CHUNK])
def test_peer_sync_protocol_on_one_gpu(cuda_device, chunk):
    """mvl_decode_step_peers_sync / mvl_peer_wait with two emulated ranks on this GPU, one launch after
    the other (never a wait before the rank it waits for has been launched): after both ranks' steps
    each control block holds both step counts, the waits pass, the step's buffer of each rank equals
    the unsharded result and the other buffer is untouched; then a wait for a rank that never
    arrives gives up after its time-out and names the rank."""
    import ctypes

    from minivllm_b200 import _ext
    from minivllm_b200.layers import attention as att
    inp = make_inputs("ragged_bf16")
    world = 2
    full = _local_decode(inp, shard_plan(inp["H"], inp["KVH"], 1, 0), cuda_device)
    B, H, D = inp["q"].shape
    bufs = torch.full((world, 2, B, H, D), float("nan"), dtype=inp["q"].dtype, device=cuda_device)
    ctrl = torch.zeros(world, 16, dtype=torch.int32, device=cuda_device)
    cptr = (ctypes.c_void_p * world)(*[ctrl[r].data_ptr() for r in range(world)])
    bt, cl = inp["block_tables"].to(cuda_device), inp["context_lens"].to(cuda_device)
    lib = _ext.lib()
    stream = torch.cuda.current_stream().cuda_stream
    shards = []
    for r in range(world):
        sh = shard_plan(H, inp["KVH"], world, r)
        shards.append((sh, inp["q"][:, sh.q_heads].contiguous().to(cuda_device),
                       inp["k_cache"][:, :, sh.kv_heads].contiguous().to(cuda_device),
                       inp["v_cache"][:, :, sh.kv_heads].contiguous().to(cuda_device)))
    nbytes = lib.mvl_paged_decode_workspace_bytes(B, H // world, inp["KVH"] // world, D, bt.shape[1], inp["BS"],
                                                  att._dtype_code(inp["q"]), chunk)
    ws = torch.zeros(max(nbytes, 16), dtype=torch.uint8, device=cuda_device)

    def launch(r, par):
        sh, q, kc, vc = shards[r]
        optr = (ctypes.c_void_p * world)(*[bufs[d, par].data_ptr() for d in range(world)])
        _ext.check(lib.mvl_decode_step_peers_sync(
            ctypes.cast(optr, ctypes.c_void_p), ctypes.cast(cptr, ctypes.c_void_p), world, r, H, sh.q_heads.start,
            q.data_ptr(), None, None, kc.data_ptr(), vc.data_ptr(), None, bt.data_ptr(), cl.data_ptr(),
            int(cl.dtype == torch.int64), B, bt.shape[1], bt.stride(0), 0, 0, sh.num_heads, sh.num_kv_heads, D,
            inp["BS"], kc.shape[0], float(inp["scale"]), att._dtype_code(q), ws.data_ptr(), ws.numel(), chunk,
            _ext.DECODE_FUSED_COMBINE, stream))

    for step in range(3):
        par = step & 1
        for r in range(world):
            launch(r, par)
        for r in range(world):
            _ext.check(lib.mvl_peer_wait(ctrl[r].data_ptr(), world, 2000, stream))
        torch.cuda.synchronize()
        c = ctrl.cpu()
        assert c[:, :world].eq(step + 1).all() and c[:, 8].eq(step + 1).all()
        assert c[:, 9].eq(0).all() and c[:, 10].eq(0).all()
        for r in range(world):
            if chunk:
                assert torch.equal(bufs[r, par], full)
            else:
                assert (bufs[r, par].float() - full.float()).abs().max() <= 2e-2
            if step == 0:
                assert torch.isnan(bufs[r, 1].float()).all()
    # rank 1 never launches step 4: rank 0's wait must give up and say so (bounded, 50 ms)
    launch(0, 1)
    _ext.check(lib.mvl_peer_wait(ctrl[0].data_ptr(), world, 50, stream))
    torch.cuda.synchronize()
    assert int(ctrl[0, 10]) == 2 and int(ctrl[0, 8]) == 4 and int(ctrl[0, 1]) == 3


def _c_abi_nccl_gather(rank, world, dev, local, full):
    """mvl_allgather_heads with a communicator made the way a non-Python host would make it: raw NCCL
    calls (here through ctypes on the process's libnccl.so.2), the ncclComm_t handed over as void*."""
    import ctypes

    import torch.distributed as dist
    from minivllm_b200 import _ext
    from minivllm_b200.layers import attention as att
    nccl = ctypes.CDLL("libnccl.so.2")

    class UniqueId(ctypes.Structure):
        _fields_ = [("internal", ctypes.c_char * 128)]

    uid = UniqueId()
    if rank == 0:
        assert nccl.ncclGetUniqueId(ctypes.byref(uid)) == 0
    buf = torch.frombuffer(bytearray(bytes(uid)), dtype=torch.uint8).to(dev)
    dist.broadcast(buf, 0)
    ctypes.memmove(ctypes.byref(uid), bytes(buf.cpu().numpy().tobytes()), 128)
    comm = ctypes.c_void_p()
    nccl.ncclCommInitRank.argtypes = [ctypes.POINTER(ctypes.c_void_p), ctypes.c_int, UniqueId, ctypes.c_int]
    assert nccl.ncclCommInitRank(ctypes.byref(comm), world, uid, rank) == 0
    try:
        B, h, D = local.shape
        staging = torch.empty(world, B, h, D, dtype=local.dtype, device=dev)
        out = torch.full((B, world * h, D), float("nan"), dtype=local.dtype, device=dev)
        _ext.check(_ext.lib().mvl_allgather_heads(out.data_ptr(), staging.data_ptr(), local.contiguous().data_ptr(), B,
                                                  h, D, att._dtype_code(local), world, comm,
                                                  torch.cuda.current_stream().cuda_stream))
        torch.cuda.synchronize()
        return bool(torch.equal(out, full))
    finally:
        nccl.ncclCommDestroy.argtypes = [ctypes.c_void_p]
        nccl.ncclCommDestroy(comm)


def _nccl_worker(rank, world, port, ret):
    import torch.distributed as dist
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
    from minivllm_b200.head_shard import gather_heads
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dev = torch.device("cuda", rank)
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        inp = make_inputs("ragged_bf16")
        full = _local_decode(inp, shard_plan(inp["H"], inp["KVH"], 1, 0), dev)
        local = _local_decode(inp, shard_plan(inp["H"], inp["KVH"], world, rank), dev)
        got = gather_heads(local)
        torch.cuda.synchronize()
        ok = bool(torch.equal(got, full))
        # the same gather done by the decode kernels' own stores into peer-mapped buffers
        from minivllm_b200.head_shard import PeerGatheredDecode
        sh = shard_plan(inp["H"], inp["KVH"], world, rank)
        fused = PeerGatheredDecode(sh, inp["q"].shape[0], inp["H"], inp["D"], inp["q"].dtype, dev)
        q = inp["q"][:, sh.q_heads].contiguous().to(dev)
        kc = inp["k_cache"][:, :, sh.kv_heads].contiguous().to(dev)
        vc = inp["v_cache"][:, :, sh.kv_heads].contiguous().to(dev)
        bt, cl = inp["block_tables"].to(dev), inp["context_lens"].to(dev)
        for it in range(8):
            # even steps: the library's own split (persistent kernel, step announced from its epilogue);
            # odd steps: pinned chunk size (chunk grid + one-thread signal launch), bit-identical to `full`
            res = fused(q, kc, vc, bt, cl, inp["scale"], inp["BS"], chunk_tokens=(CHUNK if it % 2 else 0))
            if rank == it % world:
                torch.cuda._sleep(30_000_000)   # a slow consumer on one rank: the others run ahead
            snap = res.clone()                  # the consumer's read, behind the wait in stream order
            res.fill_(float("nan"))             # poison what was read; its next writer is step it + 2
            torch.cuda.synchronize()
            if it % 2:
                ok = ok and bool(torch.equal(snap, full))
            else:
                ok = ok and bool((snap.float() - full.float()).abs().max() <= 2e-2)
        ok = ok and fused.timed_out() == 0
        ok = ok and _c_abi_nccl_gather(rank, world, dev, local, full)
        ret[rank] = ok
    finally:
        dist.destroy_process_group()


def test_nccl_gather_on_real_ranks(cuda_device):
    world = min(torch.cuda.device_count(), 8)
    world = 1 << (world.bit_length() - 1)
    if world < 2:
        pytest.skip("needs >= 2 GPUs (covered on CPU over gloo by tests/test_head_shard_gloo.py)")
    import torch.multiprocessing as mp
    port = 29500 + os.getpid() % 2000
    with mp.Manager() as mgr:
        ret = mgr.dict()
        mp.spawn(_nccl_worker, args=(world, port, ret), nprocs=world, join=True)
        assert dict(ret) == {r: True for r in range(world)}


@pytest.mark.parametrize("world", [2, 4])
def test_sharded_prefill_equals_unsharded_bitwise(cuda_device, world):
    """The prefill branch shards the same way (heads are independent; strided head slices of q/k/v go
    straight to the kernels as views): per-rank outputs concatenated over heads == the unsharded run,
    dense and over cached pages."""
    from minivllm_b200.layers.attention import flash_attention_prefill, flash_attention_prefill_paged
    from test_oracle_golden import _paged_case
    H, KVH, D, BS = 16, 4, 128, 16
    lens_ctx, lens_new = [700, 300, 129], [700, 44, 129]
    c = _paged_case(world, lens_ctx, lens_new, H, KVH, D, BS, torch.bfloat16)
    dev = cuda_device
    q, kc, vc = c["q"].to(dev), c["kc"].to(dev), c["vc"].to(dev)
    cu_q, bt, ctx = c["cu_q"].to(dev), c["bt"].to(dev), c["ctx"].to(dev)
    scale = 1.0 / D ** 0.5
    full_p = flash_attention_prefill_paged(q, kc, vc, cu_q, bt, ctx, scale, H, KVH, D, BS)
    k_all = torch.cat(c["k_full"]).to(dev)
    v_all = torch.cat(c["v_full"]).to(dev)
    q_all = torch.randn(sum(lens_ctx), H, D, generator=torch.Generator().manual_seed(1)).to(torch.bfloat16).to(dev)
    cu_k = torch.tensor([0, 700, 1000, 1129], dtype=torch.int32, device=dev)
    full_d = flash_attention_prefill(q_all, k_all, v_all, cu_k, scale, H, KVH, D)
    parts_p, parts_d = [], []
    for r in range(world):
        sh = shard_plan(H, KVH, world, r)
        parts_p.append(flash_attention_prefill_paged(q[:, sh.q_heads], kc[:, :, sh.kv_heads].contiguous(),
                                                     vc[:, :, sh.kv_heads].contiguous(), cu_q, bt, ctx, scale,
                                                     sh.num_heads, sh.num_kv_heads, D, BS))
        parts_d.append(flash_attention_prefill(q_all[:, sh.q_heads], k_all[:, sh.kv_heads], v_all[:, sh.kv_heads],
                                               cu_k, scale, sh.num_heads, sh.num_kv_heads, D))
    assert torch.equal(torch.cat(parts_p, dim=1), full_p)
    assert torch.equal(torch.cat(parts_d, dim=1), full_d)
