"""KV-head tensor-parallel plumbing on CPU: world_size-2 gloo processes each run the chain on their
KV-head shard (the oracle stands in for the CUDA kernels -- no GPU here), all-gather the output
shards and must reproduce the single-rank result exactly."""
import os
import socket
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "cann-ops-adv-vllm_b200"), os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

import kvhead_tp as tp  # noqa: E402


def test_head_shard_ranges():
    s = [tp.head_shard(r, 4, 32, 8) for r in range(4)]
    assert [(x.kv_lo, x.kv_hi, x.q_lo, x.q_hi) for x in s] == [(0, 2, 0, 8), (2, 4, 8, 16), (4, 6, 16, 24), (6, 8, 24, 32)]
    assert tp.head_shard(7, 8, 64, 8).num_heads == 8
    with pytest.raises(ValueError):
        tp.head_shard(0, 3, 32, 8)              # 8 KV heads do not divide across 3 ranks
    with pytest.raises(ValueError):
        tp.head_shard(0, 16, 32, 8)             # beyond Nkv ranks: replicas only
    with pytest.raises(ValueError):
        tp.head_shard(2, 2, 32, 8)


def test_kv_pool_layouts_shard_the_head_axis():
    Nkv, bs, D = 4, 128, 128
    sh = tp.head_shard(1, 2, 8, Nkv)
    vllm = torch.arange(3 * bs * Nkv * D, dtype=torch.float32).view(3, bs, Nkv, D)
    assert torch.equal(tp.shard_kv_pool(vllm, sh, Nkv, bs), vllm[:, :, 2:4])
    bsh = vllm.view(3, bs, Nkv * D)
    assert torch.equal(tp.shard_kv_pool(bsh, sh, Nkv, bs).view(3, bs, 2, D), vllm[:, :, 2:4])
    bnsd = vllm.permute(0, 2, 1, 3).contiguous()
    assert torch.equal(tp.shard_kv_pool(bnsd, sh, Nkv, bs), bnsd[:, 2:4])
    aq = torch.arange(2 * Nkv * D, dtype=torch.float32).view(2, Nkv, 1, D)
    assert torch.equal(tp.shard_antiquant(aq, sh, Nkv), aq.view(2, Nkv, D)[:, 2:4])
    assert tp.shard_antiquant(torch.ones(2), sh, Nkv).numel() == 2


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, ret):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.set_num_threads(2)
        from cases import make_case
        from oracle import sparse_chain as oc
        B, Nq, Nkv, C, S, k = 2, 16, 4, 128, 4096, 16
        c = make_case(2024, B, Nq, Nkv, C, S, kv_pages=96, ragged=True)          # same seed on every rank
        sh = tp.head_shard(rank, world, Nq, Nkv)
        loc = tp.shard_inputs(sh, c["q"], c["l1"], c["block_ids"], c["key"], c["value"], Nkv, 128)
        out, pos, plen, mx = oc.sparse_paged_fusion_attention(
            loc["query"], loc["key"], loc["value"], c["block_table"], loc["l1_cent"], loc["block_ids"], c["seq"],
            loc["num_heads"], c["scale"], loc["num_key_value_heads"], 128, k, 256)
        gather = tp.HeadGather(out.shape, out.dtype, out.device)
        full = gather(out)
        full2 = gather(out, torch.empty(B, Nq, 128, dtype=out.dtype))
        pos_full = tp.HeadGather(pos.shape, pos.dtype, pos.device)(pos)
        mx = tp.allreduce_max_page_position_length(mx.clone())
        if rank == 0:
            wout, wpos, wlen, wmx = oc.sparse_paged_fusion_attention(
                c["q"], c["key"], c["value"], c["block_table"], c["l1"], c["block_ids"], c["seq"], Nq, c["scale"],
                Nkv, 128, k, 256)
            ret["out"] = bool(torch.equal(full, wout)) and bool(torch.equal(full2, wout))
            ret["pos"] = bool(torch.equal(pos_full, wpos))
            ret["mx"] = bool(torch.equal(mx, wmx))
        dist.barrier()
    finally:
        dist.destroy_process_group()


def test_two_rank_gloo_matches_single_rank():
    world = 2
    port = _free_port()
    with mp.Manager() as mgr:
        ret = mgr.dict()
        mp.spawn(_worker, args=(world, port, ret), nprocs=world, join=True)
        assert dict(ret) == {"out": True, "pos": True, "mx": True}
