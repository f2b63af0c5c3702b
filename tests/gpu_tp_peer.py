"""KV-head tensor parallelism on real GPUs (run under torchrun, one rank per GPU; tests/test_gpu_tp.py
spawns it when >= 2 GPUs are visible):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 tests/gpu_tp_peer.py

Every rank runs the fused chain on its KV-head shard with the output all-gather fused into the launch
(peer_gather.TPFusedDecode: NVLink peer stores + flag exchange) and checks, launch after launch with a
fresh query each time, that its gathered [B, Nq, D] equals (a) the NCCL all-gather of the shards and
(b) on rank 0, the CPU oracle on the unsharded problem."""
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "cann-ops-adv-vllm_b200"), os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    import kvhead_tp as tp
    from cases import make_case
    from oracle import sparse_chain as oc
    from peer_gather import TPFusedDecode
    ok = True
    for (B, Nq, Nkv, C, S, k, kv) in [(3, 32, 8, 128, 8192, 16, "f16"), (2, 64, 8, 256, 4096, 32, "i8")]:
        c = make_case(4242, B, Nq, Nkv, C, S, kv_pages=B * (S // 128) + 7, ragged=True, kv_int8=(kv == "i8"),
                      permute_table=True)
        c["seq"][B - 1] = 0                     # a sequence that attends to nothing: its zero rows must reach the peers too
        sh = tp.head_shard(rank, world, Nq, Nkv)
        loc = tp.shard_inputs(sh, c["q"], c["l1"], c["block_ids"], c["key"], c["value"], Nkv, 128)
        aqs = aqo = None
        if kv == "i8":
            aqs = tp.shard_antiquant(c["aq_scale"], sh, Nkv).contiguous().to(dev)
            aqo = tp.shard_antiquant(c["aq_offset"], sh, Nkv).contiguous().to(dev)
        g = {n: loc[n].contiguous().to(dev) for n in ("query", "key", "value", "l1_cent", "block_ids")}
        bt, seq = c["block_table"].to(dev), c["seq"].to(dev)
        nq = sh.num_heads
        bp = torch.empty(B, nq, 256, dtype=torch.int32, device=dev)
        ppl = torch.empty(B, nq, 8, dtype=torch.int32, device=dev)
        mppl = torch.empty(B, 8, dtype=torch.int64, device=dev)
        step = TPFusedDecode(sh, Nq, g["query"], g["key"], g["value"], bt, g["l1_cent"], g["block_ids"], seq, bp, ppl,
                             mppl, c["scale"], 128, k=k, antiquant_scale=aqs, antiquant_offset=aqo)
        nccl = tp.HeadGather((B, nq, 128), g["query"].dtype, dev)
        tol = 1e-3 if kv == "f16" else 2e-2
        for it in range(6):
            q_full = (torch.rand(B, Nq, 128, generator=torch.Generator().manual_seed(100 + it)) - 0.25).to(c["q"].dtype)
            g["query"].copy_(q_full[:, sh.q_lo:sh.q_hi].to(dev))
            full = step.launch()
            torch.cuda.synchronize()
            # (a) the shards, gathered by NCCL, must be bit-identical to what the peers stored
            ref = nccl(full[:, sh.q_lo:sh.q_hi].contiguous())
            same = bool(torch.equal(full, ref))
            # (b) oracle on the unsharded problem (rank 0)
            close = True
            if rank == 0:
                want = oc.sparse_paged_fusion_attention(
                    q_full, c["key"], c["value"], c["block_table"], c["l1"], c["block_ids"], c["seq"], Nq, c["scale"],
                    Nkv, 128, k, 256, c.get("aq_scale") if kv == "i8" else None, c.get("aq_offset") if kv == "i8" else None)[0]
                err = (full.float().cpu() - want.float()).abs()
                close = bool((err <= tol + tol * want.float().abs()).all())
            if not (same and close):
                ok = False
                print(f"rank {rank} {kv} launch {it}: gathered==nccl {same}, oracle {close}", flush=True)
        if step.out.timed_out():
            ok = False
            print(f"rank {rank}: flag exchange timed out", flush=True)
        dist.barrier()
        step.close()
    # ---- the caller-side backend under tensor parallelism (decode_backend.SparseDecodeBackend)
    from decode_backend import SparseDecodeBackend
    B, Nq, Nkv, C, S, k = 3, 32, 8, 128, 4096, 16
    c = make_case(777, B, Nq, Nkv, C, S, kv_pages=B * 32 + 5, ragged=True, permute_table=True)
    sh = tp.head_shard(rank, world, Nq, Nkv)
    loc = tp.shard_inputs(sh, c["q"], c["l1"], c["block_ids"], c["key"], c["value"], Nkv, 128)
    be = SparseDecodeBackend(Nq, Nkv, 128, 128, k, max_batch=4, max_blocks_per_seq=S // 128, dtype=torch.float16,
                             device=dev, scale=c["scale"], tensor_parallel=True)
    be.bind_cache(loc["key"].to(dev), loc["value"].to(dev), loc["l1_cent"].to(dev), loc["block_ids"].to(dev))
    # the second half of the steps replays the launches from CUDA graphs (one per gathered buffer, alternating)
    for it, Bs in enumerate([3, 2, 3, 3, 3, 2, 3, 3, 2]):
        q_full = torch.rand(Bs, Nq, 128, generator=torch.Generator().manual_seed(500 + it)).half()
        got = be.forward_decode(q_full[:, sh.q_lo:sh.q_hi].to(dev), c["block_table"][:Bs].to(dev), c["seq"][:Bs].to(dev),
                                use_graph=it >= 4)
        torch.cuda.synchronize()
        if rank == 0:
            want = oc.sparse_paged_fusion_attention(q_full, c["key"], c["value"], c["block_table"][:Bs], c["l1"],
                                                    c["block_ids"], c["seq"][:Bs], Nq, c["scale"], Nkv, 128, k, be.max_page_num)[0]
            err = (got.float().cpu() - want.float()).abs()
            if got.shape != want.shape or not bool((err <= 1e-3 + 1e-3 * want.float().abs()).all()):
                ok = False
                print(f"backend TP step {it} B={Bs} graph={it >= 4}: max err {err.max().item():.3e}", flush=True)
        dist.barrier()
    be.close()
    flag = torch.tensor([0 if ok else 1], device=dev)
    dist.all_reduce(flag)
    if rank == 0:
        print("TP_PEER_OK" if flag.item() == 0 else "TP_PEER_FAIL", flush=True)
    dist.destroy_process_group()
    return 0 if flag.item() == 0 else 1


if __name__ == "__main__":
    sys.exit(main())
