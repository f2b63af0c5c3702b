"""KV-head tensor parallelism for the sparse decode-attention chain (host-side plumbing only).

The chain shards with zero exchange until the output: scores use l1_cent[n2] (cent_select.cpp:189),
selection uses block_ids[n2] (cent_select.cpp:203) and attention reads only the n2 slice of every KV
page (sparse_paged_attention_allvec_new.h:1378-1380).  Rank r of G owns KV heads
[r*Nkv/G, (r+1)*Nkv/G) and their g*Nkv/G query heads; block_table / seq_len / scale are replicated.
The only collective is the all-gather of attention_out shards [B, Nq/G, D] (plus an optional
all-reduce(max) of max_page_position_length, which is a max over ALL query heads).

The reference has no distributed code on this path; this module is new.  It contains
no compute and does not import the CUDA extension, so it runs under gloo on CPU for tests and under
NCCL on the GPU box.
"""
from dataclasses import dataclass
from typing import Optional, Tuple

import torch
import torch.distributed as dist


@dataclass(frozen=True)
class HeadShard:
    rank: int
    world: int
    kv_lo: int
    kv_hi: int
    q_lo: int
    q_hi: int

    @property
    def num_kv_heads(self) -> int:
        return self.kv_hi - self.kv_lo

    @property
    def num_heads(self) -> int:
        return self.q_hi - self.q_lo


def head_shard(rank: int, world: int, num_heads: int, num_key_value_heads: int) -> HeadShard:
    """Contiguous KV-head range of `rank`; query heads follow their KV head (hq // g)."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError(f"bad rank {rank} of world {world}")
    if num_heads % num_key_value_heads:
        raise ValueError("num_heads must be a multiple of num_key_value_heads")
    if num_key_value_heads % world:
        raise ValueError(f"{num_key_value_heads} KV heads do not divide across {world} ranks "
                         "(beyond Nkv ranks the path only replicates: run data-parallel replicas)")
    g = num_heads // num_key_value_heads
    per = num_key_value_heads // world
    lo, hi = rank * per, (rank + 1) * per
    return HeadShard(rank, world, lo, hi, lo * g, hi * g)


def kv_head_dim(kv: torch.Tensor, num_key_value_heads: int, block_size: int) -> Tuple[int, int]:
    """(dim, width) of the KV-head axis for the pool layouts the op accepts
    (tests/ifa_v5_case.py:408 [blocks,bs,Nkv,D]; :218 [blocks,bs,Nkv*D]; :36 [blocks,Nkv,bs,D])."""
    if kv.dim() == 3:
        return 2, kv.shape[2] // num_key_value_heads
    if kv.dim() == 4 and kv.shape[1] == block_size and kv.shape[2] == num_key_value_heads:
        return 2, 1
    if kv.dim() == 4 and kv.shape[1] == num_key_value_heads and kv.shape[2] == block_size:
        return 1, 1
    raise ValueError(f"unrecognised KV pool layout {tuple(kv.shape)}")


def shard_kv_pool(kv: torch.Tensor, sh: HeadShard, num_key_value_heads: int, block_size: int,
                  materialize: bool = True) -> torch.Tensor:
    """Rank-local pool holding only this rank's KV heads (what a TP serving engine stores anyway)."""
    dim, width = kv_head_dim(kv, num_key_value_heads, block_size)
    out = kv.narrow(dim, sh.kv_lo * width, sh.num_kv_heads * width)
    return out.contiguous() if materialize else out


def shard_antiquant(t: Optional[torch.Tensor], sh: HeadShard, num_key_value_heads: int) -> Optional[torch.Tensor]:
    """antiquant_scale / offset [2, Nkv, (1,) D] or [2, Nkv*D] -> this rank's heads; per-tensor [2] is replicated."""
    if t is None or t.numel() <= 2:
        return t
    D = t.numel() // (2 * num_key_value_heads)
    return t.reshape(2, num_key_value_heads, D)[:, sh.kv_lo:sh.kv_hi].contiguous()


def shard_inputs(sh: HeadShard, query, l1_cent, block_ids, key, value, num_key_value_heads: int,
                 block_size: int, antiquant_scale=None, antiquant_offset=None):
    """Slices of the per-head tensors for `sh`.  Returns a dict; replicated tensors are not included."""
    return dict(
        query=query[:, sh.q_lo:sh.q_hi].contiguous(),
        l1_cent=l1_cent[sh.kv_lo:sh.kv_hi].contiguous(),
        block_ids=block_ids[sh.kv_lo:sh.kv_hi].contiguous(),
        key=shard_kv_pool(key, sh, num_key_value_heads, block_size),
        value=shard_kv_pool(value, sh, num_key_value_heads, block_size),
        antiquant_scale=shard_antiquant(antiquant_scale, sh, num_key_value_heads),
        antiquant_offset=shard_antiquant(antiquant_offset, sh, num_key_value_heads),
        num_heads=sh.num_heads, num_key_value_heads=sh.num_kv_heads)


class HeadGather:
    """All-gather of attention_out shards [B, Nq/G, ...] into [B, Nq, ...] with a persistent staging
    buffer (no allocation per step, so the step stays CUDA-graph capturable)."""

    def __init__(self, local_shape, dtype, device, group=None):
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.local_shape = tuple(local_shape)
        self.stage = torch.empty((self.world,) + self.local_shape, dtype=dtype, device=device)

    def __call__(self, out_local: torch.Tensor, out_full: Optional[torch.Tensor] = None) -> torch.Tensor:
        if tuple(out_local.shape) != self.local_shape:
            raise ValueError(f"expected a shard of shape {self.local_shape}, got {tuple(out_local.shape)}")
        if self.world == 1:
            return out_local
        if dist.get_backend(self.group) == "nccl":
            dist.all_gather_into_tensor(self.stage, out_local.contiguous(), group=self.group)
        else:                                   # gloo (CPU tests): list form
            dist.all_gather(list(self.stage.unbind(0)), out_local.contiguous(), group=self.group)
        # stage[r, b, h, ...] -> full[b, r*h_local + h, ...]
        perm = self.stage.movedim(0, 1)
        B, G, H = perm.shape[0], perm.shape[1], perm.shape[2]
        full_shape = (B, G * H) + tuple(perm.shape[3:])
        if out_full is None:
            return perm.reshape(full_shape)
        out_full.view((B, G, H) + tuple(perm.shape[3:])).copy_(perm)
        return out_full


def allreduce_max_page_position_length(mppl: torch.Tensor, group=None) -> torch.Tensor:
    """max_page_position_length[b] = block_size * max over ALL query heads (cent_select.cpp:123-160):
    the per-rank value covers only the rank's heads, so a global consumer needs the max over ranks."""
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(mppl, op=dist.ReduceOp.MAX, group=group)
    return mppl
