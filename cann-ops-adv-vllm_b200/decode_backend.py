"""vLLM-style decode-attention backend over the fused sparse chain (the caller side).

The reference's intended consumer is a vLLM-Ascend attention backend (layout comments in
tests/ifa_v5_case.py:383-394); its in-framework entry `torch_npu.npu_sparse_paged_fusion_attention`
(FlashAttentionKernelNpuOpApi.cpp:1873-1935) allocates the aux outputs internally instead of taking
them as arguments.  `SparseDecodeBackend` plays that role here for ONE attention layer:

* it owns the persistent buffers of the decode step -- static input slots (query, block_table, seq_len),
  the aux outputs (block_position, page_position_length, max_page_position_length), attention_out and
  the kernel workspace -- sized once for `max_batch` x `max_blocks_per_seq`;
* it keeps one launch-ready executor per batch size (`custom_ops.FusedDecodePlan`, no validation or
  allocation per step) and optionally replays the step from a CUDA graph per batch size, the way vLLM
  captures its decode batch sizes;
* with a tensor-parallel group it shards by KV head (kvhead_tp.head_shard) and delivers the gathered
  output of all ranks from the same launch (peer_gather.TPFusedDecode).

No compute lives here: every step is one launch of the sm_100a chain kernel.
"""
from typing import Dict, Optional

import torch
import torch.distributed as dist

from custom_ops import FusedDecodePlan
from kvhead_tp import head_shard


class SparseDecodeBackend:
    def __init__(self, num_heads: int, num_kv_heads: int, head_dim: int, block_size: int, top_k: int,
                 max_batch: int, max_blocks_per_seq: int, dtype: torch.dtype, device, scale: Optional[float] = None,
                 max_page_num: Optional[int] = None, tp_group: Optional[dist.ProcessGroup] = None,
                 tensor_parallel: bool = False):
        """num_heads / num_kv_heads are the GLOBAL head counts; with `tensor_parallel` this rank computes
        its KV-head shard and the caches bound later hold only those heads."""
        self.device = torch.device(device)
        self.dtype = dtype
        self.block_size, self.top_k, self.head_dim = block_size, top_k, head_dim
        self.scale = float(scale) if scale is not None else head_dim ** -0.5
        self.tp_group = tp_group
        self.world = dist.get_world_size(tp_group) if tensor_parallel else 1
        self.rank = dist.get_rank(tp_group) if tensor_parallel else 0
        self.shard = head_shard(self.rank, self.world, num_heads, num_kv_heads)
        self.num_heads_total = num_heads
        self.num_heads, self.num_kv_heads = self.shard.num_heads, self.shard.num_kv_heads
        self.max_batch, self.max_blocks = max_batch, max_blocks_per_seq
        # the reference fixes max_page_num = 256 (add_custom.cpp:958); a longer context needs more
        self.max_page_num = max_page_num or max(256, max_blocks_per_seq)
        dev, i32 = self.device, torch.int32
        self.query = torch.zeros(max_batch, self.num_heads, head_dim, dtype=dtype, device=dev)
        self.block_table = torch.zeros(max_batch, max_blocks_per_seq, dtype=i32, device=dev)
        self.seq_len = torch.zeros(max_batch, dtype=i32, device=dev)
        self.block_position = torch.empty(max_batch, self.num_heads, self.max_page_num, dtype=i32, device=dev)
        self.page_position_length = torch.empty(max_batch, self.num_heads, 8, dtype=i32, device=dev)
        self.max_page_position_length = torch.empty(max_batch, 8, dtype=torch.int64, device=dev)
        self.attention_out = torch.empty(max_batch, self.num_heads, head_dim, dtype=dtype, device=dev)
        self._cache = None
        self._bias = (None, None)
        self._plans: Dict[int, object] = {}
        self._graphs: Dict[int, tuple] = {}

    # ------------------------------------------------------------------------------------------
    def bind_cache(self, key_cache: torch.Tensor, value_cache: torch.Tensor, l1_cent: torch.Tensor,
                   block_ids: torch.Tensor, antiquant_scale: Optional[torch.Tensor] = None,
                   antiquant_offset: Optional[torch.Tensor] = None) -> None:
        """Paged KV pool ([blocks, bs, Nkv, D] as vLLM lays it out, fp16/bf16 or int8), the centroid table
        [Nkv, C, D] and the per-page cluster ids [Nkv, blocks] of this layer (this rank's heads).  The
        tensors are cache STATE: they are bound once and updated in place by the engine."""
        self._cache = (key_cache, value_cache, l1_cent, block_ids, antiquant_scale, antiquant_offset)
        self._close_plans()

    def bind_bias(self, pse_shift: Optional[torch.Tensor] = None, attention_mask: Optional[torch.Tensor] = None) -> None:
        """Optional per-position inputs of the attention (the reference op's pse_shift / attention_mask,
        include/spa_b200.h): pse_shift [max_batch or 1, Nq, 1, >= max_blocks_per_seq * block_size] in the query dtype
        (added to the scaled logits, e.g. a relative-position bias), attention_mask [max_batch, 1, 1, same S] bool /
        int8 / uint8 (non-zero = the position is not attended).  Positions are LOGICAL token positions of each
        sequence.  Like the caches they are state: bound once, updated in place by the engine; rows [:B] are used by a
        step of batch size B.  They select the options build of the kernel; None switches them off again."""
        if self.world > 1 and (pse_shift is not None or attention_mask is not None):
            raise NotImplementedError("pse_shift / attention_mask are not wired through the tensor-parallel plan")
        smin = self.max_blocks * self.block_size
        if pse_shift is not None and (pse_shift.dim() != 4 or pse_shift.shape[0] not in (1, self.max_batch) or
                                      pse_shift.shape[1] != self.num_heads or pse_shift.shape[3] < smin):
            raise ValueError(f"pse_shift: expected [{self.max_batch} or 1, {self.num_heads}, 1, >= {smin}]")
        if attention_mask is not None and (attention_mask.shape[0] != self.max_batch or attention_mask.shape[-1] < smin):
            raise ValueError(f"attention_mask: expected [{self.max_batch}, 1, 1, >= {smin}]")
        self._bias = (pse_shift, attention_mask)
        self._close_plans()

    def _close_plans(self):
        for p in self._plans.values():
            if hasattr(p, "close"):
                p.close()
        self._plans.clear()
        self._graphs.clear()

    def close(self):
        self._close_plans()

    def _plan(self, B: int):
        plan = self._plans.get(B)
        if plan is not None:
            return plan
        if self._cache is None:
            raise RuntimeError("bind_cache() first")
        key, value, l1, bid, aqs, aqo = self._cache
        args = (self.query[:B], key, value, self.block_table[:B], l1, bid, self.seq_len[:B], self.block_position[:B],
                self.page_position_length[:B], self.max_page_position_length[:B])
        if self.world > 1:
            from peer_gather import TPFusedDecode
            plan = TPFusedDecode(self.shard, self.num_heads_total, *args, self.scale, self.block_size, k=self.top_k,
                                 antiquant_scale=aqs, antiquant_offset=aqo, group=self.tp_group)
        else:
            pse, mask = self._bias
            plan = FusedDecodePlan(*args, self.attention_out[:B], self.num_heads, self.scale, self.num_kv_heads,
                                   self.block_size, k=self.top_k, antiquant_scale=aqs, antiquant_offset=aqo,
                                   pse_shift=None if pse is None else (pse if pse.shape[0] == 1 else pse[:B]),
                                   attention_mask=None if mask is None else mask[:B])
        self._plans[B] = plan
        return plan

    # ------------------------------------------------------------------------------------------
    def forward_decode(self, query: torch.Tensor, block_tables: torch.Tensor, seq_lens: torch.Tensor,
                       use_graph: bool = False) -> torch.Tensor:
        """One decode step of this layer for B sequences: query [B, Nq(, 1), D] (this rank's heads),
        block_tables [B, <= max_blocks_per_seq] int32, seq_lens [B] (tokens in the cache, including the
        current one).  Returns attention_out [B, Nq_total, D] -- a view of a persistent buffer, valid until
        the next step with the same batch size (next-but-one under tensor parallelism)."""
        B = query.shape[0]
        if B == 0:
            return self.attention_out[:0]
        if B > self.max_batch or block_tables.shape[1] > self.max_blocks:
            raise ValueError(f"batch {B} x {block_tables.shape[1]} blocks exceeds the backend's "
                             f"{self.max_batch} x {self.max_blocks}")
        self.query[:B].copy_(query.reshape(B, self.num_heads, self.head_dim), non_blocking=True)
        self.block_table[:B, :block_tables.shape[1]].copy_(block_tables, non_blocking=True)
        self.seq_len[:B].copy_(seq_lens, non_blocking=True)
        plan = self._plan(B)
        if not use_graph:
            out = plan.launch()
            return out if self.world > 1 else self.attention_out[:B]
        if self.world > 1:
            return plan.replay()                                # one captured graph per gathered buffer (peer_gather.py)
        entry = self._graphs.get(B)
        if entry is None:
            side = torch.cuda.Stream(self.device)
            side.wait_stream(torch.cuda.current_stream(self.device))
            with torch.cuda.stream(side):
                plan.launch()                                   # warm-up outside the capture (workspace clear)
            torch.cuda.current_stream(self.device).wait_stream(side)
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                plan.launch()
            entry = (graph,)
            self._graphs[B] = entry
        entry[0].replay()
        return self.attention_out[:B]

    def selection(self, B: int):
        """(block_position [B,Nq,MP], page_position_length [B,Nq,8], max_page_position_length [B,8]) the last
        step of batch size B left behind (the reference op returns these to its caller)."""
        return self.block_position[:B], self.page_position_length[:B], self.max_page_position_length[:B]
