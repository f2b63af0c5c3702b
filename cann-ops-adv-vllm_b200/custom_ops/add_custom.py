"""Python operator API of the sparse decode-attention chain.

Positional signatures, return arity, dtypes and shapes are those of the reference's
extension/custom_ops/add_custom.py:10-135.  Keyword-only extensions (`k`, `max_page_num`,
`block_size`) default to the constants the reference hard-codes in its binding/tiling
(add_custom.cpp:940,958,987; select_position.cpp:7) so BASELINE configs with k in {8,128} or more
than 256 selected pages are reachable without changing any existing call site.

All compute happens in `custom_ops_lib` (C++ binding -> C ABI -> CUDA kernels).  There is no CPU or
eager fallback: importing this module without the built extension raises ImportError.
"""
from typing import List, Optional, Tuple

import torch

try:
    import custom_ops_lib
except ImportError as exc:  # fail loudly: the product has no fallback path
    raise ImportError(
        "custom_ops_lib (the sm_100a CUDA extension) is not built; run "
        "`python cann-ops-adv-vllm_b200/build.py` or `__graft_entry__.build()`") from exc


def _tensor(x, dtype, device):
    if torch.is_tensor(x):
        return x
    # fused_ifa_v5_case.py:92 passes actual_seq_lengths as a python list
    return torch.as_tensor(x, dtype=dtype, device=device)


def incre_flash_attention_v4(query: torch.Tensor,
                             key_list: List[torch.Tensor],
                             value_list: List[torch.Tensor],
                             pse_shift: torch.Tensor,
                             attention_mask: torch.Tensor,
                             actual_seq_lengths: torch.Tensor,
                             dequant_scale1: torch.Tensor,
                             quant_scale1: torch.Tensor,
                             dequant_scale2: torch.Tensor,
                             quant_scale2: torch.Tensor,
                             quant_offset2: torch.Tensor,
                             antiquant_scale: torch.Tensor,
                             antiquant_offset: torch.Tensor,
                             blocktable: torch.Tensor,
                             kv_padding_size: torch.Tensor,
                             num_heads: int,
                             scale_value: float,
                             input_layout: str,
                             num_key_value_heads: int,
                             block_size: int,
                             inner_precise: int) -> torch.Tensor:
    """Dense paged decode attention (add_custom.py:10-40; the reference routes it to the stock CANN
    IncreFlashAttention).  Served by the sparse kernel over every page of the block table (identity
    block_position), clipped by actual_seq_lengths; paged layout only, the stock op's other options are refused."""
    actual_seq_lengths = _tensor(actual_seq_lengths, torch.int64, query.device)
    return custom_ops_lib.incre_flash_attention_v4(
        query, key_list, value_list, pse_shift, attention_mask, actual_seq_lengths,
        dequant_scale1, quant_scale1, dequant_scale2, quant_scale2, quant_offset2,
        antiquant_scale, antiquant_offset, blocktable, kv_padding_size,
        num_heads, scale_value, input_layout, num_key_value_heads, block_size, inner_precise)


def sparse_paged_attention(query: torch.Tensor,
                           key: torch.Tensor,
                           value: torch.Tensor,
                           pse_shift: torch.Tensor,
                           attention_mask: torch.Tensor,
                           actual_seq_lengths: torch.Tensor,
                           dequant_scale1: torch.Tensor,
                           quant_scale1: torch.Tensor,
                           dequant_scale2: torch.Tensor,
                           quant_scale2: torch.Tensor,
                           quant_offset2: torch.Tensor,
                           antiquant_scale: torch.Tensor,
                           antiquant_offset: torch.Tensor,
                           blocktable: torch.Tensor,
                           kv_padding_size: torch.Tensor,
                           blockposition: torch.Tensor,
                           num_heads: int,
                           scale_value: float,
                           input_layout: str,
                           num_key_value_heads: int,
                           block_size: int,
                           inner_precise: int,
                           *, softmax_lse: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Paged flash-decode over the pages listed in `blockposition` (add_custom.py:41-75).  key / value: the query
    dtype, int8 with `antiquant_scale` (+ `antiquant_offset`), or torch.float8_e4m3fn (scales optional).

    `pse_shift` [B or 1, Nq, 1, S] (query dtype) is added to the scaled logits, `attention_mask` [B, (1, 1,) S]
    (bool / int8 / uint8, non-zero = masked) sets them to -1e12, both at the LOGICAL token position
    page * block_size + t (S >= blocktable.shape[1] * block_size); `quant_scale2` (+ `quant_offset2`), float32 or
    bfloat16, 1 element or [Nq, D]: the result is int8 = sat(rint(o * scale2 + offset2)).  Empty tensors = absent,
    as in the reference's calls.  (include/spa_b200.h has the reference citations.)

    `softmax_lse`: optional caller-allocated float32 [B, Nq], filled with ln(sum exp(scale * q.k)) per row
    (the reference kernels' softmaxLse output); fp16/bf16 and int8 KV."""
    actual_seq_lengths = _tensor(actual_seq_lengths, torch.int64, query.device)
    return custom_ops_lib.sparse_paged_attention_ex(
        query, [key], [value], pse_shift, attention_mask, actual_seq_lengths,
        dequant_scale1, quant_scale1, dequant_scale2, quant_scale2, quant_offset2,
        antiquant_scale, antiquant_offset, blocktable, kv_padding_size,
        blockposition, int(num_heads), float(scale_value), input_layout, int(num_key_value_heads),
        int(block_size), int(inner_precise), softmax_lse)


def compute_cent(query: torch.Tensor, l1_cent: torch.Tensor, *, k: int = 64) -> torch.Tensor:
    """query . l1_cent^T + TopK over clusters -> int32 [B, Nq, k] (add_custom.py:77-81)."""
    return custom_ops_lib.compute_cent_k(query, l1_cent, int(k))


def select_position(block_ids: torch.Tensor, block_table: torch.Tensor, seq_len: torch.Tensor,
                    indices: torch.Tensor, *, max_page_num: int = 256,
                    block_size: int = 128) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """-> (page_position, page_position_length, block_table_gather) (add_custom.py:83-87)."""
    return custom_ops_lib.select_position_ex(block_ids, block_table, seq_len, indices,
                                             int(max_page_num), int(block_size))


def cent_select(query: torch.Tensor, l1_cent: torch.Tensor, block_ids: torch.Tensor,
                block_table: torch.Tensor, seq_len: torch.Tensor, *, k: int = 64,
                max_page_num: int = 256, block_size: int = 128, sink_pages: int = 0, recent_pages: int = 0,
                max_fraction_den: int = 0) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """-> (page_position, page_position_length, max_page_position_length) (add_custom.py:89-93).

    `sink_pages`, `recent_pages`, `max_fraction_den`: what the reference KERNEL hard-codes as 1, 2 and 8
    (anchor pages always selected, at most P/8 pages per head; cent_select.cpp:218,350-362,410-414) but its
    golden lacks; off by default (= the golden)."""
    return custom_ops_lib.cent_select_ex(query, l1_cent, block_ids, block_table, seq_len, int(k),
                                         int(max_page_num), int(block_size), int(sink_pages), int(recent_pages),
                                         int(max_fraction_den))


def sparse_paged_fusion_attention(query: torch.Tensor,
                                  key: torch.Tensor,
                                  value: torch.Tensor,
                                  pse_shift: torch.Tensor,
                                  attention_mask: torch.Tensor,
                                  actual_seq_lengths: torch.Tensor,
                                  dequant_scale1: torch.Tensor,
                                  quant_scale1: torch.Tensor,
                                  dequant_scale2: torch.Tensor,
                                  quant_scale2: torch.Tensor,
                                  quant_offset2: torch.Tensor,
                                  antiquant_scale: torch.Tensor,
                                  antiquant_offset: torch.Tensor,
                                  blocktable: torch.Tensor,
                                  kv_padding_size: torch.Tensor,
                                  l1_cent: torch.Tensor,
                                  block_ids: torch.Tensor,
                                  total_seq_len: torch.Tensor,
                                  block_position: torch.Tensor,
                                  page_position_length: torch.Tensor,
                                  max_page_position_length: torch.Tensor,
                                  num_heads: int,
                                  scale_value: float,
                                  input_layout: str,
                                  num_key_value_heads: int,
                                  block_size: int,
                                  inner_precise: int,
                                  *, k: int = 64, sink_pages: int = 0, recent_pages: int = 0,
                                  max_fraction_den: int = 0,
                                  softmax_lse: Optional[torch.Tensor] = None) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """CentSelect + sparse paged attention in ONE launch (add_custom.py:96-135).

    `block_position`, `page_position_length`, `max_page_position_length` are caller-allocated and
    written in place; returns (attention_out, block_position, max_page_position_length) exactly as
    the reference binding does (add_custom.cpp:931).  max_page_num is `block_position.shape[2]`
    (sparse_paged_fusion_attention_tiling.cc:1606).
    """
    actual_seq_lengths = _tensor(actual_seq_lengths, torch.int64, query.device)
    return custom_ops_lib.sparse_paged_fusion_attention_ex(
        query, [key], [value], pse_shift, attention_mask, actual_seq_lengths,
        dequant_scale1, quant_scale1, dequant_scale2, quant_scale2, quant_offset2,
        antiquant_scale, antiquant_offset, blocktable, kv_padding_size,
        l1_cent, block_ids, total_seq_len,
        block_position, page_position_length, max_page_position_length,
        int(num_heads), float(scale_value), input_layout, int(num_key_value_heads), int(block_size),
        int(inner_precise), int(k), int(sink_pages), int(recent_pages), int(max_fraction_den), softmax_lse)


def assign_page_clusters(key: torch.Tensor, l1_cent: torch.Tensor, page_ids: torch.Tensor, block_ids: torch.Tensor,
                         page_tokens: Optional[torch.Tensor] = None, *, block_size: int = 128,
                         return_mean: bool = False, antiquant_scale: Optional[torch.Tensor] = None,
                         antiquant_offset: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Centroid maintenance, the step before the chain (no reference counterpart: its tests draw block_ids at
    random, tests/cent_select.py:49-50).  For every physical page in `page_ids` [n] int32 and every KV head,
    pools the page's keys and writes the index of the centroid with the largest inner product into
    block_ids[n2, page] (in place; ties -> lowest index).  Returns block_ids, or with `return_mean` the fp32
    page means [n, Nkv, D] (input of a k-means centroid update).  An int8 pool takes the attention ops'
    `antiquant_scale` / `antiquant_offset` ([2, Nkv, D] or [2]; the K row is used)."""
    return custom_ops_lib.assign_page_clusters(key, l1_cent, page_ids, page_tokens, block_ids, int(block_size),
                                               bool(return_mean), antiquant_scale, antiquant_offset)


class FusedDecodePlan:
    """Launch-ready fused executor bound to fixed tensors (extension; the reference keeps its aclnn
    executor repeatable the same way, add_custom.cpp:847).  `launch()` performs no host-side
    validation or allocation, so it is CUDA-graph capturable and is what bench.py times as the
    device-resident number."""

    def __init__(self, query, key, value, blocktable, l1_cent, block_ids, total_seq_len,
                 block_position, page_position_length, max_page_position_length, attention_out,
                 num_heads, scale_value, num_key_value_heads, block_size=128, *, k=64,
                 antiquant_scale=None, antiquant_offset=None, input_layout="BNSD", inner_precise=1,
                 sink_pages=0, recent_pages=0, max_fraction_den=0, softmax_lse=None,
                 pse_shift=None, attention_mask=None, quant_scale2=None, quant_offset2=None):
        empty = torch.empty(0, dtype=query.dtype, device=query.device)
        self.attention_out = attention_out
        self._plan = custom_ops_lib.make_fusion_plan(
            query, key, value, empty if antiquant_scale is None else antiquant_scale,
            empty if antiquant_offset is None else antiquant_offset, blocktable, l1_cent, block_ids,
            total_seq_len, block_position, page_position_length, max_page_position_length,
            attention_out, int(num_heads), float(scale_value), input_layout,
            int(num_key_value_heads), int(block_size), int(inner_precise), int(k), int(sink_pages),
            int(recent_pages), int(max_fraction_den), softmax_lse, pse_shift, attention_mask, quant_scale2,
            quant_offset2)

    @property
    def workspace_bytes(self) -> int:
        return self._plan.workspace_bytes

    @property
    def num_launches(self) -> int:
        return self._plan.num_launches

    @property
    def grid_size(self) -> int:
        return self._plan.grid_size

    def trace_phases(self) -> torch.Tensor:
        """Run one traced launch; returns int64 [grid, 16] of per-CTA %globaltimer ns at the phase
        boundaries (include/spa_b200.h: spaExecutorSetTrace)."""
        buf = torch.zeros(self.grid_size, 32, dtype=torch.int64, device=self.attention_out.device)
        self._plan.set_trace(buf)
        try:
            custom_ops_lib.launch_fusion_plan(self._plan)
            torch.cuda.synchronize()
        finally:
            self._plan.set_trace(torch.empty(0, dtype=torch.int64, device=buf.device))
        return buf

    def launch(self) -> torch.Tensor:
        custom_ops_lib.launch_fusion_plan(self._plan)
        return self.attention_out
