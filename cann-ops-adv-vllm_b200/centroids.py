"""Centroid / block_ids maintenance around the chain.

The chain consumes `l1_cent [Nkv, C, D]` and `block_ids [Nkv, kvPageLen]`; the reference never builds them
(its tests draw both at random, tests/cent_select.py:49-50).  The hot part -- pooling every page's keys from
HBM and ranking them against the centroids -- is the CUDA kernel behind `custom_ops.assign_page_clusters`
(csrc/spa_assign.cuh).  What is left of a Lloyd iteration is bookkeeping on [n, Nkv, D] page means, done here
with torch ops on the device: it is not on the decode path and runs once per centroid refresh.
"""
from typing import Optional, Tuple

import torch

import custom_ops


def assign_all_pages(key_cache: torch.Tensor, l1_cent: torch.Tensor, block_ids: torch.Tensor, num_pages: int,
                     **kw) -> torch.Tensor:
    """(Re)assign physical pages [0, num_pages) of the pool, e.g. after the centroids changed."""
    ids = torch.arange(num_pages, dtype=torch.int32, device=key_cache.device)
    return custom_ops.assign_page_clusters(key_cache, l1_cent, ids, block_ids, **kw)


def kmeans_step(key_cache: torch.Tensor, l1_cent: torch.Tensor, block_ids: torch.Tensor, page_ids: torch.Tensor,
                page_tokens: Optional[torch.Tensor] = None, *, block_size: int = 128,
                antiquant_scale: Optional[torch.Tensor] = None,
                antiquant_offset: Optional[torch.Tensor] = None) -> Tuple[torch.Tensor, torch.Tensor]:
    """One Lloyd iteration under the chain's own assignment rule (largest inner product with the pooled key):

    E-step  block_ids[n2, page] <- assign_page_clusters (CUDA kernel; also returns the fp32 page means);
    M-step  centroid[n2, c] <- token-weighted mean of the keys of the pages assigned to c (a cluster that lost
            all its pages keeps its centroid).

    `page_ids` must not repeat.  Returns (new l1_cent in the input dtype, pages per cluster [Nkv, C] int64)."""
    Nkv, C, D = l1_cent.shape
    means = custom_ops.assign_page_clusters(key_cache, l1_cent, page_ids, block_ids, page_tokens, block_size=block_size,
                                            return_mean=True, antiquant_scale=antiquant_scale,
                                            antiquant_offset=antiquant_offset)                       # [n, Nkv, D] fp32
    n = page_ids.numel()
    ids = page_ids.long()
    ok = (ids >= 0) & (ids < min(key_cache.shape[0], block_ids.shape[1]))
    tok = torch.full((n,), float(block_size), device=means.device) if page_tokens is None else \
        page_tokens.to(means.device).clamp(0, block_size).float()
    tok = tok * ok.float()
    assign = block_ids[:, ids.clamp(0, block_ids.shape[1] - 1)].long()                                # [Nkv, n]
    sums = torch.zeros(Nkv, C, D, dtype=torch.float32, device=means.device)
    wsum = torch.zeros(Nkv, C, dtype=torch.float32, device=means.device)
    cnt = torch.zeros(Nkv, C, dtype=torch.int64, device=means.device)
    weighted = means.permute(1, 0, 2) * tok.view(1, n, 1)                                             # page key sums
    sums.scatter_add_(1, assign.unsqueeze(-1).expand(Nkv, n, D), weighted)
    wsum.scatter_add_(1, assign, tok.view(1, n).expand(Nkv, n).contiguous())
    cnt.scatter_add_(1, assign, ok.long().view(1, n).expand(Nkv, n).contiguous())
    new = torch.where(wsum.unsqueeze(-1) > 0, sums / wsum.clamp(min=1.0).unsqueeze(-1), l1_cent.float())
    return new.to(l1_cent.dtype), cnt
