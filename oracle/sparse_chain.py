"""CPU oracle for the centroid-sparse paged decode-attention chain.

TEST INFRASTRUCTURE ONLY.  Nothing in the product path (`cann-ops-adv-vllm_b200/`) imports this
module; only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s CPU-baseline / `--impl reference`
legs may.  It is a numpy/torch restatement of the reference's algorithm, each function citing the
reference file:line it follows (paths relative to the reference checkout).

Parity status
-------------
* Selection (compute_cent / select_position / cent_select): PINNED.  `tests/golden/*.npz` were
  produced by importing the reference's own golden (`tests/utils.py:torch_compute_cent`,
  `torch_select_position`) in the build container (`tests/golden/make_golden.py`), and
  `tests/test_oracle_golden.py` checks this restatement against them bit-for-bit.
* Attention: the reference ships NO torch/numpy attention golden (only constant-value known-answer
  tests, `tests/ifa_v5_case.py:584-765`).  The restatement below follows the kernel sources and is
  pinned only by those known-answer vectors (restated in tests/test_gpu_parity.py::test_attention_known_answer)
  plus a float64 cross-check;
  beyond that it is "parity unpinned" and the north-star tolerances (1e-3 fp16, 2e-2 bf16/int8)
  define agreement.
"""
from __future__ import annotations

import math
from typing import List, Optional, Sequence, Tuple

import numpy as np
import torch

INT32_MAX = 0x7FFFFFFF
PAD = 8  # page_position_length / max_page_position_length replication (add_custom.cpp:963,993-994)


# --------------------------------------------------------------------------------------------
# A1 + A2: scores and TopK
# --------------------------------------------------------------------------------------------
def cent_scores(query: torch.Tensor, l1_cent: torch.Tensor) -> np.ndarray:
    """score[b,n1,c] = sum_d q[b,n1,d] * cent[n1//g,c,d]   (tests/utils.py:104-118;
    compute_cent.cpp:277-343 accumulates the half x half products in fp32).

    The reference leaves the summation order to the hardware (torch CPU GEMV vs. the NPU's
    16-lane partial sums), so its fp32 scores are order-dependent in the last ulp.  The oracle
    therefore evaluates the *exact* sum: fp16/bf16 products are exact, and a float64 accumulation
    of <=2^12 such products of magnitude < 2^16 is exact for all inputs whose exponent spread fits
    53 bits (always true for the reference's `torch.rand` / `randn` test inputs).  Every fp32
    evaluation order rounds towards this value, so it is the tie-free centre of all of them.
    """
    q = query.detach().cpu().to(torch.float64).numpy()
    cent = l1_cent.detach().cpu().to(torch.float64).numpy()
    if q.ndim == 4:  # [B, Nq, 1, D]
        q = q.reshape(q.shape[0], q.shape[1], q.shape[3])
    B, n1, D = q.shape
    n2, C, _ = cent.shape
    assert n1 % n2 == 0
    g = n1 // n2
    qg = q.reshape(B, n2, g, D)
    # [B, n2, g, C]
    s = np.einsum("bngd,ncd->bngc", qg, cent, optimize=False)
    return s.reshape(B, n1, C)


def topk_indices(scores: np.ndarray, k: int) -> np.ndarray:
    """k largest per row, sorted descending, ties broken by LOWEST index first.
    (torch.topk at tests/utils.py:121; AscendC::TopK isLargest=true at compute_cent.cpp:216-227.
    Neither pins a tie order; the build rule is (score desc, index asc) -- SURVEY 8(a) A2.)"""
    C = scores.shape[-1]
    assert 0 < k <= C, f"k={k} must be in (0, C={C}]"
    order = np.argsort(-scores, axis=-1, kind="stable")  # stable => lowest index first on ties
    return order[..., :k].astype(np.int32)


def compute_cent(query: torch.Tensor, l1_cent: torch.Tensor, k: int = 64) -> torch.Tensor:
    """custom_ops.compute_cent (add_custom.py:77-81; k hard-coded 64 at add_custom.cpp:940)."""
    return torch.from_numpy(topk_indices(cent_scores(query, l1_cent), k))


# --------------------------------------------------------------------------------------------
# A3: select_position
# --------------------------------------------------------------------------------------------
def select_position(
    block_ids: torch.Tensor,     # [Nkv, kvPageLen] int32
    block_table: torch.Tensor,   # [maxB, maxP] int32
    seq_len: torch.Tensor,       # [B] int32
    indices: torch.Tensor,       # [B, Nq, k] int
    block_size: int = 128,
    max_page_num: int = 256,
    sink_pages: int = 0,
    recent_pages: int = 0,
    max_fraction_den: int = 0,
) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    """Follows tests/utils.py:4-84 (golden) with the op's output shapes (add_custom.cpp:955-977):

    page_position        [B, Nq, max_page_num] int32, ascending selected logical pages, pad INT32_MAX
    page_position_length [B, Nq, 8] int32, count replicated x8
    block_table_gather   [B, Nq, maxP] int32, c[i] = block_ids[n2, block_table[b,i]] for i < pl, else 0

    The table value is clamped into [0, kvPageLen-1] for the block_ids lookup exactly as the golden
    does (utils.py:62).  Selections beyond max_page_num are dropped (the golden's width is
    max_b(pl); the op's is the fixed 256, add_custom.cpp:958).

    sink_pages / recent_pages / max_fraction_den restate what the reference KERNEL adds to the golden
    (cent_select.cpp:350-362: mask |= block_table == block_table[0], [pl-2], [pl-1], compared by VALUE;
    :218,410-414: length = min(count, pl // 8), lowest positions first).  0 = off = the golden.
    """
    bid = block_ids.detach().cpu().numpy().astype(np.int64)
    bt = block_table.detach().cpu().numpy().astype(np.int64)
    sl = seq_len.detach().cpu().numpy().astype(np.int64)
    idx = indices.detach().cpu().numpy().astype(np.int64)
    B = sl.shape[0]
    Nkv, kvPageLen = bid.shape
    Nq = idx.shape[1]
    assert Nq % Nkv == 0
    g = Nq // Nkv
    maxP = bt.shape[1]
    pos = np.full((B, Nq, max_page_num), INT32_MAX, dtype=np.int32)
    plen = np.zeros((B, Nq, PAD), dtype=np.int32)
    gather = np.zeros((B, Nq, maxP), dtype=np.int32)
    for b in range(B):
        pl = int((sl[b] + block_size - 1) // block_size)
        pl = min(pl, maxP)
        if pl <= 0:
            continue
        page_idx = np.clip(bt[b, :pl], 0, kvPageLen - 1)
        for h in range(Nq):
            c = bid[h // g, page_idx]                      # [pl]
            gather[b, h, :pl] = c
            mask = np.isin(c, idx[b, h])
            if sink_pages or recent_pages:
                anchors = np.concatenate([bt[b, :min(sink_pages, pl)], bt[b, max(pl - recent_pages, 0):pl]])
                mask |= np.isin(bt[b, :pl], anchors)
            cap = min(max_page_num, pl // max_fraction_den) if max_fraction_den > 0 else max_page_num
            p = np.nonzero(mask)[0][:cap]
            pos[b, h, : p.size] = p
            plen[b, h, :] = p.size
    return torch.from_numpy(pos), torch.from_numpy(plen), torch.from_numpy(gather)


def max_page_position_length(page_position_length: torch.Tensor, block_size: int = 128) -> torch.Tensor:
    """[B,8] int64 = block_size * max_h count (cent_select.cpp:123-160; golden expression at
    tests/cent_select.py:96-98)."""
    m = page_position_length[:, :, 0].to(torch.int64).max(dim=1)[0] * block_size
    return m.unsqueeze(1).repeat(1, PAD).contiguous()


def cent_select(query, l1_cent, block_ids, block_table, seq_len, k: int = 64,
                block_size: int = 128, max_page_num: int = 256, sink_pages: int = 0, recent_pages: int = 0,
                max_fraction_den: int = 0):
    """custom_ops.cent_select (add_custom.py:89-93; add_custom.cpp:979-1008).  Batch is taken from
    seq_len (add_custom.cpp:988), so query may carry more rows than seq_len."""
    B = seq_len.shape[0]
    indices = compute_cent(query[:B], l1_cent, k)
    pos, plen, _ = select_position(block_ids, block_table, seq_len, indices, block_size, max_page_num,
                                   sink_pages, recent_pages, max_fraction_den)
    return pos, plen, max_page_position_length(plen, block_size)


# --------------------------------------------------------------------------------------------
# A5-A8: sparse paged attention
# --------------------------------------------------------------------------------------------
def _kv_view(kv: torch.Tensor, Nkv: int, block_size: int, D: int) -> torch.Tensor:
    """Return a [blocks, bs, Nkv, D] view of the pool for the layouts the reference accepts
    (tests/ifa_v5_case.py:36 BNSD-paged, :218 BSH, :408 vLLM)."""
    if kv.dim() == 3:                                   # [blocks, bs, Nkv*D]
        return kv.view(kv.shape[0], kv.shape[1], Nkv, D)
    assert kv.dim() == 4
    if kv.shape[1] == Nkv and kv.shape[2] == block_size and not (
            kv.shape[1] == block_size and kv.shape[2] == Nkv):
        return kv.permute(0, 2, 1, 3)                   # [blocks, Nkv, bs, D] -> [blocks, bs, Nkv, D]
    return kv


def _antiquant(scale: Optional[torch.Tensor], offset: Optional[torch.Tensor], Nkv: int, D: int):
    """Per-channel antiquant params -> float32 [2, Nkv, D]  (shapes accepted by
    sparse_paged_attention_tiling_check.cc:497-525; per-tensor [2] at ...tiling.cc:781-790)."""
    def norm(t, fill):
        if t is None or t.numel() == 0:
            return torch.full((2, Nkv, D), fill, dtype=torch.float32)
        t = t.detach().cpu().to(torch.float32)
        if t.numel() == 2:
            return t.reshape(2, 1, 1).expand(2, Nkv, D).contiguous()
        return t.reshape(2, Nkv, D).contiguous()
    return norm(scale, 1.0), norm(offset, 0.0)


def attend_lists(
    query: torch.Tensor, key: torch.Tensor, value: torch.Tensor, block_table: torch.Tensor,
    lists: Sequence[Sequence[Sequence[Tuple[int, int]]]],   # lists[b][hq] = [(logical_page, ntok), ...]
    num_key_value_heads: int, scale_value: float, block_size: int,
    antiquant_scale: Optional[torch.Tensor] = None, antiquant_offset: Optional[torch.Tensor] = None,
    round_p: bool = True, dtype64: bool = False, lse_out: Optional[torch.Tensor] = None,
    pse_shift: Optional[torch.Tensor] = None, attention_mask: Optional[torch.Tensor] = None,
    quant_scale2: Optional[torch.Tensor] = None, quant_offset2: Optional[torch.Tensor] = None,
) -> torch.Tensor:
    """Attention over explicit per-(b, q-head) page lists.

    pse_shift [B or 1, Nq, 1, S] is added to the scaled logits, then attention_mask [B, (1, 1,) S] (non-zero =
      masked) SETS a logit to -1e12 (..._Us2.h:2862-2899, :757).  Both are indexed by the LOGICAL token position
      page * block_size + t (include/spa_b200.h: the reference, whose code is inherited from the dense op, indexes
      them by the position in the walked sequence -- the same thing for a dense walk).
    quant_scale2 (+ quant_offset2), [1] or [Nq, D]: int8 output sat(rint(o * scale2 + offset2)) (:3513-3600);
      rows that attended to nothing are zeros.

    lse_out: optional float32 [B, Nq], filled with the softmax log-sum-exp ln(sum_i exp(s_i)) = ln(l) + m
      (the reference's softmaxLse output, ..._Us2.h:2601-2623); -inf for an empty list.

    Addressing: phys = block_table[b, logical]; row r of that block, KV head n2 = hq // g
      (sparse_paged_attention_allvec_new.h:1338-1391 K, :1489-1538 V;
       sparse_paged_attention_split_Bbn2s2_Us2.h:241-279).
    Math: S = (Q K^T) in fp32, then *scale (..._Us2.h:2854-2860); softmax fp32 (:2903-2914);
      P rounded to the matmul dtype before PV (:3062-3063); O = P V / l (:3467-3494);
      int8 KV: x = (x_i8 + offset) * scale per channel (:2142, :3655-3670);
      output cast to the query dtype (:2971-2975).
    """
    qd = query.dtype
    q = query.detach().cpu()
    shape = q.shape
    if q.dim() == 4:
        q = q.reshape(shape[0], shape[1], shape[3])
    B, Nq, D = q.shape
    Nkv = num_key_value_heads
    g = Nq // Nkv
    K = _kv_view(key.detach().cpu(), Nkv, block_size, D)
    V = _kv_view(value.detach().cpu(), Nkv, block_size, D)
    bt = block_table.detach().cpu().to(torch.int64)
    # 8-bit KV: int8 with antiquant (x = (x_i8 + offset) * scale), or FP8 e4m3 (no reference analogue: numbers
    # already; scale / offset optional, same formula)
    is_i8 = K.dtype in (torch.int8, torch.float8_e4m3fn)
    ft = torch.float64 if dtype64 else torch.float32
    if is_i8:
        sc, of = _antiquant(antiquant_scale, antiquant_offset, Nkv, D)
    out = torch.zeros(B, Nq, D, dtype=ft)
    if lse_out is not None:
        lse_out.fill_(-float("inf"))
    nblk = K.shape[0]
    pse = pse_shift.detach().cpu() if pse_shift is not None and pse_shift.numel() else None
    msk = attention_mask.detach().cpu() if attention_mask is not None and attention_mask.numel() else None
    if msk is not None:
        msk = msk.reshape(msk.shape[0], msk.shape[-1]) != 0
    attended = torch.zeros(B, Nq, dtype=torch.bool)
    for b in range(len(lists)):
        for hq in range(Nq):
            ent = lists[b][hq]
            if len(ent) == 0:
                continue
            n2 = hq // g
            pages = torch.tensor([e[0] for e in ent], dtype=torch.int64)
            ntok = torch.tensor([e[1] for e in ent], dtype=torch.int64)
            phys = bt[b, pages]
            ok = (phys >= 0) & (phys < nblk)          # a block id outside the pool is skipped (include/spa_b200.h)
            physc = phys.clamp(0, nblk - 1)
            Kp = K[physc, :, n2, :].to(ft)            # [n, bs, D]
            Vp = V[physc, :, n2, :].to(ft)
            if is_i8:
                Kp = (Kp + of[0, n2].to(ft)) * sc[0, n2].to(ft)
                Vp = (Vp + of[1, n2].to(ft)) * sc[1, n2].to(ft)
            tokmask = (torch.arange(block_size).view(1, -1) < ntok.view(-1, 1)) & ok.view(-1, 1)   # [n, bs]
            Kf = Kp.reshape(-1, D)
            Vf = Vp.reshape(-1, D)
            s = (Kf @ q[b, hq].to(ft)) * float(scale_value)
            if pse is not None or msk is not None:
                pos = (pages.view(-1, 1) * block_size + torch.arange(block_size).view(1, -1)).reshape(-1)
                if pse is not None:
                    s = s + pse[b if pse.shape[0] > 1 else 0, hq, 0, pos].to(ft)
                if msk is not None:
                    s = torch.where(msk[b, pos], torch.full_like(s, -1e12), s)
            s = torch.where(tokmask.reshape(-1), s, torch.full_like(s, -float("inf")))
            m = s.max()
            if not torch.isfinite(m):
                continue
            attended[b, hq] = True
            p = torch.exp(s - m)
            l = p.sum()
            if lse_out is not None:
                lse_out[b, hq] = float(torch.log(l) + m)
            if round_p and not dtype64:
                p = p.to(qd if is_i8 else K.dtype).to(ft)
            o = (p @ Vf) / l
            out[b, hq] = o
    if quant_scale2 is not None and quant_scale2.numel():
        sc2 = quant_scale2.detach().cpu().to(ft).reshape(-1)
        of2 = quant_offset2.detach().cpu().to(ft).reshape(-1) if quant_offset2 is not None and quant_offset2.numel() else torch.zeros(1, dtype=ft)
        sc2 = sc2.view(1, Nq, D) if sc2.numel() > 1 else sc2.view(1, 1, 1)
        of2 = of2.view(1, Nq, D) if of2.numel() > 1 else of2.view(1, 1, 1)
        qv = torch.round(out * sc2 + of2).clamp(-128, 127)
        qv = torch.where(attended.view(B, Nq, 1), qv, torch.zeros_like(qv))
        return qv.to(torch.int8).reshape(shape)
    if dtype64:
        return out.reshape(shape)
    return out.to(qd).reshape(shape)


def lists_from_block_position(block_position: torch.Tensor, actual_seq_lengths, Nq: int, Nkv: int,
                              block_size: int, num_logical_pages: int):
    """Per-(b,hq) page lists for the SEPARATE op `sparse_paged_attention`.

    block_position is indexed per KV head in the kernel (..._Us2.h:258,277-279) but the fused tests
    hand it per Q head (fused_ifa_v5_case.py:76,246): dim1 == Nkv -> shared by the GQA group,
    dim1 == Nq -> per Q head.  Entry j covers sparse-sequence tokens [j*bs, (j+1)*bs) clipped to
    actual_seq_lengths[b] (..._Us2.h:241-279); the list ends at the first 0x7fffffff sentinel
    (..._Us2.h:293, fused: :311 `break`).  Reference quirk NOT replicated: heads shorter than the
    batch max are run over stale rows.
    """
    bp = block_position.detach().cpu().numpy()
    B, H, MP = bp.shape
    if torch.is_tensor(actual_seq_lengths):
        asl = actual_seq_lengths.detach().cpu().numpy().astype(np.int64).reshape(-1)
    else:
        asl = np.asarray(actual_seq_lengths, dtype=np.int64).reshape(-1)
    g = Nq // Nkv
    lists = []
    for b in range(B):
        a = int(asl[b] if asl.size > 1 else asl[0])
        per = []
        for hq in range(Nq):
            row = bp[b, hq if H == Nq else hq // g]
            ent = []
            for j in range(MP):
                if j * block_size >= a:
                    break
                lp = int(row[j])
                if lp == INT32_MAX or lp < 0 or lp >= num_logical_pages:
                    break
                ent.append((lp, min(block_size, a - j * block_size)))
            per.append(ent)
        lists.append(per)
    return lists


def sparse_paged_attention(query, key, value, actual_seq_lengths, block_table, block_position,
                           num_heads: int, scale_value: float, num_key_value_heads: int,
                           block_size: int, antiquant_scale=None, antiquant_offset=None,
                           round_p: bool = True, dtype64: bool = False, lse_out=None, pse_shift=None,
                           attention_mask=None, quant_scale2=None, quant_offset2=None) -> torch.Tensor:
    """custom_ops.sparse_paged_attention (add_custom.py:41-75)."""
    lists = lists_from_block_position(block_position, actual_seq_lengths, num_heads,
                                      num_key_value_heads, block_size, block_table.shape[1])
    return attend_lists(query, key, value, block_table, lists, num_key_value_heads, scale_value,
                        block_size, antiquant_scale, antiquant_offset, round_p, dtype64, lse_out,
                        pse_shift, attention_mask, quant_scale2, quant_offset2)


def lists_from_selection(page_position: torch.Tensor, page_position_length: torch.Tensor,
                         total_seq_len: torch.Tensor, block_size: int):
    """Per-(b,hq) page lists for the FUSED op: the golden chain's per-Q-head selections with their
    true lengths; the context's last logical page is clipped to total_seq_len (SURVEY 8(a) A5
    'correct semantics for the build')."""
    pp = page_position.detach().cpu().numpy()
    ln = page_position_length.detach().cpu().numpy()
    sl = total_seq_len.detach().cpu().numpy().astype(np.int64)
    B, Nq, _ = pp.shape
    lists = []
    for b in range(B):
        pl = int((sl[b] + block_size - 1) // block_size)
        per = []
        for hq in range(Nq):
            n = int(ln[b, hq, 0])
            ent = []
            for j in range(n):
                lp = int(pp[b, hq, j])
                nt = block_size
                if lp == pl - 1:
                    nt = int(sl[b] - (pl - 1) * block_size)
                ent.append((lp, nt))
            per.append(ent)
        lists.append(per)
    return lists


def sparse_paged_fusion_attention(query, key, value, block_table, l1_cent, block_ids, total_seq_len,
                                  num_heads: int, scale_value: float, num_key_value_heads: int,
                                  block_size: int = 128, k: int = 64, max_page_num: int = 256,
                                  antiquant_scale=None, antiquant_offset=None,
                                  round_p: bool = True, dtype64: bool = False, sink_pages: int = 0,
                                  recent_pages: int = 0, max_fraction_den: int = 0, lse_out=None,
                                  pse_shift=None, attention_mask=None, quant_scale2=None, quant_offset2=None):
    """custom_ops.sparse_paged_fusion_attention (add_custom.py:96-135): cent_select followed by the
    attention over the selected pages (sparse_paged_fusion_attention.cpp:56-76).  Returns
    (attention_out, block_position, page_position_length, max_page_position_length)."""
    B = total_seq_len.shape[0]
    pos, plen, mx = cent_select(query, l1_cent, block_ids, block_table, total_seq_len, k,
                                block_size, max_page_num, sink_pages, recent_pages, max_fraction_den)
    lists = lists_from_selection(pos, plen, total_seq_len, block_size)
    out = attend_lists(query[:B], key, value, block_table, lists, num_key_value_heads, scale_value,
                       block_size, antiquant_scale, antiquant_offset, round_p, dtype64, lse_out,
                       pse_shift, attention_mask, quant_scale2, quant_offset2)
    return out, pos, plen, mx


# --------------------------------------------------------------------------------------------
# Measurement helper: algorithmic bytes of one fused step
# --------------------------------------------------------------------------------------------
def algorithmic_bytes(page_position: torch.Tensor, page_position_length: torch.Tensor, Nkv: int,
                      C: int, D: int, block_size: int, kv_elem: int, q_elem: int, P: int,
                      max_page_num: int) -> dict:
    """bytes_alg = union-of-group selected K+V page slices + centroids + index tensors + q/out."""
    pp = page_position.detach().cpu().numpy()
    ln = page_position_length.detach().cpu().numpy()[:, :, 0]
    B, Nq, _ = pp.shape
    g = Nq // Nkv
    union = 0
    for b in range(B):
        for n2 in range(Nkv):
            s = set()
            for h in range(n2 * g, (n2 + 1) * g):
                s.update(pp[b, h, : ln[b, h]].tolist())
            union += len(s)
    page_bytes = block_size * D * 2 * kv_elem
    bytes_kv = union * page_bytes
    bytes_logical = int(ln.sum()) * page_bytes
    bytes_sel = Nkv * C * D * q_elem + Nkv * B * P * 4 + B * P * 4 + B * Nq * (max_page_num + 8) * 4 + B * 8 * 8
    bytes_io = 2 * B * Nq * D * q_elem
    return {"bytes_alg": bytes_kv + bytes_sel + bytes_io, "bytes_kv_union": bytes_kv,
            "bytes_logical": bytes_logical, "union_pages": union, "selected_pages": int(ln.sum())}


# --------------------------------------------------------------------------------------------
# centroid maintenance (page -> cluster assignment).  No reference counterpart (its tests
# draw block_ids at random, tests/cent_select.py:49-50); this restates the arithmetic the CUDA kernel
# defines (csrc/spa_assign.cuh) with the metric of ComputeCent (compute_cent.cpp:277-343).
# --------------------------------------------------------------------------------------------
def assign_page_clusters(key: torch.Tensor, l1_cent: torch.Tensor, page_ids: torch.Tensor,
                         block_ids: torch.Tensor, page_tokens: Optional[torch.Tensor] = None,
                         block_size: int = 128, antiquant_scale: Optional[torch.Tensor] = None,
                         antiquant_offset: Optional[torch.Tensor] = None) -> Tuple[torch.Tensor, torch.Tensor]:
    """Returns (block_ids updated copy, page_mean [n, Nkv, D] float32).

    pooled[d] = (s0 + s1) + (s2 + s3), s_w = float32 sum in token order of the valid tokens of the w-th
    quarter of the page; score[c] = float64 sum over d in order of pooled[d] * cent[c, d];
    block_ids[n2, page] = argmax_c score[c], ties -> lowest c.  Pages outside the pool are skipped.
    int8 pool: pooled = (integer sum + float32(ntok) * offset_k) * scale_k in float32 (x = (x_i8 + offset) * scale,
    ..._Us2.h:2142), antiquant tensors [2, Nkv, D] or [2] as the attention ops take them (row 0 = K)."""
    Nkv, C, D = l1_cent.shape
    K = _kv_view(key.detach().cpu(), Nkv, block_size, D)                      # [blocks, bs, Nkv, D]
    cent = l1_cent.detach().cpu().to(torch.float64).numpy()
    out = block_ids.detach().cpu().clone()
    ids = page_ids.detach().cpu().numpy().astype(np.int64).reshape(-1)
    toks = None if page_tokens is None else page_tokens.detach().cpu().numpy().astype(np.int64).reshape(-1)
    mean = np.zeros((ids.size, Nkv, D), dtype=np.float32)
    q4 = block_size // 4
    for i, page in enumerate(ids):
        if page < 0 or page >= K.shape[0] or page >= out.shape[1]:
            continue
        ntok = block_size if toks is None else int(min(max(toks[i], 0), block_size))
        if K.dtype == torch.int8:
            sc, of = _antiquant(antiquant_scale, antiquant_offset, Nkv, D)
            isum = K[page, :ntok].to(torch.int64).sum(dim=0).numpy().astype(np.float32)       # exact
            pooled = (isum + np.float32(ntok) * of[0].numpy().astype(np.float32)) * sc[0].numpy().astype(np.float32)
            pooled = pooled.astype(np.float32)
        else:
            kp = K[page].to(torch.float32).numpy()                            # [bs, Nkv, D]
            part = np.zeros((4, Nkv, D), dtype=np.float32)
            for w in range(4):
                for t in range(w * q4, min((w + 1) * q4, ntok)):
                    part[w] = part[w] + kp[t]                                 # float32, token order
            pooled = (part[0] + part[1]) + (part[2] + part[3])                # float32
        mean[i] = pooled / np.float32(ntok) if ntok > 0 else 0.0
        p64 = pooled.astype(np.float64)
        for n2 in range(Nkv):
            acc = np.zeros(C, dtype=np.float64)
            for d in range(D):
                acc = acc + p64[n2, d] * cent[n2, :, d]                       # sequential in d; products exact
            out[n2, page] = int(np.argmax(acc))                               # first maximum = lowest index
    return out, torch.from_numpy(mean)
