"""Seeded synthetic inputs for the sparse decode-attention chain (SURVEY 8(d) 'Synthetic inputs')."""
import math

import torch


def make_case(seed, B, Nq, Nkv, C, S, kv_pages=None, maxP=None, maxB=None, D=128, bs=128,
              qdtype=torch.float16, kv_int8=False, cent_dist="rand", ragged=False, permute_table=False,
              MP=256, device="cpu", kv_fp8=False):
    """Shapes follow tests/cent_select.py:17-53 and fused_ifa_v5_case.py:21-78 of the reference:
    q ~ U[0,1), centroids ~ U[0,1) (or N(0,1)), block_ids ~ randint(0,C), block_table ~
    randint(0,kvPageLen) (collisions allowed) or a per-sequence permutation, K/V ~ U[1e-3,1)."""
    g = torch.Generator().manual_seed(seed)
    P = (S + bs - 1) // bs
    maxP = maxP or P
    maxB = maxB or B
    kv_pages = kv_pages or max(B * P, 8)
    q = torch.rand(B, Nq, D, generator=g).to(qdtype)
    if cent_dist == "rand":
        l1 = torch.rand(Nkv, C, D, generator=g).to(qdtype)
    else:
        l1 = torch.randn(Nkv, C, D, generator=g).to(qdtype)
    block_ids = torch.randint(0, C, (Nkv, kv_pages), generator=g, dtype=torch.int32)
    if permute_table:
        assert kv_pages >= B * maxP
        block_table = torch.randperm(kv_pages, generator=g)[: maxB * maxP].view(maxB, maxP).to(torch.int32)
    else:
        block_table = torch.randint(0, kv_pages, (maxB, maxP), generator=g, dtype=torch.int32)
    if ragged:
        seq = torch.tensor([max(1, S - 37 * b - 1) for b in range(B)], dtype=torch.int32)
    else:
        seq = torch.full((B,), S, dtype=torch.int32)
    eps = 1e-3
    if kv_int8:
        key = torch.randint(-128, 128, (kv_pages, bs, Nkv, D), generator=g, dtype=torch.int8)
        value = torch.randint(-128, 128, (kv_pages, bs, Nkv, D), generator=g, dtype=torch.int8)
        aq_scale = (torch.rand(2, Nkv, D, generator=g) * 0.02 + 0.005).to(qdtype)
        aq_offset = torch.randn(2, Nkv, D, generator=g).to(qdtype)
    elif kv_fp8:
        # FP8 e4m3 pages: signed values over a few octaves; scales optional (kv_fp8 == "scaled")
        key = (torch.randn(kv_pages, bs, Nkv, D, generator=g) * 2.0).to(torch.float8_e4m3fn)
        value = (torch.randn(kv_pages, bs, Nkv, D, generator=g) * 2.0).to(torch.float8_e4m3fn)
        if kv_fp8 == "scaled":
            aq_scale = (torch.rand(2, Nkv, D, generator=g) * 0.5 + 0.25).to(qdtype)
            aq_offset = (torch.randn(2, Nkv, D, generator=g) * 0.1).to(qdtype)
        else:
            aq_scale = torch.empty(0, dtype=qdtype)
            aq_offset = torch.empty(0, dtype=qdtype)
    else:
        key = (torch.rand(kv_pages, bs, Nkv, D, generator=g) * (1 - eps) + eps).to(qdtype)
        value = (torch.rand(kv_pages, bs, Nkv, D, generator=g) * (1 - eps) + eps).to(qdtype)
        aq_scale = torch.empty(0, dtype=qdtype)
        aq_offset = torch.empty(0, dtype=qdtype)
    case = dict(q=q, l1=l1, block_ids=block_ids, block_table=block_table, seq=seq, key=key, value=value,
                aq_scale=aq_scale, aq_offset=aq_offset, B=B, Nq=Nq, Nkv=Nkv, C=C, S=S, D=D, bs=bs, MP=MP,
                scale=1.0 / math.sqrt(D))
    if device != "cpu":
        for k_, v in list(case.items()):
            if torch.is_tensor(v):
                case[k_] = v.to(device)
    return case


def to_device(case, device):
    return {k: (v.to(device) if torch.is_tensor(v) else v) for k, v in case.items()}


def fused_call(ops, c, k, garbage=True):
    """Call custom_ops.sparse_paged_fusion_attention the way the reference test does
    (fused_ifa_v5_case.py:175-190): aux outputs pre-allocated by the caller (pre-filled with garbage
    to prove they are fully overwritten)."""
    dev = c["q"].device
    B, Nq, MP = c["seq"].shape[0], c["Nq"], c["MP"]
    fill = 0x5A5A5A5A if garbage else 0
    bp = torch.full((B, Nq, MP), fill, dtype=torch.int32, device=dev)
    ppl = torch.full((B, Nq, 8), fill, dtype=torch.int32, device=dev)
    mppl = torch.full((B, 8), fill, dtype=torch.int64, device=dev)
    e = torch.empty(0, dtype=c["q"].dtype, device=dev)
    e64 = torch.empty(0, dtype=torch.int64, device=dev)
    out, bp2, mppl2 = ops.sparse_paged_fusion_attention(
        c["q"], c["key"], c["value"], e, e, [int(c["S"])], e, e, e, e, e, c["aq_scale"], c["aq_offset"],
        c["block_table"], e64, c["l1"], c["block_ids"], c["seq"], bp, ppl, mppl,
        Nq, c["scale"], "BNSD", c["Nkv"], c["bs"], 1, k=k)
    return out, bp, ppl, mppl, (bp2, mppl2)


def attn_call(ops, c, block_position, actual_seq_lengths):
    dev = c["q"].device
    e = torch.empty(0, dtype=c["q"].dtype, device=dev)
    e64 = torch.empty(0, dtype=torch.int64, device=dev)
    return ops.sparse_paged_attention(
        c["q"], c["key"], c["value"], e, e, actual_seq_lengths, e, e, e, e, e, c["aq_scale"], c["aq_offset"],
        c["block_table"], e64, block_position, c["Nq"], c["scale"], "BNSD", c["Nkv"], c["bs"], 1)
