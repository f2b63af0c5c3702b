"""Vectorised CPU variant of the fused chain: the "fair CPU" baseline of BASELINE.md section 3 / SURVEY 8(d).

TEST / BASELINE INFRASTRUCTURE ONLY (like oracle/sparse_chain.py): nothing in the product path imports it; only
`tests/` and bench.py's `cpu_baseline` / `--impl reference` legs do.

`oracle/sparse_chain.py` restates the reference golden as written -- O(B * Nq) Python loops
(tests/utils.py:4-123) -- which is what the reference itself would time on a CPU.  This module computes the same
chain with batched torch ops on all host threads (loops only over the B * Nkv work items of the attention), so
that "the reference as written" and "a CPU given a fair chance" are two separate numbers.  Selection follows the
golden's arithmetic (fp32 matmul + torch.topk, tests/utils.py:104-123), attention the kernel's math order
(sparse_paged_attention_split_Bbn2s2_Us2.h:2854-2914, 3062, 3467-3494); tests/test_oracle_golden.py pins it to
the loop oracle.
"""
import torch

INT32_MAX = 0x7FFFFFFF


def fused_chain(query, key, value, block_table, l1_cent, block_ids, seq_len, num_heads, scale_value,
                num_kv_heads, block_size, k, max_page_num, antiquant_scale=None, antiquant_offset=None):
    """-> (attention_out [B,Nq,D] in the query dtype, page_position [B,Nq,MP] i32, page_position_length [B,Nq,8]
    i32, max_page_position_length [B,8] i64).  KV pool layout [blocks, bs, Nkv, D]."""
    B = seq_len.shape[0]
    Nq, Nkv, bs, MP = num_heads, num_kv_heads, block_size, max_page_num
    g = Nq // Nkv
    D = query.shape[-1]
    q = query.reshape(query.shape[0], Nq, D)[:B]
    C = l1_cent.shape[1]
    P = block_table.shape[1]
    # ---- scores + top-k (tests/utils.py:104-123: fp32 matmul, torch.topk)
    s = torch.einsum("bngd,ncd->bngc", q.float().view(B, Nkv, g, D), l1_cent.float())
    idx = s.topk(k, dim=-1).indices
    sel = torch.zeros(B, Nkv, g, C, dtype=torch.bool).scatter_(-1, idx, True)
    # ---- membership of every logical page (tests/utils.py:57-80)
    pl = ((seq_len.long() + bs - 1) // bs).clamp(max=P)
    bt = block_table[:B].long()
    cl = block_ids.long()[:, bt.clamp(0, block_ids.shape[1] - 1)].permute(1, 0, 2)      # [B, Nkv, P]
    member = sel.gather(-1, cl.unsqueeze(2).expand(B, Nkv, g, P))
    member &= (torch.arange(P).view(1, 1, 1, P) < pl.view(B, 1, 1, 1))
    member &= member.cumsum(-1) <= MP                                                  # first max_page_num pages
    ln = member.sum(-1)                                                                # [B, Nkv, g]
    keyed = torch.where(member, torch.arange(P).view(1, 1, 1, P), torch.full((1,), INT32_MAX, dtype=torch.long))
    pos = keyed.sort(dim=-1).values[..., :MP]
    if pos.shape[-1] < MP:
        pos = torch.cat([pos, torch.full(pos.shape[:-1] + (MP - pos.shape[-1],), INT32_MAX, dtype=torch.long)], -1)
    page_position = pos.reshape(B, Nq, MP).to(torch.int32)
    ppl = ln.reshape(B, Nq, 1).expand(B, Nq, 8).to(torch.int32).contiguous()
    mppl = (ln.reshape(B, Nq).amax(dim=1) * bs).view(B, 1).expand(B, 8).to(torch.int64).contiguous()
    # ---- attention over the union of the group's pages, one (b, kv head) item at a time
    i8 = key.dtype in (torch.int8, torch.float8_e4m3fn)
    if i8 and (antiquant_scale is None or antiquant_scale.numel() == 0):
        sc, of = torch.ones(2, 1, 1), torch.zeros(2, 1, 1)                              # FP8 pages without scales
    elif i8:
        sc = antiquant_scale.float().reshape(2, -1, D) if antiquant_scale.numel() > 2 else antiquant_scale.float().view(2, 1, 1)
        of = (antiquant_offset.float().reshape(2, -1, D) if antiquant_offset.numel() > 2 else antiquant_offset.float().view(2, 1, 1)) \
            if antiquant_offset is not None and antiquant_offset.numel() else torch.zeros_like(sc)
    pdt = query.dtype if i8 else key.dtype
    out = torch.zeros(B, Nq, D)
    nblk = key.shape[0]
    tok = torch.arange(bs)
    for b in range(B):
        S = int(seq_len[b])
        for n2 in range(Nkv):
            m = member[b, n2]                                                          # [g, P]
            pages = m.any(0).nonzero().flatten()
            phys = bt[b, pages]
            okp = (phys >= 0) & (phys < nblk)
            pages, phys = pages[okp], phys[okp]
            if pages.numel() == 0:
                continue
            Kp = key[phys, :, n2, :].float()                                           # [n, bs, D]
            Vp = value[phys, :, n2, :].float()
            if i8:
                Kp = (Kp + of[0, n2 if of.shape[1] > 1 else 0]) * sc[0, n2 if sc.shape[1] > 1 else 0]
                Vp = (Vp + of[1, n2 if of.shape[1] > 1 else 0]) * sc[1, n2 if sc.shape[1] > 1 else 0]
            lg = torch.einsum("gd,ntd->gnt", q[b, n2 * g:(n2 + 1) * g].float(), Kp) * float(scale_value)
            live = m[:, pages].unsqueeze(-1) & ((pages.view(-1, 1) * bs + tok.view(1, -1)) < S).unsqueeze(0)
            lg = torch.where(live, lg, torch.full((1,), -float("inf")))
            mx = lg.flatten(1).amax(dim=1).view(g, 1, 1)
            p = torch.where(live, torch.exp(lg - torch.where(torch.isfinite(mx), mx, torch.zeros(1))), torch.zeros(1))
            l = p.flatten(1).sum(dim=1)
            p = p.to(pdt).float()                                                      # P rounded before PV (:3062)
            o = torch.einsum("gnt,ntd->gd", p, Vp)
            out[b, n2 * g:(n2 + 1) * g] = torch.where(l.view(g, 1) > 0, o / l.clamp_min(1e-38).view(g, 1), torch.zeros(1))
    return out.to(query.dtype), page_position, ppl, mppl
