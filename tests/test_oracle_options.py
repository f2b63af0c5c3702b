"""CPU checks of the oracle's inherited-IFA options (pse_shift, attention_mask, int8 post-quant;
oracle/sparse_chain.py: attend_lists) against plain torch, so that the GPU parity tests of those options
(tests/test_gpu_options.py) compare with a checker that has itself been checked.  No reference test passes any of the
options (tests/ifa_v5_case.py:133-146), so the anchor is the formula the reference kernel applies
(sparse_paged_attention_split_Bbn2s2_Us2.h:2854-2899: scale, + pse, masked -> -1e12, softmax; :3513-3600 post-quant)
evaluated by torch.nn.functional.scaled_dot_product_attention on a dense walk, where position indexing is unambiguous.
"""
import torch

from cases import make_case
from oracle import sparse_chain as oc


def dense_case(seed=5, B=2, Nq=8, Nkv=2, S=640, tail=37, qdtype=torch.float16):
    P = S // 128
    c = make_case(seed, B, Nq, Nkv, 16, S, kv_pages=2 * B * P, permute_table=True, qdtype=qdtype)
    bp = torch.arange(P, dtype=torch.int32).expand(B, Nkv, P).contiguous()
    asl = torch.full((B,), S - tail, dtype=torch.int64)
    gen = torch.Generator().manual_seed(seed + 1)
    mask = torch.rand(B, 1, 1, S, generator=gen) < 0.3
    pse = torch.randn(B, Nq, 1, S, generator=gen).to(qdtype)
    return c, bp, asl, mask, pse, P


def sdpa(c, P, S, valid, bias):
    B, Nq, Nkv = c["seq"].shape[0], c["Nq"], c["Nkv"]
    bt = c["block_table"].long()
    K = c["key"][bt[:B, :P]].reshape(B, S, Nkv, 128).permute(0, 2, 1, 3).float()
    V = c["value"][bt[:B, :P]].reshape(B, S, Nkv, 128).permute(0, 2, 1, 3).float()
    q = c["q"].float().view(B, Nkv, Nq // Nkv, 128)
    bias = bias.clone()
    bias[..., valid:] = -float("inf")
    return torch.nn.functional.scaled_dot_product_attention(q, K, V, attn_mask=bias, scale=c["scale"]).reshape(B, Nq, 128)


def test_oracle_mask_and_pse_match_sdpa_on_a_dense_walk():
    c, bp, asl, mask, pse, P = dense_case()
    B, Nq, Nkv, S = 2, c["Nq"], c["Nkv"], P * 128
    valid = int(asl[0])
    for use_pse, use_mask in ((True, True), (True, False), (False, True)):
        got = oc.sparse_paged_attention(c["q"], c["key"], c["value"], asl, c["block_table"], bp, Nq, c["scale"], Nkv, 128,
                                        pse_shift=pse if use_pse else None, attention_mask=mask if use_mask else None,
                                        round_p=False)
        bias = pse.float().view(B, Nkv, Nq // Nkv, S).clone() if use_pse else torch.zeros(B, Nkv, Nq // Nkv, S)
        if use_mask:
            bias = bias.masked_fill(mask.view(B, 1, 1, S), -1e12)
        want = sdpa(c, P, S, valid, bias)
        err = (got.float() - want).abs().max().item()
        assert err < 2e-3, f"pse={use_pse} mask={use_mask}: max err {err:.3e}"    # (the oracle's output is rounded to fp16)


def test_oracle_fully_masked_row_averages_its_tokens():
    """A masked logit is -1e12, not -inf (..._Us2.h:757): a row whose tokens are all masked is their plain average."""
    c, bp, asl, mask, pse, P = dense_case(seed=9)
    B, Nq, Nkv, S = 2, c["Nq"], c["Nkv"], P * 128
    mask[1] = True
    got = oc.sparse_paged_attention(c["q"], c["key"], c["value"], asl, c["block_table"], bp, Nq, c["scale"], Nkv, 128,
                                    attention_mask=mask, round_p=False)
    bt = c["block_table"].long()
    V = c["value"][bt[1, :P]].reshape(S, Nkv, 128)[: int(asl[1])].float().mean(0)     # [Nkv, D]
    want = V.repeat_interleave(Nq // Nkv, dim=0)
    assert (got[1].float() - want).abs().max().item() < 2e-3


def test_oracle_pse_batch_one_is_shared_and_mask_accepts_integer_dtypes():
    c, bp, asl, mask, pse, P = dense_case(seed=11)
    Nq, Nkv = c["Nq"], c["Nkv"]
    a = oc.sparse_paged_attention(c["q"], c["key"], c["value"], asl, c["block_table"], bp, Nq, c["scale"], Nkv, 128,
                                  pse_shift=pse[:1].expand(2, -1, -1, -1).contiguous(), attention_mask=mask)
    b = oc.sparse_paged_attention(c["q"], c["key"], c["value"], asl, c["block_table"], bp, Nq, c["scale"], Nkv, 128,
                                  pse_shift=pse[:1], attention_mask=mask.to(torch.uint8).view(2, -1))
    assert torch.equal(a, b)


def test_oracle_post_quant_formula_and_empty_rows():
    c, bp, asl, mask, pse, P = dense_case(seed=13)
    Nq, Nkv = c["Nq"], c["Nkv"]
    asl[1] = 0                                                        # nothing attended: zeros, whatever the offset
    plain = oc.sparse_paged_attention(c["q"], c["key"], c["value"], asl, c["block_table"], bp, Nq, c["scale"], Nkv, 128,
                                      dtype64=True)
    gen = torch.Generator().manual_seed(2)
    sc2 = 200.0 * (0.5 + torch.rand(Nq, 128, generator=gen))
    of2 = 5.0 * torch.randn(Nq, 128, generator=gen)
    got = oc.sparse_paged_attention(c["q"], c["key"], c["value"], asl, c["block_table"], bp, Nq, c["scale"], Nkv, 128,
                                    quant_scale2=sc2, quant_offset2=of2)
    assert got.dtype == torch.int8 and got.shape == c["q"].shape
    want = torch.round(plain.float() * sc2 + of2).clamp(-128, 127)
    d = (got[0].float() - want[0]).abs()
    assert d.max().item() <= 1 and (d == 0).float().mean().item() > 0.95     # (fp32 vs fp64 softmax at rounding boundaries)
    assert int(got[1].abs().max()) == 0
    assert int(got[0].int().abs().max()) == 128 or int(got[0].int().abs().max()) > 100     # the scale fills the int8 range
    one = oc.sparse_paged_attention(c["q"], c["key"], c["value"], asl, c["block_table"], bp, Nq, c["scale"], Nkv, 128,
                                    quant_scale2=torch.tensor([150.0]))
    w1 = torch.round(plain.float() * 150.0).clamp(-128, 127)
    assert (one[0].float() - w1[0]).abs().max().item() <= 1
