"""GPU parity tests of the inherited IFA options on the sparse kernels: attenMask, pseShift and
the int8 post-quant output (quantScale2 / quantOffset2) -- the optional inputs of the reference's sparse op
definition that can be non-empty (sparse_paged_attention_def.cpp:70-190; shape rules
sparse_paged_attention_tiling_check.cc:737-850, sparse_paged_attention_tiling.cc:566-647; kernel math
sparse_paged_attention_split_Bbn2s2_Us2.h:2862-2899, :3513-3600).  No reference test passes any of them
(tests/ifa_v5_case.py:133-146 and fused_ifa_v5_case.py:83-84 pass empty tensors), so the checks are: the oracle
(oracle/sparse_chain.py, same options), and -- independent of the oracle -- torch SDPA with an additive mask on a
dense walk, where the reference's position indexing and ours coincide (include/spa_b200.h).

They run in the options build of the chain kernel (csrc/spa_kernels_opt.cu); the calls go through the same
public surface as every other test: custom_ops -> custom_ops_lib -> C ABI -> sm_100a kernels.
"""
import pytest
import torch

from cases import make_case, to_device
from oracle import sparse_chain as oc

pytestmark = pytest.mark.gpu
DEV = "cuda"


@pytest.fixture(scope="module")
def ops():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    import custom_ops
    return custom_ops


def assert_close(got, want, tol, what=""):
    got, want = got.float().cpu(), want.float().cpu()
    assert not torch.isnan(got).any(), f"{what}: NaN in output"
    err = (got - want).abs()
    bad = err > tol + tol * want.abs()
    assert not bad.any(), f"{what}: {int(bad.sum())} elements outside atol=rtol={tol}; max_abs={err.max().item():.3e}"


def empties(g):
    return torch.empty(0, dtype=g["q"].dtype, device=g["q"].device), torch.empty(0, dtype=torch.int64, device=g["q"].device)


def fused(ops, g, k, pse=None, mask=None, q2=None, qo2=None, lse=None):
    e, e64 = empties(g)
    dev = g["q"].device
    B, Nq = g["seq"].shape[0], g["Nq"]
    bp = torch.full((B, Nq, 256), 7, dtype=torch.int32, device=dev)
    ppl = torch.full((B, Nq, 8), 7, dtype=torch.int32, device=dev)
    mppl = torch.full((B, 8), 7, dtype=torch.int64, device=dev)
    o = lambda t: e if t is None else t
    out, _, _ = ops.sparse_paged_fusion_attention(
        g["q"], g["key"], g["value"], o(pse), o(mask), e64, e, e, e, o(q2), o(qo2), g["aq_scale"], g["aq_offset"],
        g["block_table"], e64, g["l1"], g["block_ids"], g["seq"], bp, ppl, mppl, Nq, g["scale"], "BNSD", g["Nkv"],
        g["bs"], 1, k=k, softmax_lse=lse)
    return out, bp, ppl, mppl


def separate(ops, g, bp, asl, pse=None, mask=None, q2=None, qo2=None):
    e, e64 = empties(g)
    o = lambda t: e if t is None else t
    return ops.sparse_paged_attention(
        g["q"], g["key"], g["value"], o(pse), o(mask), asl, e, e, e, o(q2), o(qo2), g["aq_scale"], g["aq_offset"],
        g["block_table"], e64, bp, g["Nq"], g["scale"], "BNSD", g["Nkv"], g["bs"], 1)


def make_mask_pse(c, seed, mask_dtype, pse_batch, frac=0.3, pad=0):
    """mask [B, 1, 1, S] (non-zero = masked, `frac` of the positions) and pse [B or 1, Nq, 1, S] ~ N(0, 1);
    S = maxBlocksPerBatch * blockSize + pad."""
    gen = torch.Generator().manual_seed(seed)
    B, Nq = c["seq"].shape[0], c["Nq"]
    smax = c["block_table"].shape[1] * c["bs"] + pad
    mask = (torch.rand(B, 1, 1, smax, generator=gen) < frac).to(mask_dtype)
    pse = torch.randn(pse_batch if pse_batch else B, Nq, 1, smax, generator=gen).to(c["q"].dtype)
    return mask, pse


def test_mask_and_pse_on_a_dense_walk_match_sdpa(ops):
    """Identity block_position: every page in order -- the reference's indexing (position in the walked sequence)
    and ours (logical position) are the same thing, and torch SDPA with an additive mask is an independent check."""
    B, Nq, Nkv, S = 2, 32, 8, 4096
    P = S // 128
    c = make_case(193, B, Nq, Nkv, 64, S, kv_pages=2 * B * P, permute_table=True)
    g = to_device(c, DEV)
    mask, pse = make_mask_pse(c, 5, torch.bool, 0)
    bp = torch.arange(P, dtype=torch.int32).expand(B, Nkv, P).contiguous()
    asl = torch.full((B,), S - 77, dtype=torch.int64)                 # a partly filled last page as well
    got = separate(ops, g, bp.to(DEV), asl.to(DEV), pse=pse.to(DEV), mask=mask.to(DEV))
    want = oc.sparse_paged_attention(c["q"], c["key"], c["value"], asl, c["block_table"], bp, Nq, c["scale"], Nkv, 128,
                                     pse_shift=pse, attention_mask=mask)
    assert_close(got, want, 1e-3, "dense walk, mask + pse")
    bt = c["block_table"].long()
    K = c["key"][bt[:B, :P]].reshape(B, S, Nkv, 128).permute(0, 2, 1, 3).float()
    V = c["value"][bt[:B, :P]].reshape(B, S, Nkv, 128).permute(0, 2, 1, 3).float()
    qf = c["q"].float().view(B, Nkv, Nq // Nkv, 128)
    bias = pse.float().view(B, Nkv, Nq // Nkv, S).clone()
    bias.masked_fill_(mask.view(B, 1, 1, S), -1e12)
    bias[..., S - 77:] = -float("inf")
    ref = torch.nn.functional.scaled_dot_product_attention(qf, K, V, attn_mask=bias, scale=c["scale"]).reshape(B, Nq, 128)
    assert_close(got, ref, 2e-3, "dense walk vs SDPA with an additive mask")
    # the options one at a time
    got_p = separate(ops, g, bp.to(DEV), asl.to(DEV), pse=pse.to(DEV))
    want_p = oc.sparse_paged_attention(c["q"], c["key"], c["value"], asl, c["block_table"], bp, Nq, c["scale"], Nkv, 128, pse_shift=pse)
    assert_close(got_p, want_p, 1e-3, "dense walk, pse alone")
    got_m = separate(ops, g, bp.to(DEV), asl.to(DEV), mask=mask.to(DEV))
    want_m = oc.sparse_paged_attention(c["q"], c["key"], c["value"], asl, c["block_table"], bp, Nq, c["scale"], Nkv, 128, attention_mask=mask)
    assert_close(got_m, want_m, 1e-3, "dense walk, mask alone")


@pytest.mark.parametrize("kind,mask_dtype,pse_batch", [("f16", torch.bool, 0), ("bf16", torch.uint8, 1),
                                                        ("i8", torch.int8, 0), ("g16", torch.bool, 1), ("fp8", torch.uint8, 0)])
def test_mask_and_pse_fused_and_separate_vs_oracle(ops, kind, mask_dtype, pse_batch):
    """Fused op and separate chain with attenMask + pseShift at LOGICAL positions, ragged lengths, items cut into
    chunks (partials + combine): selection outputs bit-exact (the options do not touch the selection), attention
    within the dtype's tolerance, softmax LSE included.  One sequence is fully masked: its rows average the
    attended tokens (a masked logit is -1e12, not -inf)."""
    qdtype = torch.bfloat16 if kind in ("bf16", "i8") else torch.float16
    B, Nq, Nkv, C, S, k = (2, 32, 2, 64, 6000, 8) if kind == "g16" else (3, 32, 8, 128, 9000, 16)
    c = make_case(707, B, Nq, Nkv, C, S, kv_pages=B * 72 + 3, ragged=True, permute_table=True, qdtype=qdtype,
                  kv_int8=kind == "i8", kv_fp8=kind == "fp8")
    if kind == "fp8":
        c["scale"] = 0.02
    tol = 1e-3 if kind in ("f16", "g16") else 2e-2
    g = to_device(c, DEV)
    mask, pse = make_mask_pse(c, 11, mask_dtype, pse_batch)
    mask[B - 1] = 1                                                   # the last sequence: everything masked
    aq = (c["aq_scale"], c["aq_offset"])
    lse_w = torch.empty(B, Nq, dtype=torch.float32)
    want = oc.sparse_paged_fusion_attention(c["q"], c["key"], c["value"], c["block_table"], c["l1"], c["block_ids"],
                                            c["seq"], Nq, c["scale"], Nkv, 128, k, 256, *aq, lse_out=lse_w,
                                            pse_shift=pse, attention_mask=mask)
    lse = torch.empty(B, Nq, dtype=torch.float32, device=DEV)
    out, bp, ppl, mppl = fused(ops, g, k, pse=pse.to(DEV), mask=mask.to(DEV), lse=lse)
    assert torch.equal(bp.cpu(), want[1]) and torch.equal(ppl.cpu(), want[2]) and torch.equal(mppl.cpu(), want[3])
    assert_close(out, want[0], tol, f"fused, mask + pse, {kind}")
    fin = torch.isfinite(lse_w)
    assert torch.equal(torch.isfinite(lse.cpu()), fin)
    rel = (lse.cpu()[fin] - lse_w[fin]).abs() / lse_w[fin].abs().clamp_min(1.0)
    assert bool((rel < (2e-3 if tol == 1e-3 else 2e-2)).all()), f"softmax LSE with mask + pse, {kind}: {rel.max().item():.3e}"
    asl = mppl[:, 0].contiguous()
    out_s = separate(ops, g, bp, asl, pse=pse.to(DEV), mask=mask.to(DEV))
    want_s = oc.sparse_paged_attention(c["q"], c["key"], c["value"], asl.cpu(), c["block_table"], bp.cpu(), Nq, c["scale"],
                                       Nkv, 128, *aq, pse_shift=pse, attention_mask=mask)
    assert_close(out_s, want_s, tol, f"separate, mask + pse, {kind}")
    # and the plain build on the same inputs is untouched by the options' existence
    plain, bp0, _, _ = fused(ops, g, k)
    want0 = oc.sparse_paged_fusion_attention(c["q"], c["key"], c["value"], c["block_table"], c["l1"], c["block_ids"],
                                             c["seq"], Nq, c["scale"], Nkv, 128, k, 256, *aq)
    assert torch.equal(bp0.cpu(), want0[1])
    assert_close(plain, want0[0], tol, f"plain build next to the options build, {kind}")


def test_mask_and_pse_rows_not_16_byte_aligned(ops):
    """S = maxBlocksPerBatch * blockSize + 3: the rows of both tensors start at odd offsets, the producer cannot stage
    them with bulk copies, and the consumers read them from global memory instead (one page per step)."""
    B, Nq, Nkv, C, S, k = 3, 32, 8, 128, 9000, 16
    c = make_case(747, B, Nq, Nkv, C, S, kv_pages=B * 72 + 3, ragged=True, permute_table=True)
    g = to_device(c, DEV)
    mask, pse = make_mask_pse(c, 19, torch.uint8, 0, pad=3)
    want = oc.sparse_paged_fusion_attention(c["q"], c["key"], c["value"], c["block_table"], c["l1"], c["block_ids"],
                                            c["seq"], Nq, c["scale"], Nkv, 128, k, 256, pse_shift=pse, attention_mask=mask)
    out, bp, ppl, _ = fused(ops, g, k, pse=pse.to(DEV), mask=mask.to(DEV))
    assert torch.equal(bp.cpu(), want[1]) and torch.equal(ppl.cpu(), want[2])
    assert_close(out, want[0], 1e-3, "unaligned mask / pse rows")


@pytest.mark.parametrize("bs", [16, 64])
def test_mask_and_pse_with_small_blocks(ops, bs):
    """Positions are page * block_size + t for every supported block size."""
    B, Nq, Nkv, C, k = 2, 16, 4, 64, 16
    S = 40 * bs + 5
    c = make_case(717 + bs, B, Nq, Nkv, C, S, kv_pages=2 * B * 41 + 5, ragged=True, permute_table=True, bs=bs)
    g = to_device(c, DEV)
    mask, pse = make_mask_pse(c, 13, torch.bool, 1)
    want = oc.sparse_paged_fusion_attention(c["q"], c["key"], c["value"], c["block_table"], c["l1"], c["block_ids"],
                                            c["seq"], Nq, c["scale"], Nkv, bs, k, 256, pse_shift=pse, attention_mask=mask)
    out, bp, ppl, _ = fused(ops, g, k, pse=pse.to(DEV), mask=mask.to(DEV))
    assert torch.equal(bp.cpu(), want[1]) and torch.equal(ppl.cpu(), want[2])
    assert_close(out, want[0], 1e-3, f"block size {bs}, mask + pse")


@pytest.mark.parametrize("kind", ["f16_tensor", "bf16_channel", "i8_channel_f32"])
def test_post_quant_int8_output(ops, kind):
    """quant_scale2 (+ quant_offset2): attention_out is int8 = sat(rint(o * scale2 + offset2)), parameters per tensor
    or per output channel [Nq, D], float32 or (bf16 query) bfloat16.  Compared with the oracle's quantisation of its
    float32 result: identical up to one step where o * scale2 + offset2 lies at a rounding boundary."""
    qdtype = torch.float16 if kind.startswith("f16") else torch.bfloat16
    B, Nq, Nkv, C, S, k = 3, 32, 8, 128, 9000, 16
    c = make_case(727, B, Nq, Nkv, C, S, kv_pages=B * 72 + 3, ragged=True, permute_table=True, qdtype=qdtype,
                  kv_int8=kind.startswith("i8"))
    g = to_device(c, DEV)
    gen = torch.Generator().manual_seed(3)
    aq = (c["aq_scale"], c["aq_offset"])
    plain = oc.sparse_paged_fusion_attention(c["q"], c["key"], c["value"], c["block_table"], c["l1"], c["block_ids"],
                                             c["seq"], Nq, c["scale"], Nkv, 128, k, 256, *aq)[0].float()
    span = 100.0 / plain.abs().max().item()                          # spreads the outputs over most of the int8 range
    if kind == "f16_tensor":
        q2, qo2 = torch.tensor([span], dtype=torch.float32), torch.tensor([-3.0], dtype=torch.float32)
    elif kind == "bf16_channel":
        q2 = (span * (0.5 + torch.rand(Nq, 128, generator=gen))).to(torch.bfloat16)
        qo2 = (8 * torch.randn(Nq, 128, generator=gen)).to(torch.bfloat16)
    else:
        q2, qo2 = (span * (0.5 + torch.rand(Nq, 128, generator=gen))).float(), None
    want = oc.sparse_paged_fusion_attention(c["q"], c["key"], c["value"], c["block_table"], c["l1"], c["block_ids"],
                                            c["seq"], Nq, c["scale"], Nkv, 128, k, 256, *aq, quant_scale2=q2, quant_offset2=qo2)
    dq = lambda t: None if t is None else t.to(DEV)
    out, bp, ppl, mppl = fused(ops, g, k, q2=dq(q2), qo2=dq(qo2))
    assert out.dtype == torch.int8 and out.shape == c["q"].shape
    assert torch.equal(bp.cpu(), want[1]) and torch.equal(ppl.cpu(), want[2])
    d = (out.cpu().int() - want[0].int()).abs()
    # fp16 / bf16 arithmetic differences of the attention itself move o * scale2 by up to tol * 100 steps
    lim = 1 if kind == "f16_tensor" else 3
    assert int(d.max()) <= lim, f"post-quant {kind}: max step difference {int(d.max())}"
    assert float((d == 0).float().mean()) > (0.9 if kind == "f16_tensor" else 0.5), f"post-quant {kind}: exact share {(d == 0).float().mean():.3f}"
    assert int(out.cpu().int().abs().max()) > 60                     # the range is used: not a trivially small output
    asl = mppl[:, 0].contiguous()
    out_s = separate(ops, g, bp, asl, q2=dq(q2), qo2=dq(qo2))
    want_s = oc.sparse_paged_attention(c["q"], c["key"], c["value"], asl.cpu(), c["block_table"], bp.cpu(), Nq, c["scale"],
                                       Nkv, 128, *aq, quant_scale2=q2, quant_offset2=qo2)
    ds = (out_s.cpu().int() - want_s.int()).abs()
    assert out_s.dtype == torch.int8 and int(ds.max()) <= lim


def test_post_quant_with_mask_and_pse_and_empty_rows(ops):
    """All three options in one call; a sequence of length 0 gives all-zero int8 rows."""
    B, Nq, Nkv, C, S, k = 3, 16, 4, 64, 5000, 8
    c = make_case(737, B, Nq, Nkv, C, S, kv_pages=B * 40 + 3, permute_table=True)
    c["seq"][1] = 0
    g = to_device(c, DEV)
    mask, pse = make_mask_pse(c, 17, torch.bool, 0)
    q2, qo2 = torch.tensor([150.0]), torch.tensor([5.0])
    want = oc.sparse_paged_fusion_attention(c["q"], c["key"], c["value"], c["block_table"], c["l1"], c["block_ids"],
                                            c["seq"], Nq, c["scale"], Nkv, 128, k, 256, pse_shift=pse, attention_mask=mask,
                                            quant_scale2=q2, quant_offset2=qo2)
    out, bp, ppl, _ = fused(ops, g, k, pse=pse.to(DEV), mask=mask.to(DEV), q2=q2.to(DEV), qo2=qo2.to(DEV))
    assert torch.equal(bp.cpu(), want[1]) and torch.equal(ppl.cpu(), want[2])
    assert int(out[1].cpu().int().abs().max()) == 0, "rows of an empty sequence must be zeros"
    d = (out.cpu().int() - want[0].int()).abs()
    assert int(d.max()) <= 1 and float((d == 0).float().mean()) > 0.9


def test_decode_backend_with_bias_and_mask_and_graph_replay(ops):
    """SparseDecodeBackend.bind_bias: the persistent pse_shift / attention_mask of a layer go into every step's launch
    (eager and CUDA-graph replay, two batch sizes), and unbinding returns to the plain kernel."""
    from decode_backend import SparseDecodeBackend
    B, Nq, Nkv, C, S, k = 4, 16, 4, 64, 5000, 8
    c = make_case(757, B, Nq, Nkv, C, S, kv_pages=B * 40 + 3, ragged=True, permute_table=True)
    g = to_device(c, DEV)
    P = c["block_table"].shape[1]
    be = SparseDecodeBackend(Nq, Nkv, 128, 128, k, max_batch=B, max_blocks_per_seq=P, dtype=torch.float16, device=DEV,
                             scale=c["scale"], max_page_num=256)
    be.bind_cache(g["key"], g["value"], g["l1"], g["block_ids"])
    mask, pse = make_mask_pse(c, 23, torch.bool, 1)
    be.bind_bias(pse.to(DEV), mask.to(DEV))
    for nb, graph in ((B, False), (2, False), (B, True), (B, True)):
        out = be.forward_decode(g["q"][:nb], g["block_table"][:nb], g["seq"][:nb], use_graph=graph)
        want = oc.sparse_paged_fusion_attention(c["q"][:nb], c["key"], c["value"], c["block_table"][:nb], c["l1"], c["block_ids"],
                                                c["seq"][:nb], Nq, c["scale"], Nkv, 128, k, 256, pse_shift=pse,
                                                attention_mask=mask[:nb])
        assert torch.equal(be.selection(nb)[0].cpu(), want[1])
        assert_close(out, want[0], 1e-3, f"backend with bias + mask, batch {nb}, graph={graph}")
    be.bind_bias(None, None)
    out = be.forward_decode(g["q"], g["block_table"], g["seq"])
    want = oc.sparse_paged_fusion_attention(c["q"], c["key"], c["value"], c["block_table"], c["l1"], c["block_ids"],
                                            c["seq"], Nq, c["scale"], Nkv, 128, k, 256)
    assert_close(out, want[0], 1e-3, "backend after unbinding the bias")
    be.close()
