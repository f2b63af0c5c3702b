"""GPU parity tests (run on the B200 box: `pytest -m gpu`).

Every test calls the product through its public surface -- `custom_ops.*` -> `custom_ops_lib`
(C++ binding) -> C ABI (`libspa_b200.so`) -> sm_100a kernels -- and compares with the CPU oracle
(`oracle/sparse_chain.py`) or the committed reference-generated fixtures (`tests/golden/*.npz`).
Bars: indices / page_position / lengths bit-exact; attention_out atol=rtol=1e-3 (fp16),
2e-2 (bf16 or int8 KV)  -- BASELINE.json north_star.
"""
import os

import numpy as np
import pytest
import torch

from cases import attn_call, fused_call, make_case, to_device
from golden_check import check_indices, rows_with_defined_positions
from oracle import sparse_chain as oc

pytestmark = pytest.mark.gpu
DEV = "cuda"


@pytest.fixture(scope="module")
def ops():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    import custom_ops
    return custom_ops


def load(golden_dir, name):
    z = np.load(os.path.join(golden_dir, name + ".npz"))
    return {k: (torch.from_numpy(z[k]) if z[k].ndim > 0 else z[k].item()) for k in z.files}


def crop_pad(pos, width):
    B, H, W = pos.shape
    out = torch.full((B, H, width), oc.INT32_MAX, dtype=torch.int32)
    w = min(W, width)
    out[:, :, :w] = pos[:, :, :w]
    return out


def assert_close(got, want, tol, what=""):
    got = got.float().cpu()
    want = want.float().cpu()
    assert not torch.isnan(got).any(), f"{what}: NaN in output"
    err = (got - want).abs()
    bound = tol + tol * want.abs()
    bad = err > bound
    assert not bad.any(), (f"{what}: {int(bad.sum())} elements outside atol=rtol={tol}; "
                           f"max_abs={err.max().item():.3e}")


# ------------------------------------------------------------------------------------------
# selection ops against the reference-generated fixtures (k = 64, the reference constant)
# ------------------------------------------------------------------------------------------
# chain_refshape: the reference's own test shape in full (tests/compute_cent.py:16-22, tests/cent_select.py:17-39)
CHAINS = ["chain_small", "chain_32k", "chain_ragged", "chain_randn", "chain_refshape"]


@pytest.mark.parametrize("name", CHAINS)
def test_compute_cent_golden(ops, golden_dir, name):
    f = load(golden_dir, name)
    got = ops.compute_cent(f["q"].to(DEV), f["l1"].to(DEV))
    assert got.dtype == torch.int32 and tuple(got.shape) == tuple(f["ref_indices"].shape)
    check_indices(got, f, f"compute_cent {name}")          # bit-exact except on rows the fixture flags ambiguous


@pytest.mark.parametrize("name", CHAINS + ["select_dupes"])
def test_select_position_golden(ops, golden_dir, name):
    f = load(golden_dir, name)
    Bs = f["seq_len"].shape[0]
    idx = (f["indices"] if "indices" in f else f["ref_indices"][:Bs]).to(DEV)
    pp, ppl, gather = ops.select_position(f["block_ids"].to(DEV), f["block_table"].to(DEV),
                                          f["seq_len"].to(DEV), idx)
    assert tuple(pp.shape) == (Bs, idx.shape[1], 256) and tuple(ppl.shape) == (Bs, idx.shape[1], 8)
    assert torch.equal(pp.cpu(), crop_pad(f["ref_pos"], 256))
    assert torch.equal(ppl.cpu()[:, :, 0], f["ref_len"])
    assert bool((ppl == ppl[:, :, :1]).all())
    W = f["ref_cfull"].shape[2]
    assert torch.equal(gather.cpu()[:, :, :W], f["ref_cfull"])


@pytest.mark.parametrize("name", CHAINS)
def test_cent_select_golden(ops, golden_dir, name):
    f = load(golden_dir, name)
    pp, ppl, mppl = ops.cent_select(f["q"].to(DEV), f["l1"].to(DEV), f["block_ids"].to(DEV),
                                    f["block_table"].to(DEV), f["seq_len"].to(DEV))
    ok = rows_with_defined_positions(f, f["seq_len"].shape[0])
    assert torch.equal(pp.cpu()[ok], crop_pad(f["ref_pos"], 256)[ok])
    assert torch.equal(ppl.cpu()[:, :, 0][ok], f["ref_len"][ok])
    want = f["ref_len"].to(torch.int64).max(dim=1)[0] * f["block_size"]        # tests/cent_select.py:96-98
    bok = ok.all(dim=1)
    assert mppl.dtype == torch.int64 and torch.equal(mppl.cpu()[bok, 0], want[bok])
    assert bool((mppl == mppl[:, :1]).all())


# ------------------------------------------------------------------------------------------
# selection ops against the oracle on seeded cases (other k, C, group sizes, bf16)
# ------------------------------------------------------------------------------------------
SEL_CASES = [
    # B, Nq, Nkv, C, S, k, qdtype, cent_dist
    (1, 32, 8, 64, 4096, 8, torch.float16, "rand"),          # BASELINE configs[0]
    (4, 32, 8, 512, 32768, 64, torch.float16, "rand"),       # configs[1] shape, reduced batch
    (2, 64, 8, 2048, 131072, 128, torch.float16, "rand"),    # configs[3] shape, reduced batch
    (3, 32, 8, 512, 32768 - 1000, 64, torch.bfloat16, "randn"),
    (2, 8, 8, 100, 5000, 100, torch.float16, "rand"),        # g=1, k == C, C not a power of two
    (2, 16, 1, 256, 2048, 1, torch.float16, "randn"),        # g=16, k=1
]


@pytest.mark.parametrize("B,Nq,Nkv,C,S,k,qdtype,dist", SEL_CASES)
def test_selection_chain_oracle(ops, B, Nq, Nkv, C, S, k, qdtype, dist):
    P = (S + 127) // 128
    c = make_case(7 + C + k, B, Nq, Nkv, C, S, kv_pages=max(P + 7, 64), qdtype=qdtype, cent_dist=dist,
                  ragged=True, MP=max(256, P))
    g = to_device(c, DEV)
    MP = c["MP"]
    want_idx = oc.compute_cent(c["q"], c["l1"], k)
    got_idx = ops.compute_cent(g["q"], g["l1"], k=k)
    assert torch.equal(got_idx.cpu(), want_idx)
    wp, wl, wg = oc.select_position(c["block_ids"], c["block_table"], c["seq"], want_idx, 128, MP)
    gp, gl, gg = ops.select_position(g["block_ids"], g["block_table"], g["seq"], got_idx, max_page_num=MP)
    assert torch.equal(gp.cpu(), wp) and torch.equal(gl.cpu(), wl) and torch.equal(gg.cpu(), wg)
    wp2, wl2, wm2 = oc.cent_select(c["q"], c["l1"], c["block_ids"], c["block_table"], c["seq"], k, 128, MP)
    gp2, gl2, gm2 = ops.cent_select(g["q"], g["l1"], g["block_ids"], g["block_table"], g["seq"], k=k,
                                    max_page_num=MP)
    assert torch.equal(gp2.cpu(), wp2) and torch.equal(gl2.cpu(), wl2) and torch.equal(gm2.cpu(), wm2)


def test_topk_ties_lowest_index_first(ops):
    """All-equal scores: the k lowest cluster ids must be chosen, in ascending order."""
    q = torch.full((1, 4, 128), 0.5, dtype=torch.float16, device=DEV)
    l1 = torch.full((2, 300, 128), 0.25, dtype=torch.float16, device=DEV)
    got = ops.compute_cent(q, l1, k=37).cpu()
    assert torch.equal(got, torch.arange(37, dtype=torch.int32).expand(1, 4, 37))
    l1[:, 100:120] = 0.5                                           # two tied plateaus
    got = ops.compute_cent(q, l1, k=30).cpu()
    want = torch.cat([torch.arange(100, 120), torch.arange(0, 10)]).to(torch.int32).expand(1, 4, 30)
    assert torch.equal(got, want)


@pytest.mark.parametrize("qdtype", [torch.float16, torch.bfloat16])
def test_topk_near_ties_resolved_exactly(ops, qdtype):
    """Half of the clusters are copies of one centroid perturbed by single ulps in a dimension where q
    is tiny: their scores differ by ~5e-7, far below fp32 resolution at the score magnitude (ulp(32) =
    3.8e-6), and the k-th boundary is placed inside that group.  Tensor-core fp32 scores cannot order
    them; the selection must still equal the exact-arithmetic oracle (order included)."""
    gen = torch.Generator().manual_seed(77)
    B, Nq, Nkv, C, D = 2, 8, 2, 512, 128
    q = torch.rand(B, Nq, D, generator=gen)
    q[:, :, 5] = 2.0 ** -10
    q = q.to(qdtype)
    l1 = torch.rand(Nkv, C, D, generator=gen)
    ulp = 2.0 ** -11 if qdtype == torch.float16 else 2.0 ** -8          # ulp at 0.5
    for n in range(Nkv):
        base = l1[n, 0].clone()
        base[5] = 0.5
        steps = torch.randperm(C // 2, generator=gen).float()
        tie = base.unsqueeze(0).repeat(C // 2, 1)
        tie[:, 5] = 0.5 + steps * ulp
        where = torch.randperm(C, generator=gen)[: C // 2]
        l1[n, where] = tie
        l1[n, where[:7]] = tie[0]                                          # a few exact duplicates too
    l1 = l1.to(qdtype)
    exact = oc.cent_scores(q, l1)                                          # [B, Nq, C] float64
    # k such that the boundary falls inside the near-tie group for head (0, 0)
    tied_score = exact[0, 0][((l1[0, :, 5].float() >= 0.5) & (l1[0, :, 5].float() < 0.5 + C * ulp)).numpy()].mean()
    k = int((exact[0, 0] > tied_score + 1e-2).sum()) + 100
    assert 100 < k < C
    want = oc.compute_cent(q, l1, k)
    got = ops.compute_cent(q.to(DEV), l1.to(DEV), k=k)
    assert torch.equal(got.cpu(), want)


def test_topk_huge_plateau_uses_scratch(ops):
    """More than 512 exact ties around the boundary (e.g. a zero-initialised centroid table): the
    exact-score scratch in the workspace takes over; lowest indices win."""
    q = torch.rand(2, 8, 128, dtype=torch.float16, device=DEV)
    l1 = torch.zeros(2, 2048, 128, dtype=torch.float16, device=DEV)
    got = ops.compute_cent(q, l1, k=128).cpu()
    assert torch.equal(got, torch.arange(128, dtype=torch.int32).expand(2, 8, 128))
    l1[:, 1500:1540] = 1.0
    got = ops.compute_cent(q, l1, k=64).cpu()
    want = torch.cat([torch.arange(1500, 1540), torch.arange(0, 24)]).to(torch.int32)
    assert torch.equal(got, want.expand(2, 8, 64))


def test_select_truncates_at_max_page_num(ops):
    """More selected pages than max_page_num: list is cut, length is the kept count."""
    c = make_case(11, 1, 4, 2, 16, 64 * 128, kv_pages=80)
    g = to_device(c, DEV)
    idx = torch.arange(16, dtype=torch.int32).expand(1, 4, 16).contiguous()       # every cluster selected
    wp, wl, _ = oc.select_position(c["block_ids"], c["block_table"], c["seq"], idx, 128, 32)
    gp, gl, _ = ops.select_position(g["block_ids"], g["block_table"], g["seq"], idx.to(DEV), max_page_num=32)
    assert torch.equal(gp.cpu(), wp) and torch.equal(gl.cpu(), wl) and int(gl.max()) == 32


# ------------------------------------------------------------------------------------------
# attention: known-answer vectors restated from the reference tests
# ------------------------------------------------------------------------------------------
def build_kat(num_heads, block_size, total_kv, sel_tokens, layout, batch=1, head_delta=0.0):
    """tests/ifa_v5_case.py:584-765 (vLLM layout) / :12-191 (BNSD-paged): selected pages hold
    0.1 (+ h*head_delta), every other page 9.9, q = 0.1.  Head h uses physical blocks
    (3h + 2i) mod block_num; block_table packs the heads one after another and
    block_position[b,h,i] is the running index.  Expected output: the per-head constant."""
    D = 128
    block_num = total_kv // block_size
    per_head = sel_tokens // block_size
    q = torch.full((batch, num_heads, D), 0.1, dtype=torch.float16)
    shape = (block_num, block_size, num_heads, D) if layout == "vllm" else (block_num, num_heads, block_size, D)
    key = torch.full(shape, 9.9, dtype=torch.float16)
    value = torch.full(shape, 9.9, dtype=torch.float16)
    bt = torch.full((batch, num_heads * per_head), 0x7FFFFFFF, dtype=torch.int32)
    bp = torch.full((batch, num_heads, per_head), 0x7FFFFFFF, dtype=torch.int32)
    run = 0
    for h in range(num_heads):
        val = 0.1 + h * head_delta
        for i in range(per_head):
            blk = (3 * h + 2 * i) % block_num
            bt[:, run] = blk
            bp[:, h, i] = run
            run += 1
            if layout == "vllm":
                key[blk, :, h, :] = val
                value[blk, :, h, :] = val
            else:
                key[blk, h, :, :] = val
                value[blk, h, :, :] = val
    asl = torch.tensor([per_head * block_size] * batch, dtype=torch.int64)
    expect = torch.tensor([0.1 + h * head_delta for h in range(num_heads)]).to(torch.float16)
    return q, key, value, bt, bp, asl, expect


@pytest.mark.parametrize("layout,delta,bs,sel", [("vllm", 0.0, 128, 10 * 1024), ("vllm", 0.001, 128, 10 * 1024),
                                                 ("bnsd", 0.001, 128, 10 * 1024),
                                                 # tests/ifa_v5_case.py:12-191 as written: BNSD-paged pool, block 16, 256 pages per head
                                                 ("bnsd", 0.0, 16, 4 * 1024), ("bnsd", 0.001, 16, 4 * 1024)])
def test_attention_known_answer(ops, layout, delta, bs, sel):
    H = 32
    q, key, value, bt, bp, asl, expect = build_kat(H, bs, 32 * 1024, sel, layout, head_delta=delta)
    e = torch.empty(0, device=DEV)
    scale = np.float32(1.0) if bs == 128 else np.float32(1.0 / np.sqrt(128.0))      # :41 / ifa_v5_case.py:730
    out = ops.sparse_paged_attention(q.to(DEV), key.to(DEV), value.to(DEV), e, e, asl.to(DEV), e, e, e, e, e, e, e,
                                     bt.to(DEV), e, bp.to(DEV), H, scale, "BNSD", H, bs, 1)
    out = out.cpu()
    assert out.shape == q.shape and out.dtype == torch.float16
    assert not ((out.float() - 9.9).abs() < 1.0).any()            # ifa_v5_case.py:176-178 pass criterion
    want = expect.view(1, H, 1).expand_as(out)
    assert_close(out, want, 1e-3, "KAT")


# ------------------------------------------------------------------------------------------
# attention against the oracle
# ------------------------------------------------------------------------------------------
ATTN_CASES = [
    # B, Nq, Nkv, C, S, k, qdtype, int8, tol
    (1, 32, 8, 64, 4096, 8, torch.float16, False, 1e-3),         # configs[0]
    (3, 32, 8, 512, 32768, 64, torch.float16, False, 1e-3),      # configs[1] shape
    (2, 32, 8, 512, 32768 - 555, 64, torch.bfloat16, False, 2e-2),
    (3, 32, 8, 512, 32768, 64, torch.bfloat16, True, 2e-2),      # configs[2] shape: int8 KV, bf16 q
    (2, 32, 8, 256, 16384 - 77, 32, torch.float16, True, 2e-2),
    (1, 64, 8, 2048, 131072, 128, torch.float16, False, 1e-3),   # configs[3] shape
    (2, 8, 8, 128, 8192, 16, torch.float16, False, 1e-3),        # g = 1 (MHA)
    (1, 32, 2, 128, 8192 - 3, 16, torch.float16, False, 1e-3),   # g = 16 (two head tiles)
]


@pytest.mark.parametrize("B,Nq,Nkv,C,S,k,qdtype,int8,tol", ATTN_CASES)
def test_fused_vs_oracle(ops, B, Nq, Nkv, C, S, k, qdtype, int8, tol):
    P = (S + 127) // 128
    c = make_case(31 + C + k + B, B, Nq, Nkv, C, S, kv_pages=max(2 * P, 64), qdtype=qdtype, kv_int8=int8,
                  ragged=(S % 128 != 0), MP=max(256, P))
    g = to_device(c, DEV)
    aqs = c["aq_scale"] if int8 else None
    aqo = c["aq_offset"] if int8 else None
    wout, wpos, wlen, wmx = oc.sparse_paged_fusion_attention(
        c["q"], c["key"], c["value"], c["block_table"], c["l1"], c["block_ids"], c["seq"], Nq, c["scale"], Nkv,
        128, k, c["MP"], aqs, aqo)
    out, bp, ppl, mppl, ret = fused_call(ops, g, k)
    assert torch.equal(bp.cpu(), wpos) and torch.equal(ppl.cpu(), wlen) and torch.equal(mppl.cpu(), wmx)
    assert ret[0].data_ptr() == bp.data_ptr() and ret[1].data_ptr() == mppl.data_ptr()   # written in place
    assert out.dtype == qdtype and out.shape == c["q"].shape
    assert_close(out, wout, tol, "fused attention_out")


@pytest.mark.parametrize("B,Nq,Nkv,C,S,k,qdtype,int8,tol", ATTN_CASES[:5])
def test_separate_chain_matches_fused_and_oracle(ops, B, Nq, Nkv, C, S, k, qdtype, int8, tol):
    """cent_select -> sparse_paged_attention (fused_ifa_v5_case.py:228-252) must agree with the
    fused op: block_position exactly, attention_out within tolerance (S a multiple of the block
    size, otherwise the separate op cannot know the context's last partial page)."""
    S = (S + 127) // 128 * 128
    P = S // 128
    c = make_case(57 + C + k, B, Nq, Nkv, C, S, kv_pages=max(2 * P, 64), qdtype=qdtype, kv_int8=int8, MP=max(256, P))
    g = to_device(c, DEV)
    out_f, bp_f, ppl_f, mppl_f, _ = fused_call(ops, g, k)
    pp, ppl, mppl = ops.cent_select(g["q"], g["l1"], g["block_ids"], g["block_table"], g["seq"], k=k,
                                    max_page_num=c["MP"])
    assert torch.equal(pp, bp_f) and torch.equal(mppl, mppl_f)
    out_s = attn_call(ops, g, pp, mppl[:, 0].contiguous())
    assert_close(out_s, out_f, tol, "separate vs fused")
    want = oc.sparse_paged_attention(c["q"], c["key"], c["value"], mppl[:, 0].cpu(), c["block_table"], pp.cpu(),
                                     Nq, c["scale"], Nkv, 128, c["aq_scale"] if int8 else None,
                                     c["aq_offset"] if int8 else None)
    assert_close(out_s, want, tol, "separate vs oracle")


def test_attention_block_position_per_kv_head_and_partial_tail(ops):
    """block_position given per KV head (kernel indexing, ..._Us2.h:258) with actual_seq_lengths not
    a multiple of the block size: the last slot is clipped."""
    B, Nq, Nkv, S = 2, 16, 4, 8192
    c = make_case(91, B, Nq, Nkv, 64, S, kv_pages=128)
    g = to_device(c, DEV)
    gen = torch.Generator().manual_seed(5)
    bp = torch.full((B, Nkv, 256), oc.INT32_MAX, dtype=torch.int32)
    n = 11
    for b in range(B):
        for h in range(Nkv):
            bp[b, h, :n] = torch.randperm(S // 128, generator=gen)[:n].to(torch.int32)   # unsorted on purpose
    asl = torch.tensor([n * 128 - 50, (n - 3) * 128 + 1], dtype=torch.int64)
    want = oc.sparse_paged_attention(c["q"], c["key"], c["value"], asl, c["block_table"], bp, Nq, c["scale"], Nkv, 128)
    got = attn_call(ops, g, bp.to(DEV), asl.to(DEV))
    assert_close(got, want, 1e-3, "per-kv-head block_position")


def test_dense_mode_identity_block_position(ops):
    """BASELINE configs[4]: the dense comparison runs the same kernel with an identity block_position."""
    B, Nq, Nkv, S = 2, 32, 8, 4096
    P = S // 128
    c = make_case(93, B, Nq, Nkv, 64, S, kv_pages=2 * B * P, permute_table=True)
    g = to_device(c, DEV)
    bp = torch.arange(P, dtype=torch.int32).expand(B, Nkv, P).contiguous()
    asl = torch.full((B,), S, dtype=torch.int64)
    want = oc.sparse_paged_attention(c["q"], c["key"], c["value"], asl, c["block_table"], bp, Nq, c["scale"], Nkv, 128)
    got = attn_call(ops, g, bp.to(DEV), asl.to(DEV))
    assert_close(got, want, 1e-3, "dense mode")
    # cross-check against torch SDPA on the gathered dense KV (independent of the oracle)
    bt = c["block_table"].long()
    K = c["key"][bt[:B, :P]].reshape(B, S, Nkv, 128).permute(0, 2, 1, 3).float()
    V = c["value"][bt[:B, :P]].reshape(B, S, Nkv, 128).permute(0, 2, 1, 3).float()
    qf = c["q"].float().view(B, Nkv, Nq // Nkv, 128)
    ref = torch.nn.functional.scaled_dot_product_attention(qf, K, V, scale=c["scale"]).reshape(B, Nq, 128)
    assert_close(got, ref, 2e-3, "dense mode vs SDPA")


def test_empty_and_degenerate_inputs(ops):
    B, Nq, Nkv = 2, 8, 2
    c = make_case(95, B, Nq, Nkv, 64, 1024, kv_pages=64)
    g = to_device(c, DEV)
    # a sequence of length 0 selects nothing: lengths 0, padding everywhere, zero output
    g["seq"] = torch.tensor([0, 1024], dtype=torch.int32, device=DEV)
    c["seq"] = g["seq"].cpu()
    wout, wpos, wlen, wmx = oc.sparse_paged_fusion_attention(
        c["q"], c["key"], c["value"], c["block_table"], c["l1"], c["block_ids"], c["seq"], Nq, c["scale"], Nkv, 128, 8, 256)
    out, bp, ppl, mppl, _ = fused_call(ops, g, 8)
    assert torch.equal(bp.cpu(), wpos) and torch.equal(ppl.cpu(), wlen) and torch.equal(mppl.cpu(), wmx)
    assert bool((out[0] == 0).all())
    assert_close(out, wout, 1e-3, "seq_len 0")
    # all-sentinel block_position -> zeros
    bp0 = torch.full((B, Nkv, 256), oc.INT32_MAX, dtype=torch.int32, device=DEV)
    got = attn_call(ops, g, bp0, torch.tensor([1024, 1024], device=DEV))
    assert bool((got == 0).all())


def test_query_4d_and_noncontiguous(ops):
    B, Nq, Nkv, S, k = 2, 32, 8, 8192, 16
    c = make_case(97, B, Nq, Nkv, 128, S, kv_pages=256)
    g = to_device(c, DEV)
    wout, *_ = oc.sparse_paged_fusion_attention(c["q"], c["key"], c["value"], c["block_table"], c["l1"],
                                                c["block_ids"], c["seq"], Nq, c["scale"], Nkv, 128, k, 256)
    big = torch.zeros(B, Nq, 3, 128, dtype=torch.float16, device=DEV)
    big[:, :, 1] = g["q"]
    g4 = dict(g)
    g4["q"] = big[:, :, 1:2, :]                                   # [B,Nq,1,D], strided
    out, *_ = fused_call(ops, g4, k)
    assert out.shape == (B, Nq, 1, 128)
    assert_close(out.view(B, Nq, 128), wout, 1e-3, "4-D strided query")


def test_unsupported_arguments_raise(ops):
    c = make_case(99, 1, 8, 2, 64, 1024, kv_pages=64)
    g = to_device(c, DEV)
    e = torch.empty(0, device=DEV)
    bp = torch.zeros(1, 2, 256, dtype=torch.int32, device=DEV)
    asl = torch.tensor([128], device=DEV)

    def call(pse=e, mask=e, heads=8, bs=128):
        return ops.sparse_paged_attention(g["q"], g["key"], g["value"], pse, mask, asl, e, e, e, e, e, e, e,
                                          g["block_table"], e, bp, heads, 1.0, "BNSD", 2, bs, 1)

    with pytest.raises(RuntimeError, match="pse_shift"):
        call(pse=torch.zeros(1, device=DEV))
    with pytest.raises(RuntimeError, match="attention_mask"):
        call(mask=torch.zeros(1, device=DEV))
    with pytest.raises(RuntimeError, match="num_heads"):
        call(heads=16)
    with pytest.raises(RuntimeError, match="block_size"):
        call(bs=48)
    with pytest.raises(RuntimeError, match="CUDA"):
        ops.compute_cent(c["q"], c["l1"])
    with pytest.raises(RuntimeError, match="exceeds cluster count"):
        ops.compute_cent(g["q"], g["l1"], k=65)
    with pytest.raises(RuntimeError, match="pse_shift"):          # the stock IFA's other options stay refused
        ops.incre_flash_attention_v4(g["q"], [g["key"]], [g["value"]], torch.zeros(1, device=DEV), e, asl, e, e, e, e, e, e, e,
                                     g["block_table"], e, 8, 1.0, "BNSD", 2, 128, 1)


def test_incre_flash_attention_v4_is_dense_paged_decode(ops):
    """The sixth reference entry (add_custom.py:10-40, stock dense IFA) served by the same kernel over every
    page: must equal torch SDPA on the gathered KV, ragged lengths included."""
    B, Nq, Nkv, S = 3, 32, 8, 4096
    P = S // 128
    c = make_case(611, B, Nq, Nkv, 64, S, kv_pages=B * P + 9, permute_table=True)
    g = to_device(c, DEV)
    lens = torch.tensor([S, S - 300, 129], dtype=torch.int64)
    e = torch.empty(0, dtype=torch.float16, device=DEV)
    e64 = torch.empty(0, dtype=torch.int64, device=DEV)
    got = ops.incre_flash_attention_v4(g["q"], [g["key"]], [g["value"]], e, e, lens.to(DEV), e, e, e, e, e, e, e,
                                       g["block_table"], e64, Nq, c["scale"], "BNSD", Nkv, 128, 1)
    bt = c["block_table"].long()
    for b in range(B):
        n = int(lens[b])
        K = c["key"][bt[b]].reshape(P * 128, Nkv, 128)[:n].permute(1, 0, 2).float()
        V = c["value"][bt[b]].reshape(P * 128, Nkv, 128)[:n].permute(1, 0, 2).float()
        qf = c["q"][b].float().view(Nkv, Nq // Nkv, 128)
        ref = torch.nn.functional.scaled_dot_product_attention(qf, K, V, scale=c["scale"]).reshape(Nq, 128)
        assert_close(got[b], ref, 2e-3, f"dense IFA b={b}")


def test_torch_ops_namespace_and_plan(ops):
    """torch.ops.myops.* keeps the reference schema (registration.cpp:43-57); FusedDecodePlan replays."""
    c = make_case(101, 2, 32, 8, 128, 8192, kv_pages=256)
    g = to_device(c, DEV)
    a = torch.ops.myops.compute_cent(g["q"], g["l1"])
    b = ops.compute_cent(g["q"], g["l1"])
    assert torch.equal(a, b)
    out, bp, ppl, mppl, _ = fused_call(ops, g, 64)
    bp2, ppl2, mppl2 = torch.empty_like(bp), torch.empty_like(ppl), torch.empty_like(mppl)
    out2 = torch.empty_like(out)
    plan = ops.FusedDecodePlan(g["q"], g["key"], g["value"], g["block_table"], g["l1"], g["block_ids"], g["seq"],
                               bp2, ppl2, mppl2, out2, 32, c["scale"], 8, 128, k=64)
    assert plan.num_launches == 1
    for _ in range(3):
        plan.launch()
    torch.cuda.synchronize()
    # selection is bit-exact; attention_out may differ in the last bits between launches because the
    # page queue splits the items across CTAs dynamically (different summation order)
    assert torch.equal(bp2, bp) and torch.equal(mppl2, mppl)
    torch.testing.assert_close(out2.float(), out.float(), rtol=1e-3, atol=1e-3)
    # CUDA-graph capture of the fused launch
    gr = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        plan.launch()
    torch.cuda.current_stream().wait_stream(s)
    with torch.cuda.graph(gr):
        plan.launch()
    out2.zero_()
    gr.replay()
    torch.cuda.synchronize()
    torch.testing.assert_close(out2.float(), out.float(), rtol=1e-3, atol=1e-3)


# ------------------------------------------------------------------------------------------
# the vLLM-style caller (persistent buffers, per-batch-size executors, CUDA graphs)
# ------------------------------------------------------------------------------------------
def test_decode_backend_matches_oracle_across_batch_sizes_and_graph_replay(ops):
    from decode_backend import SparseDecodeBackend
    Nq, Nkv, C, S, k = 32, 8, 128, 8192, 16
    P = S // 128
    c = make_case(311, 4, Nq, Nkv, C, S, kv_pages=4 * P + 5, ragged=True, permute_table=True)
    g = to_device(c, DEV)
    be = SparseDecodeBackend(Nq, Nkv, 128, 128, k, max_batch=4, max_blocks_per_seq=P + 3, dtype=torch.float16,
                             device=DEV, scale=c["scale"])
    be.bind_cache(g["key"], g["value"], g["l1"], g["block_ids"])
    for B, graph in [(4, False), (2, False), (3, True), (3, True), (4, True)]:
        q = (torch.rand(B, Nq, 128, generator=torch.Generator().manual_seed(B * 7 + graph)) * 0.9).half()
        want = oc.sparse_paged_fusion_attention(q, c["key"], c["value"], c["block_table"][:B], c["l1"], c["block_ids"],
                                                c["seq"][:B], Nq, c["scale"], Nkv, 128, k, be.max_page_num)
        out = be.forward_decode(q.to(DEV).unsqueeze(2), g["block_table"][:B], g["seq"][:B], use_graph=graph)
        torch.cuda.synchronize()
        assert out.shape == (B, Nq, 128)
        assert_close(out, want[0], 1e-3, f"backend B={B} graph={graph}")
        bp, ppl, mx = be.selection(B)
        assert torch.equal(bp.cpu(), want[1]) and torch.equal(ppl.cpu(), want[2]) and torch.equal(mx.cpu(), want[3])
    with pytest.raises(ValueError):
        be.forward_decode(torch.zeros(5, Nq, 128, dtype=torch.float16, device=DEV), g["block_table"][:4], g["seq"][:4])


# ------------------------------------------------------------------------------------------
# reference-kernel compatibility options (anchor pages, P/8 clamp)
# ------------------------------------------------------------------------------------------
@pytest.mark.parametrize("opts", [dict(sink_pages=1, recent_pages=2), dict(sink_pages=1, recent_pages=2, max_fraction_den=8),
                                  dict(max_fraction_den=4), dict(sink_pages=3)])
def test_anchor_pages_and_clamp_options(ops, opts):
    B, Nq, Nkv, C, S, k = 3, 32, 8, 128, 8192, 8
    c = make_case(401, B, Nq, Nkv, C, S, kv_pages=160, ragged=True)       # block_table with collisions
    g = to_device(c, DEV)
    wpos, wlen, wmx = oc.cent_select(c["q"], c["l1"], c["block_ids"], c["block_table"], c["seq"], k, 128, 256, **opts)
    pos, ln, mx = ops.cent_select(g["q"], g["l1"], g["block_ids"], g["block_table"], g["seq"], k=k, **opts)
    assert torch.equal(pos.cpu(), wpos) and torch.equal(ln.cpu(), wlen) and torch.equal(mx.cpu(), wmx)
    want = oc.sparse_paged_fusion_attention(c["q"], c["key"], c["value"], c["block_table"], c["l1"], c["block_ids"],
                                            c["seq"], Nq, c["scale"], Nkv, 128, k, 256, **opts)
    dev = g["q"].device
    bp = torch.full((B, Nq, 256), 7, dtype=torch.int32, device=dev)
    ppl = torch.full((B, Nq, 8), 7, dtype=torch.int32, device=dev)
    mppl = torch.full((B, 8), 7, dtype=torch.int64, device=dev)
    e = torch.empty(0, dtype=torch.float16, device=dev)
    e64 = torch.empty(0, dtype=torch.int64, device=dev)
    out, _, _ = ops.sparse_paged_fusion_attention(
        g["q"], g["key"], g["value"], e, e, e64, e, e, e, e, e, e, e, g["block_table"], e64, g["l1"], g["block_ids"],
        g["seq"], bp, ppl, mppl, Nq, c["scale"], "BNSD", Nkv, 128, 1, k=k, **opts)
    assert torch.equal(bp.cpu(), want[1]) and torch.equal(ppl.cpu(), want[2]) and torch.equal(mppl.cpu(), want[3])
    assert_close(out, want[0], 1e-3, f"fused with {opts}")


# ------------------------------------------------------------------------------------------
# softmax log-sum-exp output
# ------------------------------------------------------------------------------------------
@pytest.mark.parametrize("qdtype,tol", [(torch.float16, 1e-3), (torch.bfloat16, 2e-2)])
def test_softmax_lse_output_fused_and_separate(ops, qdtype, tol):
    B, Nq, Nkv, C, S, k = 3, 32, 8, 128, 8192, 16
    c = make_case(501, B, Nq, Nkv, C, S, kv_pages=B * 64 + 3, ragged=True, permute_table=True, qdtype=qdtype)
    c["seq"][1] = 0                                          # a row that attends to nothing: lse = -inf, out = 0
    g = to_device(c, DEV)
    want_lse = torch.empty(B, Nq, dtype=torch.float32)
    want = oc.sparse_paged_fusion_attention(c["q"], c["key"], c["value"], c["block_table"], c["l1"], c["block_ids"],
                                            c["seq"], Nq, c["scale"], Nkv, 128, k, 256, lse_out=want_lse)
    dev = g["q"].device
    bp = torch.empty(B, Nq, 256, dtype=torch.int32, device=dev)
    ppl = torch.empty(B, Nq, 8, dtype=torch.int32, device=dev)
    mppl = torch.empty(B, 8, dtype=torch.int64, device=dev)
    lse = torch.full((B, Nq), 123.0, dtype=torch.float32, device=dev)
    e = torch.empty(0, dtype=qdtype, device=dev)
    e64 = torch.empty(0, dtype=torch.int64, device=dev)
    out, _, _ = ops.sparse_paged_fusion_attention(
        g["q"], g["key"], g["value"], e, e, e64, e, e, e, e, e, e, e, g["block_table"], e64, g["l1"], g["block_ids"],
        g["seq"], bp, ppl, mppl, Nq, c["scale"], "BNSD", Nkv, 128, 1, k=k, softmax_lse=lse)
    assert_close(out, want[0], tol, "fused out with lse")
    got = lse.cpu()
    assert bool((got[1] == -float("inf")).all()) and bool((want_lse[1] == -float("inf")).all())
    fin = torch.isfinite(want_lse)
    assert bool(torch.isfinite(got)[fin].all())
    assert float((got[fin] - want_lse[fin]).abs().max()) < 2e-3
    # separate op on the selection the fused op produced: same lse
    lse2 = torch.zeros(B, Nq, 1, dtype=torch.float32, device=dev)
    asl = (ppl[:, :, 0].max(dim=1)[0].long() * 128)
    out2 = ops.sparse_paged_attention(g["q"], g["key"], g["value"], e, e, asl, e, e, e, e, e, e, e, g["block_table"], e64, bp,
                                      Nq, c["scale"], "BNSD", Nkv, 128, 1, softmax_lse=lse2)
    want_lse2 = torch.empty(B, Nq, dtype=torch.float32)
    want2 = oc.sparse_paged_attention(c["q"], c["key"], c["value"], asl.cpu(), c["block_table"], bp.cpu(), Nq, c["scale"],
                                      Nkv, 128, lse_out=want_lse2)
    assert_close(out2, want2, tol, "separate out with lse")
    got2 = lse2.cpu().view(B, Nq)
    fin2 = torch.isfinite(want_lse2)
    assert bool((got2[~fin2] == -float("inf")).all())
    assert float((got2[fin2] - want_lse2[fin2]).abs().max()) < 2e-3
    # int8 KV: the per-head constant the softmax cancels (K offset, conversion bias) is put back into the lse
    Bi, Nqi, Nkvi = 2, 16, 4
    ci = make_case(502, Bi, Nqi, Nkvi, 64, 4096, kv_pages=80, kv_int8=True, qdtype=qdtype, ragged=True, permute_table=True)
    gi = to_device(ci, DEV)
    want_lse_i = torch.empty(Bi, Nqi, dtype=torch.float32)
    want_i = oc.sparse_paged_fusion_attention(ci["q"], ci["key"], ci["value"], ci["block_table"], ci["l1"], ci["block_ids"],
                                              ci["seq"], Nqi, ci["scale"], Nkvi, 128, 8, 256, ci["aq_scale"], ci["aq_offset"],
                                              lse_out=want_lse_i)
    lse_i = torch.zeros(Bi, Nqi, dtype=torch.float32, device=dev)
    out_i, _, _ = ops.sparse_paged_fusion_attention(
        gi["q"], gi["key"], gi["value"], e, e, e64, e, e, e, e, e, gi["aq_scale"], gi["aq_offset"], gi["block_table"], e64,
        gi["l1"], gi["block_ids"], gi["seq"], torch.empty(Bi, Nqi, 256, dtype=torch.int32, device=dev),
        torch.empty(Bi, Nqi, 8, dtype=torch.int32, device=dev), torch.empty(Bi, 8, dtype=torch.int64, device=dev),
        Nqi, ci["scale"], "BNSD", Nkvi, 128, 1, k=8, softmax_lse=lse_i)
    assert_close(out_i, want_i[0], 2e-2, "int8 fused out with lse")
    assert float((lse_i.cpu() - want_lse_i).abs().max()) < 5e-3


# ------------------------------------------------------------------------------------------
# centroid maintenance (page -> cluster assignment)
# ------------------------------------------------------------------------------------------
@pytest.mark.parametrize("qdtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("layout", ["bsnd", "bnsd"])
def test_assign_page_clusters_bit_exact(ops, qdtype, layout):
    Nkv, C, blocks = 4, 200, 24
    g = torch.Generator().manual_seed(77)
    key = (torch.randn(blocks, 128, Nkv, 128, generator=g) * 0.5).to(qdtype)
    l1 = torch.randn(Nkv, C, 128, generator=g).to(qdtype)
    l1[:, 17] = l1[:, 5]                                       # exact ties -> lowest index
    l1[1, 150] = (l1[1, 150].float() * (1 + 2 ** -7)).to(qdtype)
    page_ids = torch.tensor([3, 0, 23, 7, 7, 11, 99, -1, 5], dtype=torch.int32)      # repeats and out-of-pool entries
    toks = torch.tensor([128, 1, 37, 128, 128, 0, 128, 128, 64], dtype=torch.int32)
    bid0 = torch.randint(0, C, (Nkv, blocks), generator=g, dtype=torch.int32)
    want, wmean = oc.assign_page_clusters(key, l1, page_ids, bid0, toks)
    kd = key.to(DEV) if layout == "bsnd" else key.permute(0, 2, 1, 3).contiguous().to(DEV)
    bid = bid0.clone().to(DEV)
    mean = ops.assign_page_clusters(kd, l1.to(DEV), page_ids.to(DEV), bid, toks.to(DEV), return_mean=True)
    assert torch.equal(bid.cpu(), want)
    live = torch.tensor([0 <= p < blocks for p in page_ids.tolist()])
    assert torch.equal(mean.cpu()[live], wmean[live])
    # full pages, no token list: every page of the pool
    allp = torch.arange(blocks, dtype=torch.int32)
    want2, _ = oc.assign_page_clusters(key, l1, allp, bid0)
    bid2 = bid0.clone().to(DEV)
    r = ops.assign_page_clusters(kd, l1.to(DEV), allp.to(DEV), bid2)
    assert r.data_ptr() == bid2.data_ptr() and torch.equal(bid2.cpu(), want2)
    # the assignment is what the chain then selects by: a query equal to a page's pooled key ranks that page's cluster first
    with pytest.raises(RuntimeError):
        ops.assign_page_clusters(kd.to(torch.int8), l1.to(DEV), allp.to(DEV), bid2)        # int8 pool without antiquant
    # int8 pool with the attention ops' antiquant tensors ([2, Nkv, D]: the K row is used) and per-tensor ([2])
    k8 = torch.randint(-128, 128, (blocks, 128, Nkv, 128), generator=g, dtype=torch.int8)
    k8d = k8.to(DEV) if layout == "bsnd" else k8.permute(0, 2, 1, 3).contiguous().to(DEV)
    for shape in [(2, Nkv, 128), (2,)]:
        aqs = (torch.rand(*shape, generator=g) * 0.02 + 0.005).to(qdtype)
        aqo = torch.randn(*shape, generator=g).to(qdtype)
        want8, wmean8 = oc.assign_page_clusters(k8, l1, page_ids, bid0, toks, antiquant_scale=aqs, antiquant_offset=aqo)
        bid8 = bid0.clone().to(DEV)
        mean8 = ops.assign_page_clusters(k8d, l1.to(DEV), page_ids.to(DEV), bid8, toks.to(DEV), return_mean=True,
                                         antiquant_scale=aqs.to(DEV), antiquant_offset=aqo.to(DEV))
        assert torch.equal(bid8.cpu(), want8), f"int8 pool, antiquant {shape}"
        assert torch.equal(mean8.cpu()[live], wmean8[live])


# ------------------------------------------------------------------------------------------
# The page stream has two schedules: little work -> every CTA streams whole items (no partials, no combine);
# otherwise the queue is cut into static shares + claimed runs with partials and last-arriver combines.
# Small test shapes would only ever see the first, so the main cases run under both (SPA_SMALL_PAGES=0
# switches the whole-item schedule off; the variable is read when an executor is planned).
# ------------------------------------------------------------------------------------------
@pytest.mark.parametrize("small_pages", ["0", "64"])
@pytest.mark.parametrize("kind", ["f16", "bf16", "i8", "g16"])
def test_fused_chain_under_both_stream_schedules(ops, monkeypatch, small_pages, kind):
    monkeypatch.setenv("SPA_SMALL_PAGES", small_pages)
    if kind == "g16":
        B, Nq, Nkv, C, S, k = 3, 32, 2, 128, 8192, 16
    else:
        B, Nq, Nkv, C, S, k = 5, 32, 8, 128, 8192, 16
    qd = torch.bfloat16 if kind in ("bf16", "i8") else torch.float16
    c = make_case(701, B, Nq, Nkv, C, S, kv_pages=B * 64 + 11, ragged=True, permute_table=True, qdtype=qd,
                  kv_int8=(kind == "i8"))
    c["seq"][B - 1] = 0
    g = to_device(c, DEV)
    i8 = kind == "i8"
    want = oc.sparse_paged_fusion_attention(c["q"], c["key"], c["value"], c["block_table"], c["l1"], c["block_ids"],
                                            c["seq"], Nq, c["scale"], Nkv, 128, k, 256,
                                            c["aq_scale"] if i8 else None, c["aq_offset"] if i8 else None)
    for rep in range(2):                                      # twice: the workspace must be left clean
        out, bp, ppl, mppl, _ = fused_call(ops, g, k)
        assert torch.equal(bp.cpu(), want[1]) and torch.equal(ppl.cpu(), want[2]) and torch.equal(mppl.cpu(), want[3])
        assert_close(out, want[0], 1e-3 if kind in ("f16", "g16") else 2e-2, f"{kind} small_pages={small_pages} rep {rep}")


# ------------------------------------------------------------------------------------------
# BASELINE.json's FULL sizes, through size-independent properties (the CPU oracle takes minutes there):
# exact top-k property in float64, structural invariants of the selection outputs, fused == separate chain,
# attention against a plain torch fp32 reference over the selected pages, invariance under a relabelling
# of the physical pages.
# ------------------------------------------------------------------------------------------
@pytest.mark.parametrize("workload", ["c2", "c3", "c4"])
def test_full_size_properties(ops, workload):
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import bench
    w = bench.WORKLOADS[workload]
    dev = torch.device(DEV, 0)
    inp = bench.make_inputs(w, w["B"], 0, 1, dev)
    B, Nq, Nkv, C, k, P, MP = inp["B"], inp["Nq"], inp["Nkv"], inp["C"], inp["k"], inp["P"], 256
    g = Nq // Nkv
    i8 = w["kv"] == "i8"
    tol = 2e-2 if (i8 or w["q"] == "bf16") else 1e-3
    e = torch.empty(0, dtype=inp["qd"], device=dev)
    e64 = torch.empty(0, dtype=torch.int64, device=dev)

    def fused(bt, bid, key, value):
        bp = torch.full((B, Nq, MP), 5, dtype=torch.int32, device=dev)
        ppl = torch.full((B, Nq, 8), 5, dtype=torch.int32, device=dev)
        mppl = torch.full((B, 8), 5, dtype=torch.int64, device=dev)
        out, _, _ = ops.sparse_paged_fusion_attention(
            inp["q"], key, value, e, e, e64, e, e, e, e, e, inp["aq_scale"], inp["aq_offset"], bt, e64, inp["l1"], bid,
            inp["seq"], bp, ppl, mppl, Nq, inp["scale"], "BNSD", Nkv, 128, 1, k=k)
        return out, bp, ppl, mppl

    out, bp, ppl, mppl = fused(inp["block_table"], inp["block_ids"], inp["key"], inp["value"])
    # (1) exact top-k property: float64 scores of fp16/bf16 operands are exact
    idx = ops.compute_cent(inp["q"], inp["l1"], k=k).long()                          # [B, Nq, k]
    sc = torch.einsum("bngd,ncd->bngc", inp["q"].double().view(B, Nkv, g, 128), inp["l1"].double()).reshape(B, Nq, C)
    assert bool((idx.sort(dim=-1)[0].diff(dim=-1) > 0).all()), "duplicate cluster in a top-k row"
    sel = torch.zeros(B, Nq, C, dtype=torch.bool, device=dev).scatter_(2, idx, True)
    lo_sel = torch.where(sel, sc, torch.full_like(sc, float("inf"))).amin(dim=-1)
    hi_rest = torch.where(sel, torch.full_like(sc, -float("inf")), sc).amax(dim=-1)
    assert bool((lo_sel >= hi_rest).all()), "a selected cluster scores below an unselected one"
    got_sc = sc.gather(2, idx)
    assert bool((got_sc.diff(dim=-1) <= 0).all()), "indices are not sorted by descending score"
    # (2) structure of the selection outputs + membership: page i selected <=> its cluster is in the top-k set
    ln = ppl[:, :, 0].long()
    assert bool((ppl == ppl[:, :, :1]).all()) and bool((mppl == (ln.amax(dim=1) * 128).view(B, 1)).all())
    ar = torch.arange(MP, device=dev).view(1, 1, MP)
    assert bool((bp[ar.expand(B, Nq, MP) >= ln.unsqueeze(-1)] == oc.INT32_MAX).all())
    cl = inp["block_ids"].long()[:, inp["block_table"].long()]                       # [Nkv, B, P] cluster of every logical page
    cl = cl.permute(1, 0, 2).repeat_interleave(g, dim=1)                             # [B, Nq, P]
    member = sel.gather(2, cl)                                                       # [B, Nq, P]
    want_len = member.sum(dim=-1).clamp(max=MP)
    assert torch.equal(ln, want_len)
    onehot = torch.zeros(B, Nq, P + 1, dtype=torch.bool, device=dev)
    onehot.scatter_(2, bp.long().clamp(max=P), torch.ones_like(bp, dtype=torch.bool))
    full = want_len == member.sum(dim=-1)                                            # rows not truncated at max_page_num
    assert torch.equal(onehot[:, :, :P][full], member[full])
    valid = ar < ln.unsqueeze(-1)
    assert bool(((bp[:, :, 1:] > bp[:, :, :-1]) | ~valid[:, :, 1:]).all()), "page_position not ascending"
    # (3) fused == separate chain
    pos2, len2, mx2 = ops.cent_select(inp["q"], inp["l1"], inp["block_ids"], inp["block_table"], inp["seq"], k=k)
    assert torch.equal(pos2, bp) and torch.equal(len2, ppl) and torch.equal(mx2, mppl)
    out2 = ops.sparse_paged_attention(inp["q"], inp["key"], inp["value"], e, e, mppl[:, 0].contiguous(), e, e, e, e, e,
                                      inp["aq_scale"], inp["aq_offset"], inp["block_table"], e64, bp, Nq, inp["scale"],
                                      "BNSD", Nkv, 128, 1)
    assert_close(out2, out, tol, "separate vs fused")
    # (4) plain torch fp32 reference over the selected pages, every (b, head) row
    ref = torch.zeros(B, Nq, 128, dtype=torch.float32, device=dev)
    for b in range(B):
        for n2 in range(Nkv):
            for hh in range(g):
                h = n2 * g + hh
                pages = bp[b, h, :ln[b, h]].long()
                phys = inp["block_table"][b].long()[pages]
                K = inp["key"][phys, :, n2, :].reshape(-1, 128).float()
                V = inp["value"][phys, :, n2, :].reshape(-1, 128).float()
                if i8:
                    K = (K + inp["aq_offset"][0, n2].float()) * inp["aq_scale"][0, n2].float()
                    V = (V + inp["aq_offset"][1, n2].float()) * inp["aq_scale"][1, n2].float()
                pr = torch.softmax((K @ inp["q"][b, h].float()) * inp["scale"], dim=0)
                ref[b, h] = pr @ V
    assert_close(out, ref, tol, "fused vs torch fp32 over the selected pages")
    # (5) relabelling the physical pages changes nothing: selection bit-identical, output within tolerance
    perm = torch.randperm(inp["key"].shape[0], device=dev, generator=torch.Generator(device=dev).manual_seed(3))
    inv = torch.empty_like(perm)
    inv[perm] = torch.arange(perm.numel(), device=dev)
    key_p, val_p = inp["key"][perm], inp["value"][perm]                              # new page j holds old page perm[j]
    bt_p = inv[inp["block_table"].long()].to(torch.int32)
    bid_p = inp["block_ids"][:, perm].contiguous()
    out_p, bp_p, ppl_p, mppl_p = fused(bt_p, bid_p, key_p, val_p)
    assert torch.equal(bp_p, bp) and torch.equal(ppl_p, ppl) and torch.equal(mppl_p, mppl)
    assert_close(out_p, out, tol, "relabelled pages")


def test_kmeans_step_m_step_and_consistency(ops):
    """centroids.kmeans_step: E-step by the CUDA kernel (bit-exact vs the oracle), M-step = token-weighted mean
    of the member pages' keys, empty clusters keep their centroid."""
    import centroids
    Nkv, C, blocks = 2, 16, 40
    g = torch.Generator().manual_seed(21)
    key = torch.randn(blocks, 128, Nkv, 128, generator=g).half()
    l1 = torch.randn(Nkv, C, 128, generator=g).half()
    ids = torch.randperm(blocks, generator=g)[:30].to(torch.int32)
    toks = torch.randint(1, 129, (30,), generator=g, dtype=torch.int32)
    bid0 = torch.zeros(Nkv, blocks, dtype=torch.int32)
    want_bid, want_mean = oc.assign_page_clusters(key, l1, ids, bid0, toks)
    bid = bid0.clone().to(DEV)
    new, cnt = centroids.kmeans_step(key.to(DEV), l1.to(DEV), bid, ids.to(DEV), toks.to(DEV))
    assert torch.equal(bid.cpu(), want_bid)
    assert int(cnt.sum()) == 30 * Nkv
    ref = l1.float().clone()
    for n2 in range(Nkv):
        for c in range(C):
            m = [i for i in range(30) if int(want_bid[n2, ids[i]]) == c]
            assert int(cnt[n2, c]) == len(m)
            if m:
                w = toks[m].float()
                ref[n2, c] = (want_mean[m, n2] * w.view(-1, 1)).sum(0) / w.sum()
    assert torch.allclose(new.float().cpu(), ref.half().float(), rtol=2e-3, atol=2e-3)


# ------------------------------------------------------------------------------------------
# The pieces together: a short decode loop of one layer -- append a token's K/V to the paged cache, keep
# block_ids current with assign_page_clusters (F1), decode through SparseDecodeBackend (F3) -- against the
# same loop on the CPU oracle.
# ------------------------------------------------------------------------------------------
def test_decode_loop_with_cache_append_and_page_assignment(ops):
    from decode_backend import SparseDecodeBackend
    Nq, Nkv, C, k, blocks, maxP = 16, 4, 32, 4, 40, 8
    gen = torch.Generator().manual_seed(909)
    key = torch.zeros(blocks, 128, Nkv, 128, dtype=torch.float16)
    value = torch.zeros_like(key)
    l1 = torch.randn(Nkv, C, 128, generator=gen).half()
    lens = [3 * 128 + 126, 2 * 128 + 5]                               # sequence 0 is about to fill its 4th page
    free = list(range(blocks - 1, -1, -1))
    table = torch.zeros(2, maxP, dtype=torch.int32)
    for b, n in enumerate(lens):
        for j in range((n + 127) // 128):
            pg = free.pop()
            table[b, j] = pg
            cnt = min(128, n - j * 128)
            key[pg, :cnt] = (torch.randn(cnt, Nkv, 128, generator=gen) * 0.5).half()
            value[pg, :cnt] = torch.rand(cnt, Nkv, 128, generator=gen).half()
    bid = torch.zeros(Nkv, blocks, dtype=torch.int32)
    dkey, dvalue, dl1, dbid = key.to(DEV), value.to(DEV), l1.to(DEV), bid.clone().to(DEV)

    def assign(pages, toks):
        nonlocal bid
        ids = torch.tensor(pages, dtype=torch.int32)
        tk = torch.tensor(toks, dtype=torch.int32)
        bid, _ = oc.assign_page_clusters(key, l1, ids, bid, tk)
        ops.assign_page_clusters(dkey, dl1, ids.to(DEV), dbid, tk.to(DEV))

    pages, toks = [], []
    for b, n in enumerate(lens):
        for j in range((n + 127) // 128):
            pages.append(int(table[b, j])); toks.append(min(128, n - j * 128))
    assign(pages, toks)
    assert torch.equal(dbid.cpu(), bid)
    be = SparseDecodeBackend(Nq, Nkv, 128, 128, k, max_batch=2, max_blocks_per_seq=maxP, dtype=torch.float16, device=DEV)
    be.bind_cache(dkey, dvalue, dl1, dbid)
    for step in range(4):
        touched, tcount = [], []
        for b in range(2):                                            # append this step's token to the cache
            pos = lens[b]
            if pos % 128 == 0:
                table[b, pos // 128] = free.pop()
            pg = int(table[b, pos // 128])
            kv_k = (torch.randn(Nkv, 128, generator=gen) * 0.5).half()
            kv_v = torch.rand(Nkv, 128, generator=gen).half()
            key[pg, pos % 128], value[pg, pos % 128] = kv_k, kv_v
            dkey[pg, pos % 128], dvalue[pg, pos % 128] = kv_k.to(DEV), kv_v.to(DEV)
            lens[b] = pos + 1
            touched.append(pg); tcount.append(pos % 128 + 1)
        assign(touched, tcount)                                       # the page's cluster follows its pooled key
        assert torch.equal(dbid.cpu(), bid), f"block_ids differ at step {step}"
        q = torch.rand(2, Nq, 128, generator=gen).half()
        seq = torch.tensor(lens, dtype=torch.int32)
        want = oc.sparse_paged_fusion_attention(q, key, value, table, l1, bid, seq, Nq, 128 ** -0.5, Nkv, 128, k, be.max_page_num)
        out = be.forward_decode(q.to(DEV), table.to(DEV), seq.to(DEV), use_graph=(step >= 2))
        torch.cuda.synchronize()
        bp, ppl, mx = be.selection(2)
        assert torch.equal(bp.cpu(), want[1]) and torch.equal(ppl.cpu(), want[2]), f"selection differs at step {step}"
        assert_close(out, want[0], 1e-3, f"decode loop step {step}")
    assert lens[0] == 3 * 128 + 130                                   # sequence 0 crossed into a fifth page


# ------------------------------------------------------------------------------------------
# the reference's own fused-vs-separate test, at its own shape and with its own criterion
# ------------------------------------------------------------------------------------------
def test_fused_vs_separate_at_the_reference_shape(ops):
    """src/transformer/sparse_paged_fusion_attention/test/fused_ifa_v5_case.py:21-78: B = 1, 32/8 heads, C = 512,
    S = 128k, kvPageLen 1280, block_table [256, 1024] with collisions, randn centroids, q/K/V ~ U[1e-3, 1),
    scale 1.0.  Its criterion (:228-276): block_position of the fused op == cent_select's, and attention_out of
    the fused op vs cent_select -> sparse_paged_attention with mean relative difference < 3e-2.  On top of that:
    both against the oracle within the fp16 tolerance."""
    B, Nq, Nkv, C, S, k = 1, 32, 8, 512, 128 * 1024, 64
    c = make_case(9090, B, Nq, Nkv, C, S, kv_pages=1280, maxP=1024, maxB=256, cent_dist="randn")
    c["scale"] = 1.0
    g = to_device(c, DEV)
    out_f, bp_f, ppl_f, mppl_f, _ = fused_call(ops, g, k)
    pp, ppl, mppl = ops.cent_select(g["q"], g["l1"], g["block_ids"], g["block_table"], g["seq"])
    assert torch.equal(bp_f, pp), "fused block_position != cent_select's"
    assert torch.equal(mppl_f, mppl) and torch.equal(ppl_f, ppl)
    out_s = attn_call(ops, g, pp, mppl[:, 0].contiguous())                       # :252 sep_actual_seq_lengths
    f, s_ = out_f.float().cpu(), out_s.float().cpu()
    mean_rel = ((f - s_).abs() / s_.abs().clamp_min(1e-6)).mean().item()
    assert mean_rel < 3e-2, f"mean_rel_diff {mean_rel:.3e} (reference criterion: < 3e-2)"
    want = oc.sparse_paged_fusion_attention(c["q"], c["key"], c["value"], c["block_table"], c["l1"], c["block_ids"],
                                            c["seq"], Nq, 1.0, Nkv, 128, k, 256)
    assert torch.equal(bp_f.cpu(), want[1]) and torch.equal(ppl_f.cpu(), want[2]) and torch.equal(mppl_f.cpu(), want[3])
    assert_close(out_f, want[0], 1e-3, "fused vs oracle, reference fused shape")
    assert_close(out_s, want[0], 1e-3, "separate vs oracle, reference fused shape")


def test_topk_bf16_wide_dynamic_range(ops):
    """bf16 has an 8-bit exponent: tables whose rows mix very different magnitudes.  Inside the documented range
    (include/spa_b200.h: per dot product the non-zero |q_d * cent_d| within 2^40 of each other) float64
    accumulation is exact, so kernel and oracle must agree bit for bit whatever their summation orders."""
    B, Nq, Nkv, C, k = 2, 16, 4, 512, 64
    gen = torch.Generator().manual_seed(31337)
    qe = torch.randint(-6, 7, (B, Nq, 128), generator=gen).float()
    ce = torch.randint(-12, 13, (Nkv, C, 128), generator=gen).float()           # products span 2^-18 .. 2^19 (+ mantissas)
    q = (torch.randn(B, Nq, 128, generator=gen) * torch.exp2(qe)).to(torch.bfloat16)
    l1 = (torch.randn(Nkv, C, 128, generator=gen) * torch.exp2(ce)).to(torch.bfloat16)
    got = ops.compute_cent(q.to(DEV), l1.to(DEV), k=k)
    assert torch.equal(got.cpu(), oc.compute_cent(q, l1, k))
    # far outside the range (1e-30 next to 1e+3) exactness is not promised, but the selected SET must still be
    # right wherever the scores are separated by more than float64 rounding: distinct random scores are
    l2 = l1.clone()
    l2[:, ::7, :] = (l2[:, ::7, :].float() * 1e-30).to(torch.bfloat16)
    got2 = ops.compute_cent(q.to(DEV), l2.to(DEV), k=k).cpu()
    want2 = oc.compute_cent(q, l2, k)
    assert torch.equal(got2.sort(dim=-1)[0], want2.sort(dim=-1)[0])


# ------------------------------------------------------------------------------------------
# co-residency: the chain kernel while part of the device is held by another kernel
# ------------------------------------------------------------------------------------------
@pytest.mark.parametrize("coop", ["1", "0"])
def test_fused_op_while_other_kernel_holds_sms(ops, monkeypatch, coop):
    """Every CTA of the chain kernel waits for words other CTAs publish, so all must become resident.  While a
    spinning kernel holds 16 SMs for 50 ms on another stream the launch must neither hang nor compute garbage:
    the cooperative launch (default) starts once the whole grid fits; the ordinary launch (SPA_COOP=0) starts
    on the free SMs and finishes when the held ones are released."""
    import ctypes
    import subprocess
    import sys
    # a fresh process: the launch mode is read once per process, and a hang must not take the test session down
    code = r"""
import ctypes, os, sys, torch
sys.path[:0] = [os.path.join(ROOT, "tests"), os.path.join(ROOT, "cann-ops-adv-vllm_b200"), ROOT]
import custom_ops
from cases import fused_call, make_case, to_device
from oracle import sparse_chain as oc
L = ctypes.CDLL(os.path.join(ROOT, "cann-ops-adv-vllm_b200", "libspa_b200.so"))
c = make_case(4711, 6, 32, 8, 128, 16384, kv_pages=6 * 128 + 5, ragged=True, permute_table=True)
g = to_device(c, "cuda:0")
want = oc.sparse_paged_fusion_attention(c["q"], c["key"], c["value"], c["block_table"], c["l1"], c["block_ids"],
                                        c["seq"], 32, c["scale"], 8, 128, 16, 256)
fused_call(custom_ops, g, 16); torch.cuda.synchronize()                      # warm up (module load)
side = torch.cuda.Stream()
ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
rc = L.spaDebugHoldSMs(16, ctypes.c_uint64(50_000_000), ctypes.c_void_p(side.cuda_stream))
assert rc == 0
ev0.record()
out, bp, ppl, mppl, _ = fused_call(custom_ops, g, 16)
ev1.record()
torch.cuda.synchronize()
assert torch.equal(bp.cpu(), want[1]) and torch.equal(ppl.cpu(), want[2]) and torch.equal(mppl.cpu(), want[3])
err = (out.float().cpu() - want[0].float()).abs()
assert bool((err <= 1e-3 + 1e-3 * want[0].float().abs()).all()), err.max().item()
print("OK ms_behind_holder=%.1f" % ev0.elapsed_time(ev1))
"""
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, SPA_COOP=coop)
    r = subprocess.run([sys.executable, "-c", f"ROOT = {root!r}\n" + code], env=env, capture_output=True, text=True,
                       timeout=240)
    assert r.returncode == 0 and "OK" in r.stdout, f"SPA_COOP={coop}: rc={r.returncode}\n{r.stdout[-2000:]}\n{r.stderr[-3000:]}"
    print(f"SPA_COOP={coop}: {r.stdout.strip().splitlines()[-1]}")


@pytest.mark.parametrize("bs", [16, 32, 64])
@pytest.mark.parametrize("int8", [False, True])
def test_block_sizes_below_128(ops, bs, int8):
    """block_size 16 / 32 / 64 (the reference's sweep script varies it, sparse_paged_fusion_attention/op_eval/run.sh:26-48):
    fused op and the separate chain against the oracle, ragged contexts (partial last page)."""
    B, Nq, Nkv, C, S, k = 3, 16, 4, 64, 40 * bs + 5, 8
    qd = torch.bfloat16 if int8 else torch.float16
    tol = 2e-2 if int8 else 1e-3
    c = make_case(4000 + bs, B, Nq, Nkv, C, S, kv_pages=3 * 48 + 7, bs=bs, ragged=True, permute_table=True, qdtype=qd,
                  kv_int8=int8)
    g = to_device(c, DEV)
    aq = (c["aq_scale"], c["aq_offset"]) if int8 else (None, None)
    want = oc.sparse_paged_fusion_attention(c["q"], c["key"], c["value"], c["block_table"], c["l1"], c["block_ids"],
                                            c["seq"], Nq, c["scale"], Nkv, bs, k, 256, *aq)
    out, bp, ppl, mppl, _ = fused_call(ops, g, k)
    assert torch.equal(bp.cpu(), want[1]) and torch.equal(ppl.cpu(), want[2]) and torch.equal(mppl.cpu(), want[3])
    assert_close(out, want[0], tol, f"fused, block_size {bs}")
    pp, ppl2, mppl2 = ops.cent_select(g["q"], g["l1"], g["block_ids"], g["block_table"], g["seq"], k=k, block_size=bs)
    assert torch.equal(pp, bp) and torch.equal(mppl2, mppl)
    asl = mppl2[:, 0].contiguous()
    out_s = attn_call(ops, g, pp, asl)
    want_s = oc.sparse_paged_attention(c["q"], c["key"], c["value"], asl.cpu(), c["block_table"], pp.cpu(), Nq, c["scale"],
                                       Nkv, bs, *aq)
    assert_close(out_s, want_s, tol, f"separate, block_size {bs}")


# ------------------------------------------------------------------------------------------
# FP8 (e4m3) KV pages: SURVEY 8(f) F4's B200-native option (no reference analogue)
# ------------------------------------------------------------------------------------------
@pytest.mark.parametrize("qdtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("scaled", [False, "scaled"])
@pytest.mark.parametrize("g16", [False, True])
def test_fp8_kv_fused_and_separate(ops, qdtype, scaled, g16):
    """float8_e4m3fn K/V pools through the same ops: x = (fp8 + offset) * scale with optional [2, Nkv, D] scale /
    offset (absent: the stored numbers are the values).  Fused op and the separate chain against the oracle, which
    dequantises in float32; tolerance of the 8-bit paths (2e-2).  Also the softmax LSE."""
    B, Nq, Nkv, C, S, k = (2, 32, 2, 64, 6000, 8) if g16 else (3, 32, 8, 128, 9000, 16)
    c = make_case(606, B, Nq, Nkv, C, S, kv_pages=B * 72 + 3, ragged=True, permute_table=True, qdtype=qdtype,
                  kv_fp8=scaled or True)
    c["scale"] = 0.02                                   # keeps the logits of N(0, 2) keys in a softmax-friendly range
    g = to_device(c, DEV)
    aq = (c["aq_scale"], c["aq_offset"]) if scaled else (None, None)
    lse_w = torch.empty(B, Nq, dtype=torch.float32)
    want = oc.sparse_paged_fusion_attention(c["q"], c["key"], c["value"], c["block_table"], c["l1"], c["block_ids"],
                                            c["seq"], Nq, c["scale"], Nkv, 128, k, 256, *aq, lse_out=lse_w)
    dev = g["q"].device
    bp = torch.full((B, Nq, 256), 7, dtype=torch.int32, device=dev)
    ppl = torch.full((B, Nq, 8), 7, dtype=torch.int32, device=dev)
    mppl = torch.full((B, 8), 7, dtype=torch.int64, device=dev)
    lse = torch.empty(B, Nq, dtype=torch.float32, device=dev)
    e = torch.empty(0, dtype=qdtype, device=dev)
    e64 = torch.empty(0, dtype=torch.int64, device=dev)
    out, _, _ = ops.sparse_paged_fusion_attention(
        g["q"], g["key"], g["value"], e, e, e64, e, e, e, e, e, g["aq_scale"], g["aq_offset"], g["block_table"], e64,
        g["l1"], g["block_ids"], g["seq"], bp, ppl, mppl, Nq, c["scale"], "BNSD", Nkv, 128, 1, k=k, softmax_lse=lse)
    assert torch.equal(bp.cpu(), want[1]) and torch.equal(ppl.cpu(), want[2]) and torch.equal(mppl.cpu(), want[3])
    assert_close(out, want[0], 2e-2, "fused, FP8 KV")
    fin = torch.isfinite(lse_w)
    assert torch.equal(torch.isfinite(lse.cpu()), fin)
    assert bool(((lse.cpu()[fin] - lse_w[fin]).abs() < 2e-2).all()), "softmax LSE, FP8 KV"
    asl = mppl[:, 0].contiguous()
    out_s = attn_call(ops, g, bp, asl)
    want_s = oc.sparse_paged_attention(c["q"], c["key"], c["value"], asl.cpu(), c["block_table"], bp.cpu(), Nq, c["scale"],
                                       Nkv, 128, *aq)
    assert_close(out_s, want_s, 2e-2, "separate, FP8 KV")
