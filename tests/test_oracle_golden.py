"""Pin the CPU oracle (oracle/sparse_chain.py) against fixtures produced by the reference's own
golden functions (tests/golden/make_golden.py imports /root/reference/tests/utils.py)."""
import os

import numpy as np
import pytest
import torch

from oracle import sparse_chain as oc

from golden_check import check_indices, rows_with_defined_positions

# chain_refshape: the reference's own test shape in full (tests/compute_cent.py:16-22, tests/cent_select.py:17-39)
CHAINS = ["chain_small", "chain_32k", "chain_ragged", "chain_randn", "chain_refshape"]


def load(golden_dir, name):
    z = np.load(os.path.join(golden_dir, name + ".npz"))
    return {k: (torch.from_numpy(z[k]) if z[k].ndim > 0 else z[k].item()) for k in z.files}


def crop_pad(pos, width):
    """The golden's width is max_b(pl) (utils.py:45-48), the op's is max_page_num: pad/crop with
    INT32_MAX before comparing."""
    B, H, W = pos.shape
    out = torch.full((B, H, width), oc.INT32_MAX, dtype=torch.int32)
    w = min(W, width)
    out[:, :, :w] = pos[:, :, :w]
    return out


@pytest.mark.parametrize("name", CHAINS)
def test_compute_cent_matches_reference_golden(golden_dir, name):
    f = load(golden_dir, name)
    got = oc.compute_cent(f["q"], f["l1"], k=64)
    assert got.dtype == torch.int32
    check_indices(got, f, f"oracle compute_cent {name}")   # order included, except on rows the fixture flags


@pytest.mark.parametrize("name", CHAINS)
def test_select_position_matches_reference_golden(golden_dir, name):
    f = load(golden_dir, name)
    Bs = f["seq_len"].shape[0]
    pos, plen, gather = oc.select_position(f["block_ids"], f["block_table"], f["seq_len"],
                                           f["ref_indices"][:Bs], f["block_size"], 256)
    assert torch.equal(pos, crop_pad(f["ref_pos"], 256))
    assert torch.equal(plen[:, :, 0], f["ref_len"])
    assert bool((plen == plen[:, :, :1]).all())       # x8 replication
    W = f["ref_cfull"].shape[2]
    assert torch.equal(gather[:, :, :W], f["ref_cfull"])


@pytest.mark.parametrize("name", CHAINS)
def test_cent_select_chain_matches_reference_golden(golden_dir, name):
    f = load(golden_dir, name)
    pos, plen, mx = oc.cent_select(f["q"], f["l1"], f["block_ids"], f["block_table"], f["seq_len"],
                                   k=64, block_size=f["block_size"], max_page_num=256)
    ok = rows_with_defined_positions(f, f["seq_len"].shape[0])
    assert torch.equal(pos[ok], crop_pad(f["ref_pos"], 256)[ok])
    # tests/cent_select.py:96-98
    want = f["ref_len"].to(torch.int64).max(dim=1)[0] * f["block_size"]
    assert mx.dtype == torch.int64 and mx.shape == (f["seq_len"].shape[0], 8)
    bok = ok.all(dim=1)
    assert torch.equal(mx[bok, 0], want[bok]) and bool((mx == mx[:, :1]).all())


def test_select_position_duplicate_indices(golden_dir):
    f = load(golden_dir, "select_dupes")               # tests/select_position.py:42-45
    pos, plen, gather = oc.select_position(f["block_ids"], f["block_table"], f["seq_len"],
                                           f["indices"], f["block_size"], 256)
    assert torch.equal(pos, crop_pad(f["ref_pos"], 256))
    assert torch.equal(plen[:, :, 0], f["ref_len"])
    W = f["ref_cfull"].shape[2]
    assert torch.equal(gather[:, :, :W], f["ref_cfull"])


def test_topk_tie_break_lowest_index():
    s = np.zeros((1, 1, 16))
    s[0, 0, [3, 7, 9]] = 1.0
    got = oc.topk_indices(s, 6)
    assert got.tolist() == [[[3, 7, 9, 0, 1, 2]]]


def test_reference_kernel_anchor_and_clamp_options():
    """sink / recent anchor pages and the P/8 clamp of the reference kernels
    (cent_select.cpp:218,350-362,410-414) as oracle options; off = the golden."""
    import torch
    from oracle import sparse_chain as oc
    Nkv, Nq, C, P = 2, 4, 16, 40
    g = torch.Generator().manual_seed(5)
    block_ids = torch.randint(0, C, (Nkv, 64), generator=g, dtype=torch.int32)
    block_table = torch.randperm(64, generator=g)[:P].view(1, P).to(torch.int32)
    seq = torch.tensor([P * 128 - 5], dtype=torch.int32)
    idx = torch.randint(0, C, (1, Nq, 2), generator=g, dtype=torch.int32)
    p0, l0, _ = oc.select_position(block_ids, block_table, seq, idx)
    p1, l1, _ = oc.select_position(block_ids, block_table, seq, idx, sink_pages=1, recent_pages=2)
    for h in range(Nq):
        a = set(p0[0, h, :l0[0, h, 0]].tolist())
        b = p1[0, h, :l1[0, h, 0]].tolist()
        assert b == sorted(a | {0, P - 2, P - 1})
    p2, l2, _ = oc.select_position(block_ids, block_table, seq, idx, sink_pages=1, recent_pages=2, max_fraction_den=8)
    for h in range(Nq):
        n = min(int(l1[0, h, 0]), P // 8)
        assert int(l2[0, h, 0]) == n and p2[0, h, :n].tolist() == p1[0, h, :n].tolist()
        assert (p2[0, h, n:] == oc.INT32_MAX).all()
    # value comparison: a colliding block_table entry drags its position in, as the kernel's CompareScalar does
    bt2 = block_table.clone()
    bt2[0, 7] = bt2[0, 0]
    p3, l3, _ = oc.select_position(block_ids, bt2, seq, idx, sink_pages=1)
    assert all(7 in p3[0, h, :l3[0, h, 0]].tolist() for h in range(Nq))


def test_assign_page_clusters_oracle_agrees_with_plain_float64():
    """the page -> cluster assignment picks the centroid with the largest inner product with
    the page's pooled key (metric of ComputeCent); partial pages pool only their valid tokens."""
    import torch
    from oracle import sparse_chain as oc
    g = torch.Generator().manual_seed(9)
    Nkv, C, blocks = 2, 12, 6
    key = torch.randn(blocks, 128, Nkv, 128, generator=g).half()
    l1 = torch.randn(Nkv, C, 128, generator=g).half()
    ids = torch.tensor([4, 1, 5, 9], dtype=torch.int32)
    toks = torch.tensor([128, 50, 1, 128], dtype=torch.int32)
    bid0 = torch.full((Nkv, blocks), -7, dtype=torch.int32)
    out, mean = oc.assign_page_clusters(key, l1, ids, bid0, toks)
    for i, (pg, nt) in enumerate(zip(ids.tolist(), toks.tolist())):
        if pg >= blocks:
            continue
        pooled = key[pg, :nt].double().sum(0)                      # [Nkv, D]
        want = torch.einsum("nd,ncd->nc", pooled, l1.double()).argmax(dim=1)
        assert out[:, pg].tolist() == want.tolist()
        assert torch.allclose(mean[i].double(), pooled / nt, rtol=1e-5, atol=1e-6)
    untouched = [p for p in range(blocks) if p not in ids.tolist()]
    assert bool((out[:, untouched] == -7).all())


@pytest.mark.parametrize("int8", [False, True])
def test_vectorised_cpu_baseline_agrees_with_the_loop_oracle(int8):
    """oracle/fast_chain.py (the "fair CPU" baseline bench.py reports beside the loop port) computes the same
    chain: selection identical (tie-free seeded case), attention within the oracle-vs-kernel tolerance."""
    from cases import make_case
    from oracle import fast_chain
    qd = torch.bfloat16 if int8 else torch.float16
    c = make_case(99, 3, 16, 4, 64, 3000, kv_pages=40, ragged=True, qdtype=qd, kv_int8=int8)
    c["seq"][2] = 0
    want = oc.sparse_paged_fusion_attention(c["q"], c["key"], c["value"], c["block_table"], c["l1"], c["block_ids"],
                                            c["seq"], 16, c["scale"], 4, 128, 8, 256,
                                            c["aq_scale"] if int8 else None, c["aq_offset"] if int8 else None)
    got = fast_chain.fused_chain(c["q"], c["key"], c["value"], c["block_table"], c["l1"], c["block_ids"], c["seq"],
                                 16, c["scale"], 4, 128, 8, 256, c["aq_scale"] if int8 else None,
                                 c["aq_offset"] if int8 else None)
    assert torch.equal(got[1], want[1]) and torch.equal(got[2], want[2]) and torch.equal(got[3], want[3])
    tol = 2e-2 if int8 else 1e-3
    err = (got[0].float() - want[0].float()).abs()
    assert bool((err <= tol + tol * want[0].float().abs()).all()), err.max().item()
