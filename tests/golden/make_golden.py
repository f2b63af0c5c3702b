"""Generate the golden fixtures in this directory from the REFERENCE's own golden functions.

Run in the build container only (needs /root/reference, which does not exist on the GPU box):

    python tests/golden/make_golden.py

It imports `/root/reference/tests/utils.py` unmodified (torch_compute_cent, torch_select_position),
feeds it seeded inputs shaped like the reference's own tests (tests/compute_cent.py:16-50,
tests/select_position.py:16-45, tests/cent_select.py:17-53) and stores inputs + reference outputs as
compressed .npz.  Nothing here is a copy of reference source; only its outputs are stored.

Ambiguous rows: the reference's scores are fp32 with an unspecified summation order (tests/utils.py:117), so
rows where two of the top-(k+1) scores collide within fp32 rounding have no well-defined answer (torch's
CPU GEMV, the NPU's 16-lane partial sums and exact arithmetic can each order them differently).  Seeds are
NOT walked to avoid such rows: every fixture is generated from its base seed and stores `amb_rows`
[B, Nq] bool -- "the reference's fp32 answer differs from the exact-arithmetic ordering on this row" -- and
`amb_set` (the selected SET differs too).  Tests demand bit equality on the other rows, and on the flagged
ones what is still well defined (the same set if only the order moved; a set that differs from the golden's
only by clusters whose exact scores tie with the cut within fp32 rounding otherwise) and print the count.
At the reference's own shape (tests/compute_cent.py:16-22) about 0.2 % of the rows are flagged.
"""
import os
import sys

import numpy as np
import torch

REF = "/root/reference/tests"
sys.path.insert(0, REF)
import utils as ref_utils  # noqa: E402  (the reference's tests/utils.py)

HERE = os.path.dirname(os.path.abspath(__file__))
K = 64  # hard-coded in the reference golden (tests/utils.py:101)


def ambiguity(q, l1, ref_idx):
    """Per (b, head): does the reference's fp32 top-K differ from the exact-arithmetic one in order / as a set?"""
    s = torch.einsum("bngd,ncd->bngc", q.double().view(q.shape[0], l1.shape[0], -1, q.shape[2]),
                     l1.double()).reshape(q.shape[0], q.shape[1], -1)
    exact = torch.sort(s, dim=-1, descending=True, stable=True)[1][..., :K]
    order = (exact != ref_idx).any(dim=-1)
    sets = (torch.sort(exact, dim=-1)[0] != torch.sort(ref_idx, dim=-1)[0]).any(dim=-1)
    return order, sets


def gen_chain(name, seed, B, Nq, Nkv, C, D, S, kvPageLen, maxB, maxP, bs=128, ragged=False,
              cent_dist="rand", seq_batch=None):
    g = torch.Generator().manual_seed(seed)
    q = torch.rand(B, Nq, D, generator=g).to(torch.float16)
    if cent_dist == "rand":
        l1 = torch.rand(Nkv, C, D, generator=g).to(torch.float16)
    else:
        l1 = torch.randn(Nkv, C, D, generator=g).to(torch.float16)
    block_ids = torch.randint(0, C, (Nkv, kvPageLen), generator=g, dtype=torch.int32)
    block_table = torch.randint(0, kvPageLen, (maxB, maxP), generator=g, dtype=torch.int32)
    Bs = seq_batch if seq_batch is not None else B
    if ragged:
        seq_len = torch.tensor([max(1, S - 37 * b - 1) for b in range(Bs)], dtype=torch.int32)
    else:
        seq_len = torch.full((Bs,), S, dtype=torch.int32)
    ref_idx = ref_utils.torch_compute_cent(q, l1)                       # int64 [B,Nq,64]
    amb_rows, amb_set = ambiguity(q, l1, ref_idx)
    ref_pos, ref_len, ref_cfull = ref_utils.torch_select_position(
        block_ids, block_table, seq_len, ref_idx[:Bs], block_size=bs)
    np.savez_compressed(
        os.path.join(HERE, name + ".npz"), seed=seed, block_size=bs,
        q=q.numpy(), l1=l1.numpy(), block_ids=block_ids.numpy(), block_table=block_table.numpy(),
        seq_len=seq_len.numpy(), ref_indices=ref_idx.numpy().astype(np.int32),
        ref_pos=ref_pos.numpy(), ref_len=ref_len.numpy(), ref_cfull=ref_cfull.numpy(),
        amb_rows=amb_rows.numpy(), amb_set=amb_set.numpy())
    print(name, "seed", seed, "pos", tuple(ref_pos.shape), "len range", int(ref_len.min()), int(ref_len.max()),
          "ambiguous rows (order / set):", int(amb_rows.sum()), "/", int(amb_set.sum()), "of", amb_rows.numel())


def gen_select(name, seed, B, Nq, Nkv, C, k, S, kvPageLen, maxB, maxP, bs=128):
    """tests/select_position.py:42-45: indices are random (duplicates allowed)."""
    g = torch.Generator().manual_seed(seed)
    block_ids = torch.randint(0, C, (Nkv, kvPageLen), generator=g, dtype=torch.int32)
    block_table = torch.randint(0, kvPageLen, (maxB, maxP), generator=g, dtype=torch.int32)
    seq_len = torch.tensor([S - 129 * b for b in range(B)], dtype=torch.int32)
    indices = torch.randint(0, C, (B, Nq, k), generator=g, dtype=torch.int32)
    ref_pos, ref_len, ref_cfull = ref_utils.torch_select_position(
        block_ids, block_table, seq_len, indices, block_size=bs)
    np.savez_compressed(
        os.path.join(HERE, name + ".npz"), seed=seed, block_size=bs,
        block_ids=block_ids.numpy(), block_table=block_table.numpy(), seq_len=seq_len.numpy(),
        indices=indices.numpy(), ref_pos=ref_pos.numpy(), ref_len=ref_len.numpy(),
        ref_cfull=ref_cfull.numpy())
    print(name, "seed", seed, "pos", tuple(ref_pos.shape))


if __name__ == "__main__":
    torch.set_num_threads(1)  # keep the reference's fp32 GEMV order fixed while generating
    # c1-like: the reference's CPU-runnable case (BASELINE.json configs[0]) needs k=8, which the
    # reference golden cannot produce (k fixed at 64) -> use C=128 so k=64 is legal.
    gen_chain("chain_small", 1001, B=2, Nq=8, Nkv=2, C=128, D=128, S=4096, kvPageLen=64, maxB=4, maxP=64)
    # tests/cent_select.py shape family (C=512, seq 32k, kvPageLen 1280, table [*,1024]); batch and
    # heads reduced to keep the fixture small; seq batch = B-2 as in tests/cent_select.py:37.
    gen_chain("chain_32k", 2002, B=4, Nq=8, Nkv=2, C=512, D=128, S=32 * 1024, kvPageLen=1280,
              maxB=8, maxP=1024, seq_batch=2)
    gen_chain("chain_ragged", 3003, B=3, Nq=8, Nkv=4, C=256, D=128, S=16 * 1024, kvPageLen=300,
              maxB=4, maxP=256, ragged=True)
    # fused_ifa_v5_case.py:55 uses randn centroids
    gen_chain("chain_randn", 4004, B=2, Nq=16, Nkv=2, C=512, D=128, S=32 * 1024, kvPageLen=1280,
              maxB=4, maxP=1024, cent_dist="randn")
    # the reference's own shape, in full: tests/compute_cent.py:16-22 (batch 8, 32/8 heads, C = 512) with the tables
    # of tests/cent_select.py:17-39 (kvPageLen 1280, block_table [256, 1024], seq 32k, seq batch = batch - 2)
    gen_chain("chain_refshape", 6006, B=8, Nq=32, Nkv=8, C=512, D=128, S=32 * 1024, kvPageLen=1280,
              maxB=256, maxP=1024, seq_batch=6)
    gen_select("select_dupes", 5005, B=3, Nq=8, Nkv=2, C=512, k=64, S=32 * 1024, kvPageLen=1280,
               maxB=4, maxP=1024)
