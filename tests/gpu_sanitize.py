"""Small, fast pass over every kernel (diagnostic, not a pytest test): python tests/gpu_sanitize.py

Written for `compute-sanitizer --tool memcheck`; that tool is closed on this GPU pool (round 1: "runs under it
have left GPUs needing a reset"), so the run serves as a 5-second all-kernel consistency check instead.

One call of each op on small shapes -- both stream schedules of the chain kernel, fp16 and int8 KV, the
stand-alone selection ops, the separate attention op with LSE, the page -> cluster assignment (fp16 and int8
pools) -- each checked against the oracle so that a silent corruption would also show."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "cann-ops-adv-vllm_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import custom_ops  # noqa: E402
from cases import fused_call, make_case, to_device  # noqa: E402
from oracle import sparse_chain as oc  # noqa: E402

DEV = "cuda:0"


def main():
    ok = True
    for small_pages, int8 in [("64", False), ("0", False), ("0", True)]:
        os.environ["SPA_SMALL_PAGES"] = small_pages
        B, Nq, Nkv, C, S, k = 3, 16, 4, 64, 4096, 8
        c = make_case(11, B, Nq, Nkv, C, S, kv_pages=B * 32 + 3, ragged=True, permute_table=True, kv_int8=int8,
                      qdtype=torch.bfloat16 if int8 else torch.float16)
        g = to_device(c, DEV)
        want = oc.sparse_paged_fusion_attention(c["q"], c["key"], c["value"], c["block_table"], c["l1"], c["block_ids"],
                                                c["seq"], Nq, c["scale"], Nkv, 128, k, 256,
                                                c["aq_scale"] if int8 else None, c["aq_offset"] if int8 else None)
        out, bp, ppl, mppl, _ = fused_call(custom_ops, g, k)
        torch.cuda.synchronize()
        tol = 2e-2 if int8 else 1e-3
        good = torch.equal(bp.cpu(), want[1]) and bool(((out.float().cpu() - want[0].float()).abs() <= tol + tol * want[0].float().abs()).all())
        print(f"fused small_pages={small_pages} int8={int8}: {'ok' if good else 'MISMATCH'}", flush=True)
        ok &= good
    del os.environ["SPA_SMALL_PAGES"]
    c = make_case(12, 2, 16, 4, 64, 4096, kv_pages=70)
    g = to_device(c, DEV)
    idx = custom_ops.compute_cent(g["q"], g["l1"], k=8)
    pos, ln, _ = custom_ops.select_position(g["block_ids"], g["block_table"], g["seq"], idx)
    pos2, ln2, mx2 = custom_ops.cent_select(g["q"], g["l1"], g["block_ids"], g["block_table"], g["seq"], k=8)
    e = torch.empty(0, dtype=torch.float16, device=DEV)
    e64 = torch.empty(0, dtype=torch.int64, device=DEV)
    lse = torch.empty(2, 16, dtype=torch.float32, device=DEV)
    out = custom_ops.sparse_paged_attention(g["q"], g["key"], g["value"], e, e, mx2[:, 0].contiguous(), e, e, e, e, e, e, e,
                                            g["block_table"], e64, pos2, 16, c["scale"], "BNSD", 4, 128, 1, softmax_lse=lse)
    torch.cuda.synchronize()
    good = torch.equal(pos.cpu(), pos2.cpu()) and bool(torch.isfinite(out.float()).all())
    print(f"selection ops + separate attention: {'ok' if good else 'MISMATCH'}", flush=True)
    ok &= good
    gen = torch.Generator().manual_seed(5)
    key = torch.randn(12, 128, 4, 128, generator=gen).half()
    l1 = torch.randn(4, 72, 128, generator=gen).half()
    ids = torch.tensor([0, 11, 5, 5, 99], dtype=torch.int32)
    toks = torch.tensor([128, 3, 64, 64, 128], dtype=torch.int32)      # a repeated page must repeat its token count
    bid0 = torch.zeros(4, 12, dtype=torch.int32)
    want_bid, _ = oc.assign_page_clusters(key, l1, ids, bid0, toks)
    bid = bid0.clone().to(DEV)
    custom_ops.assign_page_clusters(key.to(DEV), l1.to(DEV), ids.to(DEV), bid, toks.to(DEV), return_mean=True)
    k8 = torch.randint(-128, 128, (12, 128, 4, 128), generator=gen, dtype=torch.int8)
    aqs = (torch.rand(2, 4, 128, generator=gen) * 0.02 + 0.005).half()
    aqo = torch.randn(2, 4, 128, generator=gen).half()
    want8, _ = oc.assign_page_clusters(k8, l1, ids, bid0, toks, antiquant_scale=aqs, antiquant_offset=aqo)
    bid8 = bid0.clone().to(DEV)
    custom_ops.assign_page_clusters(k8.to(DEV), l1.to(DEV), ids.to(DEV), bid8, toks.to(DEV), antiquant_scale=aqs.to(DEV),
                                    antiquant_offset=aqo.to(DEV))
    torch.cuda.synchronize()
    good = torch.equal(bid.cpu(), want_bid) and torch.equal(bid8.cpu(), want8)
    print(f"assign_page_clusters fp16 + int8: {'ok' if good else 'MISMATCH'}", flush=True)
    ok &= good
    print("SANITIZE_RUN_OK" if ok else "SANITIZE_RUN_MISMATCH", flush=True)
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
