#!/usr/bin/env python
"""BASELINE configs[4]: sparsity and context sweep at the c2 head shape (B=16, 32 Q / 8 KV heads,
D=128, fp16), fused SparsePagedFusionAttention vs the DENSE path on the same kernel
(SparsePagedAttention with an identity block_position, i.e. every page of the context).

    python tools/sweep.py [--steps 20] [--out profiles/rNN_sweep_c5.md]

k/C in {1/64 .. 1/4} (C = 512, k = 8 .. 128) x S in {8k .. 128k}.  Every cell is timed like bench.py's
device-resident number: CUDA events on the launching stream, L2 flushed (write 256 MiB + read 256 MiB)
before every timed launch, warm-up first.  Columns: ms and GB/s on the algorithmic bytes
of each path, and the speed-up of the sparse step over the dense one.
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for _p in (ROOT, os.path.join(ROOT, "cann-ops-adv-vllm_b200")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import torch  # noqa: E402

import bench  # noqa: E402


def time_fn(fn, flush, steps, warmup=3):
    for _ in range(warmup):
        flush()
        fn()
    torch.cuda.synchronize()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    for a, b in ev:
        flush()
        a.record()
        fn()
        b.record()
    torch.cuda.synchronize()
    t = sorted(x.elapsed_time(y) for x, y in ev)
    return sum(t) / len(t), t[len(t) // 2]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--out", default="")
    ap.add_argument("--seqs", default="8192,16384,32768,65536,131072")
    ap.add_argument("--ks", default="8,16,32,64,128")
    args = ap.parse_args()
    assert torch.cuda.is_available(), "needs a CUDA device (no CPU fallback exists)"
    import custom_ops
    dev = torch.device("cuda", 0)
    base = bench.WORKLOADS["c2"]
    flush = bench.L2Flush(dev)
    rows = []
    for S in [int(x) for x in args.seqs.split(",")]:
        w = dict(base, S=S)
        inp = bench.make_inputs(w, w["B"], 0, 1, dev)
        B, nq, nkv, P = inp["B"], inp["Nq"], inp["Nkv"], inp["P"]
        qd = inp["qd"]
        out = torch.empty(B, nq, bench.D, dtype=qd, device=dev)
        e = torch.empty(0, dtype=qd, device=dev)
        e64 = torch.empty(0, dtype=torch.int64, device=dev)
        # dense: identity block_position per KV head, every page of the context
        ident = torch.arange(P, dtype=torch.int32, device=dev).expand(B, nkv, P).contiguous()
        asl = torch.full((B,), S, dtype=torch.int64, device=dev)

        def dense():
            return custom_ops.sparse_paged_attention(
                inp["q"], inp["key"], inp["value"], e, e, asl, e, e, e, e, e, e, e, inp["block_table"], e64,
                ident, nq, inp["scale"], "BNSD", nkv, bench.BS, 1)

        d_ms, d_med = time_fn(dense, flush, args.steps)
        dense_bytes = B * nkv * P * bench.BS * bench.D * 2 * 2
        for k in [int(x) for x in args.ks.split(",")]:
            MP = max(256, P)
            bp = torch.empty(B, nq, MP, dtype=torch.int32, device=dev)
            ppl = torch.empty(B, nq, 8, dtype=torch.int32, device=dev)
            mppl = torch.empty(B, 8, dtype=torch.int64, device=dev)
            plan = custom_ops.FusedDecodePlan(
                inp["q"], inp["key"], inp["value"], inp["block_table"], inp["l1"], inp["block_ids"], inp["seq"],
                bp, ppl, mppl, out, nq, inp["scale"], nkv, bench.BS, k=k)
            s_ms, s_med = time_fn(plan.launch, flush, args.steps)
            inp["MP"] = MP
            ab = bench.algorithmic_bytes(inp, bp, ppl)
            rows.append(dict(S=S, C=base["C"], k=k, ratio=f"1/{base['C'] // k}", sparse_ms=s_ms, dense_ms=d_ms,
                             sparse_gbs=ab["bytes_alg"] / s_ms / 1e6, dense_gbs=dense_bytes / d_ms / 1e6,
                             union_frac=ab["bytes_kv_union"] / ab["bytes_dense"],
                             speedup=d_ms / s_ms, tokens_per_s=B / (s_ms * 1e-3)))
            print(json.dumps(rows[-1]), flush=True)
            del plan
        del inp
        torch.cuda.empty_cache()
    lines = ["| S | k/C | union pages / all pages | sparse ms | sparse GB/s (alg.) | dense ms | dense GB/s | dense/sparse | tokens/s |",
             "|---|---|---|---|---|---|---|---|---|"]
    for r in rows:
        lines.append(f"| {r['S'] // 1024}k | {r['ratio']} | {r['union_frac']:.3f} | {r['sparse_ms']:.4f} | {r['sparse_gbs']:.0f} | "
                     f"{r['dense_ms']:.4f} | {r['dense_gbs']:.0f} | {r['speedup']:.2f}x | {r['tokens_per_s']:.0f} |")
    text = "\n".join(lines)
    print(text)
    if args.out:
        with open(args.out, "w") as f:
            f.write("# Sparsity x context sweep (BASELINE configs[4]): B=16, 32Q/8KV, D=128, C=512, fp16, 1 x B200\n\n"
                    "`python tools/sweep.py` -- fused SparsePagedFusionAttention vs the dense path on the same kernel\n"
                    "(SparsePagedAttention with identity block_position); CUDA events, L2 flushed before every launch.\n\n")
            f.write(text + "\n")


if __name__ == "__main__":
    main()
