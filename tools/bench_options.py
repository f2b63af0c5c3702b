#!/usr/bin/env python
"""Device time of the fused op with the inherited IFA options (attenMask / pseShift / int8 post-quant)
next to the plain call, on a bench.py workload (default c2).  The options run in the options build of the chain
kernel (csrc/spa_kernels_opt.cu): one page per consumer step, the mask bytes and pse values read from global memory
at their logical positions.

    python tools/bench_options.py [--workload c2] [--steps 30]      -> one JSON line per variant

Same rules as bench.py: inputs resident in HBM, L2 flushed between steps, CUDA events on the launching stream,
5 warm-up steps.  The op is called through custom_ops.sparse_paged_fusion_attention (plan cache on: a repeated call
is one launch).  Algorithmic bytes = bench.py's for the selection produced + the mask / pse bytes of the union pages.
"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="c2")
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    args = ap.parse_args()
    import custom_ops                                                # (bench.py put the package on sys.path)
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    w = bench.WORKLOADS[args.workload]
    inp = bench.make_inputs(w, w["B"], 0, 1, dev)
    B, nq, nkv, qd, P = inp["B"], inp["Nq"], inp["Nkv"], inp["qd"], inp["P"]
    smax = P * bench.BS
    gen = torch.Generator(device=dev).manual_seed(7)
    mask = torch.rand(B, 1, 1, smax, generator=gen, device=dev) < 0.3
    pse = torch.randn(B, nq, 1, smax, generator=gen, device=dev).to(qd)
    q2 = torch.full((1,), 100.0, dtype=torch.float32, device=dev)
    e = torch.empty(0, dtype=qd, device=dev)
    e64 = torch.empty(0, dtype=torch.int64, device=dev)
    bp = torch.empty(B, nq, inp["MP"], dtype=torch.int32, device=dev)
    ppl = torch.empty(B, nq, 8, dtype=torch.int32, device=dev)
    mppl = torch.empty(B, 8, dtype=torch.int64, device=dev)
    i8 = w["kv"] == "i8"
    flush = bench.L2Flush(dev)
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            peak = float(json.load(f).get("hbm_gbs", 6650.0))
    except OSError:
        peak = 6650.0
    variants = [("plain", {}), ("mask", dict(mask=mask)), ("mask_pse", dict(mask=mask, pse=pse)),
                ("post_quant", dict(q2=q2)), ("mask_pse_post_quant", dict(mask=mask, pse=pse, q2=q2))]
    for name, o in variants:
        def step():
            return custom_ops.sparse_paged_fusion_attention(
                inp["q"], inp["key"], inp["value"], o.get("pse", e), o.get("mask", e), e64, e, e, e, o.get("q2", e), e,
                inp["aq_scale"] if i8 else e, inp["aq_offset"] if i8 else e, inp["block_table"], e64, inp["l1"],
                inp["block_ids"], inp["seq"], bp, ppl, mppl, nq, inp["scale"], "BNSD", nkv, bench.BS, 1, k=inp["k"])
        for _ in range(args.warmup):
            flush()
            step()
        torch.cuda.synchronize()
        ab = bench.algorithmic_bytes(inp, bp, ppl)
        extra = 0
        if "mask" in o:
            extra += ab["union_pages"] * bench.BS                       # one byte per position of a union page (per kv head item)
        if "pse" in o:
            extra += ab["union_pages"] * bench.BS * (nq // nkv) * 2     # g heads x 2 bytes per position
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
        for s0, s1 in ev:
            flush()
            s0.record()
            step()
            s1.record()
        torch.cuda.synchronize()
        ms = sorted(a.elapsed_time(b) for a, b in ev)
        mean = sum(ms) / len(ms)
        gbs = (ab["bytes_alg"] + extra) / (mean * 1e-3) / 1e9
        rec = {"variant": name, "workload": w["desc"], "ms_per_step": round(mean, 4), "ms_median": round(ms[len(ms) // 2], 4),
               "tokens_per_s": round(B / (mean * 1e-3), 1), "algorithmic_bytes": ab["bytes_alg"] + extra,
               "option_bytes": extra, "achieved_GBps": round(gbs, 1), "steps": args.steps, "warmup": args.warmup,
               "l2": "flushed between steps", "api": "custom_ops.sparse_paged_fusion_attention (plan cache)"}
        if peak:
            rec["frac_of_measured_hbm_peak"] = round(gbs / peak, 3)
        print(json.dumps(rec), flush=True)


if __name__ == "__main__":
    main()
