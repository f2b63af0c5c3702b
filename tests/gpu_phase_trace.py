"""Per-phase timing of the fused chain kernel from its in-kernel %globaltimer trace
(diagnostic, not a test):  python tests/gpu_phase_trace.py [workload]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "cann-ops-adv-vllm_b200"))
import bench  # noqa: E402
import custom_ops  # noqa: E402


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "c2"
    w = bench.WORKLOADS[name]
    dev = torch.device("cuda", 0)
    inp = bench.make_inputs(w, w["B"], 0, 1, dev)
    B, nq, nkv = inp["B"], inp["Nq"], inp["Nkv"]
    bp = torch.empty(B, nq, inp["MP"], dtype=torch.int32, device=dev)
    ppl = torch.empty(B, nq, 8, dtype=torch.int32, device=dev)
    mppl = torch.empty(B, 8, dtype=torch.int64, device=dev)
    out = torch.empty(B, nq, 128, dtype=inp["qd"], device=dev)
    i8 = w["kv"] == "i8"
    plan = custom_ops.FusedDecodePlan(inp["q"], inp["key"], inp["value"], inp["block_table"], inp["l1"],
                                      inp["block_ids"], inp["seq"], bp, ppl, mppl, out, nq, inp["scale"], nkv, 128,
                                      k=inp["k"], antiquant_scale=inp["aq_scale"] if i8 else None,
                                      antiquant_offset=inp["aq_offset"] if i8 else None)
    flush = bench.L2Flush(dev)
    for _ in range(3):
        plan.launch()
    torch.cuda.synchronize()
    names = ["A select", "sync1", "B stream", "sync2", "C combine"]
    for rep in range(3):
        if not os.environ.get('NOFLUSH'):
            flush()
        t = plan.trace_phases().cpu().double()
        t0 = t[:, 0].min()
        rel = (t - t0) / 1e3                                     # us
        print(f"[{name}] rep {rep}: kernel span {rel[:, 5].max():.1f} us; start skew {rel[:, 0].max():.1f} us")
        for i, nm in enumerate(names):
            d = rel[:, i + 1] - rel[:, i]
            print(f"   {nm:10s} per-CTA min {d.min():7.1f}  mean {d.mean():7.1f}  max {d.max():7.1f} us | "
                  f"phase ends at max {rel[:, i + 1].max():7.1f} us")
        if rep == 2:
            sub = ["gathers issued", "scores(mma)", "grp sync", "topk", "grp sync", "page masks", "compaction", "union"]
            have = t[:, 15] > 0
            prev = t[have, 0]
            for i, nm in enumerate(sub):
                cur = t[have, 8 + i]
                d = (cur - prev) / 1e3
                print(f"      A.{nm:15s} mean {d.mean():6.2f} max {d.max():6.2f} us (thread 0 of the CTA, last item)")
                prev = cur


if __name__ == "__main__":
    main()
