"""Per-phase timing of the fused chain kernel from its in-kernel %globaltimer trace
(diagnostic, not a test):  python tests/gpu_phase_trace.py [workload]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.environ.get("SPA_PKG_DIR") or os.path.join(ROOT, "cann-ops-adv-vllm_b200"))
import bench  # noqa: E402
import custom_ops  # noqa: E402


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "c2"
    w = bench.WORKLOADS[name]
    dev = torch.device("cuda", 0)
    inp = bench.make_inputs(w, w["B"], 0, 1, dev, cent_dist=os.environ.get("CENT_DIST", "rand"))
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
    spans = [("A select", 0, 1), ("B stream+combine", 1, 3), ("exit", 3, 5)]
    for rep in range(3):
        if not os.environ.get('NOFLUSH'):
            flush()
        t = plan.trace_phases().cpu().double()
        t0 = t[:, 0].min()
        rel = (t - t0) / 1e3                                     # us
        print(f"[{name}] rep {rep}: kernel span {rel[:, 5].max():.1f} us; start skew {rel[:, 0].max():.1f} us")
        for nm, a, b in spans:
            d = rel[:, b] - rel[:, a]
            print(f"   {nm:18s} per-CTA min {d.min():7.1f}  mean {d.mean():7.1f}  max {d.max():7.1f} us | "
                  f"ends at min {rel[:, b].min():7.1f} max {rel[:, b].max():7.1f} us")
        order = torch.argsort(rel[:, 3], descending=True)[:6]
        for c in order.tolist():
            print(f"      late CTA {c:3d}: A end {rel[c, 1]:6.1f}  last claim {rel[c, 2]:6.1f}  producer exit {rel[c, 7]:6.1f}  "
                  f"B end {rel[c, 3]:6.1f} us  pages {int(t[c, 4])}  runs {int(t[c, 6])}")
            if os.environ.get("SPA_DBG_REPEAT"):
                print(f"           sentinel seen {rel[c, 8]:6.1f}  last flush done {rel[c, 9]:6.1f}  combines(after sentinel) {int(t[c, 10])} "
                      f"taking {t[c, 11] / 1e3:6.1f} us")
        print(f"      pages per CTA min {int(t[:, 4].min())} mean {t[:, 4].mean():.1f} max {int(t[:, 4].max())}; "
              f"runs per CTA mean {t[:, 6].mean():.1f} max {int(t[:, 6].max())}")
        if rep == 2 and os.environ.get("SPA_SMID"):
            smid = t[:, 12].long(); pages = t[:, 4].long()
            idx = torch.argsort(smid)
            print("      smid:pages " + " ".join(f"{int(smid[i])}:{int(pages[i])}" for i in idx.tolist()))
        if rep == 2:
            sub = ["set-up", "scores(mma)", "hist", "bin", "classify", "resolve", "page masks", "compaction"]
            have = t[:, 15] > 0
            prev = t[have, 0]
            for i, nm in enumerate(sub):
                cur = t[have, 8 + i]
                d = (cur - prev) / 1e3
                print(f"      A.{nm:15s} mean {d.mean():6.2f} max {d.max():6.2f} us (thread 0 of the CTA, last item)")
                prev = cur
            ck = t[have][:, 16:32]
            names2 = ["S0: issue copies, init, wait block_table", "issue block_ids gather, q l1 norms", "centroid tiles (mma)",
                      "score reductions", "sync", "hist", "sync", "bin", "sync", "classify", "sync", "resolve",
                      "wait gather, masks, sync", "compaction (pass1, prefix, pass2, lengths)", "publish"]
            for i, nm in enumerate(names2):
                d = ck[:, i + 1] - ck[:, i]
                print(f"      clk {nm:45s} mean {d.mean():8.0f} max {d.max():8.0f} cycles")

if __name__ == "__main__":
    main()
