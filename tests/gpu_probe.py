"""Diagnostic run of every op against the oracle (prints, does not assert).  Used while bringing
the kernels up on the GPU box:  python tests/gpu_probe.py"""
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "cann-ops-adv-vllm_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import custom_ops  # noqa: E402
from cases import attn_call, fused_call, make_case, to_device  # noqa: E402
from oracle import sparse_chain as oc  # noqa: E402


def report(name, got, want, exact=True):
    got = got.cpu()
    if exact:
        ok = torch.equal(got, want)
        nbad = int((got != want).sum()) if got.shape == want.shape else -1
        print(f"  {name}: {'OK' if ok else 'MISMATCH'} (bad={nbad}/{want.numel()})", flush=True)
        if not ok and got.shape == want.shape:
            idx = (got != want).nonzero()[:5]
            for i in idx:
                t = tuple(i.tolist())
                print("     at", t, "got", got[t].item(), "want", want[t].item())
    else:
        d = (got.float() - want.float()).abs()
        rel = d / want.float().abs().clamp_min(1e-3)
        print(f"  {name}: max_abs={d.max().item():.3e} max_rel={rel.max().item():.3e} "
              f"nan={int(torch.isnan(got.float()).sum())}", flush=True)
        if d.max() > 1e-2 or torch.isnan(got.float()).any():
            bad = (d > 1e-2) | torch.isnan(got.float())
            print("     bad per (b,h):", bad.reshape(bad.shape[0], bad.shape[1], -1).sum(-1).tolist())


def main():
    dev = "cuda"
    torch.manual_seed(0)
    for (B, Nq, Nkv, C, S, k) in [(2, 8, 2, 128, 4096, 8), (2, 32, 8, 512, 32768, 64), (1, 16, 2, 256, 8192 - 77, 16), (1, 32, 2, 128, 8192 - 3, 16)]:
        print(f"== case B={B} Nq={Nq} Nkv={Nkv} C={C} S={S} k={k}", flush=True)
        c = make_case(100 + B + Nq, B, Nq, Nkv, C, S, kv_pages=max(64, 2 * ((S + 127) // 128)), ragged=(S % 128 != 0))
        g = to_device(c, dev)
        t0 = time.time()
        want_idx = oc.compute_cent(c["q"], c["l1"], k)
        got_idx = custom_ops.compute_cent(g["q"], g["l1"], k=k)
        torch.cuda.synchronize()
        report("compute_cent", got_idx, want_idx)
        wp, wl, wg = oc.select_position(c["block_ids"], c["block_table"], c["seq"], want_idx, 128, 256)
        gp, gl, gg = custom_ops.select_position(g["block_ids"], g["block_table"], g["seq"], got_idx)
        torch.cuda.synchronize()
        report("select.page_position", gp, wp)
        report("select.length", gl, wl)
        report("select.gather", gg, wg)
        wp, wl, wm = oc.cent_select(c["q"], c["l1"], c["block_ids"], c["block_table"], c["seq"], k, 128, 256)
        gp, gl, gm = custom_ops.cent_select(g["q"], g["l1"], g["block_ids"], g["block_table"], g["seq"], k=k)
        torch.cuda.synchronize()
        report("cent_select.page_position", gp, wp)
        report("cent_select.length", gl, wl)
        report("cent_select.max", gm, wm)
        # separate attention on the oracle's selection
        want = oc.sparse_paged_attention(c["q"], c["key"], c["value"], wm[:, 0], c["block_table"], wp, Nq,
                                         c["scale"], Nkv, 128)
        got = attn_call(custom_ops, g, wp.to(dev), wm[:, 0].to(dev))
        torch.cuda.synchronize()
        report("sparse_paged_attention", got, want, exact=False)
        # fused
        wout, wpos, wlen, wmx = oc.sparse_paged_fusion_attention(
            c["q"], c["key"], c["value"], c["block_table"], c["l1"], c["block_ids"], c["seq"], Nq, c["scale"],
            Nkv, 128, k, 256)
        out, bp, ppl, mppl, _ = fused_call(custom_ops, g, k)
        torch.cuda.synchronize()
        report("fused.block_position", bp, wpos)
        report("fused.length", ppl, wlen)
        report("fused.max", mppl, wmx)
        report("fused.attention_out", out, wout, exact=False)
        print(f"  ({time.time() - t0:.1f}s)", flush=True)


if __name__ == "__main__":
    main()
