#!/usr/bin/env python
"""HBM roofline of the page -> cluster assignment kernel: python tools/bench_assign.py

Assigns every page of a c2-sized K pool (4096 pages x 8 KV heads, fp16, C=512) in one launch.  Algorithmic
bytes = pages * Nkv * bs * D * 2 (each K page-head slice once; the 1 MiB centroid table stays in L2);
CUDA events, L2 flushed before every timed launch."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for _p in (ROOT, os.path.join(ROOT, "cann-ops-adv-vllm_b200")):
    sys.path.insert(0, _p)
import torch  # noqa: E402

import bench  # noqa: E402


def main():
    import custom_ops
    dev = torch.device("cuda", 0)
    pages, Nkv, C = 4096, 8, 512
    g = torch.Generator(device=dev).manual_seed(1)
    key = torch.empty(pages, 128, Nkv, 128, dtype=torch.float16, device=dev)
    for p0 in range(0, pages, 512):
        key[p0:p0 + 512] = torch.randn(512, 128, Nkv, 128, generator=g, device=dev).half()
    l1 = torch.randn(Nkv, C, 128, generator=g, device=dev).half()
    ids = torch.arange(pages, dtype=torch.int32, device=dev)
    bid = torch.zeros(Nkv, pages, dtype=torch.int32, device=dev)
    flush = bench.L2Flush(dev)
    for _ in range(3):
        flush()
        custom_ops.assign_page_clusters(key, l1, ids, bid)
    torch.cuda.synchronize()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(20)]
    for a, b in ev:
        flush()
        a.record()
        custom_ops.assign_page_clusters(key, l1, ids, bid)
        b.record()
    torch.cuda.synchronize()
    ms = sum(a.elapsed_time(b) for a, b in ev) / len(ev)
    nbytes = pages * Nkv * 128 * 128 * 2
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
    gbs = nbytes / ms / 1e6
    print(json.dumps({"kernel": "spa_assign_kernel", "pages": pages, "kv_heads": Nkv, "clusters": C, "ms": ms,
                      "pages_per_s": pages / (ms * 1e-3), "bytes_alg": nbytes,
                      "roofline": {"bound": "hbm", "achieved": gbs, "peak": peak, "unit": "GB/s", "frac": gbs / peak}}))


if __name__ == "__main__":
    main()
