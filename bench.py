#!/usr/bin/env python
"""Benchmark of the fused centroid-sparse paged decode-attention step (SparsePagedFusionAttention).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--workload c2]

One "step" = one fused decode-attention call for one layer: ComputeCent + SelectPosition + sparse
paged flash-decode for B sequences (BASELINE.json metric: decode attn tokens/s + achieved HBM GB/s
on the selected-KV bytes).  Prints ONE JSON line on rank 0.

  value     tokens/s with every input resident in HBM, launch-ready executor (FusedDecodePlan),
            CUDA events on the launching stream, L2 flushed before every timed step.
  e2e       same metric through `custom_ops.sparse_paged_fusion_attention` (the reference-facing
            plugin call: validation, descriptors, workspace allocation) with the step's inputs
            (query, block_table, total_seq_len) copied from pinned host memory and attention_out
            copied back to pinned host memory inside the timed region.  The paged KV pool, the
            centroids and block_ids are the resident cache state of the serving engine (they live
            in device memory in the reference too) and are not per-step inputs.
  roofline  algorithmic bytes of one launch (union-of-group selected K+V page slices +
            centroids + index tensors + q/out, computed from the selection the kernel produced)
            / CUDA-event duration of that launch, against MEASURED_PEAKS.json hbm_gbs.  `traffic` is
            the ncu-measured DRAM bytes of one launch (profiles/ncu_traffic.json), reported only while
            that capture belongs to the kernel sources being timed (content hash), else null.
  cpu_baseline  the CPU oracle (port of the reference golden chain as written + attention
            restatement) timed on this box's host cores on the same workload (whole batch at c1/c2,
            else a bounded sample); cpu_baseline_vectorised = the same chain with batched torch ops
            (oracle/fast_chain.py), the "fair CPU" number of BASELINE.md section 3.
  extra_workloads (N=1, default run)  the same measurement for BASELINE configs[2] (c3: int8 KV),
            configs[3] (c4: 70B shape), c2 with randn centroids (independent heads: larger union), c3 with
            contiguous page-head slices / FP8 pages, and c2 with attention_mask + pse_shift (options build).
  strong_c4 (N>1)  STRONG scaling of configs[3] (c4, B=32 fixed, KV heads split over the ranks):
            ms/step at N, at N=1 (rank 0 alone, same run), speed-up and efficiency.
  gather_verified (N>1)  the fused peer-store all-gather compared, outside the timed region, bit for
            bit with an NCCL all-gather of the same shards and with the oracle on one sequence; the
            run fails if it differs.

Multi-GPU (torchrun, one rank per GPU): KV-head tensor parallelism -- rank r owns KV heads
[r*Nkv/N, (r+1)*Nkv/N) with their query heads, centroids, block_ids and a rank-local KV pool; the
only exchange is the all-gather of attention_out, fused into the launch (NVLink peer stores) or by
NCCL (--gather nccl).  Default scaling is "weak": the batch grows with N (B = 16*N) so the bytes
streamed per GPU stay fixed; `--scaling strong` keeps the global workload fixed.
"""
import argparse
import json
import math
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
# SPA_PKG_DIR: another build of the package (same-box A/B of two kernel builds, tools/gpu_batch.sh)
for _p in (ROOT, os.environ.get("SPA_PKG_DIR") or os.path.join(ROOT, "cann-ops-adv-vllm_b200")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import torch  # noqa: E402

BS = 128
D = 128


def kernel_source_sha():
    """Content hash of the kernel sources + the C ABI header: identifies the build an ncu capture belongs to
    (git is not available on the GPU box: .git does not travel with the snapshot)."""
    import glob
    import hashlib
    h = hashlib.sha256()
    for f in sorted(glob.glob(os.path.join(ROOT, "cann-ops-adv-vllm_b200", "csrc", "*.cu*"))) + \
            [os.path.join(ROOT, "include", "spa_b200.h")]:
        with open(f, "rb") as fh:
            h.update(fh.read())
    return h.hexdigest()[:16]


def measured_traffic(workload):
    """dram__bytes_read.sum + dram__bytes_write.sum of ONE chain-kernel launch from the committed `ncu --set full`
    capture (profiles/ncu_traffic.json, written by tools/ncu_traffic.py next to the summary it was read from).
    Reported only while the capture belongs to the kernel sources being timed; otherwise null."""
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
            t = json.load(f)
    except (OSError, ValueError):
        return None, "no capture committed (profiles/ncu_traffic.json)"
    rec = t.get("workloads", {}).get(workload)
    if rec is None:
        return None, f"no capture of {workload}"
    if rec.get("src_sha") != kernel_source_sha():
        return None, f"capture {rec.get('summary')} belongs to kernel sources {rec.get('src_sha')}, not the ones timed"
    return int(rec["dram_read"]) + int(rec["dram_write"]), f"ncu --set full, one launch ({rec.get('summary')})"

WORKLOADS = {
    # BASELINE.json configs[0..3]
    "c1": dict(B=1, Nq=32, Nkv=8, S=4096, C=64, k=8, q="f16", kv="f16",
               desc="CPU golden case: B=1 32Q/8KV D=128 seq 4096 C=64 top-k=8 fp16"),
    "c2": dict(B=16, Nq=32, Nkv=8, S=32768, C=512, k=64, q="f16", kv="f16",
               desc="Llama-3-8B decode: B=16 32Q/8KV D=128 seq 32k C=512 top-k=64 fp16 KV, fused"),
    "c3": dict(B=64, Nq=32, Nkv=8, S=32768, C=512, k=64, q="bf16", kv="i8",
               desc="Llama-3-8B decode, int8 KV (antiquant), bf16 query: B=64 seq 32k C=512 top-k=64, fused"),
    "c3_fp8": dict(B=64, Nq=32, Nkv=8, S=32768, C=512, k=64, q="bf16", kv="f8",
                   desc="c3 with FP8 (e4m3) KV pages instead of int8 + antiquant: B=64 seq 32k C=512 top-k=64, bf16 query, fused"),
    # the inherited IFA options: attention_mask (30 % of the positions) + pse_shift [B, Nq, 1, S] on c2;
    # runs the options build of the chain kernel (csrc/spa_kernels_opt.cu)
    "c2_mask_pse": dict(B=16, Nq=32, Nkv=8, S=32768, C=512, k=64, q="f16", kv="f16", opts="mask_pse",
                        desc="c2 with attention_mask + pse_shift (options build): B=16 32Q/8KV D=128 seq 32k C=512 top-k=64 fp16 KV, fused"),
    "c4": dict(B=32, Nq=64, Nkv=8, S=131072, C=2048, k=128, q="f16", kv="f16",
               desc="Llama-3-70B decode: B=32 64Q/8KV D=128 seq 128k C=2048 top-k=128 fp16 KV, fused"),
    # corners of the configs[4] sweep (tools/sweep.py), for traces
    "c5_8k_64": dict(B=16, Nq=32, Nkv=8, S=8192, C=512, k=8, q="f16", kv="f16",
                     desc="sweep corner: B=16 32Q/8KV seq 8k C=512 top-k=8 (k/C = 1/64) fp16"),
    "c5_128k_64": dict(B=16, Nq=32, Nkv=8, S=131072, C=512, k=8, q="f16", kv="f16",
                       desc="sweep corner: B=16 32Q/8KV seq 128k C=512 top-k=8 (k/C = 1/64) fp16"),
}
METRIC = "decode_attn_tokens_per_s"
UNIT = "tokens/s"


def qdtype(w):
    return torch.float16 if w["q"] == "f16" else torch.bfloat16


# ------------------------------------------------------------------------------------------------
# synthetic inputs: generated on the device, seeded; block_table is a per-sequence
# permutation of distinct physical pages so aliasing does not inflate L2 hits
# ------------------------------------------------------------------------------------------------
def make_inputs(w, B, rank, world, device, seed=1234, kv_layout="bsnd", cent_dist="rand"):
    """kv_layout: "bsnd" = vLLM pool [blocks, bs, Nkv, D] (primary, tests/ifa_v5_case.py:608-610);
    "bnsd" = [blocks, Nkv, bs, D] (the reference's BNSD-paged KAT layout, tests/ifa_v5_case.py:36)."""
    Nq, Nkv, S, C = w["Nq"], w["Nkv"], w["S"], w["C"]
    assert Nkv % world == 0, "KV heads must divide across ranks"
    nkv = Nkv // world
    nq = Nq // world
    P = (S + BS - 1) // BS
    pages = B * P
    qd = qdtype(w)
    g_common = torch.Generator(device=device).manual_seed(seed)
    g_rank = torch.Generator(device=device).manual_seed(seed + 1000 * (rank + 1))
    block_table = torch.randperm(pages, generator=g_common, device=device).to(torch.int32).view(B, P)
    seq = torch.full((B,), S, dtype=torch.int32, device=device)
    q = torch.rand(B, nq, D, generator=g_rank, device=device).to(qd)
    # "rand": selection tests of the reference (tests/cent_select.py:49); "randn": its fused test (fused_ifa_v5_case.py:55)
    # -- with zero-mean centroids the heads of a GQA group rank the clusters independently (larger union)
    l1 = (torch.rand if cent_dist == "rand" else torch.randn)(nkv, C, D, generator=g_rank, device=device).to(qd)
    block_ids = torch.randint(0, C, (nkv, pages), generator=g_rank, device=device, dtype=torch.int32)
    pool_shape = (pages, BS, nkv, D) if kv_layout == "bsnd" else (pages, nkv, BS, D)
    if w["kv"] == "f8":
        key = torch.empty(pool_shape, dtype=torch.float8_e4m3fn, device=device)
        value = torch.empty_like(key)
        aq_scale = torch.empty(0, dtype=qd, device=device)
        aq_offset = torch.empty(0, dtype=qd, device=device)
    elif w["kv"] == "i8":
        key = torch.empty(pool_shape, dtype=torch.int8, device=device)
        value = torch.empty_like(key)
        aq_scale = (torch.rand(2, nkv, D, generator=g_rank, device=device) * 0.02 + 0.005).to(qd)
        aq_offset = torch.randn(2, nkv, D, generator=g_rank, device=device).to(qd)
    else:
        key = torch.empty(pool_shape, dtype=qd, device=device)
        value = torch.empty_like(key)
        aq_scale = torch.empty(0, dtype=qd, device=device)
        aq_offset = torch.empty(0, dtype=qd, device=device)
    chunk = max(1, (64 << 20) // (BS * nkv * D))              # ~64 Mi elements per chunk
    for t in (key, value):
        for p0 in range(0, pages, chunk):
            n = min(chunk, pages - p0)
            if w["kv"] == "f8":
                t[p0:p0 + n] = torch.randn((n,) + pool_shape[1:], generator=g_rank, device=device).to(torch.float8_e4m3fn)
            elif w["kv"] == "i8":
                t[p0:p0 + n] = torch.randint(-128, 128, (n,) + pool_shape[1:], generator=g_rank, device=device,
                                             dtype=torch.int8)
            else:
                t[p0:p0 + n] = (torch.rand((n,) + pool_shape[1:], generator=g_rank, device=device) * (1 - 1e-3) + 1e-3).to(qd)
    return dict(q=q, l1=l1, block_ids=block_ids, block_table=block_table, seq=seq, key=key, value=value,
                aq_scale=aq_scale, aq_offset=aq_offset, B=B, Nq=nq, Nkv=nkv, C=C, S=S, P=P, k=w["k"],
                scale=1.0 / math.sqrt(D), MP=256, qd=qd, kv_layout=kv_layout)


def algorithmic_bytes(inp, page_position, page_position_length):
    """compulsory bytes of one fused launch, from the selection actually produced."""
    B, nq, nkv, P, MP = inp["B"], inp["Nq"], inp["Nkv"], inp["P"], inp["MP"]
    g = nq // nkv
    ln = page_position_length[:, :, 0].long()
    valid = torch.arange(MP, device=ln.device).view(1, 1, MP) < ln.unsqueeze(-1)
    pp = page_position.long().clamp(0, P - 1)
    onehot = torch.zeros(B, nq, P, dtype=torch.bool, device=ln.device)
    onehot.scatter_(2, pp, valid)
    union_pages = int(onehot.view(B, nkv, g, P).any(dim=2).sum().item())
    selected = int(ln.sum().item())
    kv_elem = inp["key"].element_size()
    q_elem = inp["q"].element_size()
    page_bytes = BS * D * 2 * kv_elem                        # K + V slice of one page for one kv head
    bytes_kv = union_pages * page_bytes
    bytes_sel = (nkv * inp["C"] * D * q_elem + nkv * B * P * 4 + B * P * 4 + B * nq * (MP + 8) * 4 + B * 8 * 8)
    bytes_io = 2 * B * nq * D * q_elem
    if kv_elem == 1 and inp["aq_scale"].numel():
        bytes_sel += 2 * 2 * nkv * D * q_elem
    return dict(bytes_alg=bytes_kv + bytes_sel + bytes_io, bytes_kv_union=bytes_kv,
                bytes_logical=selected * page_bytes, union_pages=union_pages, selected_pages=selected,
                bytes_dense=B * nkv * P * page_bytes)


class L2Flush:
    """Evict the 126 MB L2 between timed steps: write a 256 MiB buffer, then read a second 256 MiB
    buffer so that the lines left in L2 are CLEAN -- after a write-only flush the timed kernel would
    pay for the write-back of ~126 MB of dirty lines, a bench artefact rather than kernel traffic."""

    def __init__(self, device, nbytes=256 << 20):
        self.a = torch.empty(nbytes, dtype=torch.uint8, device=device)
        self.b = torch.zeros(nbytes // 4, dtype=torch.int32, device=device)
        self.sink = torch.zeros(1, dtype=torch.int64, device=device)

    def __call__(self):
        self.a.zero_()
        torch.sum(self.b, dim=0, keepdim=True, out=self.sink)


# ------------------------------------------------------------------------------------------------
# clocks sampler (B200_PROFILING.md "clocks DURING the timed region")
# ------------------------------------------------------------------------------------------------
class ClockSampler:
    FIELDS = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu = gpu_index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.gpu}", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for nm, v in zip(names, f[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------------------------------------
# CPU baseline (oracle port) on a bounded sample
# ------------------------------------------------------------------------------------------------
def cpu_sample(inp, nseq):
    """First `nseq` sequences of the workload, their pages compacted into a small CPU pool."""
    bt = inp["block_table"][:nseq].long()                          # [nseq, P]
    flat = bt.reshape(-1)
    c = dict(q=inp["q"][:nseq].cpu(), l1=inp["l1"].cpu(), seq=inp["seq"][:nseq].cpu(),
             key=inp["key"][flat].cpu(), value=inp["value"][flat].cpu(),
             block_ids=inp["block_ids"][:, flat].cpu().contiguous(),
             block_table=torch.arange(flat.numel(), dtype=torch.int32).view(nseq, -1),
             aq_scale=inp["aq_scale"].cpu(), aq_offset=inp["aq_offset"].cpu())
    return c


def run_oracle(c, inp, fast=False):
    from oracle import sparse_chain as oc
    i8 = c["key"].dtype == torch.int8
    if fast:
        from oracle import fast_chain
        return fast_chain.fused_chain(
            c["q"], c["key"], c["value"], c["block_table"], c["l1"], c["block_ids"], c["seq"], inp["Nq"],
            inp["scale"], inp["Nkv"], BS, inp["k"], inp["MP"], c["aq_scale"] if i8 else None,
            c["aq_offset"] if i8 else None)
    return oc.sparse_paged_fusion_attention(
        c["q"], c["key"], c["value"], c["block_table"], c["l1"], c["block_ids"], c["seq"], inp["Nq"],
        inp["scale"], inp["Nkv"], BS, inp["k"], inp["MP"], c["aq_scale"] if i8 else None,
        c["aq_offset"] if i8 else None)


def cpu_nseq(w):
    """Sequences of the workload the CPU legs run per step: the whole batch where the host copy of their pages
    stays modest (c1, c2: <= 2 GiB), else a bounded sample."""
    P = (w["S"] + BS - 1) // BS
    page_bytes = BS * w["Nkv"] * D * 2 * (1 if w["kv"] in ("i8", "f8") else 2)
    return max(1, min(w["B"], (2 << 30) // (P * page_bytes)))


def time_cpu_baseline(inp, w, budget_s=12.0):
    """The oracle port (reference golden as written: Python loops) and the vectorised variant, on the host cores."""
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    nseq = cpu_nseq(w)
    c = cpu_sample(inp, nseq)
    res = {}
    for fast in (False, True):
        best, reps, t_all = float("inf"), 0, time.perf_counter()
        while reps < 2 or (time.perf_counter() - t_all < budget_s and reps < 8):
            t0 = time.perf_counter()
            run_oracle(c, inp, fast)
            best = min(best, time.perf_counter() - t0)
            reps += 1
        what = "oracle/fast_chain.py (vectorised torch, loops over B*Nkv items only)" if fast else \
            "oracle/sparse_chain.py (the reference golden as written: Python loops over B*Nq rows)"
        res[fast] = {"value": nseq / best, "unit": UNIT, "cores": cores, "kind": "port",
                     "sample": f"{'all' if nseq == inp['B'] else 'first'} {nseq} of {inp['B']} sequences of the workload "
                               f"(all {inp['Nq']} q heads), fused chain by {what}, best of {reps}, torch threads={cores}"}
    return res[False], res[True]


# ------------------------------------------------------------------------------------------------
def bench_workload(args, name, ctx, steps, warmup, scaling="weak", cent_dist="rand", do_e2e=True, verify=False,
                   sample_clocks=True, kv_layout=None):
    """Time `steps` fused decode steps of workload `name` on the ranks of ctx (device-resident, then end to end).
    Returns the measurement record on rank 0 of ctx, None elsewhere."""
    import custom_ops
    rank, world, dev, dist = ctx["rank"], ctx["world"], ctx["dev"], ctx["dist"]
    w = WORKLOADS[name]
    B = w["B"] * world if scaling == "weak" else w["B"]
    kv_layout = kv_layout or args.kv_layout
    inp = make_inputs(w, B, rank, world, dev, kv_layout=kv_layout, cent_dist=cent_dist)
    nq, nkv, qd = inp["Nq"], inp["Nkv"], inp["qd"]
    bp = torch.empty(B, nq, inp["MP"], dtype=torch.int32, device=dev)
    ppl = torch.empty(B, nq, 8, dtype=torch.int32, device=dev)
    mppl = torch.empty(B, 8, dtype=torch.int64, device=dev)
    out = torch.empty(B, nq, D, dtype=qd, device=dev)

    # The step's inputs are staged the way a serving engine stages them: ONE pinned host blob (query | block_table |
    # seq_len, 256-byte aligned sections) -> one H2D copy per step into a device blob the op's tensors are views of.
    def _sec(n):
        return (n + 255) // 256 * 256
    nb_q, nb_bt, nb_seq = inp["q"].numel() * inp["q"].element_size(), inp["block_table"].numel() * 4, inp["seq"].numel() * 4
    o_bt, o_seq, nb_all = _sec(nb_q), _sec(nb_q) + _sec(nb_bt), _sec(nb_q) + _sec(nb_bt) + _sec(nb_seq)
    hblob = torch.zeros(nb_all, dtype=torch.uint8).pin_memory()
    dblob = torch.zeros(nb_all, dtype=torch.uint8, device=dev)
    hblob[:nb_q].view(qd).view(inp["q"].shape).copy_(inp["q"])
    hblob[o_bt:o_bt + nb_bt].view(torch.int32).view(inp["block_table"].shape).copy_(inp["block_table"])
    hblob[o_seq:o_seq + nb_seq].view(torch.int32).view(inp["seq"].shape).copy_(inp["seq"])
    sq = dblob[:nb_q].view(qd).view(inp["q"].shape)
    sbt = dblob[o_bt:o_bt + nb_bt].view(torch.int32).view(inp["block_table"].shape)
    sseq = dblob[o_seq:o_seq + nb_seq].view(torch.int32).view(inp["seq"].shape)
    dblob.copy_(hblob)
    gather = out_full = None
    i8 = w["kv"] == "i8"
    peer = world > 1 and args.gather == "peer"
    aq = dict(antiquant_scale=inp["aq_scale"] if i8 else None, antiquant_offset=inp["aq_offset"] if i8 else None)
    opt_mask = opt_pse = None
    if w.get("opts") == "mask_pse":
        assert world == 1, "the options workloads are single-GPU lines"
        g_opt = torch.Generator(device=dev).manual_seed(77)
        smax = inp["P"] * BS
        opt_mask = torch.rand(B, 1, 1, smax, generator=g_opt, device=dev) < 0.3
        opt_pse = torch.randn(B, nq, 1, smax, generator=g_opt, device=dev).to(qd)
    if world > 1:
        from kvhead_tp import HeadGather, head_shard
        gather = HeadGather(out.shape, qd, dev)                      # NCCL all-gather of the output shards
        out_full = torch.empty(B, nq * world, D, dtype=qd, device=dev)
    if peer:
        # the launch itself delivers every rank's rows to every rank (peer stores + flag exchange)
        from peer_gather import TPFusedDecode
        shard = head_shard(rank, world, nq * world, nkv * world)
        plan = TPFusedDecode(shard, nq * world, sq, inp["key"], inp["value"], sbt, inp["l1"], inp["block_ids"], sseq,
                             bp, ppl, mppl, inp["scale"], BS, k=inp["k"], **aq)
    else:
        plan = custom_ops.FusedDecodePlan(inp["q"], inp["key"], inp["value"], inp["block_table"], inp["l1"],
                                          inp["block_ids"], inp["seq"], bp, ppl, mppl, out, nq, inp["scale"], nkv, BS,
                                          k=inp["k"], pse_shift=opt_pse, attention_mask=opt_mask, **aq)
    flush = None if args.no_flush else L2Flush(dev)

    def step():
        r = plan.launch()
        if world > 1 and args.gather == "nccl":
            return gather(out, out_full)
        return r

    # N>1: the ranks are coupled by the output exchange at the end of every step, so a rank whose L2 flush ran a few
    # microseconds longer makes all the others wait inside THEIR timed region.  A one-element all-reduce after the
    # flush (outside the timed region) lines the ranks up again, as the collective in front of the attention does
    # in a tensor-parallel decoder layer.
    align_buf = torch.zeros(1, device=dev) if world > 1 else None

    def pre_step():
        if flush is not None:
            flush()
        if align_buf is not None and not args.no_align:
            dist.all_reduce(align_buf)

    graph = None
    if args.graph and peer:
        plan.capture()                                               # one graph per gathered buffer, replayed alternately
    elif args.graph:
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            step()
        torch.cuda.current_stream().wait_stream(s)
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            step()

    def run_step():
        if args.graph and peer:
            return plan.replay()
        return graph.replay() if graph is not None else step()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    for _ in range(warmup):
        pre_step()
        run_step()
    barrier()

    # ---------------- N>1: is the gathered output right?  (outside the timed region) ----------------
    gather_verified = None
    if verify and world > 1:
        got = step().clone()                                         # gathered [B, Nq_total, D] of this rank
        torch.cuda.synchronize()
        lo = rank * nq
        mine = got[:, lo:lo + nq].contiguous()
        ref_full = gather(mine, torch.empty_like(got))               # the same shards through NCCL
        ok = torch.equal(got, ref_full)                              # bit for bit: the exchange moved every row
        c1 = cpu_sample(inp, 1)                                      # this rank's heads of sequence 0 vs the oracle
        wo = run_oracle(c1, inp)[0].float()
        tol = 2e-2 if (w["kv"] != "f16" or w["q"] == "bf16") else 1e-3
        ok = ok and bool(((mine[:1].float().cpu() - wo).abs() <= tol + tol * wo.abs()).all())
        flag = torch.tensor([1 if ok else 0], device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        gather_verified = bool(flag.item())
        barrier()

    # ---------------- device-resident timing ----------------
    ab = algorithmic_bytes(inp, bp, ppl)
    if opt_mask is not None:
        # + one mask byte and g pse values (2 bytes) per position of every union page
        ab["bytes_alg"] += ab["union_pages"] * BS * (1 + 2 * (nq // nkv))
    sampler = ClockSampler(ctx["local_rank"]) if (rank == 0 and sample_clocks) else None
    if sampler:
        sampler.start()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    barrier()
    xch0 = (plan.out.exchange_ns(), plan.out.exchange_send_ns()) if peer else (0, 0)
    for s0, s1 in ev:
        pre_step()
        s0.record()
        run_step()
        s1.record()
    barrier()
    xch_us = None
    if peer:
        mine = torch.tensor([(plan.out.exchange_ns() - xch0[0]) / steps / 1e3,
                             (plan.out.exchange_send_ns() - xch0[1]) / steps / 1e3], dtype=torch.float64, device=dev)
        allx = [torch.zeros_like(mine) for _ in range(world)]
        dist.all_gather(allx, mine)
        xch_us = {"total_per_rank": [round(x[0].item(), 2) for x in allx],
                  "fence_and_flags_per_rank": [round(x[1].item(), 2) for x in allx]}
    times = [a.elapsed_time(b) for a, b in ev]                      # ms
    total_ms = torch.tensor([sum(times)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(total_ms, op=dist.ReduceOp.MAX)
    ms_per_step = total_ms.item() / steps
    st = sorted(times)

    # ---------------- end-to-end through the reference-facing call ----------------
    e2e = None
    if do_e2e:
        # N>1: a tensor-parallel engine consumes the gathered output on the DEVICE; what returns to the host here
        # is this rank's shard of it (the whole-job D2H volume is then one full output, not N of them)
        hout = torch.empty(B, nq, D, dtype=qd).pin_memory()
        dshard = torch.empty(B, nq, D, dtype=qd, device=dev) if world > 1 else None
        e = torch.empty(0, dtype=qd, device=dev)
        e64 = torch.empty(0, dtype=torch.int64, device=dev)

        def e2e_step():
            dblob.copy_(hblob, non_blocking=True)                    # query | block_table | seq_len: one H2D copy
            if peer:
                # same public surface as the TP serving step: the plan is bound to views of dblob, the launch
                # returns the gathered output of all ranks
                full = plan.launch()
                dshard.copy_(full[:, rank * nq:(rank + 1) * nq])
                hout.copy_(dshard, non_blocking=True)
                return
            o, _, _ = custom_ops.sparse_paged_fusion_attention(
                sq, inp["key"], inp["value"], e if opt_pse is None else opt_pse, e if opt_mask is None else opt_mask,
                e64, e, e, e, e, e, inp["aq_scale"], inp["aq_offset"], sbt, e64,
                inp["l1"], inp["block_ids"], sseq, bp, ppl, mppl, nq, inp["scale"], "BNSD", nkv, BS, 1, k=inp["k"])
            if world > 1 and args.gather == "nccl":
                gather(o, out_full)
            hout.copy_(o, non_blocking=True)

        for _ in range(warmup):
            e2e_step()
        barrier()
        ev2 = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        for s0, s1 in ev2:
            pre_step()
            s0.record()
            e2e_step()
            s1.record()
            s1.synchronize()                                         # the caller consumes hout every step
        barrier()
        tot2 = torch.tensor([sum(a.elapsed_time(b) for a, b in ev2)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tot2, op=dist.ReduceOp.MAX)
        e2e_ms = tot2.item() / steps
        e2e = {"value": B / (e2e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": nb_all,
               "d2h_bytes_per_step": hout.numel() * hout.element_size(), "ms_per_step": e2e_ms,
               "api": ("peer_gather.TPFusedDecode.launch (gathered output stays on the device; each rank returns its "
                       "own shard to the host; bytes are per rank)") if peer else
                      "custom_ops.sparse_paged_fusion_attention" + (" + kvhead_tp.HeadGather (NCCL); each rank returns its own shard; bytes are per rank" if world > 1 else "")}
    clocks = sampler.stop() if sampler else None
    if peer and plan.out.timed_out():
        raise SystemExit(f"rank {rank}: a flag exchange of the fused output gather timed out (a peer did not launch)")

    agg = torch.tensor([ab["bytes_alg"], ab["bytes_kv_union"], ab["bytes_logical"], ab["bytes_dense"]],
                       dtype=torch.float64, device=dev)             # whole job
    if world > 1:
        dist.all_reduce(agg, op=dist.ReduceOp.SUM)
    rec = None
    if rank == 0:
        achieved = ab["bytes_alg"] / (ms_per_step * 1e-3) / 1e9      # per-GPU kernel (rank 0's launch)
        traffic, traffic_src = measured_traffic(name) if (world == 1 and cent_dist == "rand" and kv_layout == "bsnd" and not w.get("opts")) else (None, "not captured for this configuration")
        rec = {
            "workload": name, "B": B, "ms_per_step": ms_per_step, "value": B / (ms_per_step * 1e-3),
            "achieved_hbm_gbs_selected_kv": agg[1].item() / (ms_per_step * 1e-3) / 1e9,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": ctx["hbm_peak"], "unit": "GB/s",
                         "frac": achieved / ctx["hbm_peak"], "traffic": traffic, "traffic_source": traffic_src,
                         "peak_source": ctx["peak_src"],
                         "kernel": "spa_chain_kernel (fused: select + flash-decode + combine)",
                         "bytes_alg_per_launch": ab["bytes_alg"], "bytes_kv_union": ab["bytes_kv_union"],
                         "bytes_logical_per_q_head": ab["bytes_logical"], "bytes_dense": ab["bytes_dense"],
                         "union_pages": ab["union_pages"], "selected_pages": ab["selected_pages"],
                         "union_over_selected": ab["union_pages"] / max(1, ab["selected_pages"])},
            "timing": {"min_ms": st[0], "median_ms": st[len(st) // 2], "p90_ms": st[int(len(st) * 0.9)], "max_ms": st[-1]},
            "e2e": e2e, "launches_per_step": plan.num_launches, "tp_flag_exchange_us_per_step": xch_us,
            "clocks": clocks, "gather_verified": gather_verified, "cent_dist": cent_dist,
            "kv_layout": "[blocks,bs,Nkv,D]" if kv_layout == "bsnd" else "[blocks,Nkv,bs,D]",
            "l2": ("flushed before every timed step (256 MiB write + 256 MiB read of a second buffer, so L2 holds clean lines); " if flush is not None else "not flushed; ") +
                  f"per-step read set {ab['bytes_kv_union'] / 2**20:.0f} MiB > 126 MB L2" +
                  ("; ranks aligned by a 1-element all-reduce before each timed step (outside the timed region)" if world > 1 and not args.no_align else ""),
            "launch": "cuda graph replay" if (graph is not None or (args.graph and peer)) else ("TPFusedDecode.launch()" if peer else "FusedDecodePlan.launch()"),
            "_inp": inp,
        }
    if peer:
        plan.close()
    return rec


# ------------------------------------------------------------------------------------------------
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS))
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"])
    ap.add_argument("--cent-dist", default="rand", choices=["rand", "randn"],
                    help="centroid distribution: rand = U[0,1) (tests/cent_select.py:49), randn (fused_ifa_v5_case.py:55)")
    ap.add_argument("--no-flush", action="store_true", help="do not flush L2 between timed steps")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extra", action="store_true",
                    help="main workload only: no extra_workloads (N=1) / strong_c4 (N>1) records")
    ap.add_argument("--no-align", action="store_true", help="N>1: no rank-aligning all-reduce before the timed steps")
    ap.add_argument("--graph", action="store_true", help="replay the step from a CUDA graph")
    ap.add_argument("--gather", default="peer", choices=["peer", "nccl", "none"],
                    help="N>1: output all-gather fused into the launch (NVLink peer stores, default) or NCCL after it")
    ap.add_argument("--kv-layout", default="bsnd", choices=["bsnd", "bnsd"],
                    help="paged KV pool layout: bsnd = [blocks,bs,Nkv,D] (vLLM, default), bnsd = [blocks,Nkv,bs,D]")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    w = WORKLOADS[args.workload]
    peaks = {}
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            peaks = json.load(f)
    except OSError:
        pass
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback 6650 GB/s"

    if args.impl == "reference":
        if rank != 0:
            return 0
        return reference_arm(args, w)

    assert torch.cuda.is_available(), "bench.py needs a CUDA device (no CPU fallback exists)"
    import custom_ops  # noqa: F401  (fails loudly if the extension is not built)
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
    ctx = dict(rank=rank, local_rank=local_rank, world=world, dev=dev, dist=dist, hbm_peak=hbm_peak, peak_src=peak_src)
    solo = dict(ctx, rank=0, world=1, dist=None)                    # one rank on its own (no collectives)

    main_rec = bench_workload(args, args.workload, ctx, args.steps, args.warmup, scaling=args.scaling,
                              cent_dist=args.cent_dist, verify=(world > 1 and args.gather != "none"))
    cpu = cpu_fair = None
    if rank == 0:
        inp = main_rec.pop("_inp")
        if world == 1 and not args.no_cpu_baseline:
            cpu, cpu_fair = time_cpu_baseline(inp, w)
        del inp
    torch.cuda.empty_cache()

    def brief(r):
        r.pop("_inp", None)
        return {"workload": WORKLOADS[r["workload"]]["desc"], "global_batch": r["B"], "cent_dist": r["cent_dist"],
                "kv_layout": r["kv_layout"],
                "ms_per_step": r["ms_per_step"], "value": r["value"], "unit": UNIT, "roofline": r["roofline"],
                "e2e": r["e2e"], "clocks": r["clocks"], "timing": r["timing"]}

    # ---- N=1: the other single-GPU BASELINE configs and the randn-centroid variant, in the same line, so that
    # their numbers are driver-visible too (each is one more fused launch per step, timed exactly like the headline)
    extra = None
    if world == 1 and not args.no_extra and args.workload == "c2" and args.cent_dist == "rand":
        extra = {}
        xs = max(10, min(args.steps, 30))
        # (c3_bnsd: int8 KV with contiguous page-head slices -- in the vLLM layout an int8 row is 128 B at a 1 KiB stride)
        for nm, cd, lay in (("c3", "rand", None), ("c4", "rand", None), ("c2", "randn", None), ("c3", "rand", "bnsd"),
                            ("c3_fp8", "rand", None), ("c2_mask_pse", "rand", None)):
            r = bench_workload(args, nm, ctx, xs, args.warmup, cent_dist=cd, kv_layout=lay)
            extra[nm + ("" if cd == "rand" else f"_{cd}") + (f"_{lay}" if lay else "")] = dict(brief(r), steps=xs)
            del r
            torch.cuda.empty_cache()

    # ---- N>1: STRONG scaling of BASELINE configs[3] (c4, B = 32 fixed, KV heads split over the ranks) next to the
    # weak-scaling headline: rank 0 first times the whole problem alone, then all ranks time their shards
    strong = None
    if world > 1 and not args.no_extra and args.scaling == "weak" and WORKLOADS["c4"]["Nkv"] % world == 0:
        xs = max(10, min(args.steps, 30))
        n1 = None
        if rank == 0:
            r1 = bench_workload(args, "c4", solo, xs, args.warmup, do_e2e=False, sample_clocks=False)
            n1 = r1["ms_per_step"]
            del r1
            torch.cuda.empty_cache()
        dist.barrier()
        rN = bench_workload(args, "c4", ctx, xs, args.warmup, scaling="strong", do_e2e=False, sample_clocks=False,
                            verify=args.gather != "none")
        if rank == 0:
            rN.pop("_inp", None)
            strong = {"workload": WORKLOADS["c4"]["desc"], "global_batch": rN["B"], "n_gpus": world, "steps": xs,
                      "ms_per_step": rN["ms_per_step"], "n1_ms_per_step": n1, "speedup_vs_n1": n1 / rN["ms_per_step"],
                      "efficiency": n1 / rN["ms_per_step"] / world, "value": rN["value"], "unit": UNIT,
                      "roofline_frac_per_gpu": rN["roofline"]["frac"], "gather_verified": rN["gather_verified"],
                      "tp_flag_exchange_us_per_step": rN["tp_flag_exchange_us_per_step"],
                      "n1_measured": "rank 0 alone, same run, same kernel, whole c4 problem on one GPU"}
        torch.cuda.empty_cache()

    if rank == 0:
        r = main_rec
        peer = world > 1 and args.gather == "peer"
        line = {
            "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True,
            "scaling": args.scaling, "vs_baseline": None, "dtype": "f16" if w["q"] == "f16" else "bf16",
            "data": "synthetic (seeded torch.rand/randint on device; block_table = random permutation of distinct pages)",
            "config": {"workload": f"{args.workload}: {w['desc']}", "global_batch": r["B"], "q_heads": w["Nq"],
                       "kv_heads": w["Nkv"], "head_dim": D, "seq_len": w["S"], "block_size": BS, "clusters": w["C"],
                       "top_k": w["k"], "kv_dtype": w["kv"], "q_dtype": w["q"], "cent_dist": args.cent_dist,
                       "kv_layout": "[blocks,bs,Nkv,D]" if args.kv_layout == "bsnd" else "[blocks,Nkv,bs,D]",
                       "parallelism": (f"kv-head tp{world}, output all-gather " + ("fused into the launch (NVLink peer stores)" if peer else "by NCCL")) if world > 1 else "single gpu",
                       "l2": r["l2"], "launch": r["launch"]},
            "achieved_hbm_gbs_selected_kv": r["achieved_hbm_gbs_selected_kv"],
            "roofline": r["roofline"], "timing": r["timing"], "e2e": r["e2e"],
            "gpu_launches": r["launches_per_step"] * args.steps,
            "tp_flag_exchange_us_per_step": r["tp_flag_exchange_us_per_step"],
            "clocks": r["clocks"], "kernel_src_sha": kernel_source_sha(),
        }
        if world > 1:
            line["gather_verified"] = r["gather_verified"]
        if cpu is not None:
            line["cpu_baseline"] = cpu
            line["cpu_baseline_vectorised"] = cpu_fair
        if extra is not None:
            line["extra_workloads"] = extra
        if strong is not None:
            line["strong_c4"] = strong
        print(json.dumps(line), flush=True)
    bad = world > 1 and rank == 0 and (main_rec["gather_verified"] is False or (strong and strong["gather_verified"] is False))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if bad:
        raise SystemExit("gathered output differs from the NCCL gather / the oracle: run REJECTED")
    return 0


def reference_arm(args, w):
    """The reference's CPU implementation of the path (its torch golden for the selection + the CPU restatement of
    the attention kernel), i.e. the oracle port as the reference wrote it (Python loops), on this box's host cores,
    on the b200 arm's config: the whole batch where the host copy of the pages stays modest (c1, c2), else a
    bounded sample.  The reference's kernels are AscendC and cannot be built here."""
    dev = torch.device("cuda", 0) if torch.cuda.is_available() else torch.device("cpu")
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    nseq = cpu_nseq(w)
    # inputs: same generator as the b200 arm when a GPU is present, else a CPU-generated sample
    if dev.type == "cuda":
        inp = make_inputs(w, w["B"], 0, 1, dev, cent_dist=args.cent_dist)
    else:
        small = dict(w); small["B"] = nseq
        inp = make_inputs(small, nseq, 0, 1, dev, cent_dist=args.cent_dist)
    c = cpu_sample(inp, nseq)
    warm = min(args.warmup, 3)
    for _ in range(warm):
        run_oracle(c, inp)
    steps = max(1, min(args.steps, 20))
    t0 = time.perf_counter()
    for _ in range(steps):
        run_oracle(c, inp)
    dt = (time.perf_counter() - t0) / steps
    v = nseq / dt
    t1 = time.perf_counter()
    run_oracle(c, inp, fast=True)
    vfast = nseq / (time.perf_counter() - t1)
    sample = (f"each step = {'all' if nseq == w['B'] else 'the first'} {nseq} of {w['B']} sequences of the workload (all {w['Nq']} q heads); "
              f"oracle/sparse_chain.py fused chain (reference golden as written); torch threads={cores}")
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": steps,
            "warmup": warm, "ms_per_step": dt * 1e3, "higher_is_better": True,
            "scaling": args.scaling, "vs_baseline": None, "dtype": "f16" if w["q"] == "f16" else "bf16",
            "data": "synthetic",
            "config": {"workload": f"{args.workload}: {w['desc']}", "global_batch": nseq, "sample": sample,
                       "cent_dist": args.cent_dist},
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
            "cpu_baseline_vectorised": {"value": vfast, "unit": UNIT, "cores": cores, "kind": "port",
                                        "sample": "same sequences, oracle/fast_chain.py (vectorised torch), one pass"},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)
    return 0


if __name__ == "__main__":
    sys.exit(main())
