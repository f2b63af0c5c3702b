"""GPU test of the allocation-free entry point `spaSparsePagedFusionAttentionStatic` (include/spa_b200.h).

The entry point has no counterpart in the reference (it was added for
callers that cannot call cudaMalloc); it is checked against the EXISTING two-phase fused entry point
(`spaSparsePagedFusionAttentionGetWorkspaceSize` + `spaSparsePagedFusionAttention`) on the same inputs, both
called through the C ABI with ctypes: selection outputs bit-exact, attention_out within the fp16 tolerance the
parity suite uses (atol = rtol = 1e-3, BASELINE.json north_star), on two calls in a row with different inputs
(the shared scratch keeps the first call's contents).

The file name sorts last on purpose: a failure here hides no other test.
It passes alone in a fresh process and after the other GPU tests.
"""
import ctypes
import os

import pytest
import torch

from cases import make_case, to_device

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "cann-ops-adv-vllm_b200", "libspa_b200.so")
SPA_OK, SPA_ERR_WORKSPACE = 0, -6
DTYPES = {torch.float16: 0, torch.bfloat16: 1, torch.float32: 2, torch.int8: 3, torch.int32: 4, torch.int64: 5}
TOL = 1e-3                                                    # fp16 (tests/test_gpu_parity.py)


class SpaTensor(ctypes.Structure):
    _fields_ = [("data", ctypes.c_void_p), ("dtype", ctypes.c_int32), ("ndim", ctypes.c_int32),
                ("sizes", ctypes.c_int64 * 8), ("strides", ctypes.c_int64 * 8)]


def desc(t):
    """spaTensor descriptor of a torch tensor (None / empty -> NULL = absent)."""
    if t is None or t.numel() == 0:
        return None
    d = SpaTensor()
    d.data = t.data_ptr()
    d.dtype = DTYPES[t.dtype]
    d.ndim = t.dim()
    for i in range(t.dim()):
        d.sizes[i] = t.shape[i]
        d.strides[i] = t.stride(i)
    return d


def fake(sizes, dtype, data=0x10000):
    """Descriptor with a made-up (aligned) device address: for calls that must be refused before any launch."""
    d = SpaTensor()
    d.data, d.dtype, d.ndim = data, dtype, len(sizes)
    s = 1
    for i in reversed(range(len(sizes))):
        d.sizes[i], d.strides[i] = sizes[i], s
        s *= sizes[i]
    return d


def ref(d):
    return ctypes.byref(d) if d is not None else None


@pytest.fixture(scope="module")
def lib():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    if not os.path.exists(LIB):
        raise RuntimeError("libspa_b200.so is not built (run __graft_entry__.build()); there is no fallback")
    L = ctypes.CDLL(LIB)
    L.spaGetLastError.restype = ctypes.c_char_p
    L.spaStaticScratchBytes.restype = ctypes.c_uint64
    return L


def fused_args(c, k, bp, ppl, mppl, out):
    """The 29 tensor / attribute arguments shared by both entry points, in the header's order."""
    keep = [desc(c["q"]), desc(c["key"]), desc(c["value"]), desc(c["aq_scale"]), desc(c["aq_offset"]),
            desc(c["block_table"]), desc(c["l1"]), desc(c["block_ids"]), desc(c["seq"]),
            desc(bp), desc(ppl), desc(mppl), desc(out)]
    q, key, value, aqs, aqo, bt, l1, bid, seq, dbp, dppl, dmppl, dout = keep
    args = [ref(q), ref(key), ref(value), None, None, None, None, None, None, None, None, ref(aqs), ref(aqo),
            ref(bt), None, ref(l1), ref(bid), ref(seq),
            ctypes.c_int64(c["Nq"]), ctypes.c_double(c["scale"]), ctypes.c_char_p(b"BNSD"), ctypes.c_int64(c["Nkv"]),
            ctypes.c_int64(c["bs"]), ctypes.c_int64(1), ctypes.c_int64(k),
            ref(dbp), ref(dppl), ref(dmppl), ref(dout)]
    return args, keep                                          # keep: the descriptors must outlive the call


def outputs(c, fill):
    dev = c["q"].device
    B, Nq, MP = c["seq"].shape[0], c["Nq"], c["MP"]
    return (torch.full((B, Nq, MP), fill, dtype=torch.int32, device=dev),
            torch.full((B, Nq, 8), fill, dtype=torch.int32, device=dev),
            torch.full((B, 8), fill, dtype=torch.int64, device=dev),
            torch.full((B, Nq, c["D"]), float("nan"), dtype=c["q"].dtype, device=dev))


def run_two_phase(L, c, k):
    """The existing entry point: plan, caller-allocated workspace (torch), launch on the current stream."""
    bp, ppl, mppl, out = outputs(c, 0x5A5A5A5A)
    args, keep = fused_args(c, k, bp, ppl, mppl, out)
    ws, ex = ctypes.c_uint64(0), ctypes.c_void_p(None)
    rc = L.spaSparsePagedFusionAttentionGetWorkspaceSize(*args, ctypes.byref(ws), ctypes.byref(ex))
    assert rc == SPA_OK, L.spaGetLastError().decode()
    work = torch.empty(max(int(ws.value), 256), dtype=torch.uint8, device=c["q"].device)
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = L.spaSparsePagedFusionAttention(ctypes.c_void_p(work.data_ptr()), ctypes.c_uint64(work.numel()), ex, stream)
    assert rc == SPA_OK, L.spaGetLastError().decode()
    torch.cuda.synchronize()
    L.spaDestroyExecutor(ex)
    del keep
    return bp, ppl, mppl, out, int(ws.value)


def run_static(L, c, k, check_no_alloc=False):
    """The new entry point: inputs and outputs only; no workspace, no executor."""
    bp, ppl, mppl, out = outputs(c, 0x33333333)
    args, keep = fused_args(c, k, bp, ppl, mppl, out)
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    torch_before = torch.cuda.memory_allocated()
    free_before = torch.cuda.mem_get_info()[0]
    rc = L.spaSparsePagedFusionAttentionStatic(*args, stream)
    assert rc == SPA_OK, L.spaGetLastError().decode()
    torch.cuda.synchronize()
    if check_no_alloc:                                         # (the first call of a process may load the CUDA module)
        assert torch.cuda.memory_allocated() == torch_before
        assert torch.cuda.mem_get_info()[0] >= free_before, "device memory was allocated inside the call"
    del keep
    return bp, ppl, mppl, out


def compare(got, want, what):
    gbp, gppl, gmppl, gout = got
    wbp, wppl, wmppl, wout = want[:4]
    assert torch.equal(gbp, wbp), f"{what}: block_position differs from the two-phase entry point"
    assert torch.equal(gppl, wppl), f"{what}: page_position_length differs"
    assert torch.equal(gmppl, wmppl), f"{what}: max_page_position_length differs"
    g, w = gout.float(), wout.float()
    assert not torch.isnan(g).any(), f"{what}: attention_out was not fully written"
    err = (g - w).abs()
    assert bool((err <= TOL + TOL * w.abs()).all()), f"{what}: attention_out max err {err.max().item():.3e} (tol {TOL})"


def test_static_entry_point_matches_two_phase_on_two_calls_in_a_row(lib):
    k = 8
    # call 1: BASELINE configs[0] shape, B = 2 (whole items per CTA); call 2: a larger, ragged problem whose
    # items are cut into chunks (partials + combine) -- it runs on the scratch the first call left behind
    c1 = to_device(make_case(777, 2, 32, 8, 64, 4096, kv_pages=128), "cuda:0")
    c2 = to_device(make_case(778, 4, 32, 8, 128, 16384, kv_pages=1024, ragged=True, permute_table=True), "cuda:0")
    torch.cuda.set_device(0)
    want1 = run_two_phase(lib, c1, k)
    want2 = run_two_phase(lib, c2, 16)
    assert max(want1[4], want2[4]) <= lib.spaStaticScratchBytes()
    got1 = run_static(lib, c1, k)
    got2 = run_static(lib, c2, 16, check_no_alloc=True)
    compare(got1, want1, "call 1")
    compare(got2, want2, "call 2")
    # and once more on the first problem: the scratch now holds call 2's leftovers
    compare(run_static(lib, c1, k, check_no_alloc=True), want1, "call 3")


def test_static_entry_point_rejects_problems_beyond_its_scratch(lib):
    torch.cuda.set_device(0)
    B, Nq, Nkv, C, P, blocks = 256, 32, 8, 512, 1024, 4096     # 2048 work items x 9 fp32 partials: ~90 MB of workspace
    f16, i32, i64 = 0, 4, 5
    q, key, value = fake([B, Nq, 128], f16), fake([blocks, 128, Nkv, 128], f16), fake([blocks, 128, Nkv, 128], f16)
    bt, l1, bid, seq = fake([B, P], i32), fake([Nkv, C, 128], f16), fake([Nkv, blocks], i32), fake([B], i32)
    bp, ppl, mppl, out = fake([B, Nq, 256], i32), fake([B, Nq, 8], i32), fake([B, 8], i64), fake([B, Nq, 128], f16)
    rc = lib.spaSparsePagedFusionAttentionStatic(
        ref(q), ref(key), ref(value), None, None, None, None, None, None, None, None, None, None, ref(bt), None,
        ref(l1), ref(bid), ref(seq), ctypes.c_int64(Nq), ctypes.c_double(0.088), ctypes.c_char_p(b"BNSD"),
        ctypes.c_int64(Nkv), ctypes.c_int64(128), ctypes.c_int64(1), ctypes.c_int64(64),
        ref(bp), ref(ppl), ref(mppl), ref(out), ctypes.c_void_p(None))
    assert rc == SPA_ERR_WORKSPACE
    msg = lib.spaGetLastError().decode()
    assert "static scratch" in msg and str(lib.spaStaticScratchBytes()) in msg, msg
    torch.cuda.synchronize()                                   # nothing was launched: no sticky error
