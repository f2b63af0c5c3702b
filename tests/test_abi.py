"""CPU-side checks of the drop-in boundary: the C-ABI library loads without a GPU, exports every
symbol `include/spa_b200.h` declares, and its phase-1 ("GetWorkspaceSize") validation -- the place
the reference keeps its tiling checks (sparse_paged_attention_tiling_check.cc) -- rejects malformed
descriptors with the documented status codes before touching CUDA.  No compute is launched."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "spa_b200.h")
LIB = os.path.join(ROOT, "cann-ops-adv-vllm_b200", "libspa_b200.so")

SPA_OK, SPA_ERR_ARG, SPA_ERR_SHAPE, SPA_ERR_DTYPE, SPA_ERR_UNSUPPORTED = 0, -1, -2, -3, -4
SPA_F16, SPA_BF16, SPA_F32, SPA_I8, SPA_I32, SPA_I64 = range(6)


class SpaTensor(ctypes.Structure):
    _fields_ = [("data", ctypes.c_void_p), ("dtype", ctypes.c_int32), ("ndim", ctypes.c_int32),
                ("sizes", ctypes.c_int64 * 8), ("strides", ctypes.c_int64 * 8)]


def tensor(sizes, dtype, data=0x1000):
    t = SpaTensor()
    t.data = data
    t.dtype = dtype
    t.ndim = len(sizes)
    s = 1
    for i in reversed(range(len(sizes))):
        t.sizes[i] = sizes[i]
        t.strides[i] = s
        s *= sizes[i]
    return t


@pytest.fixture(scope="module")
def lib():
    if not os.path.exists(LIB):
        import importlib.util
        spec = importlib.util.spec_from_file_location("spa_build", os.path.join(ROOT, "cann-ops-adv-vllm_b200", "build.py"))
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        mod.build_kernels()
    L = ctypes.CDLL(LIB)
    L.spaGetLastError.restype = ctypes.c_char_p
    return L


def declared_symbols():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"SPA_API\s+[\w\s\*]+?\b(spa\w+)\s*\(", text)))


def test_header_declares_the_five_ops_two_phase():
    syms = declared_symbols()
    for op in ("ComputeCent", "SelectPosition", "CentSelect", "SparsePagedAttention", "SparsePagedFusionAttention"):
        assert f"spa{op}GetWorkspaceSize" in syms and f"spa{op}" in syms      # aclnn two-phase shape
    assert {"spaGetLastError", "spaDestroyExecutor", "spaGetVersion", "spaGetCompiledArch"} <= set(syms)


def test_library_exports_every_declared_symbol(lib):
    for s in declared_symbols():
        assert hasattr(lib, s), f"{s} declared in include/spa_b200.h but not exported"


def test_version_and_arch(lib):
    assert lib.spaGetVersion() >= 100
    assert lib.spaGetCompiledArch() == 100                        # sm_100a


def test_compute_cent_validation_needs_no_gpu(lib):
    ws = ctypes.c_uint64(0)
    ex = ctypes.c_void_p(0)
    q = tensor([2, 32, 128], SPA_F16)
    cent = tensor([8, 512, 128], SPA_F16)
    idx = tensor([2, 32, 64], SPA_I32)
    f = lib.spaComputeCentGetWorkspaceSize
    # absent query
    assert f(None, ctypes.byref(cent), ctypes.byref(idx), ctypes.byref(ws), ctypes.byref(ex)) == SPA_ERR_ARG
    assert b"query" in lib.spaGetLastError()
    # wrong dtype
    qf = tensor([2, 32, 128], SPA_F32)
    assert f(ctypes.byref(qf), ctypes.byref(cent), ctypes.byref(idx), ctypes.byref(ws), ctypes.byref(ex)) == SPA_ERR_DTYPE
    # head dim other than 128
    q64 = tensor([2, 32, 64], SPA_F16)
    assert f(ctypes.byref(q64), ctypes.byref(cent), ctypes.byref(idx), ctypes.byref(ws), ctypes.byref(ex)) == SPA_ERR_UNSUPPORTED
    # centroid dtype must equal the query's (cent_select_def.cpp:12-21 is fp16-only)
    cb = tensor([8, 512, 128], SPA_BF16)
    assert f(ctypes.byref(q), ctypes.byref(cb), ctypes.byref(idx), ctypes.byref(ws), ctypes.byref(ex)) == SPA_ERR_DTYPE
    # k (= indices.shape[2], compute_cent_tiling.cpp:45-50) larger than C
    big = tensor([2, 32, 513], SPA_I32)
    assert f(ctypes.byref(q), ctypes.byref(cent), ctypes.byref(big), ctypes.byref(ws), ctypes.byref(ex)) == SPA_ERR_SHAPE
    assert b"exceeds cluster count" in lib.spaGetLastError()
    # q heads not a multiple of kv heads
    q30 = tensor([2, 30, 128], SPA_F16)
    i30 = tensor([2, 30, 64], SPA_I32)
    assert f(ctypes.byref(q30), ctypes.byref(cent), ctypes.byref(i30), ctypes.byref(ws), ctypes.byref(ex)) == SPA_ERR_SHAPE
    # null out-pointers
    assert f(ctypes.byref(q), ctypes.byref(cent), ctypes.byref(idx), None, ctypes.byref(ex)) == SPA_ERR_ARG


def test_select_position_validation_needs_no_gpu(lib):
    ws = ctypes.c_uint64(0)
    ex = ctypes.c_void_p(0)
    bid = tensor([8, 1280], SPA_I32)
    bt = tensor([4, 1024], SPA_I32)
    sl = tensor([4], SPA_I32)
    idx = tensor([4, 32, 64], SPA_I32)
    pp = tensor([4, 32, 256], SPA_I32)
    ppl = tensor([4, 32, 8], SPA_I32)
    f = lib.spaSelectPositionGetWorkspaceSize
    args = lambda **kw: [ctypes.byref(kw.get("bid", bid)), ctypes.byref(kw.get("bt", bt)), ctypes.byref(kw.get("sl", sl)),
                         ctypes.byref(kw.get("idx", idx)), ctypes.c_int64(kw.get("bs", 128)),
                         ctypes.byref(kw.get("pp", pp)), ctypes.byref(kw.get("ppl", ppl)), None,
                         ctypes.byref(ws), ctypes.byref(ex)]
    assert f(*args(bid=tensor([8, 1280], SPA_I64))) == SPA_ERR_DTYPE
    assert f(*args(sl=tensor([9], SPA_I32))) == SPA_ERR_SHAPE           # more sequences than block_table rows
    assert f(*args(bs=0)) == SPA_ERR_SHAPE
    assert f(*args(ppl=tensor([4, 32, 4], SPA_I32))) == SPA_ERR_SHAPE   # x8 padding is part of the contract
    assert f(*args(idx=tensor([4, 30, 64], SPA_I32))) == SPA_ERR_SHAPE


def test_attention_validates_ifa_options_without_gpu(lib):
    """pse_shift / attention_mask / quant_scale2 (+ offset) are checked as the reference tiling checks them
    (sparse_paged_attention_tiling_check.cc:737-850, sparse_paged_attention_tiling.cc:566-647) before anything touches
    the device; the int8-query scales and kv_padding_size stay refused."""
    ws = ctypes.c_uint64(0)
    ex = ctypes.c_void_p(0)
    q = tensor([2, 32, 128], SPA_F16)
    kv = tensor([64, 128, 8, 128], SPA_F16)
    asl = tensor([2], SPA_I64)
    bt = tensor([2, 32], SPA_I32)
    bp = tensor([2, 8, 256], SPA_I32)
    out = tensor([2, 32, 128], SPA_F16)
    out8 = tensor([2, 32, 128], SPA_I8)
    extra = tensor([1], SPA_F16)
    SPA_F32, SPA_I8_, SPA_U8, SPA_BOOL = 2, 3, 7, 8
    S = 32 * 128
    f = lib.spaSparsePagedAttentionGetWorkspaceSize
    f.argtypes = [ctypes.c_void_p] * 16 + [ctypes.c_int64, ctypes.c_double, ctypes.c_char_p, ctypes.c_int64,
                                           ctypes.c_int64, ctypes.c_int64, ctypes.c_void_p, ctypes.c_void_p,
                                           ctypes.c_void_p]

    def call(pse=None, mask=None, dq1=None, q2=None, qo2=None, kvpad=None, heads=32, layout=b"BNSD", bs=128, kvh=8, o=out):
        p = lambda t: ctypes.cast(ctypes.pointer(t), ctypes.c_void_p) if t is not None else None
        return f(p(q), p(kv), p(kv), p(pse), p(mask), p(asl), p(dq1), None, None, p(q2), p(qo2), None, None, p(bt),
                 p(kvpad), p(bp), heads, 0.088, layout, kvh, bs, 1, p(o), ctypes.addressof(ws),
                 ctypes.addressof(ex))

    err = lambda: lib.spaGetLastError()
    assert call(pse=extra) == SPA_ERR_SHAPE and b"pse_shift" in err()                       # not [B or 1, Nq, 1, S]
    assert call(pse=tensor([2, 32, 1, S], SPA_BF16)) == SPA_ERR_DTYPE and b"pse_shift" in err()
    assert call(pse=tensor([2, 32, 1, S - 1], SPA_F16)) == SPA_ERR_SHAPE and b"pse_shift" in err()
    assert call(pse=tensor([3, 16, 1, S], SPA_F16)) == SPA_ERR_SHAPE
    assert call(mask=extra) == SPA_ERR_DTYPE and b"attention_mask" in err()                 # bool / int8 / uint8 only
    assert call(mask=tensor([2, 1, 1, S - 1], SPA_BOOL)) == SPA_ERR_SHAPE and b"attention_mask" in err()
    assert call(mask=tensor([1, S], SPA_U8)) == SPA_ERR_SHAPE                               # batch of the mask < B
    assert call(mask=tensor([2, 2, S], SPA_I8_)) == SPA_ERR_SHAPE                           # middle dims must be 1
    assert call(q2=tensor([1], SPA_F32)) == SPA_ERR_DTYPE and b"int8" in err()              # post-quant needs an int8 output
    assert call(o=out8) == SPA_ERR_ARG and b"quant_scale2" in err()                         # and the other way round
    assert call(q2=tensor([1], SPA_F16), o=out8) == SPA_ERR_DTYPE
    assert call(q2=tensor([7], SPA_F32), o=out8) == SPA_ERR_SHAPE
    assert call(q2=tensor([1], SPA_F32), qo2=tensor([32 * 128], SPA_BF16), o=out8) == SPA_ERR_DTYPE
    assert call(qo2=tensor([1], SPA_F32)) == SPA_ERR_ARG
    assert call(dq1=extra) == SPA_ERR_UNSUPPORTED
    assert call(kvpad=extra) == SPA_ERR_UNSUPPORTED
    assert call(heads=16) == SPA_ERR_SHAPE and b"num_heads" in lib.spaGetLastError()
    assert call(layout=b"XYZ") == SPA_ERR_ARG
    assert call(bs=48) == SPA_ERR_UNSUPPORTED
    assert call(kvh=5) == SPA_ERR_SHAPE


def test_destroy_null_executor_is_ok(lib):
    assert lib.spaDestroyExecutor(None) == SPA_OK
    assert lib.spaExecutorNumLaunches(None) == 0


def test_meta_impls_infer_output_shapes_without_gpu(lib):
    """`myops` Meta registrations (shape inference for FakeTensor / torch.compile tracing; the
    reference's are op_plugin/python/meta/_meta_registrations.py:38-93)."""
    import sys
    import torch
    sys.path.insert(0, os.path.join(ROOT, "cann-ops-adv-vllm_b200"))
    pytest.importorskip("custom_ops_lib")

    def m(*s, dt=torch.float16):
        return torch.empty(*s, dtype=dt, device="meta")

    B, Nq, Nkv, C, P = 4, 32, 8, 512, 256
    i32 = torch.int32
    idx = torch.ops.myops.compute_cent(m(B, Nq, 128), m(Nkv, C, 128))
    assert idx.shape == (B, Nq, 64) and idx.dtype == i32
    pp, ppl, ga = torch.ops.myops.select_position(m(Nkv, 1280, dt=i32), m(B, P, dt=i32), m(B, dt=i32), m(B, Nq, 64, dt=i32))
    assert pp.shape == (B, Nq, 256) and ppl.shape == (B, Nq, 8) and ga.shape == (B, Nq, P)
    pp, ppl, mx = torch.ops.myops.cent_select(m(B, Nq, 128), m(Nkv, C, 128), m(Nkv, 1280, dt=i32), m(B, P, dt=i32),
                                              m(B - 2, dt=i32))
    assert pp.shape == (B - 2, Nq, 256) and ppl.shape == (B - 2, Nq, 8) and mx.shape == (B - 2, 8) and mx.dtype == torch.int64
    e = m(0)
    kv = m(1280, 128, Nkv, 128)
    out = torch.ops.myops.sparse_paged_attention(m(B, Nq, 128), [kv], [kv], e, e, m(B, dt=torch.int64), e, e, e, e, e, e, e,
                                                 m(B, P, dt=i32), e, m(B, Nkv, 256, dt=i32), Nq, 1.0, "BNSD", Nkv, 128, 1)
    assert out.shape == (B, Nq, 128) and out.dtype == torch.float16
    out, bp, mx = torch.ops.myops.sparse_paged_fusion_attention(
        m(B, Nq, 128, dt=torch.bfloat16), [kv], [kv], e, e, m(B, dt=torch.int64), e, e, e, e, e, e, e, m(B, P, dt=i32), e,
        m(Nkv, C, 128), m(Nkv, 1280, dt=i32), m(B, dt=i32), m(B, Nq, 512, dt=i32), m(B, Nq, 8, dt=i32),
        m(B, 8, dt=torch.int64), Nq, 1.0, "BNSD", Nkv, 128, 1)
    assert out.shape == (B, Nq, 128) and out.dtype == torch.bfloat16 and bp.shape == (B, Nq, 512) and mx.shape == (B, 8)


def test_extension_entry_points_validate_without_gpu(lib):
    """The entry points added beyond the five aclnn pairs (centroid maintenance, executor options, peer
    output) reject malformed arguments with the documented codes before touching CUDA."""
    ws = ctypes.c_uint64(0)
    ex = ctypes.c_void_p()
    Nkv, C, blocks = 4, 64, 16
    key = tensor([blocks, 128, Nkv, 128], SPA_F16)
    cent = tensor([Nkv, C, 128], SPA_F16)
    ids = tensor([5], SPA_I32)
    bid = tensor([Nkv, blocks], SPA_I32)
    f = lib.spaAssignPageClustersGetWorkspaceSize
    args = lambda k=key, c=cent, i=ids, b=bid, bs=128, aq=None: (ctypes.byref(k), ctypes.byref(c), ctypes.byref(i), None,
                                                                 aq, None, ctypes.c_int64(bs), ctypes.byref(b), None,
                                                                 ctypes.byref(ws), ctypes.byref(ex))
    assert f(*args(k=tensor([blocks, 128, Nkv, 128], SPA_I8))) == SPA_ERR_ARG          # int8 pool without antiquant scale
    assert b"antiquantScale" in lib.spaGetLastError()
    assert f(*args(k=tensor([blocks, 128, Nkv, 128], SPA_BF16))) == SPA_ERR_DTYPE       # neither int8 nor the centroid dtype
    assert f(*args(aq=ctypes.byref(tensor([2, Nkv, 128], SPA_F16)))) == SPA_ERR_ARG     # antiquant for a non-int8 pool
    assert f(*args(k=tensor([blocks, 128, Nkv, 128], SPA_I8), aq=ctypes.byref(tensor([2, Nkv, 64], SPA_F16)))) == SPA_ERR_SHAPE
    assert f(*args(b=tensor([Nkv + 1, blocks], SPA_I32))) == SPA_ERR_SHAPE
    assert f(*args(c=tensor([Nkv, 9000, 128], SPA_F16))) == SPA_ERR_UNSUPPORTED
    assert f(*args(k=tensor([blocks, 64, Nkv, 128], SPA_F16))) == SPA_ERR_SHAPE        # pool does not match blockSize
    assert f(*args(i=tensor([5], SPA_I64))) == SPA_ERR_DTYPE
    # an empty page list plans (and launches) nothing and needs no device
    empty = tensor([0], SPA_I32)
    assert f(*args(i=empty)) == SPA_OK and ex.value
    assert lib.spaAssignPageClusters(None, ctypes.c_uint64(0), ex, None) == SPA_OK
    # executor setters: type-checked against the op the executor was planned for
    assert lib.spaExecutorSetSelectOptions(ex, 1, 2, 8) == SPA_ERR_UNSUPPORTED
    assert lib.spaExecutorSetSoftmaxLse(ex, None) == SPA_ERR_UNSUPPORTED
    assert lib.spaExecutorSetPeerOutput(ex, 0, 1, None, None) == SPA_OK                 # world 1: off
    assert lib.spaDestroyExecutor(ex) == SPA_OK
    assert lib.spaExecutorSetSelectOptions(None, 0, 0, 0) == SPA_ERR_ARG
    assert lib.spaExecutorSetPeerOutput(None, 0, 2, None, None) == SPA_ERR_ARG
    assert lib.spaPeerBufferOpen(None, None) == SPA_ERR_ARG
    assert lib.spaPeerBufferCreate(ctypes.c_uint64(0), None, None) == SPA_ERR_ARG
