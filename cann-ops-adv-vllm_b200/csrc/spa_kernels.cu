// sm_100a kernels + C ABI (include/spa_b200.h) of the centroid-sparse paged decode-attention chain.
//
// Replaces, for this path only, the reference's device kernels
//   src/transformer/{compute_cent,select_position,cent_select,sparse_paged_attention,
//                    sparse_paged_fusion_attention}/*.cpp|*.h
// and the host "ophost" layer (op defs, tiling, aclnn wrappers) under src/transformer/*/ophost/.
#include <cooperative_groups.h>
#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <mutex>
#include <new>

#include "../../include/spa_b200.h"
#include "spa_assign.cuh"
#include "spa_attn.cuh"
#include "spa_chain.cuh"
#include "spa_common.cuh"
#include "spa_select.cuh"

namespace cg = cooperative_groups;
using namespace spa;

// =============================================================================================
// kernels
// =============================================================================================
// Stand-alone selection ops (ComputeCent / SelectPosition / CentSelect share the device code).
template <typename QT, int FLAGS>
__global__ void __launch_bounds__(kThreads, 1) spa_select_kernel(const __grid_constant__ ChainParams p,
                                                                 const __grid_constant__ CUtensorMap tmC) {
    extern __shared__ unsigned char smem_raw[];
    unsigned char *smem = aligned_smem(smem_raw);
    select_phase<QT, FLAGS>(p, &tmC, smem, nullptr);
}

// Test hook (spaDebugHoldSMs): CTAs that claim a whole SM each (shared memory) and spin for a given time.
__global__ void spa_hold_sm_kernel(unsigned long long ns) {
    extern __shared__ unsigned char hold_smem[];
    const unsigned long long t0 = globaltimer_ns();
    if (threadIdx.x == 0) hold_smem[0] = 1;
    while (globaltimer_ns() - t0 < ns) __nanosleep(1000);
}

__global__ void spa_max_ppl_kernel(const __grid_constant__ ChainParams p) {
    for (int b = blockIdx.x; b < p.B; b += gridDim.x) write_max_ppl(p, b);
}

// =============================================================================================
// host side: validation ("tiling checks"), plans, launches
// =============================================================================================
static thread_local char g_err[512] = "";

static int fail(int code, const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}
#define CUDA_TRY(expr)                                                                         \
    do {                                                                                       \
        cudaError_t e__ = (expr);                                                              \
        if (e__ != cudaSuccess)                                                                \
            return fail(SPA_ERR_CUDA, "%s failed: %s", #expr, cudaGetErrorString(e__));        \
    } while (0)

enum OpKind { OP_COMPUTE_CENT = 1, OP_SELECT_POSITION, OP_CENT_SELECT, OP_ATTENTION, OP_FUSION, OP_ASSIGN };

struct spaExecutor {
    int op;
    ChainParams p;
    CUtensorMap tmK, tmV, tmC;
    int flags;            // selection flags
    int grid;
    int q_bf16, kv_i8, nt_tiles;       // kv_i8: 8-bit KV pages (int8 or FP8)
    int kv_f8;                         // ... of which FP8 e4m3
    int opt;                           // options build of the chain kernel (attenMask / pseShift / post-quant), spa_chain.cuh
    uint64_t ws_bytes;
    // workspace offsets (bytes)
    uint64_t off_ctrl, off_flat, off_desc, off_info, off_ippl, off_po, off_pml, off_xs;
    uint64_t ctrl_bytes;
    AssignParams ap;       // OP_ASSIGN
    int persistent_ws;     // see spaExecutorSetPersistentWorkspace
    void *clean_ws;        // workspace whose queue words are known to be zero
    int launches;
};

static bool absent(const spaTensor *t) {
    if (!t || !t->data) return true;
    if (t->ndim <= 0) return false;
    for (int i = 0; i < t->ndim; ++i)
        if (t->sizes[i] == 0) return true;
    return false;
}
static int64_t numel(const spaTensor *t) {
    int64_t n = 1;
    for (int i = 0; i < t->ndim; ++i) n *= t->sizes[i];
    return n;
}
static bool contiguous(const spaTensor *t) {
    int64_t s = 1;
    for (int i = t->ndim - 1; i >= 0; --i) {
        if (t->sizes[i] != 1 && t->strides[i] != s) return false;
        s *= t->sizes[i];
    }
    return true;
}
static const char *dtype_name(int d) {
    switch (d) {
        case SPA_F16: return "float16"; case SPA_BF16: return "bfloat16"; case SPA_F32: return "float32";
        case SPA_I8: return "int8"; case SPA_I32: return "int32"; case SPA_I64: return "int64";
        case SPA_F8E4M3: return "float8_e4m3"; case SPA_U8: return "uint8"; case SPA_BOOL: return "bool";
    }
    return "?";
}
#define REQUIRE_PRESENT(t, name)                                                                \
    if (absent(t)) return fail(SPA_ERR_ARG, "%s: required tensor is absent", name)
#define REQUIRE_DTYPE(t, dt, name)                                                              \
    if ((t)->dtype != (dt))                                                                     \
    return fail(SPA_ERR_DTYPE, "%s: expected %s, got %s", name, dtype_name(dt), dtype_name((t)->dtype))
#define REQUIRE_CONTIG(t, name)                                                                 \
    if (!contiguous(t)) return fail(SPA_ERR_SHAPE, "%s: must be contiguous", name)

static uint64_t align_up(uint64_t x, uint64_t a) { return (x + a - 1) / a * a; }

struct DevInfo { int sms; int smem_optin; int coop; };
static int dev_info(DevInfo *d) {
    int dev = 0;
    CUDA_TRY(cudaGetDevice(&dev));
    CUDA_TRY(cudaDeviceGetAttribute(&d->sms, cudaDevAttrMultiProcessorCount, dev));
    CUDA_TRY(cudaDeviceGetAttribute(&d->smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
    CUDA_TRY(cudaDeviceGetAttribute(&d->coop, cudaDevAttrCooperativeLaunch, dev));
    return SPA_OK;
}


// Raise a kernel's dynamic shared-memory limit to everything the device allows next to the kernel's static
// shared memory.  Done at plan time with the DEVICE's maximum rather than the plan's need: the attribute is
// global to the function, so plans with different needs (other threads, other executors) must never lower
// it under each other's launches.
static int raise_smem_limit(const void *func, int optin) {
    cudaFuncAttributes fa;
    CUDA_TRY(cudaFuncGetAttributes(&fa, func));
    CUDA_TRY(cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, optin - (int)fa.sharedSizeBytes));
    return SPA_OK;
}

// ---- query: [B,Nq,D] or [B,Nq,1,D]
static int parse_query(const spaTensor *q, const char *name, int *B, int *Nq, long long *sb,
                       long long *sh, int *is_bf16) {
    REQUIRE_PRESENT(q, name);
    if (q->dtype != SPA_F16 && q->dtype != SPA_BF16)
        return fail(SPA_ERR_DTYPE, "%s: expected float16/bfloat16, got %s", name, dtype_name(q->dtype));
    if (!(q->ndim == 3 || (q->ndim == 4 && q->sizes[2] == 1)))
        return fail(SPA_ERR_SHAPE, "%s: expected [B,N,D] or [B,N,1,D]", name);
    const int dl = q->ndim - 1;
    if (q->sizes[dl] != kD) return fail(SPA_ERR_UNSUPPORTED, "%s: head dim %lld, this build supports D=%d", name, (long long)q->sizes[dl], kD);
    if (q->strides[dl] != 1) return fail(SPA_ERR_SHAPE, "%s: last dim must be contiguous", name);
    if ((q->strides[0] & 3) || (q->strides[1] & 3) || (reinterpret_cast<uintptr_t>(q->data) & 7))
        return fail(SPA_ERR_SHAPE, "%s: rows must be 8-byte aligned", name);
    *B = (int)q->sizes[0]; *Nq = (int)q->sizes[1]; *sb = q->strides[0]; *sh = q->strides[1];
    *is_bf16 = q->dtype == SPA_BF16;
    return SPA_OK;
}

static int parse_cent(const spaTensor *c, int qdtype, int *Nkv, int *C) {
    REQUIRE_PRESENT(c, "l1_cent");
    REQUIRE_DTYPE(c, qdtype, "l1_cent");
    if (c->ndim != 3 || c->sizes[2] != kD) return fail(SPA_ERR_SHAPE, "l1_cent: expected [Nkv,C,%d]", kD);
    REQUIRE_CONTIG(c, "l1_cent");
    if (reinterpret_cast<uintptr_t>(c->data) & 15) return fail(SPA_ERR_SHAPE, "l1_cent: must be 16-byte aligned");
    *Nkv = (int)c->sizes[0]; *C = (int)c->sizes[1];
    if (*C > 65535) return fail(SPA_ERR_UNSUPPORTED, "l1_cent: C=%d > 65535", *C);
    return SPA_OK;
}

static int parse_tables(const spaTensor *bid, const spaTensor *bt, const spaTensor *sl, ChainParams *p) {
    REQUIRE_PRESENT(bid, "block_ids"); REQUIRE_DTYPE(bid, SPA_I32, "block_ids");
    REQUIRE_PRESENT(bt, "block_table"); REQUIRE_DTYPE(bt, SPA_I32, "block_table");
    REQUIRE_PRESENT(sl, "seq_len"); REQUIRE_DTYPE(sl, SPA_I32, "seq_len");
    if (bid->ndim != 2 || bid->strides[1] != 1) return fail(SPA_ERR_SHAPE, "block_ids: expected [Nkv,kvPageLen], row-contiguous");
    if (bt->ndim != 2 || bt->strides[1] != 1) return fail(SPA_ERR_SHAPE, "block_table: expected [maxB,maxP], row-contiguous");
    if (sl->ndim != 1 || !contiguous(sl)) return fail(SPA_ERR_SHAPE, "seq_len: expected contiguous [B]");
    p->block_ids = (const int *)bid->data; p->bid_stride = bid->strides[0]; p->kvPageLen = (int)bid->sizes[1];
    p->block_table = (const int *)bt->data; p->bt_stride = bt->strides[0]; p->maxP = (int)bt->sizes[1];
    p->seq_len = (const int *)sl->data;
    if (sl->sizes[0] > bt->sizes[0]) return fail(SPA_ERR_SHAPE, "seq_len batch %lld exceeds block_table rows %lld", (long long)sl->sizes[0], (long long)bt->sizes[0]);
    return SPA_OK;
}

// Phase A shared memory: [ngroups x sel_ns centroid-tile stages][ngroups x group state].  Chooses the
// group width (two 4-warp groups when g <= 4 and there are more items than CTAs) and the stage count.
static int plan_select_smem(ChainParams *p, int flags, int grid, int budget, int *bytes) {
    const int Ceff = std::max(p->C, 1), keff = std::max(p->k, 1);
    const bool scores = (flags & kSelScores) != 0;
    const int ntiles = (Ceff + kCentTileRows - 1) / kCentTileRows;
    SelSmem L(p->g, Ceff, keff, p->maxP, flags);
    int gw = (p->g > 4 || p->items <= grid) ? 8 : 4;
    if (const char *s = getenv("SPA_SEL_GW")) { const int v = atoi(s); if (v == 4 || v == 8) gw = v; }
    for (; gw <= 8; gw *= 2) {
        const int ngroups = kConsumerWarps / gw;
        const long long state = (long long)L.total * ngroups;
        int ns = 0;
        if (scores) {
            ns = (int)std::min<long long>(std::min(ntiles, 4), (budget - state) / ((long long)ngroups * kCentTileBytes));
            if (ns < 1) continue;
        } else if (state > budget) continue;
        p->sel_gw = gw; p->sel_group_bytes = L.total; p->sel_ns = ns;
        *bytes = (int)(state + (long long)ngroups * ns * kCentTileBytes);
        return SPA_OK;
    }
    return fail(SPA_ERR_UNSUPPORTED, "selection working set (g=%d, C=%d, k=%d, maxP=%d) exceeds shared memory", p->g, p->C, p->k, p->maxP);
}

// ---------------------------------------------------------------------------------------------
// TMA tensor maps over the paged K / V pools
// ---------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                  CUtensorMapFloatOOBfill);
static EncodeTiledFn get_encode_fn() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void *ptr = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(ptr);
    }
    return fn;
}

struct KvLayout { long long nblocks, s_blk, s_row, s_head; };   // strides in elements

// Accepted pool layouts (tests/ifa_v5_case.py:408 vLLM [blocks,bs,Nkv,D]; :218 [blocks,bs,Nkv*D];
// :36 BNSD-paged [blocks,Nkv,bs,D]).
static int parse_kv(const spaTensor *t, const char *name, int Nkv, int bs, KvLayout *L) {
    REQUIRE_PRESENT(t, name);
    if (t->ndim == 3) {
        if (t->sizes[1] != bs || t->sizes[2] != (int64_t)Nkv * kD || t->strides[2] != 1)
            return fail(SPA_ERR_SHAPE, "%s: expected [blocks,%d,%d]", name, bs, Nkv * kD);
        *L = {t->sizes[0], t->strides[0], t->strides[1], kD};
    } else if (t->ndim == 4 && t->sizes[3] == kD && t->strides[3] == 1) {
        if (t->sizes[1] == bs && t->sizes[2] == Nkv) *L = {t->sizes[0], t->strides[0], t->strides[1], t->strides[2]};
        else if (t->sizes[1] == Nkv && t->sizes[2] == bs) *L = {t->sizes[0], t->strides[0], t->strides[2], t->strides[1]};
        else return fail(SPA_ERR_SHAPE, "%s: expected [blocks,%d,%d,%d] or [blocks,%d,%d,%d]", name, bs, Nkv, kD, Nkv, bs, kD);
    } else {
        return fail(SPA_ERR_SHAPE, "%s: expected a 3-D or 4-D paged pool with head dim %d", name, kD);
    }
    return SPA_OK;
}

static int make_kv_map(CUtensorMap *map, const spaTensor *t, const KvLayout &L, int Nkv, int bs,
                       int *ax_row, int *ax_head, int *ax_blk, const char *name) {
    EncodeTiledFn enc = get_encode_fn();
    if (!enc) return fail(SPA_ERR_CUDA, "cuTensorMapEncodeTiled entry point not available");
    const int esz = (t->dtype == SPA_I8 || t->dtype == SPA_F8E4M3) ? 1 : 2;
    struct Ax { long long size, stride; int id; };
    Ax ax[3] = {{bs, L.s_row * esz, 0}, {Nkv, L.s_head * esz, 1}, {L.nblocks, L.s_blk * esz, 2}};
    // size-1 axes carry arbitrary strides: push them outermost and give them a consistent stride
    std::stable_sort(ax, ax + 3, [](const Ax &a, const Ax &b) {
        if ((a.size == 1) != (b.size == 1)) return b.size == 1;
        return a.stride < b.stride;
    });
    long long prev_stride = (long long)kD * esz, prev_size = 1;
    for (int i = 0; i < 3; ++i) {
        if (ax[i].size == 1) ax[i].stride = prev_stride * prev_size;
        prev_stride = ax[i].stride; prev_size = ax[i].size;
    }
    cuuint64_t gdim[4] = {(cuuint64_t)kD, (cuuint64_t)ax[0].size, (cuuint64_t)ax[1].size, (cuuint64_t)ax[2].size};
    cuuint64_t gstr[3] = {(cuuint64_t)ax[0].stride, (cuuint64_t)ax[1].stride, (cuuint64_t)ax[2].stride};
    cuuint32_t box[4] = {(cuuint32_t)(128 / esz), 1, 1, 1};
    cuuint32_t estr[4] = {1, 1, 1, 1};
    for (int i = 0; i < 3; ++i) {
        if (gstr[i] % 16) return fail(SPA_ERR_SHAPE, "%s: strides must be multiples of 16 bytes", name);
        if (ax[i].id == 0) { box[i + 1] = (cuuint32_t)bs; *ax_row = i + 1; }
        if (ax[i].id == 1) *ax_head = i + 1;
        if (ax[i].id == 2) *ax_blk = i + 1;
    }
    if (reinterpret_cast<uintptr_t>(t->data) & 15) return fail(SPA_ERR_SHAPE, "%s: must be 16-byte aligned", name);
    CUtensorMapDataType dt = (t->dtype == SPA_I8 || t->dtype == SPA_F8E4M3) ? CU_TENSOR_MAP_DATA_TYPE_UINT8
                             : t->dtype == SPA_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16
                                                    : CU_TENSOR_MAP_DATA_TYPE_FLOAT16;
    // L2 fetch granularity: a token row of one kv head is kD*esz contiguous bytes.  With int8 KV in the vLLM
    // layout that is 128 B and the neighbouring 128 B belong to another head (another CTA, another time):
    // 256 B promotion would pull them from DRAM for nothing (ncu, c3: 1.55 GB read for 1.07 GB needed).
    const bool rows_contiguous = (L.s_row == kD);
    const CUtensorMapL2promotion promo = ((long long)kD * esz >= 256 || rows_contiguous)
                                             ? CU_TENSOR_MAP_L2_PROMOTION_L2_256B : CU_TENSOR_MAP_L2_PROMOTION_L2_128B;
    CUresult r = enc(map, dt, 4, t->data, gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, promo, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(SPA_ERR_CUDA, "%s: cuTensorMapEncodeTiled failed (%d)", name, (int)r);
    return SPA_OK;
}

// 2-D map over the centroid table [Nkv*C, 128]; box = 64 columns x 128 rows, 128B swizzle (the same
// smem image as one half of a K page, so phase A reuses the ldmatrix addressing of phase B).
static int make_cent_map(CUtensorMap *map, const void *cent, int rows, int bf16) {
    EncodeTiledFn enc = get_encode_fn();
    if (!enc) return fail(SPA_ERR_CUDA, "cuTensorMapEncodeTiled entry point not available");
    cuuint64_t gdim[2] = {(cuuint64_t)kD, (cuuint64_t)std::max(rows, 1)};
    cuuint64_t gstr[1] = {(cuuint64_t)kD * 2};
    cuuint32_t box[2] = {64, (cuuint32_t)kCentTileRows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(map, bf16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2,
                     const_cast<void *>(cent), gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(SPA_ERR_CUDA, "l1_cent: cuTensorMapEncodeTiled failed (%d)", (int)r);
    return SPA_OK;
}

// ---------------------------------------------------------------------------------------------
// kernel dispatch tables
// ---------------------------------------------------------------------------------------------
static ChainKernel chain_kernel_for(int bf16, int kvk, int nt, int opt) {
    return opt ? opt_chain_kernel_for(bf16, kvk, nt) : chain_kernel_table<0>(bf16, kvk, nt);
}
typedef void (*SelectKernel)(const ChainParams, const CUtensorMap);
static SelectKernel select_kernel_for(int bf16, int flags) {
    switch (flags) {
        case kSelScores | kSelWriteIndices:
            return bf16 ? spa_select_kernel<__nv_bfloat16, kSelScores | kSelWriteIndices> : spa_select_kernel<__half, kSelScores | kSelWriteIndices>;
        case kSelScores | kSelSelect:
            return bf16 ? spa_select_kernel<__nv_bfloat16, kSelScores | kSelSelect> : spa_select_kernel<__half, kSelScores | kSelSelect>;
        case kSelSelect:
            return spa_select_kernel<__half, kSelSelect>;
        case kSelSelect | kSelGather:
            return spa_select_kernel<__half, kSelSelect | kSelGather>;
    }
    return nullptr;
}

// =============================================================================================
// C ABI
// =============================================================================================
extern "C" {

const char *spaGetLastError(void) { return g_err; }
int spaGetVersion(void) { return 100; }
int spaGetCompiledArch(void) { return 100; }
int spaDestroyExecutor(spaExecutor *e) { delete e; return SPA_OK; }
int spaExecutorNumLaunches(const spaExecutor *e) { return e ? e->launches : 0; }
int spaExecutorGridSize(const spaExecutor *e) { return e ? e->grid : 0; }
int spaExecutorSetPersistentWorkspace(spaExecutor *e, int enable) {
    if (!e) return fail(SPA_ERR_ARG, "bad executor");
    e->persistent_ws = enable ? 1 : 0;
    e->clean_ws = nullptr;
    return SPA_OK;
}
int spaExecutorSetSelectOptions(spaExecutor *e, int sinkPages, int recentPages, int maxFractionDen) {
    if (!e) return fail(SPA_ERR_ARG, "bad executor");
    if (e->op != OP_CENT_SELECT && e->op != OP_FUSION)
        return fail(SPA_ERR_UNSUPPORTED, "select options apply to CentSelect and SparsePagedFusionAttention");
    if (sinkPages < 0 || recentPages < 0 || maxFractionDen < 0 || sinkPages > 16 || recentPages > 16)
        return fail(SPA_ERR_ARG, "select options: sink %d recent %d (0..16), fraction 1/%d", sinkPages, recentPages, maxFractionDen);
    e->p.sink_pages = sinkPages; e->p.recent_pages = recentPages; e->p.cap_den = maxFractionDen;
    return SPA_OK;
}
int spaExecutorSetSoftmaxLse(spaExecutor *e, const spaTensor *lse) {
    if (!e) return fail(SPA_ERR_ARG, "bad executor");
    if (e->op != OP_ATTENTION && e->op != OP_FUSION) return fail(SPA_ERR_UNSUPPORTED, "softmax lse: only the attention ops");
    if (absent(lse)) { e->p.lse = nullptr; return SPA_OK; }
    if (lse->dtype != SPA_F32) return fail(SPA_ERR_DTYPE, "softmax lse must be float32, got %s", dtype_name(lse->dtype));
    if (lse->ndim < 2 || lse->sizes[0] != e->p.B || lse->sizes[1] != e->p.Nq || numel(lse) != (int64_t)e->p.B * e->p.Nq)
        return fail(SPA_ERR_SHAPE, "softmax lse must be [B=%d, Nq=%d(, 1)]", e->p.B, e->p.Nq);
    e->p.lse = (float *)lse->data; e->p.lse_sb = lse->strides[0]; e->p.lse_sh = lse->strides[1];
    return SPA_OK;
}
int spaExecutorSetTrace(spaExecutor *e, void *deviceBuffer) {
    if (!e) return fail(SPA_ERR_ARG, "bad executor");
    e->p.trace = (unsigned long long *)deviceBuffer;
    e->p.dbg = getenv("SPA_DBG_REPEAT") ? 1 : 0;
    return SPA_OK;
}

// ---------------------------------------------------------------------------------------------
// KV-head tensor parallelism: peer buffers and the fused output all-gather
// ---------------------------------------------------------------------------------------------
int spaPeerBufferCreate(uint64_t bytes, void **devPtr, unsigned char *ipcHandle) {
    if (!devPtr || !ipcHandle || bytes == 0) return fail(SPA_ERR_ARG, "spaPeerBufferCreate: bad argument");
    static_assert(sizeof(cudaIpcMemHandle_t) == SPA_IPC_HANDLE_BYTES, "IPC handle size");
    void *ptr = nullptr;
    CUDA_TRY(cudaMalloc(&ptr, bytes));
    cudaError_t rc = cudaMemset(ptr, 0, bytes);
    cudaIpcMemHandle_t h;
    if (rc == cudaSuccess) rc = cudaIpcGetMemHandle(&h, ptr);
    if (rc != cudaSuccess) { cudaFree(ptr); return fail(SPA_ERR_CUDA, "spaPeerBufferCreate: %s", cudaGetErrorString(rc)); }
    memcpy(ipcHandle, &h, sizeof(h));
    *devPtr = ptr;
    return SPA_OK;
}
int spaPeerBufferOpen(const unsigned char *ipcHandle, void **devPtr) {
    if (!devPtr || !ipcHandle) return fail(SPA_ERR_ARG, "spaPeerBufferOpen: bad argument");
    cudaIpcMemHandle_t h;
    memcpy(&h, ipcHandle, sizeof(h));
    CUDA_TRY(cudaIpcOpenMemHandle(devPtr, h, cudaIpcMemLazyEnablePeerAccess));
    return SPA_OK;
}
int spaPeerBufferClose(void *devPtr) {
    if (devPtr) CUDA_TRY(cudaIpcCloseMemHandle(devPtr));
    return SPA_OK;
}
int spaPeerBufferDestroy(void *devPtr) {
    if (devPtr) CUDA_TRY(cudaFree(devPtr));
    return SPA_OK;
}
int spaExecutorSetPeerOutput(spaExecutor *e, int rank, int world, void *const *gatheredOut, void *const *signals) {
    if (!e) return fail(SPA_ERR_ARG, "bad executor");
    ChainParams &p = e->p;
    p.n_peer = 0; p.tp_rank = 0; p.sig_local = nullptr;
    for (int i = 0; i < kMaxPeers; ++i) { p.out_peer[i] = nullptr; p.sig_peer[i] = nullptr; p.peer_rank[i] = 0; }
    if (world <= 1 || !gatheredOut) return SPA_OK;
    if (e->op != OP_FUSION && e->op != OP_ATTENTION) return fail(SPA_ERR_UNSUPPORTED, "peer output: only the attention ops write attentionOut");
    if (e->p.pq_scale != nullptr) return fail(SPA_ERR_UNSUPPORTED, "peer output: not with an int8 (post-quant) attentionOut");
    if (world > kMaxPeers || rank < 0 || rank >= world || !signals) return fail(SPA_ERR_ARG, "peer output: rank %d of world %d (max %d)", rank, world, kMaxPeers);
    for (int r = 0; r < world; ++r)
        if (!gatheredOut[r] || !signals[r]) return fail(SPA_ERR_ARG, "peer output: null buffer for rank %d", r);
    const ptrdiff_t delta = (const char *)p.out - (const char *)gatheredOut[rank];
    if (delta < 0) return fail(SPA_ERR_ARG, "peer output: attentionOut is not a view into the local gathered buffer");
    int n = 0;
    for (int r = 0; r < world; ++r) {
        if (r == rank) continue;
        p.out_peer[n] = (char *)gatheredOut[r] + delta;
        p.sig_peer[n] = (unsigned *)signals[r];
        p.peer_rank[n] = r;
        ++n;
    }
    p.sig_local = (unsigned *)signals[rank];
    p.n_peer = n; p.tp_rank = rank;
    return SPA_OK;
}

// ---------------------------------------------------------------------------------------------
// centroid maintenance: page -> cluster assignment
// ---------------------------------------------------------------------------------------------
int spaAssignPageClustersGetWorkspaceSize(const spaTensor *key, const spaTensor *l1Cent, const spaTensor *pageIds,
                                          const spaTensor *pageTokens, const spaTensor *antiquantScale,
                                          const spaTensor *antiquantOffset, int64_t blockSize, const spaTensor *blockIds,
                                          const spaTensor *pageMeanOut, uint64_t *ws, spaExecutor **executor) {
    if (!ws || !executor) return fail(SPA_ERR_ARG, "null output pointer");
    *executor = nullptr; *ws = 0;
    REQUIRE_PRESENT(l1Cent, "l1Cent"); REQUIRE_PRESENT(blockIds, "blockIds");
    if (l1Cent->ndim != 3 || l1Cent->sizes[2] != kD) return fail(SPA_ERR_SHAPE, "l1Cent: expected [Nkv,C,%d]", kD);
    REQUIRE_CONTIG(l1Cent, "l1Cent");
    if (l1Cent->dtype != SPA_F16 && l1Cent->dtype != SPA_BF16) return fail(SPA_ERR_DTYPE, "l1Cent: expected float16/bfloat16");
    const int Nkv = (int)l1Cent->sizes[0], C = (int)l1Cent->sizes[1];
    if (C < 1 || C > 8192) return fail(SPA_ERR_UNSUPPORTED, "clusters %d outside 1..8192", C);
    if (blockSize < 4 || blockSize > 1024 || (blockSize & 3)) return fail(SPA_ERR_UNSUPPORTED, "block size %lld", (long long)blockSize);
    REQUIRE_PRESENT(key, "key");
    const bool i8 = key->dtype == SPA_I8;
    if (!i8 && key->dtype != l1Cent->dtype) return fail(SPA_ERR_DTYPE, "key must be int8 or have the centroid dtype %s, got %s",
                                                        dtype_name(l1Cent->dtype), dtype_name(key->dtype));
    int aq_per_tensor = 0;
    if (i8) {
        // K antiquant: x = (x_i8 + offset) * scale per channel (antiquant tensors [2, Nkv, D]: row 0 = K), or one value
        if (absent(antiquantScale)) return fail(SPA_ERR_ARG, "int8 key pool needs antiquantScale");
        for (const spaTensor *t : {antiquantScale, antiquantOffset}) {
            if (absent(t)) continue;
            if (t->dtype != l1Cent->dtype) return fail(SPA_ERR_DTYPE, "antiquant tensors must have the centroid dtype");
            if (!contiguous(t)) return fail(SPA_ERR_SHAPE, "antiquant tensors must be contiguous");
            const int64_t n = numel(t);
            if (n != 2 && n != 2ll * Nkv * kD) return fail(SPA_ERR_SHAPE, "antiquant tensors: expected [2,%d,%d] or [2]", Nkv, kD);
            if (t == antiquantScale) aq_per_tensor = n == 2;
            else if ((n == 2) != (aq_per_tensor != 0)) return fail(SPA_ERR_SHAPE, "antiquant scale and offset must have the same shape");
        }
    } else if (!absent(antiquantScale) || !absent(antiquantOffset)) {
        return fail(SPA_ERR_ARG, "antiquant tensors given for a non-int8 key pool");
    }
    KvLayout L;
    int rc = parse_kv(key, "key", Nkv, (int)blockSize, &L);
    if (rc) return rc;
    REQUIRE_DTYPE(blockIds, SPA_I32, "blockIds");
    if (blockIds->ndim != 2 || blockIds->sizes[0] != Nkv || blockIds->strides[1] != 1)
        return fail(SPA_ERR_SHAPE, "blockIds: expected [Nkv=%d, kvPageLen]", Nkv);
    spaExecutor *e = new (std::nothrow) spaExecutor();
    if (!e) return fail(SPA_ERR_ARG, "out of memory");
    e->op = OP_ASSIGN; e->launches = 1; e->q_bf16 = l1Cent->dtype == SPA_BF16; e->kv_i8 = i8;
    AssignParams &a = e->ap;
    a.n = absent(pageIds) ? 0 : (int)numel(pageIds);
    if (a.n) {
        if (pageIds->dtype != SPA_I32 || !contiguous(pageIds)) { delete e; return fail(SPA_ERR_DTYPE, "pageIds: expected contiguous int32"); }
        if (!absent(pageTokens) && (pageTokens->dtype != SPA_I32 || !contiguous(pageTokens) || numel(pageTokens) != a.n)) {
            delete e; return fail(SPA_ERR_SHAPE, "pageTokens: expected contiguous int32 [%d]", a.n);
        }
        if (!absent(pageMeanOut) && (pageMeanOut->dtype != SPA_F32 || !contiguous(pageMeanOut) ||
                                     numel(pageMeanOut) != (int64_t)a.n * Nkv * kD)) {
            delete e; return fail(SPA_ERR_SHAPE, "pageMeanOut: expected contiguous float32 [%d,%d,%d]", a.n, Nkv, kD);
        }
    }
    a.key = key->data; a.s_blk = L.s_blk; a.s_row = L.s_row; a.s_head = L.s_head;
    a.cent = l1Cent->data;
    a.page_ids = a.n ? (const int *)pageIds->data : nullptr;
    a.page_tokens = (a.n && !absent(pageTokens)) ? (const int *)pageTokens->data : nullptr;
    a.block_ids = (int *)blockIds->data; a.bid_stride = blockIds->strides[0];
    a.page_mean = (a.n && !absent(pageMeanOut)) ? (float *)pageMeanOut->data : nullptr;
    a.aq_scale = i8 ? antiquantScale->data : nullptr;
    a.aq_offset = (i8 && !absent(antiquantOffset)) ? antiquantOffset->data : nullptr;
    a.aq_per_tensor = aq_per_tensor;
    a.Nkv = Nkv; a.C = C; a.bs = (int)blockSize; a.num_blocks = (int)L.nblocks; a.kvPageLen = (int)blockIds->sizes[1];
    if (a.n) {
        DevInfo d;
        if ((rc = dev_info(&d))) { delete e; return rc; }
        const int64_t units = (int64_t)((a.n + kAssignPages - 1) / kAssignPages) * Nkv;
        const AssignSmem S(C, e->q_bf16 ? 3 : 2);
        if (S.total > d.smem_optin) { delete e; return fail(SPA_ERR_UNSUPPORTED, "clusters %d need %d bytes of shared memory", C, S.total); }
        const int per_sm = std::max(1, std::min(4, (d.smem_optin - 1024) / (S.total + 1024)));
        e->grid = (int)std::min<int64_t>(units, (int64_t)d.sms * per_sm);
        void (*kern)(AssignParams) = e->q_bf16 ? (i8 ? spa_assign_kernel<__nv_bfloat16, true> : spa_assign_kernel<__nv_bfloat16, false>)
                                               : (i8 ? spa_assign_kernel<__half, true> : spa_assign_kernel<__half, false>);
        if ((rc = raise_smem_limit((const void *)kern, d.smem_optin))) { delete e; return rc; }
    }
    *executor = e;
    return SPA_OK;
}
int spaAssignPageClusters(void *, uint64_t, spaExecutor *e, spaStream stream_) {
    if (!e || e->op != OP_ASSIGN) return fail(SPA_ERR_ARG, "bad executor");
    if (e->ap.n == 0) return SPA_OK;
    cudaStream_t stream = (cudaStream_t)stream_;
    const AssignSmem S(e->ap.C, e->q_bf16 ? 3 : 2);
    void (*kern)(AssignParams) = e->q_bf16 ? (e->kv_i8 ? spa_assign_kernel<__nv_bfloat16, true> : spa_assign_kernel<__nv_bfloat16, false>)
                                           : (e->kv_i8 ? spa_assign_kernel<__half, true> : spa_assign_kernel<__half, false>);
    kern<<<e->grid, kAssignThreads, S.total, stream>>>(e->ap);       // (smem limit raised once, at plan time)
    CUDA_TRY(cudaGetLastError());
    return SPA_OK;
}

// ---------------------------------------------------------------------------------------------
// selection ops
// ---------------------------------------------------------------------------------------------
static int finish_select_plan(spaExecutor *e, uint64_t *ws, spaExecutor **out) {
    DevInfo d;
    int rc = dev_info(&d);
    if (rc) { delete e; return rc; }
    int bytes = 0;
    // stand-alone ops: enough CTAs for one item per group, capped at a few waves
    rc = plan_select_smem(&e->p, e->flags, 0, d.smem_optin - 2048, &bytes);
    if (rc) { delete e; return rc; }
    memset(&e->tmC, 0, sizeof(e->tmC));
    if (e->flags & kSelScores) {
        rc = make_cent_map(&e->tmC, e->p.cent, e->p.Nkv * e->p.C, e->q_bf16);
        if (rc) { delete e; return rc; }
    }
    e->p.smem_bytes = bytes + 1024;
    if (SelectKernel k = select_kernel_for(e->q_bf16, e->flags)) {
        if ((rc = raise_smem_limit((const void *)k, d.smem_optin))) { delete e; return rc; }
    }
    const int ngroups = kConsumerWarps / e->p.sel_gw;
    e->grid = std::max(1, std::min((e->p.items + ngroups - 1) / ngroups, 8 * d.sms));
    // exact-score scratch for the rare case of more than kXCap near-ties around the k-th score
    e->off_xs = 0;
    e->ws_bytes = (e->flags & kSelScores) ? align_up((uint64_t)e->grid * kConsumerWarps * std::max(e->p.C, 1) * 8, 256) : 0;
    *ws = e->ws_bytes;
    *out = e;
    return SPA_OK;
}

static int launch_select(void *workspace, uint64_t wsSize, spaExecutor *e, cudaStream_t stream) {
    if (e->p.items == 0) return SPA_OK;
    if (e->ws_bytes && (!workspace || wsSize < e->ws_bytes))
        return fail(SPA_ERR_WORKSPACE, "workspace %llu bytes < required %llu", (unsigned long long)wsSize, (unsigned long long)e->ws_bytes);
    ChainParams p = e->p;
    p.exact_scratch = (double *)workspace;
    SelectKernel k = select_kernel_for(e->q_bf16, e->flags);
    if (!k) return fail(SPA_ERR_ARG, "no selection kernel for flags %d", e->flags);
    k<<<e->grid, kThreads, p.smem_bytes, stream>>>(p, e->tmC);       // (smem limit raised once, in finish_select_plan)
    CUDA_TRY(cudaGetLastError());
    if (e->op == OP_CENT_SELECT) {
        spa_max_ppl_kernel<<<std::max(1, p.B), 32, 0, stream>>>(p);
        CUDA_TRY(cudaGetLastError());
    }
    return SPA_OK;
}

int spaComputeCentGetWorkspaceSize(const spaTensor *query, const spaTensor *l1Cent,
                                   const spaTensor *indicesOut, uint64_t *ws, spaExecutor **executor) {
    if (!ws || !executor) return fail(SPA_ERR_ARG, "workspaceSize/executor must not be null");
    spaExecutor *e = new (std::nothrow) spaExecutor();
    if (!e) return fail(SPA_ERR_ARG, "out of host memory");
    memset(e, 0, sizeof(*e));
    ChainParams &p = e->p;
    int rc;
#define BAIL(x) do { rc = (x); if (rc) { delete e; return rc; } } while (0)
    BAIL(parse_query(query, "query", &p.B, &p.Nq, &p.q_sb, &p.q_sh, &e->q_bf16));
    BAIL(parse_cent(l1Cent, query->dtype, &p.Nkv, &p.C));
    // k is read from the output's last dim (compute_cent_tiling.cpp:45-50); k == 0 => nothing to do
    if (!indicesOut) { delete e; return fail(SPA_ERR_ARG, "indices: required [B,Nq,k] int32"); }
    if (indicesOut->dtype != SPA_I32) { delete e; return fail(SPA_ERR_DTYPE, "indices: expected int32"); }
    if (indicesOut->ndim != 3 || indicesOut->sizes[0] != p.B || indicesOut->sizes[1] != p.Nq || !contiguous(indicesOut)) {
        delete e; return fail(SPA_ERR_SHAPE, "indices: expected contiguous [%d,%d,k]", p.B, p.Nq);
    }
    p.k = (int)indicesOut->sizes[2];
    if (p.Nkv <= 0 || p.Nq % p.Nkv) { delete e; return fail(SPA_ERR_SHAPE, "num q heads %d not a multiple of kv heads %d", p.Nq, p.Nkv); }
    p.g = p.Nq / p.Nkv;
    if (p.g > kMaxGroup) { delete e; return fail(SPA_ERR_UNSUPPORTED, "group size %d > %d", p.g, kMaxGroup); }
    if (p.k > p.C) { delete e; return fail(SPA_ERR_SHAPE, "k=%d exceeds cluster count C=%d", p.k, p.C); }
    p.q = query->data; p.cent = l1Cent->data; p.indices_out = (int *)indicesOut->data;
    p.items = (p.k == 0) ? 0 : p.B * p.Nkv; p.maxP = 1; p.bs = kBS; p.MP = 1;
    e->op = OP_COMPUTE_CENT; e->flags = kSelScores | kSelWriteIndices; e->launches = 1;
    return finish_select_plan(e, ws, executor);
}

int spaComputeCent(void *workspace, uint64_t wsSize, spaExecutor *e, spaStream stream) {
    if (!e || e->op != OP_COMPUTE_CENT) return fail(SPA_ERR_ARG, "bad executor");
    return launch_select(workspace, wsSize, e, (cudaStream_t)stream);
}

static int parse_select_outputs(ChainParams &p, const spaTensor *pp, const spaTensor *ppl) {
    REQUIRE_PRESENT(pp, "page_position"); REQUIRE_DTYPE(pp, SPA_I32, "page_position");
    REQUIRE_PRESENT(ppl, "page_position_length"); REQUIRE_DTYPE(ppl, SPA_I32, "page_position_length");
    if (pp->ndim != 3 || pp->sizes[0] < p.B || pp->sizes[1] != p.Nq || !contiguous(pp))
        return fail(SPA_ERR_SHAPE, "page_position: expected contiguous [%d,%d,max_page_num]", p.B, p.Nq);
    if (ppl->ndim != 3 || ppl->sizes[0] < p.B || ppl->sizes[1] != p.Nq || ppl->sizes[2] != kPad || !contiguous(ppl))
        return fail(SPA_ERR_SHAPE, "page_position_length: expected contiguous [%d,%d,%d]", p.B, p.Nq, kPad);
    p.MP = (int)pp->sizes[2];
    p.page_position = (int *)pp->data; p.page_position_length = (int *)ppl->data;
    return SPA_OK;
}

int spaSelectPositionGetWorkspaceSize(const spaTensor *blockIds, const spaTensor *blockTable,
                                      const spaTensor *seqLen, const spaTensor *indices, int64_t blockSize,
                                      const spaTensor *pagePositionOut, const spaTensor *pagePositionLengthOut,
                                      const spaTensor *blockTableGatherOut, uint64_t *ws, spaExecutor **executor) {
    if (!ws || !executor) return fail(SPA_ERR_ARG, "workspaceSize/executor must not be null");
    spaExecutor *e = new (std::nothrow) spaExecutor();
    if (!e) return fail(SPA_ERR_ARG, "out of host memory");
    memset(e, 0, sizeof(*e));
    ChainParams &p = e->p;
    int rc;
    BAIL(parse_tables(blockIds, blockTable, seqLen, &p));
    if (absent(indices)) { delete e; return fail(SPA_ERR_ARG, "indices: required tensor is absent"); }
    if (indices->dtype != SPA_I32) { delete e; return fail(SPA_ERR_DTYPE, "indices: expected int32"); }
    if (indices->ndim != 3 || indices->strides[2] != 1) { delete e; return fail(SPA_ERR_SHAPE, "indices: expected [B,Nq,k]"); }
    if (blockSize <= 0) { delete e; return fail(SPA_ERR_SHAPE, "block_size must be positive"); }
    p.B = (int)seqLen->sizes[0]; p.Nq = (int)indices->sizes[1]; p.Nkv = (int)blockIds->sizes[0];
    if (indices->sizes[0] < p.B) { delete e; return fail(SPA_ERR_SHAPE, "indices batch smaller than seq_len batch"); }
    if (p.Nkv <= 0 || p.Nq % p.Nkv) { delete e; return fail(SPA_ERR_SHAPE, "num q heads %d not a multiple of kv heads %d", p.Nq, p.Nkv); }
    p.g = p.Nq / p.Nkv;
    if (p.g > kMaxGroup) { delete e; return fail(SPA_ERR_UNSUPPORTED, "group size %d > %d", p.g, kMaxGroup); }
    p.k = (int)indices->sizes[2]; p.C = 32;
    p.indices_in = (const int *)indices->data; p.idx_sb = indices->strides[0]; p.idx_sh = indices->strides[1];
    p.bs = (int)blockSize;
    BAIL(parse_select_outputs(p, pagePositionOut, pagePositionLengthOut));
    if (!absent(blockTableGatherOut)) {
        if (blockTableGatherOut->dtype != SPA_I32 || blockTableGatherOut->ndim != 3 || blockTableGatherOut->sizes[0] != p.B ||
            blockTableGatherOut->sizes[1] != p.Nq || blockTableGatherOut->sizes[2] != p.maxP || !contiguous(blockTableGatherOut)) {
            delete e; return fail(SPA_ERR_SHAPE, "block_table_gather: expected contiguous int32 [%d,%d,%d]", p.B, p.Nq, p.maxP);
        }
        p.bt_gather = (int *)blockTableGatherOut->data;
    }
    p.items = p.B * p.Nkv;
    e->op = OP_SELECT_POSITION; e->flags = kSelSelect | (p.bt_gather ? kSelGather : 0); e->launches = 1;
    return finish_select_plan(e, ws, executor);
}

int spaSelectPosition(void *workspace, uint64_t wsSize, spaExecutor *e, spaStream stream) {
    if (!e || e->op != OP_SELECT_POSITION) return fail(SPA_ERR_ARG, "bad executor");
    return launch_select(workspace, wsSize, e, (cudaStream_t)stream);
}

static int parse_cent_select_common(spaExecutor *e, const spaTensor *query, const spaTensor *l1Cent,
                                    const spaTensor *blockIds, const spaTensor *blockTable,
                                    const spaTensor *seqLen, int64_t topK, int64_t blockSize,
                                    const spaTensor *pp, const spaTensor *ppl, const spaTensor *mppl) {
    ChainParams &p = e->p;
    int rc, Bq = 0;
    if ((rc = parse_query(query, "query", &Bq, &p.Nq, &p.q_sb, &p.q_sh, &e->q_bf16))) return rc;
    if ((rc = parse_cent(l1Cent, query->dtype, &p.Nkv, &p.C))) return rc;
    if ((rc = parse_tables(blockIds, blockTable, seqLen, &p))) return rc;
    p.B = (int)seqLen->sizes[0];                                  // add_custom.cpp:988
    if (p.B > Bq) return fail(SPA_ERR_SHAPE, "seq_len batch %d exceeds query batch %d", p.B, Bq);
    if (p.Nkv <= 0 || p.Nq % p.Nkv) return fail(SPA_ERR_SHAPE, "num q heads %d not a multiple of kv heads %d", p.Nq, p.Nkv);
    if (blockIds->sizes[0] != p.Nkv) return fail(SPA_ERR_SHAPE, "block_ids rows %lld != kv heads %d", (long long)blockIds->sizes[0], p.Nkv);
    p.g = p.Nq / p.Nkv;
    if (p.g > kMaxGroup) return fail(SPA_ERR_UNSUPPORTED, "group size %d > %d", p.g, kMaxGroup);
    if (topK <= 0 || topK > p.C) return fail(SPA_ERR_SHAPE, "top-k %lld must be in [1, C=%d]", (long long)topK, p.C);
    if (blockSize <= 0) return fail(SPA_ERR_SHAPE, "block_size must be positive");
    p.k = (int)topK; p.bs = (int)blockSize;
    if ((rc = parse_select_outputs(p, pp, ppl))) return rc;
    REQUIRE_PRESENT(mppl, "max_page_position_length"); REQUIRE_DTYPE(mppl, SPA_I64, "max_page_position_length");
    if (mppl->ndim != 2 || mppl->sizes[0] < p.B || mppl->sizes[1] != kPad || !contiguous(mppl))
        return fail(SPA_ERR_SHAPE, "max_page_position_length: expected contiguous int64 [%d,%d]", p.B, kPad);
    p.max_ppl = (long long *)mppl->data;
    p.q = query->data; p.cent = l1Cent->data;
    p.items = p.B * p.Nkv;
    return SPA_OK;
}

int spaCentSelectGetWorkspaceSize(const spaTensor *query, const spaTensor *l1Cent, const spaTensor *blockIds,
                                  const spaTensor *blockTable, const spaTensor *seqLen, int64_t topK,
                                  int64_t blockSize, const spaTensor *pagePositionOut,
                                  const spaTensor *pagePositionLengthOut, const spaTensor *maxPagePositionLengthOut,
                                  uint64_t *ws, spaExecutor **executor) {
    if (!ws || !executor) return fail(SPA_ERR_ARG, "workspaceSize/executor must not be null");
    spaExecutor *e = new (std::nothrow) spaExecutor();
    if (!e) return fail(SPA_ERR_ARG, "out of host memory");
    memset(e, 0, sizeof(*e));
    int rc;
    BAIL(parse_cent_select_common(e, query, l1Cent, blockIds, blockTable, seqLen, topK, blockSize,
                                  pagePositionOut, pagePositionLengthOut, maxPagePositionLengthOut));
    e->op = OP_CENT_SELECT; e->flags = kSelScores | kSelSelect; e->launches = 2;
    return finish_select_plan(e, ws, executor);
}

int spaCentSelect(void *workspace, uint64_t wsSize, spaExecutor *e, spaStream stream) {
    if (!e || e->op != OP_CENT_SELECT) return fail(SPA_ERR_ARG, "bad executor");
    return launch_select(workspace, wsSize, e, (cudaStream_t)stream);
}

// ---------------------------------------------------------------------------------------------
// attention ops
// ---------------------------------------------------------------------------------------------
// The optional tensors of the aclnn prototypes between `value` and `antiquantScale` (+ kvPaddingSize).  None of
// them is non-empty in any reference sparse test (tests/ifa_v5_case.py:133-146 pass empty tensors).
struct AttnOptions { const spaTensor *pse, *mask, *dq1, *q1, *dq2, *q2, *qo2, *kvpad; };

// pseShift / attenMask / quantScale2 / quantOffset2 -> the options build of the chain kernel.  Shapes
// and dtypes as the reference tiling checks them (sparse_paged_attention_tiling_check.cc:737-850,
// sparse_paged_attention_tiling.cc:566-647).  Called once B, Nq, maxP and bs are known.
static int parse_attn_options(spaExecutor *e, const AttnOptions &o, const spaTensor *query) {
    ChainParams &p = e->p;
    // int8 QUERY (full-quant IFA) does not exist on the sparse path: the query is f16 / bf16
    if (!absent(o.dq1) || !absent(o.q1) || !absent(o.dq2))
        return fail(SPA_ERR_UNSUPPORTED, "dequant_scale1 / quant_scale1 / dequant_scale2 belong to int8 queries, which the sparse path does not take");
    if (!absent(o.kvpad)) return fail(SPA_ERR_UNSUPPORTED, "kv_padding_size (left padding) is not defined for paged KV");
    const long long smax = (long long)p.maxP * p.bs;
    if (!absent(o.pse)) {
        const spaTensor *t = o.pse;
        if (t->dtype != query->dtype) return fail(SPA_ERR_DTYPE, "pse_shift dtype %s must equal the query dtype %s", dtype_name(t->dtype), dtype_name(query->dtype));
        if (t->ndim != 4 || (t->sizes[0] != 1 && t->sizes[0] < p.B) || t->sizes[1] != p.Nq || t->sizes[2] != 1)
            return fail(SPA_ERR_SHAPE, "pse_shift: expected [B or 1, Nq, 1, S]");
        if (t->sizes[3] < smax) return fail(SPA_ERR_SHAPE, "pse_shift: S=%lld < maxBlocksPerBatch * blockSize = %lld", (long long)t->sizes[3], smax);
        if (t->strides[3] != 1) return fail(SPA_ERR_SHAPE, "pse_shift rows must be contiguous");
        p.pse = t->data; p.pse_sb = t->sizes[0] == 1 ? 0 : t->strides[0]; p.pse_sh = t->strides[1];
        e->opt = 1;
    }
    if (!absent(o.mask)) {
        const spaTensor *t = o.mask;
        if (t->dtype != SPA_BOOL && t->dtype != SPA_I8 && t->dtype != SPA_U8)
            return fail(SPA_ERR_DTYPE, "attention_mask: bool, int8 or uint8 (got %s)", dtype_name(t->dtype));
        if (t->ndim < 2 || t->sizes[0] < p.B) return fail(SPA_ERR_SHAPE, "attention_mask: expected [B, (1, (1,)) S]");
        for (int i = 1; i + 1 < t->ndim; ++i)
            if (t->sizes[i] != 1) return fail(SPA_ERR_SHAPE, "attention_mask: expected [B, (1, (1,)) S]");
        if (t->sizes[t->ndim - 1] < smax)
            return fail(SPA_ERR_SHAPE, "attention_mask: S=%lld < maxBlocksPerBatch * blockSize = %lld", (long long)t->sizes[t->ndim - 1], smax);
        if (t->strides[t->ndim - 1] != 1) return fail(SPA_ERR_SHAPE, "attention_mask rows must be contiguous");
        p.mask = (const unsigned char *)t->data; p.mask_sb = t->strides[0];
        e->opt = 1;
    }
    if (!absent(o.q2)) {
        const spaTensor *t = o.q2;
        if (t->dtype != SPA_F32 && !(t->dtype == SPA_BF16 && query->dtype == SPA_BF16))
            return fail(SPA_ERR_DTYPE, "quant_scale2: float32, or bfloat16 with a bfloat16 query");
        const int64_t n = numel(t);
        if (n != 1 && n != (int64_t)p.Nq * kD) return fail(SPA_ERR_SHAPE, "quant_scale2: expected 1 or Nq*D = %lld elements", (long long)p.Nq * kD);
        REQUIRE_CONTIG(t, "quant_scale2");
        p.pq_scale = t->data; p.pq_scale_chn = n != 1; p.pq_bf16 = t->dtype == SPA_BF16;
        if (!absent(o.qo2)) {
            if (o.qo2->dtype != t->dtype) return fail(SPA_ERR_DTYPE, "quant_offset2 dtype must equal quant_scale2's");
            const int64_t m = numel(o.qo2);
            if (m != 1 && m != (int64_t)p.Nq * kD) return fail(SPA_ERR_SHAPE, "quant_offset2: expected 1 or Nq*D elements");
            REQUIRE_CONTIG(o.qo2, "quant_offset2");
            p.pq_offset = o.qo2->data; p.pq_offset_chn = m != 1;
        }
        e->opt = 1;
    } else if (!absent(o.qo2)) {
        return fail(SPA_ERR_ARG, "quant_offset2 given without quant_scale2");
    }
    p.opt = e->opt;
    return SPA_OK;
}

static int plan_attention(spaExecutor *e, const spaTensor *query, const spaTensor *key, const spaTensor *value,
                          const AttnOptions &opts, const spaTensor *aqs, const spaTensor *aqo, const spaTensor *blocktable,
                          int64_t numHeads, double scaleValue, const char *inputLayout, int64_t numKV,
                          int64_t blockSize, const spaTensor *out, uint64_t *ws) {
    ChainParams &p = e->p;
    int rc;
    if (inputLayout && strcmp(inputLayout, "BNSD") && strcmp(inputLayout, "BSH") && strcmp(inputLayout, "BSND"))
        return fail(SPA_ERR_ARG, "input_layout '%s' not one of BNSD/BSH/BSND", inputLayout);
    if (p.Nq != numHeads) return fail(SPA_ERR_SHAPE, "num_heads=%lld disagrees with query heads %d", (long long)numHeads, p.Nq);
    if (numKV <= 0) numKV = numHeads;                    // reference default: 0 => MHA
    if (p.Nq % numKV) return fail(SPA_ERR_SHAPE, "num_heads %d not a multiple of num_key_value_heads %lld", p.Nq, (long long)numKV);
    p.Nkv = (int)numKV; p.g = p.Nq / p.Nkv;
    if (p.g > kMaxGroup) return fail(SPA_ERR_UNSUPPORTED, "group size %d > %d", p.g, kMaxGroup);
    // 128 (every BASELINE config) streams full stage images; 16 / 32 / 64 (the reference's BNSD known-answer test runs
    // block 16, tests/ifa_v5_case.py:22) fill the first rows of an image -- functional, not tuned
    if (blockSize != 16 && blockSize != 32 && blockSize != 64 && blockSize != kBS)
        return fail(SPA_ERR_UNSUPPORTED, "block_size %lld: this build supports 16, 32, 64 and %d", (long long)blockSize, kBS);
    p.bs = (int)blockSize;
    REQUIRE_PRESENT(blocktable, "blocktable"); REQUIRE_DTYPE(blocktable, SPA_I32, "blocktable");
    if (blocktable->ndim != 2 || blocktable->strides[1] != 1) return fail(SPA_ERR_SHAPE, "blocktable: expected [maxB,maxP]");
    if (blocktable->sizes[0] < p.B) return fail(SPA_ERR_SHAPE, "blocktable rows %lld < batch %d", (long long)blocktable->sizes[0], p.B);
    p.block_table = (const int *)blocktable->data; p.bt_stride = blocktable->strides[0]; p.maxP = (int)blocktable->sizes[1];
    if ((rc = parse_attn_options(e, opts, query))) return rc;
    // output: same shape as query (add_custom.cpp:559-560); its dtype, or int8 exactly when quant_scale2 is given
    // (sparse_paged_attention_tiling.cc:570-602)
    REQUIRE_PRESENT(out, "attention_out");
    if (p.pq_scale != nullptr) {
        if (out->dtype != SPA_I8) return fail(SPA_ERR_DTYPE, "quant_scale2 is given: attention_out must be int8 (got %s)", dtype_name(out->dtype));
    } else if (out->dtype == SPA_I8) {
        return fail(SPA_ERR_ARG, "attention_out is int8 but quant_scale2 is absent");
    } else {
        REQUIRE_DTYPE(out, query->dtype, "attention_out");
    }
    if (out->ndim != query->ndim) return fail(SPA_ERR_SHAPE, "attention_out must have the query's shape");
    for (int i = 0; i < out->ndim; ++i)
        if (out->sizes[i] != query->sizes[i] && !(i == 0 && out->sizes[0] >= p.B))
            return fail(SPA_ERR_SHAPE, "attention_out must have the query's shape");
    if (out->strides[out->ndim - 1] != 1 || (out->strides[0] & 3) || (out->strides[1] & 3) ||
        (reinterpret_cast<uintptr_t>(out->data) & 7))
        return fail(SPA_ERR_SHAPE, "attention_out rows must be contiguous and 8-byte aligned");
    REQUIRE_PRESENT(key, "key"); REQUIRE_PRESENT(value, "value");
    if (key->dtype != value->dtype) return fail(SPA_ERR_DTYPE, "key/value dtypes differ");
    e->kv_f8 = key->dtype == SPA_F8E4M3;
    e->kv_i8 = key->dtype == SPA_I8 || e->kv_f8;
    if (!e->kv_i8 && key->dtype != query->dtype)
        return fail(SPA_ERR_DTYPE, "key dtype %s must equal query dtype %s, int8 or float8_e4m3", dtype_name(key->dtype), dtype_name(query->dtype));
    KvLayout LK, LV;
    if ((rc = parse_kv(key, "key", p.Nkv, p.bs, &LK))) return rc;
    if ((rc = parse_kv(value, "value", p.Nkv, p.bs, &LV))) return rc;
    p.num_blocks = (int)LK.nblocks;
    int ar, ah, ab, vr, vh, vb;
    if ((rc = make_kv_map(&e->tmK, key, LK, p.Nkv, p.bs, &ar, &ah, &ab, "key"))) return rc;
    if ((rc = make_kv_map(&e->tmV, value, LV, p.Nkv, p.bs, &vr, &vh, &vb, "value"))) return rc;
    if (ar != vr || ah != vh || ab != vb) return fail(SPA_ERR_SHAPE, "key and value pools must share a layout");
    p.ax_row = ar; p.ax_head = ah; p.ax_blk = ab;
    // antiquant: non-empty <=> int8 KV
    if (e->kv_f8 && absent(aqs)) {
        // FP8 (e4m3) pages are numbers already: scale (and offset) are optional
        if (!absent(aqo)) return fail(SPA_ERR_ARG, "antiquant_offset given without antiquant_scale");
    } else if (e->kv_i8) {
        if (absent(aqs)) return fail(SPA_ERR_ARG, "int8 KV requires antiquant_scale");
        if (aqs->dtype != query->dtype) return fail(SPA_ERR_DTYPE, "antiquant_scale dtype must equal query dtype");
        const int64_t n = numel(aqs);
        if (n == 2) p.aq_per_tensor = 1;
        else if (n == 2LL * p.Nkv * kD && aqs->sizes[0] == 2) p.aq_per_tensor = 0;
        else return fail(SPA_ERR_SHAPE, "antiquant_scale: expected [2] or [2,Nkv,(1,)D]");
        REQUIRE_CONTIG(aqs, "antiquant_scale");
        p.aq_scale = aqs->data;
        if (!absent(aqo)) {
            if (aqo->dtype != query->dtype || numel(aqo) != n) return fail(SPA_ERR_SHAPE, "antiquant_offset must match antiquant_scale");
            REQUIRE_CONTIG(aqo, "antiquant_offset");
            p.aq_offset = aqo->data;
        }
    } else if (!absent(aqs) || !absent(aqo)) {
        return fail(SPA_ERR_ARG, "antiquant_scale/offset given but KV is neither int8 nor float8_e4m3");
    }
    p.out = out->data; p.out_sb = out->strides[0]; p.out_sh = out->strides[1];
    p.q = query->data;
    p.scale_log2 = (float)scaleValue * kLog2e;
    p.items = p.B * p.Nkv;
    if (p.items > kMaxItems) return fail(SPA_ERR_UNSUPPORTED, "B*Nkv = %d > %d work items", p.items, kMaxItems);

    // ---- launch shape
    DevInfo d;
    if ((rc = dev_info(&d))) return rc;
    if (!d.coop) return fail(SPA_ERR_CUDA, "device lacks cooperative launch");
    e->nt_tiles = (p.g + 7) / 8;
    // ---- scheduling (spa_select.cuh publish_item, spa_attn.cuh producer): a CTA keeps static_pct % of its fair
    // share as a private prefix of its own items; the rest is offered in chunks, about dyn_runs per CTA
    p.static_pct = 80; p.dyn_runs = 3; p.dyn_min = 2;
    p.small_pages = 16;    // measured crossover on the configs[4] sweep: ~16.5 pages per CTA
    if (const char *s = getenv("SPA_SMALL_PAGES")) p.small_pages = std::max(0, atoi(s));
    if (const char *s = getenv("SPA_STATIC_PCT")) p.static_pct = std::min(100, std::max(0, atoi(s)));
    if (const char *s = getenv("SPA_DYN_RUNS")) p.dyn_runs = std::max(1, atoi(s));
    if (const char *s = getenv("SPA_DYN_MIN")) p.dyn_min = std::max(1, atoi(s));
    p.chunk_num = 1; p.chunk_den = 2; p.claim_lead = 3;
    if (const char *s = getenv("SPA_CHUNK_FRAC")) { int a = 0, b = 0; if (sscanf(s, "%d/%d", &a, &b) == 2 && a > 0 && b >= a) { p.chunk_num = a; p.chunk_den = b; } }
    if (const char *s = getenv("SPA_CLAIM_LEAD")) p.claim_lead = std::max(0, atoi(s));
    p.qstride = p.maxP * (p.mode == kModeAttnOnly ? 2 : 1);     // x2: a page may be listed twice (split tail entries)
    // Partials: fixed slots, one for the owner's prefix and one per chunk of every item; chunks are not smaller
    // than 1/32 of a CTA's fair share (publish_item), so with very many items an item is never cut more than once
    // few items (tensor-parallel shards): up to 15 chunks each, so that the idle CTAs get even shares of long items
    p.pslots = ((long long)p.items >= 32ll * d.sms) ? 2 : (p.items <= d.sms ? kMaxChunks + 1 : 9);
    if (p.qstride >= 65536) return fail(SPA_ERR_UNSUPPORTED, "more than 65535 page entries per work item");
    if (p.num_blocks > kMaxBlocks) return fail(SPA_ERR_UNSUPPORTED, "more than %d KV blocks", kMaxBlocks);
    p.stage_bytes = e->kv_i8 ? 16384 : 32768;
    p.page_tx = p.bs * kD * (e->kv_i8 ? 1 : 2);
    const int budget = d.smem_optin - 1024 - 512;     // 1 KiB alignment slack + the kernels' static shared memory
    // the producer stages every run's query rows (and the K scales) in shared memory ahead of the consumers
    // (bulk copies: 16-byte aligned rows); otherwise the consumers read them from global memory per segment
    p.q_stage = e->kv_i8 && (reinterpret_cast<uintptr_t>(p.q) % 16 == 0) && p.q_sb % 8 == 0 && p.q_sh % 8 == 0 &&
                (p.aq_per_tensor || reinterpret_cast<uintptr_t>(p.aq_scale) % 16 == 0);      // (8-bit KV kernels only)
    if (const char *s = getenv("SPA_Q_STAGE")) p.q_stage = p.q_stage && atoi(s) != 0;
    // options build: the producer stages each page's mask bytes and pse rows next to its K image (bulk copies:
    // 16-byte aligned rows); otherwise the consumers read them from global memory, one page per step
    p.aux_stage = 0; p.aux_bytes = 0;
    if (p.mask != nullptr || p.pse != nullptr) {
        const bool mask_ok = p.mask == nullptr || (reinterpret_cast<uintptr_t>(p.mask) % 16 == 0 && p.mask_sb % 16 == 0);
        const bool pse_ok = p.pse == nullptr || (reinterpret_cast<uintptr_t>(p.pse) % 16 == 0 && p.pse_sb % 8 == 0 && p.pse_sh % 8 == 0);
        p.aux_stage = mask_ok && pse_ok;
        if (const char *s = getenv("SPA_AUX_STAGE")) p.aux_stage = p.aux_stage && atoi(s) != 0;
        if (p.aux_stage) p.aux_bytes = AttnSmem::aux_bytes(p.pse != nullptr ? p.g : 0, p.bs);
    }
    int nst = e->kv_i8 ? 12 : 6;
    for (; nst >= 2; --nst)
        if (AttnSmem(nst, p.stage_bytes, e->nt_tiles, p.q_stage ? p.g : 0, p.aux_bytes).total <= budget) break;
    if (nst < 2) return fail(SPA_ERR_UNSUPPORTED, "attention working set exceeds shared memory");
    p.nstages = nst;
    p.pair_pages = nst >= 4 && (p.aux_stage || (p.mask == nullptr && p.pse == nullptr));   // (mask / pse from global memory: one page per step)
    if (const char *s = getenv("SPA_PAIR")) p.pair_pages = p.pair_pages && atoi(s) != 0;
    int attn_bytes = AttnSmem(nst, p.stage_bytes, e->nt_tiles, p.q_stage ? p.g : 0, p.aux_bytes).total;
    int sel_bytes = p.maxP * 4;
    if (p.mode == kModeFused) {
        if ((rc = plan_select_smem(&p, kSelScores | kSelSelect | kSelUnion, d.sms, budget - 1024, &sel_bytes))) return rc;
        if ((rc = make_cent_map(&e->tmC, p.cent, p.Nkv * p.C, e->q_bf16))) return rc;
    }
    p.smem_bytes = std::max(attn_bytes, sel_bytes) + 1024;
    if (p.smem_bytes > d.smem_optin) return fail(SPA_ERR_UNSUPPORTED, "shared memory need %d > %d", p.smem_bytes, d.smem_optin);
    ChainKernel kern = chain_kernel_for(e->q_bf16, e->kv_f8 ? kKvF8 : (e->kv_i8 ? kKvI8 : kKvSame), e->nt_tiles, e->opt);
    if ((rc = raise_smem_limit((const void *)kern, d.smem_optin))) return rc;
    int occ = 0;
    CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, kThreads, p.smem_bytes));
    if (occ < 1) return fail(SPA_ERR_CUDA, "chain kernel does not fit on an SM (smem %d)", p.smem_bytes);
    e->grid = d.sms;                                     // one persistent CTA per SM
    // ---- workspace
    const int HP = e->nt_tiles * 8;
    uint64_t o = 0;
    // words that must be zero when a launch starts (the launch leaves them zero again)
    const uint64_t qcap = (uint64_t)p.items * p.qstride;
    if (qcap >= (uint64_t)kMaxQueue) return fail(SPA_ERR_UNSUPPORTED, "B*Nkv*maxBlocks = %llu exceeds the page array", (unsigned long long)qcap);
    e->off_ctrl = o;   o = align_up(o + (uint64_t)Ctrl::words(p.items) * 4, 256);
    e->off_info = o;   o = align_up(o + (uint64_t)p.items * 8, 256);
    e->off_ippl = o;   o = align_up(o + (uint64_t)p.items * 4, 256);
    // chunk queue: every CTA claims up to two indices past the end
    e->off_desc = o;   o = align_up(o + ((uint64_t)p.items * kMaxChunks + 2ull * d.sms + 64) * 8, 256);
    e->off_flat = o;   o = align_up(o + qcap * 8, 256);
    e->ctrl_bytes = o;
    e->off_po = o;     o = align_up(o + (uint64_t)p.items * p.pslots * HP * kD * 4, 256);
    e->off_pml = o;    o = align_up(o + (uint64_t)p.items * p.pslots * HP * 2 * 4, 256);
    e->off_xs = o;
    if (p.mode == kModeFused) o = align_up(o + (uint64_t)e->grid * kConsumerWarps * std::max(p.C, 1) * 8, 256);
    e->ws_bytes = o;
    *ws = o;
    e->launches = 1;
    return SPA_OK;
}

static int launch_chain(void *workspace, uint64_t wsSize, spaExecutor *e, cudaStream_t stream) {
    if (e->p.items == 0) return SPA_OK;
    if (!workspace || wsSize < e->ws_bytes)
        return fail(SPA_ERR_WORKSPACE, "workspace %llu bytes < required %llu", (unsigned long long)wsSize, (unsigned long long)e->ws_bytes);
    if (reinterpret_cast<uintptr_t>(workspace) & 255) return fail(SPA_ERR_WORKSPACE, "workspace must be 256-byte aligned");
    ChainParams p = e->p;
    char *w = (char *)workspace;
    p.ctrl = (int *)(w + e->off_ctrl); p.flat = (unsigned long long *)(w + e->off_flat);
    p.desc = (unsigned long long *)(w + e->off_desc);
    p.item_info = (unsigned long long *)(w + e->off_info);
    p.item_ppl = (int *)(w + e->off_ippl);
    p.part_o = (float *)(w + e->off_po); p.part_ml = (float *)(w + e->off_pml);
    p.exact_scratch = (double *)(w + e->off_xs);
    ChainKernel kern = chain_kernel_for(e->q_bf16, e->kv_f8 ? kKvF8 : (e->kv_i8 ? kKvI8 : kKvSame), e->nt_tiles, e->opt);   // (smem limit raised once, in plan_attention)
    // queue words start at zero and the kernel leaves them at zero again, but the workspace is the
    // caller's scratch: clear it unless the caller vouches for it (spaExecutorSetPersistentWorkspace).
    // "Known clean" is only recorded for a clear that really executes: under stream capture the memset is
    // merely recorded into the graph (it then runs with every replay), so the flag stays as it was.
    if (!e->persistent_ws || e->clean_ws != workspace) {
        cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
        CUDA_TRY(cudaStreamIsCapturing(stream, &cap));
        CUDA_TRY(cudaMemsetAsync(p.ctrl, 0, e->ctrl_bytes, stream));
        if (cap == cudaStreamCaptureStatusNone) e->clean_ws = e->persistent_ws ? workspace : nullptr;
    }
    // One CTA per SM, launched cooperatively: there is no grid-wide barrier, but CTAs wait for words that other
    // CTAs publish after their own selection, so all of them must be co-resident.  The cooperative launch makes
    // the runtime guarantee that -- or refuse the launch with an error (SM partitioning, MPS limits) instead of
    // leaving resident CTAs spinning for one that can never be scheduled.  SPA_COOP=0: ordinary launch.
    static const bool coop = !(getenv("SPA_COOP") && atoi(getenv("SPA_COOP")) == 0);
    if (coop) {
        void *args[] = {(void *)&p, (void *)&e->tmK, (void *)&e->tmV, (void *)&e->tmC};
        CUDA_TRY(cudaLaunchCooperativeKernel((const void *)kern, dim3(e->grid), dim3(kThreads), args,
                                             (size_t)p.smem_bytes, stream));
    } else {
        kern<<<e->grid, kThreads, p.smem_bytes, stream>>>(p, e->tmK, e->tmV, e->tmC);
        CUDA_TRY(cudaGetLastError());
    }
    return SPA_OK;
}

int spaSparsePagedAttentionGetWorkspaceSize(
    const spaTensor *query, const spaTensor *key, const spaTensor *value, const spaTensor *pseShift,
    const spaTensor *attenMask, const spaTensor *actualSeqLengths, const spaTensor *dequantScale1,
    const spaTensor *quantScale1, const spaTensor *dequantScale2, const spaTensor *quantScale2,
    const spaTensor *quantOffset2, const spaTensor *antiquantScale, const spaTensor *antiquantOffset,
    const spaTensor *blocktable, const spaTensor *kvPaddingSize, const spaTensor *blockPosition,
    int64_t numHeads, double scaleValue, const char *inputLayout, int64_t numKeyValueHeads,
    int64_t blockSize, int64_t innerPrecise, const spaTensor *attentionOut, uint64_t *ws,
    spaExecutor **executor) {
    (void)innerPrecise;   // accepted and ignored: the intermediate type is always float (…_Us2.h:69-70)
    if (!ws || !executor) return fail(SPA_ERR_ARG, "workspaceSize/executor must not be null");
    spaExecutor *e = new (std::nothrow) spaExecutor();
    if (!e) return fail(SPA_ERR_ARG, "out of host memory");
    memset(e, 0, sizeof(*e));
    ChainParams &p = e->p;
    int rc;
    const AttnOptions opts = {pseShift, attenMask, dequantScale1, quantScale1, dequantScale2, quantScale2, quantOffset2, kvPaddingSize};
    BAIL(parse_query(query, "query", &p.B, &p.Nq, &p.q_sb, &p.q_sh, &e->q_bf16));
    p.mode = kModeAttnOnly;
    if (absent(blockPosition)) { delete e; return fail(SPA_ERR_ARG, "blockposition: required tensor is absent"); }
    if (blockPosition->dtype != SPA_I32) { delete e; return fail(SPA_ERR_DTYPE, "blockposition: expected int32"); }
    if (absent(actualSeqLengths)) { delete e; return fail(SPA_ERR_ARG, "actual_seq_lengths: required tensor is absent"); }
    if (actualSeqLengths->dtype != SPA_I64) { delete e; return fail(SPA_ERR_DTYPE, "actual_seq_lengths: expected int64"); }
    BAIL(plan_attention(e, query, key, value, opts, antiquantScale, antiquantOffset, blocktable, numHeads, scaleValue,
                        inputLayout, numKeyValueHeads, blockSize, attentionOut, ws));
    if (blockPosition->ndim != 3 || blockPosition->sizes[0] < p.B || blockPosition->strides[2] != 1 ||
        (blockPosition->sizes[1] != p.Nkv && blockPosition->sizes[1] != p.Nq)) {
        delete e; return fail(SPA_ERR_SHAPE, "blockposition: expected [B, Nkv or Nq, max_page_num]");
    }
    p.bp_in = (const int *)blockPosition->data; p.bp_heads = (int)blockPosition->sizes[1];
    p.bp_sb = blockPosition->strides[0]; p.bp_sh = blockPosition->strides[1]; p.bp_len = (int)blockPosition->sizes[2];
    const int64_t na = numel(actualSeqLengths);
    if (!(na == 1 || na >= p.B) || !contiguous(actualSeqLengths)) { delete e; return fail(SPA_ERR_SHAPE, "actual_seq_lengths: expected int64 [B] or [1]"); }
    p.actual_seq = (const long long *)actualSeqLengths->data; p.actual_n = (int)na;
    e->op = OP_ATTENTION;
    *executor = e;
    return SPA_OK;
}

int spaSparsePagedAttention(void *workspace, uint64_t wsSize, spaExecutor *e, spaStream stream) {
    if (!e || e->op != OP_ATTENTION) return fail(SPA_ERR_ARG, "bad executor");
    return launch_chain(workspace, wsSize, e, (cudaStream_t)stream);
}

int spaSparsePagedFusionAttentionGetWorkspaceSize(
    const spaTensor *query, const spaTensor *key, const spaTensor *value, const spaTensor *pseShift,
    const spaTensor *attenMask, const spaTensor *actualSeqLengths, const spaTensor *dequantScale1,
    const spaTensor *quantScale1, const spaTensor *dequantScale2, const spaTensor *quantScale2,
    const spaTensor *quantOffset2, const spaTensor *antiquantScale, const spaTensor *antiquantOffset,
    const spaTensor *blocktable, const spaTensor *kvPaddingSize, const spaTensor *l1Cent,
    const spaTensor *blockIds, const spaTensor *totalSeqLen, int64_t numHeads, double scaleValue,
    const char *inputLayout, int64_t numKeyValueHeads, int64_t blockSize, int64_t innerPrecise,
    int64_t topK, const spaTensor *blockPosition, const spaTensor *pagePositionLength,
    const spaTensor *maxPagePositionLength, const spaTensor *attentionOut, uint64_t *ws,
    spaExecutor **executor) {
    (void)innerPrecise; (void)actualSeqLengths;
    if (!ws || !executor) return fail(SPA_ERR_ARG, "workspaceSize/executor must not be null");
    spaExecutor *e = new (std::nothrow) spaExecutor();
    if (!e) return fail(SPA_ERR_ARG, "out of host memory");
    memset(e, 0, sizeof(*e));
    ChainParams &p = e->p;
    int rc;
    const AttnOptions opts = {pseShift, attenMask, dequantScale1, quantScale1, dequantScale2, quantScale2, quantOffset2, kvPaddingSize};
    p.mode = kModeFused;
    BAIL(parse_cent_select_common(e, query, l1Cent, blockIds, blocktable, totalSeqLen, topK, blockSize,
                                  blockPosition, pagePositionLength, maxPagePositionLength));
    if (numKeyValueHeads > 0 && numKeyValueHeads != p.Nkv) { delete e; return fail(SPA_ERR_SHAPE, "num_key_value_heads=%lld disagrees with l1_cent heads %d", (long long)numKeyValueHeads, p.Nkv); }
    BAIL(plan_attention(e, query, key, value, opts, antiquantScale, antiquantOffset, blocktable, numHeads, scaleValue,
                        inputLayout, p.Nkv, blockSize, attentionOut, ws));
    e->op = OP_FUSION;
    *executor = e;
    return SPA_OK;
}

int spaSparsePagedFusionAttention(void *workspace, uint64_t wsSize, spaExecutor *e, spaStream stream) {
    if (!e || e->op != OP_FUSION) return fail(SPA_ERR_ARG, "bad executor");
    return launch_chain(workspace, wsSize, e, (cudaStream_t)stream);
}

// ---------------------------------------------------------------------------------------------
// Allocation-free variant of the fused op (no counterpart in the reference):
// the workspace is a file-scope device array instead of caller memory, the plan lives on
// the stack, and nothing on this path calls cudaMalloc / cudaFree or any allocating routine.
// ---------------------------------------------------------------------------------------------
static __device__ __align__(256) unsigned char g_static_scratch[SPA_STATIC_SCRATCH_BYTES];
constexpr int kMaxDevices = 64;
static void *g_static_scratch_ptr[kMaxDevices];       // device address of g_static_scratch on device i (0 = not asked yet)
static std::mutex g_static_scratch_mu;

// In host code `g_static_scratch` is only the symbol's host-side stand-in: the device address comes from
// cudaGetSymbolAddress, once per device, with that device current.
static int static_scratch_address(void **ptr) {
    int dev = -1;
    CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices) return fail(SPA_ERR_CUDA, "device ordinal %d outside 0..%d", dev, kMaxDevices - 1);
    std::lock_guard<std::mutex> lock(g_static_scratch_mu);
    if (!g_static_scratch_ptr[dev]) {
        void *addr = nullptr;
        CUDA_TRY(cudaGetSymbolAddress(&addr, g_static_scratch));
        if (!addr) return fail(SPA_ERR_CUDA, "cudaGetSymbolAddress returned a null address for the static scratch");
        g_static_scratch_ptr[dev] = addr;
    }
    *ptr = g_static_scratch_ptr[dev];
    return SPA_OK;
}

uint64_t spaStaticScratchBytes(void) { return SPA_STATIC_SCRATCH_BYTES; }

int spaDebugHoldSMs(int numSMs, uint64_t nanoseconds, spaStream stream) {
    if (numSMs <= 0) return SPA_OK;
    DevInfo d;
    int rc = dev_info(&d);
    if (rc) return rc;
    const int smem = d.smem_optin - 4096;               // more than half an SM: one such CTA per SM
    if ((rc = raise_smem_limit((const void *)spa_hold_sm_kernel, d.smem_optin))) return rc;
    spa_hold_sm_kernel<<<numSMs, 32, smem, (cudaStream_t)stream>>>((unsigned long long)nanoseconds);
    CUDA_TRY(cudaGetLastError());
    return SPA_OK;
}

int spaSparsePagedFusionAttentionStatic(
    const spaTensor *query, const spaTensor *key, const spaTensor *value, const spaTensor *pseShift,
    const spaTensor *attenMask, const spaTensor *actualSeqLengths, const spaTensor *dequantScale1,
    const spaTensor *quantScale1, const spaTensor *dequantScale2, const spaTensor *quantScale2,
    const spaTensor *quantOffset2, const spaTensor *antiquantScale, const spaTensor *antiquantOffset,
    const spaTensor *blocktable, const spaTensor *kvPaddingSize, const spaTensor *l1Cent,
    const spaTensor *blockIds, const spaTensor *totalSeqLen, int64_t numHeads, double scaleValue,
    const char *inputLayout, int64_t numKeyValueHeads, int64_t blockSize, int64_t innerPrecise,
    int64_t topK, const spaTensor *blockPosition, const spaTensor *pagePositionLength,
    const spaTensor *maxPagePositionLength, const spaTensor *attentionOut, spaStream stream) {
    (void)innerPrecise; (void)actualSeqLengths;
    spaExecutor plan;                                   // on the stack: no host allocation either
    memset(&plan, 0, sizeof(plan));
    spaExecutor *e = &plan;
    ChainParams &p = e->p;
    int rc;
    const AttnOptions opts = {pseShift, attenMask, dequantScale1, quantScale1, dequantScale2, quantScale2, quantOffset2, kvPaddingSize};
    p.mode = kModeFused;
    if ((rc = parse_cent_select_common(e, query, l1Cent, blockIds, blocktable, totalSeqLen, topK, blockSize,
                                       blockPosition, pagePositionLength, maxPagePositionLength))) return rc;
    if (numKeyValueHeads > 0 && numKeyValueHeads != p.Nkv)
        return fail(SPA_ERR_SHAPE, "num_key_value_heads=%lld disagrees with l1_cent heads %d", (long long)numKeyValueHeads, p.Nkv);
    uint64_t need = 0;
    if ((rc = plan_attention(e, query, key, value, opts, antiquantScale, antiquantOffset, blocktable, numHeads, scaleValue,
                             inputLayout, p.Nkv, blockSize, attentionOut, &need))) return rc;
    e->op = OP_FUSION;
    if (need > SPA_STATIC_SCRATCH_BYTES)
        return fail(SPA_ERR_WORKSPACE, "this problem needs %llu bytes of scratch; the static scratch of "
                    "spaSparsePagedFusionAttentionStatic holds %llu (use the two-phase entry points with a caller workspace)",
                    (unsigned long long)need, (unsigned long long)SPA_STATIC_SCRATCH_BYTES);
    void *scratch = nullptr;
    if ((rc = static_scratch_address(&scratch))) return rc;
    // The array keeps the previous call's contents: launch_chain clears the words that must start at zero
    // (cudaMemsetAsync on the stream, persistent_ws == 0) in front of every launch; the rest is written before read.
    return launch_chain(scratch, SPA_STATIC_SCRATCH_BYTES, e, (cudaStream_t)stream);
}

}  // extern "C"
