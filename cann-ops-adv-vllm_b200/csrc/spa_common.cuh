// Common definitions for the sm_100a sparse decode-attention chain kernels.
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace spa {

constexpr int kD = 128;              // head dim (all reference sparse tests)
constexpr int kBS = 128;             // tokens per KV page (PAGESIZE, select_position.cpp:7)
constexpr int kConsumerWarps = 8;    // each owns 16 of the 128 tokens of a page
constexpr int kPostWarps = 3;         // bookkeeping of partials + combine, off the consumers' path
constexpr int kThreads = (kConsumerWarps + 1 + kPostWarps) * 32;   // + TMA producer warp + post warps
constexpr int kSentinel = 0x7fffffff;
constexpr int kPad = 8;              // x8 replication of the length outputs (add_custom.cpp:963,993)
constexpr int kMaxGroup = 16;        // q heads per kv head handled by one CTA (2 n8 tiles)
constexpr float kLog2e = 1.4426950408889634f;

// One launch description, shared by every kernel of the chain.  Filled by the host plan.
constexpr int kMaxPeers = 8;      // peers of one NVSwitch domain a launch stores its output rows to
constexpr int kSigEpoch = 8;      // signal word holding this rank's launch count
constexpr int kSigError = 9;      // set when a flag exchange timed out
constexpr int kSigWaitNs = 10;    // words 10-11: uint64 total ns spent in the flag exchange (statistics)

struct ChainParams {
    // ---- problem dims
    int B, Nq, Nkv, g;               // g = Nq / Nkv
    int C, k;                        // clusters, top-k
    int MP;                          // max_page_num = page_position.shape[2]
    int maxP;                        // block_table.shape[1]
    int kvPageLen;                   // block_ids.shape[1]
    int bs;                          // block size (tokens per page): 128, or 16 / 32 / 64 (a stage then holds a partly filled image)
    int items;                       // B * Nkv work items (b, kv head)
    // ---- selection inputs
    const void *q;       long long q_sb, q_sh;          // element strides (batch, head); D contiguous
    const void *cent;                                    // [Nkv, C, D] contiguous
    const int *block_ids; long long bid_stride;
    const int *block_table; long long bt_stride;
    const int *seq_len;                                  // [B] int32 (total context length)
    const int *indices_in; long long idx_sb, idx_sh;     // select_position: given indices [B,Nq,k]
    // reference deployment options, off by default as in the golden (spaExecutorSetSelectOptions)
    int sink_pages, recent_pages;                        // always-selected first / last pages
    int cap_den;                                         // > 0: at most pl / cap_den pages per head
    // ---- selection outputs
    int *indices_out;                                    // [B,Nq,k] (compute_cent) or null
    int *page_position;                                  // [B,Nq,MP] or null
    int *page_position_length;                           // [B,Nq,8] or null
    long long *max_ppl;                                  // [B,8] or null
    int *bt_gather;                                      // [B,Nq,maxP] or null
    // ---- attention inputs
    const long long *actual_seq; int actual_n;           // separate op: device int64 [B] or [1]
    const int *bp_in; int bp_heads; long long bp_sb, bp_sh; int bp_len;   // block_position input
    const void *aq_scale, *aq_offset;                    // int8 KV: [2, Nkv, D] in q dtype (or null)
    int aq_per_tensor;
    void *out; long long out_sb, out_sh;
    float *lse; long long lse_sb, lse_sh;                // optional softmax log-sum-exp [B, Nq] (natural log), or null
    // ---- KV-head tensor parallelism: every output row is also stored into the gathered-output buffers of
    // the peer GPUs (NVLink peer stores), and the launch ends with a flag exchange -- see spaExecutorSetPeerOutput
    void *out_peer[kMaxPeers];                            // the n_peer peers' images of `out` (same strides); n_peer = 0: off
    unsigned *sig_peer[kMaxPeers];                        // the peers' signal blocks; this rank writes word [tp_rank] of each
    int peer_rank[kMaxPeers];                             // rank of peer i = word of sig_local it writes
    unsigned *sig_local;                                  // own signal block: [rank] = launch count flagged by that rank,
                                                          // [kSigEpoch] = own launch count, [kSigError] = time-out marker
    int n_peer, tp_rank;
    // ---- options build only (OPT = 1 kernels, spa_chain.cuh).  Positions are LOGICAL token positions
    // page * bs + t; page entries then carry the logical page and the producer looks the block up
    int opt;
    const unsigned char *mask; long long mask_sb;        // attenMask [B, >= maxP*bs] bytes, non-zero = masked (..._Us2.h:2876-2899)
    const void *pse; long long pse_sb, pse_sh;           // pseShift [B or 1, Nq, 1, >= maxP*bs] in the query dtype; pse_sb = 0: shared (:2862-2874)
    const void *pq_scale, *pq_offset;                    // int8 post-quant of the output (:3513-3600): [1] or [Nq, D]; `out` is int8 then
    int pq_scale_chn, pq_offset_chn, pq_bf16;            // per-channel flags; parameter dtype bf16 (else fp32)
    int aux_stage, aux_bytes;                            // the producer stages mask / pse slices next to each K page (AttnSmem::aux)
    float scale_log2;                                    // scale_value * log2(e)
    int ax_row, ax_head, ax_blk;                         // TMA coordinate slots (1..3)
    int num_blocks;
    // ---- workspace
    unsigned long long *flat;                            // [items, qstride] union pages of every item, see queue_entry(); zero = not published
    unsigned long long *desc;                            // chunk queue, see chunk_desc(); zero = not published
    unsigned long long *item_info;                       // [items] see item_record(); zero = not published
    int *item_ppl;                                       // [items] 1 + max selected pages over the item's heads; zero = not published
    int qstride;                                         // entries reserved per item in `flat`
    int *ctrl;                                           // queue counters, see Ctrl (all zero between launches)
    int pslots;                                          // partial slots per item: the owner's prefix + its chunks
    float *part_o;                                       // [items * pslots, HP, D]
    float *part_ml;                                      // [items * pslots, HP, 2]
    double *exact_scratch;                               // [grid, kConsumerWarps, C] (rare: > 512 near-ties)
    // ---- launch shape
    int sel_gw;                                          // warps per selection group (4 or 8)
    int sel_group_bytes;                                 // smem bytes of one selection group
    int sel_ns;                                          // centroid-tile stages per selection group
    int nstages;                                         // pipeline stages
    int stage_bytes;                                     // ring stride: one 128-row K or V image
    int page_tx;                                         // bytes TMA delivers per K or V page: bs * D * sizeof(kv)
    int smem_bytes;
    int mode;                                            // see Mode
    int static_pct;                                      // share of a CTA's fair share it keeps of its OWN items (streamed first), percent
    int dyn_runs, dyn_min;                               // (dyn_runs unused) chunks of the rest are >= dyn_min pages
    int chunk_num, chunk_den;                            // each chunk takes chunk_num / chunk_den of what is left of the item
    int claim_lead;                                      // a chunk is claimed when the current run has this many pages left to issue
    int small_pages;                                     // a CTA keeps its own items whole while they add up to <= this many pages (0 = off)
    int pair_pages;                                      // consumers take two pages of an item per step (needs >= 4 stages)
    int q_stage;                                         // the producer stages each run's query rows (and K scales) in shared memory
    // ---- optional phase trace: [gridDim.x, kTraceSlots] globaltimer ns (null = off)
    unsigned long long *trace;
    int dbg;                                             // trace-only experiments (env SPA_DBG_REPEAT)
};
constexpr int kTraceSlots = 32;
constexpr int kQStageStride = kD + 8;                    // halves per staged query row: +16 B keeps the 8 rows of a fragment load on distinct banks

enum Mode : int {
    kModeFused = 0,        // scores + topk + select + attention
    kModeAttnOnly = 1,     // union from block_position + attention
};

// Queue control block (ints).  Every word is zero when a launch starts and is returned to zero by
// the launch itself (per-item words by the item's combiner, the globals by the last CTA to leave).
struct Ctrl {
    // kTail (low: the largest number of chunks any item was cut into) and kItemsDone (high: items reported) form
    // one aligned 64-bit word: a single load tells a consumer whether every item has reported and, if so, where
    // the chunk queue ends (kTail * items)
    enum { kHead = 0, kTail = 2, kItemsDone = 3, kPool = 4, kFinished = 5, kError = 6, kGlobals = 16 };
    __host__ __device__ static inline int pages_done(int item) { return kGlobals + item; }
    __host__ __device__ static inline int words(int items) { return kGlobals + items; }
};

// ------------------------------------------------------------------------------------------
// small device helpers
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ int ld_acquire(const int *p) {
    int v;
    asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release(int *p, int v) {
    asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
// Page-queue entry: physical block id in the low word; high word = valid | tail page | item | head mask.
// Low word: physical block id (24 bits) | (tokens on the page if it is a tail page) - 1 (7 bits).
constexpr int kMaxBlocks = 1 << 24;
__host__ __device__ __forceinline__ unsigned long long queue_entry(int phys, unsigned mask16, bool tail, int item, int tail_cnt) {
    const unsigned hi = 0x80000000u | (tail ? 0x40000000u : 0u) | ((unsigned)item << 16) | (mask16 & 0xffffu);
    const unsigned lo = ((unsigned)phys & 0xffffffu) | (((unsigned)(tail_cnt - 1) & 0x7fu) << 24);
    return ((unsigned long long)hi << 32) | lo;
}
constexpr int kMaxItems = 1 << 14;
// Per-item record: valid | tail tokens (8 bits, 1..128) | union pages (24 bits).  The item's entries sit at the
// fixed place flat[item * qstride ...), so publishing needs no reservation.
constexpr int kMaxQueue = 1 << 28;
__host__ __device__ __forceinline__ unsigned long long item_record(int len, int tail_cnt) {
    return (1ull << 63) | ((unsigned long long)(tail_cnt & 0xff) << 52) | ((unsigned long long)(len & 0xffffff) << 28);
}
__host__ __device__ __forceinline__ int record_len(unsigned long long r) { return (int)((r >> 28) & 0xffffff); }
__host__ __device__ __forceinline__ int record_tail(unsigned long long r) { return (int)((r >> 52) & 0xff); }
// Chunk of an item's pages offered to every CTA: valid | item (14 bits) | pages (16 bits) | first entry (32 bits)
constexpr int kMaxChunks = 15;        // chunks one item is cut into at most (partial ids 0..15 share a 16-bit field)
constexpr int kMaxOwn = 8;            // own items a CTA keeps a private prefix of
__host__ __device__ __forceinline__ unsigned long long chunk_desc(int item, int start, int n) {
    return (1ull << 63) | ((unsigned long long)(item & (kMaxItems - 1)) << 48) | ((unsigned long long)(n & 0xffff) << 32) |
           (unsigned long long)(unsigned)start;
}
// Per-CTA publisher state (shared memory): what this CTA keeps of its own items
struct PubState {
    int n_own, priv_used;
    int own_item[kMaxOwn], own_len[kMaxOwn];
    // selection group -> publisher (TMA producer warp) mailbox: {item, union pages, tail tokens, max pages per head}
    volatile int req[2][4];
    volatile unsigned req_seq[2], ack_seq[2];
};
__device__ __forceinline__ unsigned long long ld_relaxed_u64(const unsigned long long *p) {
    unsigned long long v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed_u64(unsigned long long *p, unsigned long long v) {
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_u64(const unsigned long long *p) {
    unsigned long long v;
    asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_u64(unsigned long long *p, unsigned long long v) {
    asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ int ld_relaxed(const int *p) {
    int v;
    asm volatile("ld.relaxed.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ uint32_t smem_u32(const void *p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ int lane_id() { return threadIdx.x & 31; }
__device__ __forceinline__ int warp_id() { return threadIdx.x >> 5; }

template <typename T> struct QTraits;
template <> struct QTraits<__half> {
    static __device__ __forceinline__ double to_double(unsigned short bits) {
        return static_cast<double>(__half2float(__ushort_as_half(bits)));
    }
    static __device__ __forceinline__ float to_float(unsigned short bits) {
        return __half2float(__ushort_as_half(bits));
    }
    static __device__ __forceinline__ unsigned short from_float(float f) {
        return __half_as_ushort(__float2half_rn(f));
    }
    static __device__ __forceinline__ uint32_t pack2(float a, float b) {
        __half2 h = __floats2half2_rn(a, b);
        return *reinterpret_cast<uint32_t *>(&h);
    }
};
template <> struct QTraits<__nv_bfloat16> {
    static __device__ __forceinline__ double to_double(unsigned short bits) {
        return static_cast<double>(__uint_as_float(static_cast<uint32_t>(bits) << 16));
    }
    static __device__ __forceinline__ float to_float(unsigned short bits) {
        return __uint_as_float(static_cast<uint32_t>(bits) << 16);
    }
    static __device__ __forceinline__ unsigned short from_float(float f) {
        return __bfloat16_as_ushort(__float2bfloat16_rn(f));
    }
    static __device__ __forceinline__ uint32_t pack2(float a, float b) {
        __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
        return *reinterpret_cast<uint32_t *>(&h);
    }
};

// KV kind of a chain-kernel instantiation: the query dtype (fp16 / bf16), int8 with antiquant, FP8 e4m3
enum : int { kKvSame = 0, kKvI8 = 1, kKvF8 = 2 };
// Element type of the attention matmuls: the query dtype, except that 8-bit KV always runs in fp16
// (exact and cheap int8 / e4m3 -> fp16 conversion; see attention_phase).
template <typename QT, int KVK> struct MmaType { typedef __half type; };
template <typename QT> struct MmaType<QT, 0> { typedef QT type; };

// ---- mbarrier / TMA (PTX ISA 8.x, sm_90+/sm_100a) -----------------------------------------
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void fence_barrier_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async() {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
// 4-D tiled TMA load, completion on an mbarrier (SASS: UTMALDG)
__device__ __forceinline__ void tma_load_4d(void *dst, const CUtensorMap *map, uint64_t *bar, int c0,
                                            int c1, int c2, int c3) {
    asm volatile(
        "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes"
        " [%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(smem_u32(dst)),
        "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
        : "memory");
}
// 2-D tiled TMA load (centroid table tiles)
__device__ __forceinline__ void tma_load_2d(void *dst, const CUtensorMap *map, uint64_t *bar, int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes"
        " [%0], [%1, {%3, %4}], [%2];" ::"r"(smem_u32(dst)),
        "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
        : "memory");
}
// 1-D bulk copy global -> shared, completion on an mbarrier (SASS: UBLKCP); 16-byte aligned, size multiple of 16
__device__ __forceinline__ void bulk_load(void *dst, const void *src, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
// 4-byte asynchronous global -> shared copy (SASS: LDGSTS): gathers that must not hold registers
__device__ __forceinline__ void cp_async_4(void *dst, const void *src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() {
    asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;" ::: "memory");
}
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap *map) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(map)) : "memory");
}

// ---- warp-level matrix ops ------------------------------------------------------------------
__device__ __forceinline__ void ldmatrix_x4(uint32_t &r0, uint32_t &r1, uint32_t &r2, uint32_t &r3,
                                            uint32_t addr) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3)
                 : "r"(addr));
}
__device__ __forceinline__ void ldmatrix_x4_trans(uint32_t &r0, uint32_t &r1, uint32_t &r2,
                                                  uint32_t &r3, uint32_t addr) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3)
                 : "r"(addr));
}
__device__ __forceinline__ uint32_t movmatrix_trans(uint32_t a) {
    uint32_t d;
    asm volatile("movmatrix.sync.aligned.m8n8.trans.b16 %0, %1;" : "=r"(d) : "r"(a));
    return d;
}
template <typename QT>
__device__ __forceinline__ void mma_16816(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2,
                                          uint32_t a3, uint32_t b0, uint32_t b1);
template <>
__device__ __forceinline__ void mma_16816<__half>(float (&c)[4], uint32_t a0, uint32_t a1,
                                                  uint32_t a2, uint32_t a3, uint32_t b0,
                                                  uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, "
        "{%0,%1,%2,%3};"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
template <>
__device__ __forceinline__ void mma_16816<__nv_bfloat16>(float (&c)[4], uint32_t a0, uint32_t a1,
                                                         uint32_t a2, uint32_t a3, uint32_t b0,
                                                         uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, "
        "{%0,%1,%2,%3};"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

// int8 x int8 -> int32 (IMMA): A = 16 rows x 32 k (row-major bytes), B = 32 k x 8 columns
__device__ __forceinline__ void mma_16832_s8(int (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3,
                                             uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k32.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
        : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3])
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
// Two-digit int8 split of a value u with |u| <= 127: u ~ d1 + d2 / 254, d1 = rint(u), d2 = rint((u - d1) * 254);
// |u - (d1 + d2/254)| <= 1/508, i.e. 2^-16 of the range (the reference's MSD scheme for int8 KV,
// sparse_paged_attention_split_Bbn2s2_Us2.h:2051-2110, restated for two digits)
constexpr float kQDigitRange = 63.5f;     // q * scale_k is brought to [1, 2) per head, then to (-127, 127)
constexpr float kQDigit2 = 254.f;
__device__ __forceinline__ void split_digits(float u, int &d1, int &d2) {
    d1 = __float2int_rn(u);
    d2 = __float2int_rn((u - (float)d1) * kQDigit2);
}

// two FP8 (e4m3) values in the low 16 bits of v -> f16x2 (exact: every e4m3 value is an fp16 value); low byte -> low half
__device__ __forceinline__ uint32_t cvt_e4m3x2_f16x2(uint32_t v) {
    uint32_t d;
    asm("{\n.reg .b16 t;\ncvt.u16.u32 t, %1;\ncvt.rn.f16x2.e4m3x2 %0, t;\n}" : "=r"(d) : "r"(v));
    return d;
}

// ---- int8 -> fp16/bf16 pairs (exact), for the fused KV antiquant ------------------------------
__device__ __forceinline__ uint32_t prmt(uint32_t a, uint32_t b, uint32_t sel) {
    uint32_t d;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(sel));
    return d;
}
// Convert two signed bytes of r (byte indices I and J) to a packed pair {lo = byte I, hi = byte J}.
template <typename QT, int I, int J> struct I8Pair;
template <int I, int J> struct I8Pair<__half, I, J> {
    // 0x6400 | (b + 128) is the half 1024 + (b + 128); subtracting 1152 leaves b exactly
    static __device__ __forceinline__ uint32_t cvt(uint32_t r) {
        const uint32_t x = r ^ 0x80808080u;
        uint32_t h = prmt(x, 0x64646464u, 0x4040u | I | (J << 8));
        const uint32_t magic = 0x64806480u;
        __half2 v = __hsub2(*reinterpret_cast<__half2 *>(&h), *reinterpret_cast<const __half2 *>(&magic));
        return *reinterpret_cast<uint32_t *>(&v);
    }
};
// Biased variant for the attention matmuls (fp16 only): 1152 + b, i.e. the same PRMT without the subtraction.
// The constant never has to be taken out element by element: in Q.K^T it adds 1152 * sum_d q_d to every
// logit of a head, which the softmax cancels; in P.V it adds 1152 * sum_t P_t, removed once per output
// row as O/l - 1152 because l is accumulated from the same rounded P (see consume_pages / write_out_vals).
constexpr float kI8Bias = 1152.f;
template <int I, int J> struct I8PairBiased {
    static __device__ __forceinline__ uint32_t cvt(uint32_t x /* bytes already xor 0x80 */) {
        return prmt(x, 0x64646464u, 0x4040u | I | (J << 8));
    }
};
template <int I, int J> struct I8Pair<__nv_bfloat16, I, J> {
    // 0x4B000000 | (b + 128) is the float 2^23 + (b + 128); |b| <= 128 is exact in bf16
    static __device__ __forceinline__ uint32_t cvt(uint32_t r) {
        const uint32_t x = r ^ 0x80808080u;
        const float fa = __uint_as_float(prmt(x, 0x4B000000u, 0x7440u | I)) - 8388736.f;
        const float fb = __uint_as_float(prmt(x, 0x4B000000u, 0x7440u | J)) - 8388736.f;
        return prmt(__float_as_uint(fa), __float_as_uint(fb), 0x7632u);
    }
};

__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

__device__ __forceinline__ unsigned long long globaltimer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
__device__ __forceinline__ void trace_mark(const ChainParams &p, int slot) {
    if (p.trace != nullptr && threadIdx.x == 0) p.trace[(size_t)blockIdx.x * kTraceSlots + slot] = globaltimer_ns();
}

// any single thread: raw value into a slot of this CTA's trace row
__device__ __forceinline__ void trace_put(const ChainParams &p, int slot, unsigned long long v) {
    if (p.trace != nullptr) p.trace[(size_t)blockIdx.x * kTraceSlots + slot] = v;
}

// fine-grained probe: SM clock of thread 0 (slots 16..31)
__device__ __forceinline__ void trace_clk(const ChainParams &p, int slot) {
    if (p.trace != nullptr && threadIdx.x == 0) {
        long long c;
        asm volatile("mov.u64 %0, %%clock64;" : "=l"(c));
        p.trace[(size_t)blockIdx.x * kTraceSlots + slot] = (unsigned long long)c;
    }
}

__device__ __forceinline__ float fast_exp2(float x) {
    float y;
    asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// order-preserving map float -> uint32 (larger float => larger key)
__device__ __forceinline__ uint32_t float_key(float f) {
    uint32_t u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

// (score desc, index asc) strict total order: does (sa, ia) come before (sb, ib)?
__device__ __forceinline__ bool precedes(double sa, int ia, double sb, int ib) {
    return (sa > sb) || (sa == sb && ia < ib);
}

}  // namespace spa
