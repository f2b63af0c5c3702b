// The persistent chain kernel (selection -> page stream -> combine in one launch).  A header, because it is
// instantiated in two translation units: spa_kernels.cu holds the plain build (OPT = 0: what every BASELINE
// configuration runs), spa_kernels_opt.cu the options build (OPT = 1: attenMask / pseShift / int8 post-quant
// compiled in) -- the options cost registers and instructions in the page loop, and the plain
// kernels sit at the register limit of a 384-thread CTA, so they are kept apart at
// compile time; the two files also compile in parallel.
#pragma once
#include "spa_attn.cuh"
#include "spa_common.cuh"
#include "spa_select.cuh"

namespace spa {

// 1 KiB-aligned start of the dynamic shared memory (TMA 128B-swizzled tiles).  Plain pointer + offset:
// going through an integer would hide the address space from the compiler, and every access below
// would become a generic LD/ST/ATOM instead of LDS/STS/ATOMS.
__device__ __forceinline__ unsigned char *aligned_smem(unsigned char *raw) {
    return raw + ((1024u - (smem_u32(raw) & 1023u)) & 1023u);
}

// The whole chain in one persistent launch (SparsePagedFusionAttention,
// sparse_paged_fusion_attention.cpp:56-76: InitCentSelect -> ProcessCentSelect -> SyncAll ->
// Init/Process attention), or, with mode == kModeAttnOnly, SparsePagedAttention.  The reference's
// SyncAll between selection and attention becomes a producer/consumer page queue: CTAs only wait
// for entries that are not published yet, so the launch is cooperative (co-residency) but has no
// grid-wide barrier.
// End of a tensor-parallel launch (last CTA, one thread).  Every CTA fenced its peer stores at system scope
// before it reported, so after one more fence the flag stores below order behind ALL output rows of this
// rank.  Then wait until every peer has flagged the same launch count: from there on the local gathered
// buffer is complete and the stream may go on to its consumer.  The count lives in the signal words, not
// in the parameters, so a captured CUDA graph replays correctly.
static __device__ __noinline__ void peer_exchange_flags(const ChainParams &p) {
    const unsigned long long t_in = globaltimer_ns();
    const unsigned e = p.sig_local[kSigEpoch] + 1u;
    p.sig_local[kSigEpoch] = e;
    asm volatile("fence.acq_rel.sys;" ::: "memory");        // one release fence, then relaxed flag stores
    for (int i = 0; i < p.n_peer; ++i)
        asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(p.sig_peer[i] + p.tp_rank), "r"(e) : "memory");
    const unsigned long long t0 = globaltimer_ns();
    *reinterpret_cast<unsigned long long *>(p.sig_local + kSigWaitNs + 2) += t0 - t_in;    // fence + flag stores alone
    for (int i = 0; i < p.n_peer; ++i) {
        for (;;) {
            unsigned v;
            asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p.sig_local + p.peer_rank[i]) : "memory");
            if ((int)(v - e) >= 0) break;
            if (globaltimer_ns() - t0 > 2000000000ull) {
                // a peer did not launch within 2 s: the gathered buffer is incomplete and the lock-step that makes
                // the double buffer safe is broken.  Fail LOUDLY: leave the marker for the host, then trap -- the
                // stream's next synchronisation returns an error instead of a consumer reading stale rows.
                p.sig_local[kSigError] = e;
                asm volatile("fence.acq_rel.sys;" ::: "memory");
                __trap();
            }
            __nanosleep(64);
        }
    }
    // statistics: total ns this rank spent in the exchange (fence + flag stores + waiting for the peers)
    *reinterpret_cast<unsigned long long *>(p.sig_local + kSigWaitNs) += globaltimer_ns() - t_in;
}

template <typename QT, int KVK, int NT, int OPT>
__global__ void __launch_bounds__(kThreads, 1)
spa_chain_kernel(const __grid_constant__ ChainParams p, const __grid_constant__ CUtensorMap tmK,
                 const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmC) {
    extern __shared__ unsigned char smem_raw[];
    unsigned char *smem = aligned_smem(smem_raw);
    __shared__ PubState pub;

    trace_mark(p, 0);
    if (threadIdx.x == 0) {
        pub.n_own = 0; pub.priv_used = 0;
        pub.req_seq[0] = pub.req_seq[1] = 0u; pub.ack_seq[0] = pub.ack_seq[1] = 0u;
    }
    // ---- phase A: selection (or union building) per (b, kv head); every item ends by publishing its
    // pages (both branches start with a __syncthreads: the publisher state above is visible)
    if (p.mode == kModeFused)
        select_phase<QT, kSelScores | kSelSelect | kSelUnion>(p, &tmC, smem, &pub);
    else
        for (int item = blockIdx.x; item < p.items; item += gridDim.x) union_from_block_position(p, item, smem, &pub);
    trace_mark(p, 1);
    // ---- phase B + C: stream own prefixes and claimed chunks; the CTA completing an item combines it
    attention_phase<QT, KVK, NT, OPT>(p, &tmK, &tmV, smem, &pub);
    trace_mark(p, 3);
    __syncthreads();
    if (threadIdx.x == 0) {
        // the last CTA to leave returns the queue counters to zero for the next launch
        if (p.n_peer) asm volatile("fence.acq_rel.sys;" ::: "memory"); else __threadfence();
        if (atomicAdd(p.ctrl + Ctrl::kFinished, 1) == (int)gridDim.x - 1) {
            p.ctrl[Ctrl::kHead] = 0; p.ctrl[Ctrl::kTail] = 0;
            p.ctrl[Ctrl::kItemsDone] = 0; p.ctrl[Ctrl::kPool] = 0; p.ctrl[Ctrl::kFinished] = 0;
            if (p.n_peer) peer_exchange_flags(p);
        }
    }
    trace_mark(p, 5);
}

typedef void (*ChainKernel)(const ChainParams, const CUtensorMap, const CUtensorMap, const CUtensorMap);
template <int OPT>
static ChainKernel chain_kernel_table(int bf16, int kvk, int nt) {       // kvk: kKvSame / kKvI8 / kKvF8
    if (!bf16) {
        if (kvk == kKvSame) return nt == 1 ? spa_chain_kernel<__half, kKvSame, 1, OPT> : spa_chain_kernel<__half, kKvSame, 2, OPT>;
        if (kvk == kKvI8) return nt == 1 ? spa_chain_kernel<__half, kKvI8, 1, OPT> : spa_chain_kernel<__half, kKvI8, 2, OPT>;
        return nt == 1 ? spa_chain_kernel<__half, kKvF8, 1, OPT> : spa_chain_kernel<__half, kKvF8, 2, OPT>;
    }
    if (kvk == kKvSame) return nt == 1 ? spa_chain_kernel<__nv_bfloat16, kKvSame, 1, OPT> : spa_chain_kernel<__nv_bfloat16, kKvSame, 2, OPT>;
    if (kvk == kKvI8) return nt == 1 ? spa_chain_kernel<__nv_bfloat16, kKvI8, 1, OPT> : spa_chain_kernel<__nv_bfloat16, kKvI8, 2, OPT>;
    return nt == 1 ? spa_chain_kernel<__nv_bfloat16, kKvF8, 1, OPT> : spa_chain_kernel<__nv_bfloat16, kKvF8, 2, OPT>;
}
// the options build lives in spa_kernels_opt.cu
ChainKernel opt_chain_kernel_for(int bf16, int kvk, int nt);

}  // namespace spa
