// Phase B (paged flash-decode over the union of the selected pages) with the split combine folded in.
//
// Work item = (b, kv head).  All g query heads of the GQA group are processed together, so every
// selected K/V page slice is fetched from HBM once per group (the compulsory traffic)
// even though each query head has its own page set: per-page head masks switch heads off.
//
// Scheduling: phase A leaves each item's union pages at a fixed place of the page array and publishes them
// (spa_select.cuh, publish_item): a private prefix that the selecting CTA streams itself, first thing after its
// own selection -- while other CTAs may still be selecting -- and chunks of the rest in one global chunk queue
// that every CTA claims from, one atomicAdd per chunk with the next claim and the next descriptor always in
// flight.  So all SMs drain HBM until the very end and no grid-wide barrier separates selection, streaming and
// combine.  (The reference separates them with SyncAll and prototyped equal static ranges in
// tests/test_block_lookup.py:22-48.)  Pages are streamed through a TMA -> shared-memory ring; whenever the item
// changes the CTA leaves one partial (O, m, l), and the CTA that completes an item's page count merges that
// item's partials (flash-decode combine, sparse_paged_attention_split_Bbn2s2_Us2.h:2392-2495) and writes
// attention_out.
#pragma once
#include "spa_common.cuh"

namespace spa {

constexpr int kMergeHalf = 64;                          // head dims merged per round
constexpr int kMergeHeadStride = kMergeHalf + 4;        // floats; +4 padding keeps the 4 lanes of a quad on distinct banks
constexpr int kMergeWarpFloats = 8 * kMergeHeadStride;
constexpr float kMaskedLog2 = -1e12f * kLog2e;         // BOOL_ATTEN_MASK_SCALAR_VALUE (..._Us2.h:757) in the log2 domain
constexpr int kMailSlots = 4;                           // flush -> post-warp messages in flight
constexpr int kPostWarp = kConsumerWarps + 1;           // warp 8: TMA producer, warps 9..11: partial bookkeeping + combine

constexpr int kAuxMask = 128;                           // bytes reserved for a page's mask in its stage's aux block
struct AttnSmem {
    int stages, mbuf, wml, meta, bars, mail, qst, aux, total;
    __host__ __device__ static inline int align(int x, int a) { return (x + a - 1) / a * a; }
    // one of the two query staging buffers: g rows of kQStageStride halves + the kv head's K antiquant scales
    __host__ __device__ static inline int qstage_bytes(int g) { return g * kQStageStride * 2 + kD * 2; }
    // options build: per stage, the page's attenMask bytes (kAuxMask reserved) and its g pseShift rows of bs values
    __host__ __device__ static inline int aux_bytes(int g_pse, int bs) { return align(kAuxMask + g_pse * bs * 2, 16); }
    __host__ __device__ AttnSmem(int nstages, int stage_bytes, int nt_tiles, int g_staged, int aux_per_stage = 0) {
        int o = 0;
        stages = o;   o = align(o + nstages * stage_bytes, 1024);
        mbuf = o;     o = align(o + kConsumerWarps * kMergeWarpFloats * 4, 16);   // 8 heads x 64 dims per warp
        wml = o;      o = align(o + kConsumerWarps * nt_tiles * 8 * 2 * 4 + 16, 16);   // + the flush's slot word
        meta = o;     o = align(o + nstages * 16, 16);
        bars = o;     o = align(o + (nstages * 2 + kPostWarps * kMailSlots * 2) * 8, 16);
        mail = o;     o = align(o + kPostWarps * kMailSlots * 32, 16);
        qst = o;      o = align(o + 2 * qstage_bytes(g_staged), 16);      // g_staged == 0: no staging
        aux = o;      o = align(o + nstages * aux_per_stage, 16);
        total = o;
    }
};

// head-dim index held by accumulator register r of m-tile mt (rows gid / gid+8 of the mma tile)
template <int KVK>
__device__ __forceinline__ int o_dim(int mt, int gid, int r) {
    return KVK != 0 ? mt * 16 + 2 * gid + (r >> 1) : mt * 16 + gid + (r >> 1) * 8;      // 8-bit V pages: see consume_pages
}

// normalise NV consecutive head dims d0.. of one output row, V antiquant, store
template <typename QT, int KVK, int NV, int OPT>
__device__ __forceinline__ void write_out_vals(const ChainParams &p, int b, int n2, int h, int d0, const float (&acc)[NV],
                                               float Lsum) {
    const float inv = (Lsum > 0.f) ? 1.f / Lsum : 0.f;
    float r[NV];
#pragma unroll
    for (int e = 0; e < NV; ++e) r[e] = acc[e] * inv;
    if (KVK != 0 && Lsum > 0.f) {
        // V dequant applied once to the normalised sum: (O + offset_v) * scale_v (…_Us2.h:3655-3670).  int8 pages
        // entered the matmul biased by 1152 (see consume_pages); FP8 pages need no scale at all (absent = 1)
        const unsigned short *sv = (KVK == kKvI8 || p.aq_scale) ? reinterpret_cast<const unsigned short *>(p.aq_scale) +
                                   (p.aq_per_tensor ? 1 : (size_t)(p.Nkv + n2) * kD) : nullptr;
        const unsigned short *ov = p.aq_offset ? reinterpret_cast<const unsigned short *>(p.aq_offset) +
                                   (p.aq_per_tensor ? 1 : (size_t)(p.Nkv + n2) * kD) : nullptr;
#pragma unroll
        for (int e = 0; e < NV; ++e) {
            const int d = p.aq_per_tensor ? 0 : d0 + e;
            const float off = ov ? QTraits<QT>::to_float(ov[d]) : 0.f;
            const float scl = (KVK == kKvI8 || sv) ? QTraits<QT>::to_float(sv[d]) : 1.f;
            r[e] = ((KVK == 1 ? r[e] - kI8Bias : r[e]) + off) * scl;             // int8: O'/l - 1152 = O/l
        }
    }
    const size_t off = (size_t)b * p.out_sb + (size_t)(n2 * p.g + h) * p.out_sh + d0;
    if (OPT && p.pq_scale != nullptr) {
        // int8 post-quant (PostQuant, ..._Us2.h:3513-3600): q = sat_s8(rint(x * scale2 + offset2)), parameters per
        // tensor or per output channel [Nq, D], fp32 or bf16; the row is stored as bytes
        const size_t chn = (size_t)(n2 * p.g + h) * kD + d0;
        auto param = [&](const void *base, int per_chn, int e) -> float {
            const size_t i = per_chn ? chn + e : 0;
            return p.pq_bf16 ? __bfloat162float(reinterpret_cast<const __nv_bfloat16 *>(base)[i])
                             : reinterpret_cast<const float *>(base)[i];
        };
        unsigned char *ob = reinterpret_cast<unsigned char *>(p.out) + off;
#pragma unroll
        for (int e = 0; e < NV; e += 2) {
            unsigned short two = 0;
#pragma unroll
            for (int u = 0; u < 2; ++u) {
                const float x = r[e + u] * param(p.pq_scale, p.pq_scale_chn, e + u) +
                                (p.pq_offset ? param(p.pq_offset, p.pq_offset_chn, e + u) : 0.f);
                // (a row that attended to nothing is all zeros, like the rows of an item without pages)
                const int q = (Lsum > 0.f) ? max(-128, min(127, __float2int_rn(fminf(fmaxf(x, -1e9f), 1e9f)))) : 0;
                two |= (unsigned short)((unsigned)q & 0xffu) << (8 * u);
            }
            *reinterpret_cast<unsigned short *>(ob + e) = two;
        }
        return;
    }
    unsigned short *op = reinterpret_cast<unsigned short *>(p.out) + off;
    uint32_t pk[NV / 2];
#pragma unroll
    for (int e = 0; e < NV; e += 2) {
        pk[e / 2] = QTraits<QT>::pack2(r[e], r[e + 1]);
        *reinterpret_cast<uint32_t *>(op + e) = pk[e / 2];
    }
    // output all-gather fused into the epilogue: the same row goes to every peer's gathered buffer
#pragma unroll 1
    for (int i = 0; i < p.n_peer; ++i) {
        unsigned short *pp = reinterpret_cast<unsigned short *>(p.out_peer[i]) + off;
#pragma unroll
        for (int e = 0; e < NV; e += 2) *reinterpret_cast<uint32_t *>(pp + e) = pk[e / 2];
    }
}

// softmax log-sum-exp of one (b, head) row: ln(sum_i exp(s_i)) with s = scale * q.k, from the running max (log2
// domain) and the sum of 2^(s2 - max) (the reference's optional softmaxLse output, ..._Us2.h:2601-2623).
// Whole warp.  int8 KV: the kernel's logits are short of the true ones by a per-head constant -- K enters the
// matmul as k instead of k + offset_k, with q*scale_k split into two int8 digits after a power-of-two
// normalisation (see the Q fragments in attention_phase) -- which the softmax never sees but its log-sum
// does: it is rebuilt here from the same quantised operands, 4 head dims per lane.
template <typename QT, int KVK>
__device__ __forceinline__ void write_lse(const ChainParams &p, int b, int n2, int h, float m2, float lsum) {
    const int lane = lane_id();
    float shift = 0.f;
    if (KVK != 0) {
        const unsigned short *qrow = reinterpret_cast<const unsigned short *>(p.q) + (size_t)b * p.q_sb +
                                     (size_t)(n2 * p.g + h) * p.q_sh + lane * 4;
        const unsigned short *sk = (KVK == kKvI8 || p.aq_scale) ? reinterpret_cast<const unsigned short *>(p.aq_scale) + (p.aq_per_tensor ? 0 : (size_t)n2 * kD + lane * 4) : nullptr;
        const unsigned short *ok = p.aq_offset ? reinterpret_cast<const unsigned short *>(p.aq_offset) +
                                   (p.aq_per_tensor ? 0 : (size_t)n2 * kD + lane * 4) : nullptr;
        float f[4], off[4], amax = 0.f;
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            const int j = p.aq_per_tensor ? 0 : e;
            f[e] = QTraits<QT>::to_float(qrow[e]) * ((KVK == kKvI8 || sk) ? QTraits<QT>::to_float(sk[j]) : 1.f);
            off[e] = ok ? QTraits<QT>::to_float(ok[j]) : 0.f;
            amax = fmaxf(amax, fabsf(f[e]));
        }
#pragma unroll
        for (int o = 16; o; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
        const uint32_t ex = min(max((__float_as_uint(amax) >> 23) & 0xffu, 1u), 253u);
        const float up = __uint_as_float((254u - ex) << 23) * (KVK == 1 ? kQDigitRange : 1.f), down = __uint_as_float(ex << 23);
        float sum = 0.f;
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            if (KVK == 1) {
                int d1, d2;
                split_digits(f[e] * up, d1, d2);
                sum += ((float)d1 + (float)d2 * (1.f / kQDigit2)) * off[e];
            } else {
                sum += __half2float(__float2half_rn(f[e] * up)) * off[e];      // FP8: Q enters as fp16(q * scale_k * 2^n)
            }
        }
#pragma unroll
        for (int o = 16; o; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
        shift = p.scale_log2 * down * (KVK == 1 ? 1.f / kQDigitRange : 1.f) * sum;
    }
    if (lane == 0)
        p.lse[(size_t)b * p.lse_sb + (size_t)(n2 * p.g + h) * p.lse_sh] =
            (lsum > 0.f) ? (m2 + log2f(lsum) + shift) * 0.6931471805599453f : -INFINITY;
}

// ---- combine (post warp): the CTA whose partial completed the item's page count merges all partials ----
// Partial slots are fixed: slot = item * pslots + pid, pid = 0 for the owner's prefix, 1 + j for chunk j.  The
// word that counts the item's finished pages also collects one bit per partial (see post_warp_loop), so the
// last arriver knows the slots without searching.  Latency-bound (everything sits in L2): (m, l) of every
// partial and head in one round of loads, then the O rows of two heads x all partials per round trip.
template <typename QT, int KVK, int NT, int OPT>
__device__ void combine_item_warp(const ChainParams &p, int item, unsigned bits) {
    constexpr int HP = NT * 8;
    constexpr int NB = 8;                                     // partials per round of O-row loads
    const int lane = lane_id();
    const int b = item / p.Nkv, n2 = item % p.Nkv, g = p.g;
    const int n = __popc(bits);                               // <= kMaxChunks + 1 = 16
    // lane j < n holds the j-th partial: its slot and, per round of heads, its (m, l)
    int slot = -1;
    {
        unsigned rest = bits;
        for (int j = 0; j < n; ++j) {
            const int pid = __ffs(rest) - 1;
            rest &= rest - 1;
            if (lane == j) slot = item * p.pslots + pid;
        }
    }
#pragma unroll 1
    for (int h0 = 0; h0 < g; h0 += 2) {
        const int nh = min(2, g - h0);
        float2 me[2];
#pragma unroll
        for (int e = 0; e < 2; ++e)
            me[e] = (slot >= 0 && e < nh) ? __ldcg(reinterpret_cast<const float2 *>(p.part_ml + ((size_t)slot * HP + h0 + e) * 2))
                                          : make_float2(-INFINITY, 0.f);
        float acc[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
        float M[2] = {0.f, 0.f}, Ls[2] = {0.f, 0.f}, w[2] = {0.f, 0.f};
#pragma unroll 1
        for (int u0 = 0; u0 < n; u0 += NB) {
            // one round of independent L2 loads: the O rows of two heads x up to eight partials (with the (m, l)
            // loads above in the first round)
            float4 v[2][NB];
#pragma unroll
            for (int u = 0; u < NB; ++u) {
                const int sj = __shfl_sync(0xffffffffu, slot, min(u0 + u, n - 1));
#pragma unroll
                for (int e = 0; e < 2; ++e)
                    if (u0 + u < n) v[e][u] = __ldcg(reinterpret_cast<const float4 *>(p.part_o + ((size_t)sj * HP + h0 + min(e, nh - 1)) * kD) + lane);
            }
            if (u0 == 0) {
#pragma unroll
                for (int e = 0; e < 2; ++e) {
                    float cm = me[e].x;
#pragma unroll
                    for (int o = 16; o; o >>= 1) cm = fmaxf(cm, __shfl_xor_sync(0xffffffffu, cm, o));
                    M[e] = cm;
                    w[e] = (me[e].x == -INFINITY) ? 0.f : fast_exp2(me[e].x - cm);
                    float L = me[e].y * w[e];
#pragma unroll
                    for (int o = 16; o; o >>= 1) L += __shfl_xor_sync(0xffffffffu, L, o);
                    Ls[e] = L;
                }
            }
#pragma unroll
            for (int u = 0; u < NB; ++u)
#pragma unroll
                for (int e = 0; e < 2; ++e) {
                    const float we = __shfl_sync(0xffffffffu, w[e], min(u0 + u, n - 1));
                    if (u0 + u < n) {
                        acc[e][0] += v[e][u].x * we; acc[e][1] += v[e][u].y * we;
                        acc[e][2] += v[e][u].z * we; acc[e][3] += v[e][u].w * we;
                    }
                }
        }
#pragma unroll
        for (int e = 0; e < 2; ++e) {
            if (e < nh) write_out_vals<QT, KVK, 4, OPT>(p, b, n2, h0 + e, lane * 4, acc[e], Ls[e]);
            if (e < nh && p.lse != nullptr) write_lse<QT, KVK>(p, b, n2, h0 + e, M[e], Ls[e]);
        }
    }
    if (lane == 0) { p.item_info[item] = 0ull; p.ctrl[Ctrl::pages_done(item)] = 0; }
}

// ---- post warp: takes the flush messages of this CTA's consumers; publishes each partial (fence, page count +
// partial bit) and combines the items this CTA completed.  Keeps every global round trip
// of the bookkeeping off the consumers' path: under a saturated HBM each one costs microseconds.
// max_page_position_length[b, 0..7] = bs * (longest per-head page list of b) (cent_select.cpp:123-160), fused op:
// the items' publishers leave 1 + their maximum in item_ppl (non-zero once written); post warp 0 of CTA
// (b mod grid) waits for the Nkv words of b, reduces, writes, and returns the words to zero.  Nothing else in the
// launch waits for this, and the waiting warp has nothing to do until its CTA's first partial.
static __device__ __noinline__ void max_ppl_poll(const ChainParams &p) {
    const int lane = lane_id();
    for (int b = blockIdx.x; b < p.B; b += gridDim.x) {
        int m = 0;
        for (int n0 = 0; n0 < p.Nkv; n0 += 32) {
            const int n2 = n0 + lane;
            if (n2 < p.Nkv) {
                int *w = p.item_ppl + (size_t)b * p.Nkv + n2;
                int v;
                while ((v = ld_relaxed(w)) == 0) __nanosleep(100);
                *w = 0;
                m = max(m, v - 1);
            }
        }
#pragma unroll
        for (int o = 16; o; o >>= 1) m = max(m, __shfl_xor_sync(0xffffffffu, m, o));
        if (lane < kPad) p.max_ppl[(size_t)b * kPad + lane] = (long long)m * p.bs;
    }
}

template <typename QT, int KVK, int NT, int OPT>
__device__ void post_warp_loop(const ChainParams &p, const int4 *mail, uint64_t *mfull, uint64_t *mempty) {
    const int lane = lane_id();
    int mi = 0;
    uint32_t mph = 0;
    for (;;) {
        mbar_wait(&mfull[mi], mph);
        const int4 m0 = mail[mi * 2];
        __syncwarp();
        if (lane == 0) mbar_arrive(&mempty[mi]);
        if (++mi == kMailSlots) { mi = 0; mph ^= 1u; }
        const int item = m0.x, pid = m0.y, cnt = m0.z, len = m0.w;
        if (item < 0) break;
        // one word per item: finished pages in the low half, one bit per partial in the high half
        int old = 0;
        if (lane == 0) {
            __threadfence();
            old = atomicAdd(p.ctrl + Ctrl::pages_done(item), (int)((unsigned)cnt | (1u << (16 + pid))));
        }
        old = __shfl_sync(0xffffffffu, old, 0);
        if ((old & 0xffff) + cnt == len) {
            __threadfence();
            const unsigned long long tc0 = (p.dbg && lane == 0) ? globaltimer_ns() : 0ull;
            combine_item_warp<QT, KVK, NT, OPT>(p, item, ((unsigned)old >> 16) | (1u << pid));
            if (p.dbg && lane == 0 && p.trace) {
                p.trace[(size_t)blockIdx.x * kTraceSlots + 10] += 1;
                p.trace[(size_t)blockIdx.x * kTraceSlots + 11] += globaltimer_ns() - tc0;
            }
        }
    }
}

// consumers' thread 0: hand a message to the post warp
// mi counts messages: message k goes to post warp k % kPostWarps, slot (k / kPostWarps) % kMailSlots
__device__ __forceinline__ void post_message(int4 *mail, uint64_t *mfull, uint64_t *mempty, int &mi, uint32_t &mph,
                                             int4 m0, int4 m1) {
    const int w = mi % kPostWarps, r = mi / kPostWarps;
    const int sl = w * kMailSlots + (r % kMailSlots);
    const uint32_t ph = (uint32_t)(r / kMailSlots) & 1u;
    mbar_wait(&mempty[sl], ph ^ 1u);
    mail[sl * 2] = m0; mail[sl * 2 + 1] = m1;
    mbar_arrive(&mfull[sl]);
    ++mi;
    (void)mph;
}

// ---- segment flush: merge the 8 warps' running states of `item`; either finish the item (all of its
// pages were in this segment) or leave a partial and tell the post warp.  All consumers.
template <typename QT, int KVK, int NT, int OPT>
__device__ __forceinline__ void flush_segment(const ChainParams &p, int item, int pid, int cnt,
                                              unsigned long long rec, float (&o)[NT][8][4],
                                              float (&m_run)[NT][2], float (&l_run)[NT][2], float *mbuf,
                                              float *wml, int4 *mail, uint64_t *mfull, uint64_t *mempty, int &mi,
                                              uint32_t &mph) {
    constexpr int HP = NT * 8;
    constexpr int kAll = kConsumerWarps * 32;
    const int lane = lane_id(), warp = warp_id(), gid = lane >> 2, tig = lane & 3;
    const int b = item / p.Nkv, n2 = item % p.Nkv;
    // l: reduce the thread-partials over the 8 token rows held by different lanes
#pragma unroll
    for (int nt = 0; nt < NT; ++nt)
#pragma unroll
        for (int e = 0; e < 2; ++e) {
            float l = l_run[nt][e];
            l += __shfl_xor_sync(0xffffffffu, l, 4);
            l += __shfl_xor_sync(0xffffffffu, l, 8);
            l += __shfl_xor_sync(0xffffffffu, l, 16);
            if (gid == 0) {
                const int h = nt * 8 + tig * 2 + e;
                wml[(warp * HP + h) * 2 + 0] = m_run[nt][e];
                wml[(warp * HP + h) * 2 + 1] = l;
            }
        }
    while (!(rec >> 63)) rec = ld_relaxed_u64(p.item_info + item);     // published before the entries; normally valid at once
    const int len = record_len(rec);
    // uniform over the CTA.  With the optional LSE output every segment takes the partial -> combine route (the
    // combiner writes the LSE): the consumers' path carries no code for it
    const bool whole = (cnt == len) && p.lse == nullptr;
    const int slot = item * p.pslots + pid;                              // the segment's fixed partial slot
    named_bar_sync(1, kAll);
    // bring every warp's accumulators to the common max of the head
#pragma unroll
    for (int nt = 0; nt < NT; ++nt)
#pragma unroll
        for (int e = 0; e < 2; ++e) {
            const int h = nt * 8 + tig * 2 + e;
            float M = -INFINITY;
#pragma unroll
            for (int w = 0; w < kConsumerWarps; ++w) M = fmaxf(M, wml[(w * HP + h) * 2]);
            const float sc = (m_run[nt][e] == -INFINITY) ? 0.f : fast_exp2(m_run[nt][e] - M);
#pragma unroll
            for (int mt = 0; mt < 8; ++mt) { o[nt][mt][e] *= sc; o[nt][mt][2 + e] *= sc; }
        }
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
        // thread (warp, lane) owns head nt*8 + warp: its max and sum over the 8 warps
        const int h = nt * 8 + warp;
        float M = -INFINITY, Ls = 0.f;
#pragma unroll
        for (int w = 0; w < kConsumerWarps; ++w) M = fmaxf(M, wml[(w * HP + h) * 2]);
#pragma unroll
        for (int w = 0; w < kConsumerWarps; ++w) {
            const float mw = wml[(w * HP + h) * 2];
            Ls += (mw == -INFINITY) ? 0.f : wml[(w * HP + h) * 2 + 1] * fast_exp2(mw - M);
        }
#pragma unroll
        for (int half = 0; half < 2; ++half) {
            if (nt + half > 0) named_bar_sync(1, kAll);               // merge buffers are reused per round
            float *mine = mbuf + warp * kMergeWarpFloats;
#pragma unroll
            for (int mt = half * 4; mt < half * 4 + 4; ++mt)
#pragma unroll
                for (int r = 0; r < 4; ++r)
                    mine[(tig * 2 + (r & 1)) * kMergeHeadStride + o_dim<KVK>(mt, gid, r) - half * kMergeHalf] = o[nt][mt][r];
            named_bar_sync(1, kAll);
            // dims half*64 + 2*lane, +1 of head h: sum over the 8 warps
            float acc[2] = {0.f, 0.f};
#pragma unroll
            for (int w = 0; w < kConsumerWarps; ++w) {
                const float2 v = *reinterpret_cast<const float2 *>(mbuf + w * kMergeWarpFloats + warp * kMergeHeadStride + lane * 2);
                acc[0] += v.x; acc[1] += v.y;
            }
            const int d0 = half * kMergeHalf + lane * 2;
            if (whole) {
                if (h < p.g) write_out_vals<QT, KVK, 2, OPT>(p, b, n2, h, d0, acc, Ls);
            } else {
                *reinterpret_cast<float2 *>(p.part_o + ((size_t)slot * HP + h) * kD + d0) = make_float2(acc[0], acc[1]);
            }
        }
        if (!whole && lane == 0) *reinterpret_cast<float2 *>(p.part_ml + ((size_t)slot * HP + h) * 2) = make_float2(M, Ls);
    }
    named_bar_sync(1, kAll);      // every thread's partial stores precede the message; smem free for the next flush
    if (threadIdx.x == 0) {
        if (whole) p.item_info[item] = 0ull;                          // consumed: nobody else holds pages of it
        else post_message(mail, mfull, mempty, mi, mph, make_int4(item, pid, cnt, len), make_int4(0, 0, 0, 0));
    }
}

// ---- producer side ------------------------------------------------------------------------------------
// One page entry (any lane).  An item is on offer as soon as its page count is known, possibly before its
// entries are written: poll.  A consumed entry is returned to zero, so the array is clean for the next launch.
__device__ __forceinline__ unsigned long long take_entry(const ChainParams &p, int idx, unsigned long long v) {
    unsigned long long *e = p.flat + idx;
    while (!(v >> 63)) { __nanosleep(32); v = ld_relaxed_u64(e); }
    st_relaxed_u64(e, 0ull);
    return v;
}

// ---- consumers: PP (1 or 2) pages of the current item ------------------------------------------------
// Per page and warp: S^T = K Q^T for the warp's 16 tokens, online softmax in the log2 domain,
// O^T += V^T P^T.  With PP = 2 the two pages' chains are interleaved instruction by instruction, the
// running max is reduced once over both and O is rescaled once (and only when some max moved).
// ks/vs: ring stages of the K and V tiles (K full barriers already passed), vph: parities of the V stages.
template <typename QT, int KVK, int NT, int PP, int OPT>
__device__ __forceinline__ void consume_pages(const ChainParams &p, uint32_t stage_u32, const int (&ks)[PP],
                                              const int (&vs)[PP], const uint32_t (&vph)[PP],
                                              const unsigned (&metas)[PP], const int (&tails)[PP],
                                              const int (&lps)[PP], int item, const unsigned char *aux,
                                              const uint32_t (&qf)[NT][8][2], float (&o)[NT][8][4],
                                              float (&m_run)[NT][2], float (&l_run)[NT][2],
                                              const float (&hsc)[NT][2], uint64_t *full, uint64_t *empty) {
    using MT = typename MmaType<QT, KVK>::type;
    const int lane = lane_id(), warp = warp_id(), gid = lane >> 2, tig = lane & 3;
    const int t0 = warp * 16;                       // this warp's 16 tokens of every page
    // ldmatrix lane offsets (128B-swizzled tile: row r, 16B chunk c stored at chunk c ^ (r & 7))
    const int a_row = t0 + (lane & 15);             // K tile as A operand (non-transposed)
    const int a_chk = lane >> 4;
    const int v_row = t0 + (lane & 7) + ((lane >> 4) & 1) * 8;   // V tile as A^T operand (.trans)
    const int v_chk = (lane >> 3) & 1;

    // options build (attenMask / pseShift, ..._Us2.h:2862-2899): the mask bytes and pse values of this thread's two
    // tokens of each page (rows gid and gid + 8 of the warp's 16) at their LOGICAL positions page * bs + t; fetched
    // here, used after the K matmuls
    uint32_t mk[PP];                             // byte 0: row gid masked, byte 1: row gid + 8
    uint32_t psv[PP][NT][2];                     // per head: pse of (row gid, row gid + 8), query dtype
    if (OPT && !p.aux_stage) {
        // (fallback for tensors whose rows are not 16-byte aligned: read from global memory -- a round trip per page)
        const int b = item / p.Nkv, n2 = item % p.Nkv;
#pragma unroll
        for (int pp = 0; pp < PP; ++pp) {
            const size_t pos = (size_t)lps[pp] * p.bs + t0 + gid;
            const bool in0 = t0 + gid < tails[pp], in1 = t0 + gid + 8 < tails[pp];
            mk[pp] = 0u;
            if (p.mask != nullptr) {
                const unsigned char *mrow = p.mask + (size_t)b * p.mask_sb + pos;
                mk[pp] = (in0 ? (uint32_t)__ldg(mrow) : 0u) | (in1 ? (uint32_t)__ldg(mrow + 8) << 8 : 0u);
            }
#pragma unroll
            for (int nt = 0; nt < NT; ++nt)
#pragma unroll
                for (int e = 0; e < 2; ++e) {
                    const int h = nt * 8 + tig * 2 + e;
                    psv[pp][nt][e] = 0u;
                    if (p.pse != nullptr && h < p.g) {
                        const unsigned short *prow = reinterpret_cast<const unsigned short *>(p.pse) + (size_t)b * p.pse_sb +
                                                     (size_t)(n2 * p.g + h) * p.pse_sh + pos;
                        psv[pp][nt][e] = (in0 ? (uint32_t)__ldg(prow) : 0u) | (in1 ? (uint32_t)__ldg(prow + 8) << 16 : 0u);
                    }
                }
        }
    }
    // ------------------------------ K: S^T = K Q^T ------------------------------
    float s[PP][NT][4];
#pragma unroll
    for (int pp = 0; pp < PP; ++pp)
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) s[pp][nt][0] = s[pp][nt][1] = s[pp][nt][2] = s[pp][nt][3] = 0.f;
    if (KVK == kKvF8) {
        // FP8 (e4m3) page: [128 tokens][128 B], the same shared-memory image as an int8 page.  One ldmatrix.x4 covers
        // 16 tokens x 32 d; a register holds 4 consecutive d of one token, so after the (exact) conversion to fp16
        // pairs k-slot (2 tig + e) of an fp16 mma is d = 4 tig + e and slot (2 tig + 8 + e) is d = 4 tig + 2 + e --
        // the same permutation on the Q fragments (see there); two m16n8k16 mmas per 32 d.
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int chunk = j * 2 + a_chk;
            const uint32_t off = a_row * 128 + ((chunk ^ (a_row & 7)) << 4);
#pragma unroll
            for (int pp = 0; pp < PP; ++pp) {
                uint32_t r0, r1, r2, r3;
                ldmatrix_x4(r0, r1, r2, r3, stage_u32 + (uint32_t)ks[pp] * p.stage_bytes + off);
                const uint32_t k00 = cvt_e4m3x2_f16x2(r0), k01 = cvt_e4m3x2_f16x2(r1), k02 = cvt_e4m3x2_f16x2(r0 >> 16),
                               k03 = cvt_e4m3x2_f16x2(r1 >> 16), k10 = cvt_e4m3x2_f16x2(r2), k11 = cvt_e4m3x2_f16x2(r3),
                               k12 = cvt_e4m3x2_f16x2(r2 >> 16), k13 = cvt_e4m3x2_f16x2(r3 >> 16);
#pragma unroll
                for (int nt = 0; nt < NT; ++nt) {
                    mma_16816<MT>(s[pp][nt], k00, k01, k02, k03, qf[nt][2 * j][0], qf[nt][2 * j][1]);
                    mma_16816<MT>(s[pp][nt], k10, k11, k12, k13, qf[nt][2 * j + 1][0], qf[nt][2 * j + 1][1]);
                }
            }
        }
    } else if (KVK == kKvSame) {
#pragma unroll
        for (int k8 = 0; k8 < 8; ++k8) {
            const int chunk = (k8 & 3) * 2 + a_chk;
            const uint32_t off = (k8 >> 2) * (kBS * 128) + a_row * 128 + ((chunk ^ (a_row & 7)) << 4);
#pragma unroll
            for (int pp = 0; pp < PP; ++pp) {
                uint32_t a0, a1, a2, a3;
                ldmatrix_x4(a0, a1, a2, a3, stage_u32 + (uint32_t)ks[pp] * p.stage_bytes + off);
#pragma unroll
                for (int nt = 0; nt < NT; ++nt)
                    mma_16816<MT>(s[pp][nt], a0, a1, a2, a3, qf[nt][k8][0], qf[nt][k8][1]);
            }
        }
    } else {
        // int8 page: [128 tokens][128 B]; one ldmatrix.x4 covers 16 tokens x 32 d = the A fragment of an
        // int8 mma (m16n8k32) as it lies in memory: no conversion.  Q enters as two int8 digits per element
        // (see the Q fragments), the products accumulate exactly in int32, and the digits are joined once
        // per logit: 254 * S1 + S2 (|S1| < 2^21, so the sum fits).
        int s1[PP][NT][4], s2[PP][NT][4];
#pragma unroll
        for (int pp = 0; pp < PP; ++pp)
#pragma unroll
            for (int nt = 0; nt < NT; ++nt)
#pragma unroll
                for (int i = 0; i < 4; ++i) { s1[pp][nt][i] = 0; s2[pp][nt][i] = 0; }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int chunk = j * 2 + a_chk;
            const uint32_t off = a_row * 128 + ((chunk ^ (a_row & 7)) << 4);
#pragma unroll
            for (int pp = 0; pp < PP; ++pp) {
                uint32_t r0, r1, r2, r3;
                ldmatrix_x4(r0, r1, r2, r3, stage_u32 + (uint32_t)ks[pp] * p.stage_bytes + off);
#pragma unroll
                for (int nt = 0; nt < NT; ++nt) {
                    mma_16832_s8(s1[pp][nt], r0, r1, r2, r3, qf[nt][2 * j][0], qf[nt][2 * j][1]);
                    mma_16832_s8(s2[pp][nt], r0, r1, r2, r3, qf[nt][2 * j + 1][0], qf[nt][2 * j + 1][1]);
                }
            }
        }
#pragma unroll
        for (int pp = 0; pp < PP; ++pp)
#pragma unroll
            for (int nt = 0; nt < NT; ++nt)
#pragma unroll
                for (int i = 0; i < 4; ++i) s[pp][nt][i] = (float)(s1[pp][nt][i] * 254 + s2[pp][nt][i]);
    }
    if (OPT && p.aux_stage) {
        // the producer staged the page's mask bytes and pse rows next to its K image (same barrier); they are read
        // before the K stage is released
#pragma unroll
        for (int pp = 0; pp < PP; ++pp) {
            const unsigned char *ax = aux + (size_t)ks[pp] * p.aux_bytes;
            const bool in0 = t0 + gid < tails[pp], in1 = t0 + gid + 8 < tails[pp];
            mk[pp] = 0u;
            if (p.mask != nullptr) mk[pp] = (in0 ? (uint32_t)ax[t0 + gid] : 0u) | (in1 ? (uint32_t)ax[t0 + gid + 8] << 8 : 0u);
            const unsigned short *px = reinterpret_cast<const unsigned short *>(ax + kAuxMask);
#pragma unroll
            for (int nt = 0; nt < NT; ++nt)
#pragma unroll
                for (int e = 0; e < 2; ++e) {
                    const int h = nt * 8 + tig * 2 + e;
                    psv[pp][nt][e] = 0u;
                    if (p.pse != nullptr && h < p.g)
                        psv[pp][nt][e] = (in0 ? (uint32_t)px[h * p.bs + t0 + gid] : 0u) | (in1 ? (uint32_t)px[h * p.bs + t0 + gid + 8] << 16 : 0u);
                }
        }
    }
    __syncwarp();
    if (lane == 0) {
#pragma unroll
        for (int pp = 0; pp < PP; ++pp) mbar_arrive(&empty[ks[pp]]);
    }

    // ------------------------------ online softmax ------------------------------
    uint32_t pb[PP][NT][2];                      // P^T as B fragments (after transpose)
    float alpha[NT][2];
    bool moved = false;
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
        float pr[PP][4];
#pragma unroll
        for (int e = 0; e < 2; ++e) {
            const int h = nt * 8 + tig * 2 + e;
            const float sc2 = KVK != kKvSame ? hsc[nt][e] : p.scale_log2;
            float x[PP][2];
            float mx = -INFINITY;
#pragma unroll
            for (int pp = 0; pp < PP; ++pp) {
                const int ntok = ((metas[pp] >> h) & 1u) ? tails[pp] : 0;
                x[pp][0] = (t0 + gid < ntok) ? s[pp][nt][e] * sc2 : -INFINITY;
                x[pp][1] = (t0 + gid + 8 < ntok) ? s[pp][nt][2 + e] * sc2 : -INFINITY;
                if (OPT) {
                    // pse is added to the scaled logit, a masked position is then SET to -1e12 (finite, as the reference's
                    // SelectWithBytesMask does: a row whose tokens are all masked averages them)
                    x[pp][0] += QTraits<QT>::to_float(psv[pp][nt][e] & 0xffffu) * kLog2e;
                    x[pp][1] += QTraits<QT>::to_float(psv[pp][nt][e] >> 16) * kLog2e;
                    if ((mk[pp] & 0xffu) && x[pp][0] != -INFINITY) x[pp][0] = kMaskedLog2;
                    if ((mk[pp] >> 8) && x[pp][1] != -INFINITY) x[pp][1] = kMaskedLog2;
                }
                mx = fmaxf(mx, fmaxf(x[pp][0], x[pp][1]));
            }
            mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 4));
            mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 8));
            mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 16));
            const float m_new = fmaxf(m_run[nt][e], mx);
            const float m_use = (m_new == -INFINITY) ? 0.f : m_new;
            alpha[nt][e] = fast_exp2(m_run[nt][e] - m_use);
            moved |= (m_new != m_run[nt][e]);
            float psum = 0.f;
#pragma unroll
            for (int pp = 0; pp < PP; ++pp) {
                const float p0 = fast_exp2(x[pp][0] - m_use);
                const float p1 = fast_exp2(x[pp][1] - m_use);
                if (KVK != kKvI8) psum += p0 + p1;       // int8 KV: l is summed from the ROUNDED P below
                pr[pp][e] = p0; pr[pp][2 + e] = p1;
            }
            l_run[nt][e] = l_run[nt][e] * alpha[nt][e] + psum;
            m_run[nt][e] = m_new;
        }
        // P rounded to the matmul dtype before PV (…_Us2.h:3062-3063); 8x8 transposes turn
        // the accumulator layout [tok][head] into the B-operand layout [k=tok][n=head]
#pragma unroll
        for (int pp = 0; pp < PP; ++pp) {
            const uint32_t pk0 = QTraits<MT>::pack2(pr[pp][0], pr[pp][1]);
            const uint32_t pk1 = QTraits<MT>::pack2(pr[pp][2], pr[pp][3]);
            if (KVK == kKvI8) {
                // V enters the matmul as 1152 + v: O' = O + 1152 * sum(P16).  With l = sum(P16) -- the very
                // values the tensor cores multiply by -- the constant leaves as O'/l - 1152 in write_out_vals.
                // (l differs from the fp32 sum of the unrounded p by < 2^-12 relative.)
                const float2 f0 = __half22float2(*reinterpret_cast<const __half2 *>(&pk0));
                const float2 f1 = __half22float2(*reinterpret_cast<const __half2 *>(&pk1));
                l_run[nt][0] += f0.x + f1.x;
                l_run[nt][1] += f0.y + f1.y;
            }
            pb[pp][nt][0] = movmatrix_trans(pk0);
            pb[pp][nt][1] = movmatrix_trans(pk1);
        }
    }
    // rescale O only when some head's running max moved (it settles after the first pages of an item)
    if (__any_sync(0xffffffffu, moved)) {
#pragma unroll
        for (int nt = 0; nt < NT; ++nt)
#pragma unroll
            for (int mt = 0; mt < 8; ++mt) {
                o[nt][mt][0] *= alpha[nt][0]; o[nt][mt][1] *= alpha[nt][1];
                o[nt][mt][2] *= alpha[nt][0]; o[nt][mt][3] *= alpha[nt][1];
            }
    }

    // ------------------------------ V: O^T += V^T P^T ------------------------------
#pragma unroll
    for (int pp = 0; pp < PP; ++pp) mbar_wait(&full[vs[pp]], vph[pp]);
    if (KVK == kKvF8) {
        // FP8 page, transposed 8x8 b16 loads as for int8 (below): a register holds V[tok 2tig][d, d+1] and
        // V[tok 2tig+1][d, d+1]; converted pairwise (exact) and regrouped into (tok, tok+1) pairs per d
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int chunk = j * 2 + a_chk;
            const uint32_t off = a_row * 128 + ((chunk ^ (a_row & 7)) << 4);
#pragma unroll
            for (int pp = 0; pp < PP; ++pp) {
                uint32_t r[4], va[4], vb[4];
                ldmatrix_x4_trans(r[0], r[1], r[2], r[3], stage_u32 + (uint32_t)vs[pp] * p.stage_bytes + off);
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const uint32_t x = cvt_e4m3x2_f16x2(r[i]), y = cvt_e4m3x2_f16x2(r[i] >> 16);
                    va[i] = prmt(x, y, 0x5410u);          // {V[t][d],   V[t+1][d]}
                    vb[i] = prmt(x, y, 0x7632u);          // {V[t][d+1], V[t+1][d+1]}
                }
#pragma unroll
                for (int nt = 0; nt < NT; ++nt) {
                    mma_16816<MT>(o[nt][2 * j], va[0], vb[0], va[1], vb[1], pb[pp][nt][0], pb[pp][nt][1]);
                    mma_16816<MT>(o[nt][2 * j + 1], va[2], vb[2], va[3], vb[3], pb[pp][nt][0], pb[pp][nt][1]);
                }
            }
        }
    } else if (KVK == kKvSame) {
#pragma unroll
        for (int mt = 0; mt < 8; ++mt) {
            const int chunk = (mt & 3) * 2 + v_chk;
            const uint32_t off = (mt >> 2) * (kBS * 128) + v_row * 128 + ((chunk ^ (v_row & 7)) << 4);
#pragma unroll
            for (int pp = 0; pp < PP; ++pp) {
                uint32_t a0, a1, a2, a3;
                ldmatrix_x4_trans(a0, a1, a2, a3, stage_u32 + (uint32_t)vs[pp] * p.stage_bytes + off);
#pragma unroll
                for (int nt = 0; nt < NT; ++nt)
                    mma_16816<MT>(o[nt][mt], a0, a1, a2, a3, pb[pp][nt][0], pb[pp][nt][1]);
            }
        }
    } else {
        // int8 page, transposed 8x8 b16 loads: a register holds V[tok 2tig][d,d+1] and
        // V[tok 2tig+1][d,d+1] with d = 16*chunk + 2*gid; m-tile row gid <-> d, row
        // gid+8 <-> d+1 (see o_dim()).
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int chunk = j * 2 + a_chk;
            const uint32_t off = a_row * 128 + ((chunk ^ (a_row & 7)) << 4);
#pragma unroll
            for (int pp = 0; pp < PP; ++pp) {
                uint32_t r0, r1, r2, r3;
                ldmatrix_x4_trans(r0, r1, r2, r3, stage_u32 + (uint32_t)vs[pp] * p.stage_bytes + off);
                r0 ^= 0x80808080u; r1 ^= 0x80808080u; r2 ^= 0x80808080u; r3 ^= 0x80808080u;
                const uint32_t v00 = I8PairBiased<0, 2>::cvt(r0), v01 = I8PairBiased<1, 3>::cvt(r0),
                               v02 = I8PairBiased<0, 2>::cvt(r1), v03 = I8PairBiased<1, 3>::cvt(r1),
                               v10 = I8PairBiased<0, 2>::cvt(r2), v11 = I8PairBiased<1, 3>::cvt(r2),
                               v12 = I8PairBiased<0, 2>::cvt(r3), v13 = I8PairBiased<1, 3>::cvt(r3);
#pragma unroll
                for (int nt = 0; nt < NT; ++nt) {
                    mma_16816<MT>(o[nt][2 * j], v00, v01, v02, v03, pb[pp][nt][0], pb[pp][nt][1]);
                    mma_16816<MT>(o[nt][2 * j + 1], v10, v11, v12, v13, pb[pp][nt][0], pb[pp][nt][1]);
                }
            }
        }
    }
    __syncwarp();
    if (lane == 0) {
#pragma unroll
        for (int pp = 0; pp < PP; ++pp) mbar_arrive(&empty[vs[pp]]);
    }
}

template <typename QT, int KVK, int NT, int OPT>
__device__ void attention_phase(const ChainParams &p, const CUtensorMap *tmK, const CUtensorMap *tmV,
                                unsigned char *smem, PubState *pub) {
    // query staging is compiled into the int8 kernels only: the fp16 stream gains nothing from it (its pages are
    // twice as long) and the extra live state cost the fp16 kernels 1-2 us (same-box A/B, c2 / c4)
    const bool q_stage = KVK != kKvSame && p.q_stage;
    AttnSmem L(p.nstages, p.stage_bytes, NT, q_stage ? p.g : 0, (OPT && p.aux_stage) ? p.aux_bytes : 0);
    unsigned char *stage_base = smem + L.stages;
    float *mbuf = reinterpret_cast<float *>(smem + L.mbuf);
    float *wml = reinterpret_cast<float *>(smem + L.wml);
    int4 *metaS = reinterpret_cast<int4 *>(smem + L.meta);
    uint64_t *full = reinterpret_cast<uint64_t *>(smem + L.bars);
    uint64_t *empty = full + p.nstages;
    uint64_t *mfull = empty + p.nstages;
    uint64_t *mempty = mfull + kPostWarps * kMailSlots;
    int4 *mail = reinterpret_cast<int4 *>(smem + L.mail);

    const int tid = threadIdx.x, lane = lane_id(), warp = warp_id();
    const int nstages = p.nstages;

    __syncthreads();                                  // phase A of every warp group is done with the shared memory
    if (p.bs < kBS) {
        // pages shorter than the 128-row stage image (block_size 16 / 32 / 64): TMA fills the first bs rows of a
        // stage, the other rows are never written again -- zero them once, so that the matmuls meet finite values
        // there (their tokens are masked, but 0 x NaN would poison the P.V sums)
        uint4 *z = reinterpret_cast<uint4 *>(stage_base);
        for (int i = tid; i < nstages * p.stage_bytes / 16; i += blockDim.x) z[i] = make_uint4(0u, 0u, 0u, 0u);
        fence_proxy_async();
        __syncthreads();
    }
    if (tid == 0) {
        for (int s = 0; s < nstages; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], kConsumerWarps); }
        for (int s = 0; s < kPostWarps * kMailSlots; ++s) { mbar_init(&mfull[s], 1); mbar_init(&mempty[s], 1); }
        fence_barrier_init();
        fence_proxy_async();
    }
    __syncthreads();

    int stage = 0;
    uint32_t phase = 0;

    if (warp >= kPostWarp) {
        const int pw = warp - kPostWarp;
        if (pw == 0 && p.mode == kModeFused && p.max_ppl != nullptr) max_ppl_poll(p);
        post_warp_loop<QT, KVK, NT, OPT>(p, mail + pw * kMailSlots * 2, mfull + pw * kMailSlots, mempty + pw * kMailSlots);
        return;
    }
    if (warp == kConsumerWarps) {
        // =============================== TMA producer ===============================
        if (lane == 0) { tma_prefetch_desc(tmK); tma_prefetch_desc(tmV); }
        // Runs of page entries, in this order: the private prefixes of this CTA's own items (known since its own
        // selection: streaming starts while other CTAs may still be selecting), then chunks claimed from the
        // global chunk queue.  A chunk is claimed LATE -- when the current run has claim_lead pages left to
        // issue -- so that at the end of the launch no CTA sits on a chunk it has not started while others are
        // idle; the claim (atomic) and the descriptor load are issued at different page counts, so neither
        // round trip (microseconds while HBM is saturated) is waited for.
        const int nown = pub->n_own;
        int oi = 0;
        int rf = 0, rn = 0, rpid = 0;                                      // current run: flat[rf, rf + rn), its partial id
        bool rfirst = false;                                               // the next chunk handed out starts a run
        int nf = 0, claim_idx = 0;                                         // claim in flight (lane 0) / claimed index
        unsigned long long dv = 0ull;                                      // descriptor word of claim_idx as last loaded (lane 0)
        int cst = 0;                                                       // claim state: 0 none, 1 atomic in flight, 2 descriptor loading
        bool more = true;
        unsigned long long t_runs = 0, t_pages = 0;
        const int lead = p.claim_lead;
        // called per issued page with the pages the current run still has to issue
        auto advance_claim = [&](int remaining) {
            if (!more || oi < nown) return;
            if (cst == 0 && remaining <= lead) {
                if (lane == 0) nf = atomicAdd(p.ctrl + Ctrl::kHead, 1);
                cst = 1;
            } else if (cst == 1 && remaining <= (lead >> 1)) {
                claim_idx = __shfl_sync(0xffffffffu, nf, 0);
                if (lane == 0) dv = ld_relaxed_u64(p.desc + claim_idx);
                cst = 2;
            }
        };
        // next chunk (<= 32 entries of one run) -> cf, cn, last (it ends its run).  Returns 1: got one, 0: the queue
        // is exhausted, -1 (only with block == false): the next run is not known yet -- issue what you have first
        auto next_chunk = [&](int &cf, int &cn, int &cpid, bool &rstart, bool &last, bool block) -> int {
            while (rn == 0) {
                if (oi < nown) { rf = pub->own_item[oi] * p.qstride; rn = pub->own_len[oi]; rpid = 0; rfirst = true; ++oi; ++t_runs; continue; }
                if (!more) return 0;
                if (!block) return -1;
                if (cst == 0) { if (lane == 0) nf = atomicAdd(p.ctrl + Ctrl::kHead, 1); cst = 1; }
                if (cst == 1) {
                    claim_idx = __shfl_sync(0xffffffffu, nf, 0);
                    if (lane == 0) dv = ld_relaxed_u64(p.desc + claim_idx);
                    cst = 2;
                }
                unsigned long long d = 0ull;
                int verdict = 1;                                           // 1 chunk, 0 empty slot (claim again), -1 end of the queue
                if (lane == 0) {
                    d = dv;
                    while (!(d >> 63)) {
                        // not published yet -- or never: once every item has reported, a zero descriptor is an
                        // empty slot (its item has fewer chunks) and kTail * items is the end of the queue
                        const unsigned long long w = ld_acquire_u64(reinterpret_cast<const unsigned long long *>(p.ctrl + Ctrl::kTail));
                        d = ld_relaxed_u64(p.desc + claim_idx);
                        if (d >> 63) break;
                        if ((int)(w >> 32) == p.items) { verdict = (claim_idx >= (int)(unsigned)w * p.items) ? -1 : 0; break; }
                        __nanosleep(64);
                    }
                    if (d >> 63) { st_relaxed_u64(p.desc + claim_idx, 0ull); trace_put(p, 2, globaltimer_ns()); }
                }
                d = __shfl_sync(0xffffffffu, d, 0);
                verdict = __shfl_sync(0xffffffffu, verdict, 0);
                cst = 0;
                if (!(d >> 63)) {
                    if (verdict < 0) { more = false; return 0; }
                    continue;                                              // empty slot
                }
                rf = (int)((d >> 48) & (kMaxItems - 1)) * p.qstride + (int)(unsigned)d;
                rn = (int)((d >> 32) & 0xffffu);
                rpid = 1 + claim_idx / p.items;
                rfirst = true;
                ++t_runs;
            }
            cf = rf; cn = min(32, rn); cpid = rpid; rstart = rfirst;
            rfirst = false;
            rf += cn; rn -= cn;
            last = rn == 0;
            return 1;
        };
        int cf = 0, cn = 0, cpid = 0;
        bool rstart = false, last = false;
        int ok = next_chunk(cf, cn, cpid, rstart, last, true);
        // query staging (lane 0): run r's rows go to buffer r & 1.  A buffer is free again once the first stage of
        // the run that used it has been released by the consumers (they read the rows before they release it)
        int qrun = 0, pg = 0;
        int qstg0 = 0, qstg1 = 0, qpg0 = -(1 << 20), qpg1 = -(1 << 20);      // per buffer: first stage / page count of its last user
        uint32_t qph0 = 0u, qph1 = 0u;
        unsigned char *qst_base = smem + L.qst;
        const int qst_bytes = AttnSmem::qstage_bytes(p.g);
        unsigned long long ent = (ok == 1 && lane < cn) ? ld_relaxed_u64(p.flat + cf + lane) : 0ull;
        while (ok == 1) {
            // entries of the following chunk travel while this chunk is issued
            int cf2 = 0, cn2 = 0, cpid2 = 0;
            bool rstart2 = false, last2 = false;
            int ok2 = next_chunk(cf2, cn2, cpid2, rstart2, last2, false);
            unsigned long long ent2 = (ok2 == 1 && lane < cn2) ? ld_relaxed_u64(p.flat + cf2 + lane) : 0ull;
            if (lane < cn) ent = take_entry(p, cf + lane, ent);
            // options build: an entry names the LOGICAL page (mask / pse positions); every lane looks up the block of
            // its own entry -- one round trip per 32 pages, in front of the stores that wait for a free stage anyway
            int lph = 0;
            if (OPT && lane < cn)
                lph = __ldg(p.block_table + (size_t)((int)((ent >> 48) & (kMaxItems - 1)) / p.Nkv) * p.bt_stride + (int)((unsigned)ent & 0xffffffu));
            t_pages += cn;
            bool second = false;                                           // this entry closes a pair
            for (int j = 0; j < cn; ++j) {
                if (last) advance_claim(cn - j);
                const unsigned long long e = __shfl_sync(0xffffffffu, ent, j);
                const unsigned long long en = __shfl_sync(0xffffffffu, ent, min(j + 1, cn - 1));
                const int ph_opt = OPT ? __shfl_sync(0xffffffffu, lph, j) : 0;
                // two consecutive pages of one item (inside a chunk) are handed to the consumers as a pair
                const bool first = p.pair_pages && !second && j + 1 < cn && (((e ^ en) >> 48) & (kMaxItems - 1)) == 0;
                second = first;
                if (lane == 0) {
                    const unsigned hi = (unsigned)(e >> 32), lo = (unsigned)e;
                    const int ph = OPT ? ph_opt : (int)(lo & 0xffffffu), item = (int)((hi >> 16) & (kMaxItems - 1));
                    const int n2 = item % p.Nkv;
                    // stage header: head mask | tail flag << 16 | pair flag << 17 | query buffer << 18, item, partial id
                    // of the run, tokens on a tail page
                    const bool stage_q = q_stage && rstart && j == 0;
                    const int qi = qrun & 1;
                    const int4 md = make_int4((int)(hi & 0xffffu) | (int)(((hi >> 30) & 1u) << 16) | (first ? (1 << 17) : 0) | (qi << 18),
                                              item, cpid, ((int)((lo >> 24) & 0x7fu) + 1) | (OPT ? (int)(lo & 0xffffffu) << 8 : 0));
                    int c[4];
                    c[1] = (p.ax_head == 1) ? n2 : (p.ax_blk == 1 ? ph : 0);
                    c[2] = (p.ax_head == 2) ? n2 : (p.ax_blk == 2 ? ph : 0);
                    c[3] = (p.ax_head == 3) ? n2 : (p.ax_blk == 3 ? ph : 0);
#pragma unroll
                    for (int kv = 0; kv < 2; ++kv) {
                        mbar_wait(&empty[stage], phase ^ 1u);
                        metaS[stage] = md;
                        uint32_t tx = (uint32_t)p.page_tx;
                        const bool with_q = stage_q && kv == 0;
                        const bool with_s = with_q && !p.aq_per_tensor && (KVK == kKvI8 || p.aq_scale != nullptr);
                        if (with_q) {
                            // the rows of this run travel with its first K page and complete on the same barrier
                            if (2 * (pg - (qi ? qpg1 : qpg0)) <= nstages) mbar_wait(&empty[qi ? qstg1 : qstg0], qi ? qph1 : qph0);
                            tx += (uint32_t)(p.g * kD * 2) + (with_s ? (uint32_t)(kD * 2) : 0u);
                        }
                        const bool with_aux = OPT && p.aux_stage && kv == 0;
                        if (with_aux) tx += (p.mask != nullptr ? (uint32_t)p.bs : 0u) + (p.pse != nullptr ? (uint32_t)(p.g * p.bs * 2) : 0u);
                        mbar_expect_tx(&full[stage], tx);
                        unsigned char *dst = stage_base + (size_t)stage * p.stage_bytes;
                        const CUtensorMap *tm = kv ? tmV : tmK;
                        if (KVK != kKvSame) {
                            tma_load_4d(dst, tm, &full[stage], 0, c[1], c[2], c[3]);
                        } else {
                            tma_load_4d(dst, tm, &full[stage], 0, c[1], c[2], c[3]);
                            tma_load_4d(dst + p.stage_bytes / 2, tm, &full[stage], 64, c[1], c[2], c[3]);
                        }
                        if (with_aux) {
                            // the page's mask bytes and the group's pse rows at its logical position (16-byte aligned rows)
                            unsigned char *ax = smem + L.aux + (size_t)stage * p.aux_bytes;
                            const size_t b = (size_t)(item / p.Nkv), pos = (size_t)(lo & 0xffffffu) * p.bs;
                            if (p.mask != nullptr) bulk_load(ax, p.mask + b * p.mask_sb + pos, p.bs, &full[stage]);
                            if (p.pse != nullptr)
                                for (int h = 0; h < p.g; ++h)
                                    bulk_load(ax + kAuxMask + (size_t)h * p.bs * 2,
                                              reinterpret_cast<const unsigned short *>(p.pse) + b * p.pse_sb + (size_t)(n2 * p.g + h) * p.pse_sh + pos,
                                              p.bs * 2, &full[stage]);
                        }
                        if (with_q) {
                            unsigned char *qd = qst_base + (size_t)qi * qst_bytes;
                            const unsigned short *qrow = reinterpret_cast<const unsigned short *>(p.q) + (size_t)(item / p.Nkv) * p.q_sb +
                                                         (size_t)(n2 * p.g) * p.q_sh;
                            for (int h = 0; h < p.g; ++h)
                                bulk_load(qd + (size_t)h * kQStageStride * 2, qrow + (size_t)h * p.q_sh, kD * 2, &full[stage]);
                            if (with_s)
                                bulk_load(qd + (size_t)p.g * kQStageStride * 2,
                                          reinterpret_cast<const unsigned short *>(p.aq_scale) + (size_t)n2 * kD, kD * 2, &full[stage]);
                            if (qi) { qstg1 = stage; qph1 = phase; qpg1 = pg; } else { qstg0 = stage; qph0 = phase; qpg0 = pg; }
                            ++qrun;
                        }
                        if (++stage == nstages) { stage = 0; phase ^= 1u; }
                    }
                    ++pg;
                }
            }
            __syncwarp();
            if (ok2 < 0) {
                ok2 = next_chunk(cf2, cn2, cpid2, rstart2, last2, true);
                ent2 = (ok2 == 1 && lane < cn2) ? ld_relaxed_u64(p.flat + cf2 + lane) : 0ull;
            }
            ok = ok2; cf = cf2; cn = cn2; cpid = cpid2; rstart = rstart2; last = last2; ent = ent2;
        }
        if (lane == 0 && cst == 1) pub->priv_used = nf;  // a claim in flight has landed before this CTA reports "finished"
        if (lane == 0) { unsigned smid; asm volatile("mov.u32 %0, %%smid;" : "=r"(smid)); trace_put(p, 12, smid); }
        if (lane == 0) { trace_put(p, 4, t_pages); trace_put(p, 6, t_runs); trace_put(p, 7, globaltimer_ns()); }
        if (lane == 0) {                                  // end marker: a stage without data
            mbar_wait(&empty[stage], phase ^ 1u);
            metaS[stage] = make_int4(0, -2, 0, 0);       // item -2: never equals a consumer's current item
            mbar_arrive(&full[stage]);
        }
        return;
    }
    // ================================= consumers =================================
    const int gid = lane >> 2, tig = lane & 3;
    const int t0 = warp * 16;                       // this warp's 16 tokens of every page
    const unsigned short *qg = reinterpret_cast<const unsigned short *>(p.q);
    const uint32_t stage_u32 = smem_u32(stage_base);
    // ldmatrix lane offsets (128B-swizzled tile: row r, 16B chunk c stored at chunk c ^ (r & 7))
    const int a_row = t0 + (lane & 15);             // K tile as A operand (non-transposed)
    const int a_chk = lane >> 4;
    const int v_row = t0 + (lane & 7) + ((lane >> 4) & 1) * 8;   // V tile as A^T operand (.trans)
    const int v_chk = (lane >> 3) & 1;

    // running state of the current segment: O^T[d][head] fragments, m (log2 domain), l (thread-partial)
    using MT = typename MmaType<QT, KVK>::type;     // element type of the tensor-core operands
    uint32_t qf[NT][8][2];
    float o[NT][8][4];
    float m_run[NT][2], l_run[NT][2];
    float hsc[NT][2];                                // int8 KV: per-head logit scale (see the Q fragments)
    int cur = -1, cnt = 0, pid = 0;
    unsigned long long rec = 0ull;
    int mi = 0;
    uint32_t mph = 0;
    for (;;) {
        mbar_wait(&full[stage], phase);
        const int4 md = metaS[stage];
        const unsigned meta = (unsigned)md.x & 0xffffu;      // head mask of the page
        const bool tail_page = (md.x >> 16) & 1;
        const int tail_cnt = OPT ? (md.w & 0xff) : md.w;     // options build: logical page << 8 rides along
        if (md.y != cur) {
            if (p.dbg && tid == 0 && md.y < 0) { trace_put(p, 10, 0); trace_put(p, 11, 0); trace_put(p, 8, globaltimer_ns()); }
            if (cur >= 0) flush_segment<QT, KVK, NT, OPT>(p, cur, pid, cnt, rec, o, m_run, l_run, mbuf, wml, mail, mfull, mempty, mi, mph);
            if (p.dbg && tid == 0 && md.y < 0) trace_put(p, 9, globaltimer_ns());
            if (md.y < 0) break;
            cur = md.y; cnt = 0; pid = md.z;
            rec = ld_relaxed_u64(p.item_info + cur);             // first used at the flush
            const int b = cur / p.Nkv, n2 = cur % p.Nkv;
            // The item's query rows (and K scales): staged in shared memory by the producer together with this
            // run's first K page (the stage barrier just passed covers them), else read from global memory --
            // a round trip of microseconds under a saturated HBM, per segment and warp (ncu, c3: 23 % of all
            // stall samples sat on these loads before the staging).
            const unsigned short *qbase = qg + (size_t)b * p.q_sb + (size_t)(n2 * p.g) * p.q_sh;
            size_t qstride = (size_t)p.q_sh;
            const unsigned short *sk_staged = nullptr;
            if (q_stage) {
                qbase = reinterpret_cast<const unsigned short *>(smem + L.qst + (size_t)((md.x >> 18) & 1) * AttnSmem::qstage_bytes(p.g));
                qstride = kQStageStride;
                sk_staged = qbase + (size_t)p.g * kQStageStride;
            }
            // ---- Q fragments (B operand of S^T = K Q^T): b0 = Q[head][d0,d0+1], b1 = Q[head][d0+8,+9]
#pragma unroll
            for (int nt = 0; nt < NT; ++nt) {
                const int h = nt * 8 + gid;
                const unsigned short *qrow = qbase + (size_t)h * qstride;
                if (KVK == kKvSame) {
#pragma unroll
                    for (int ks = 0; ks < 8; ++ks) {
                        const int d0 = ks * 16 + tig * 2;
                        uint32_t v0 = 0, v1 = 0;
                        if (h < p.g) {
                            v0 = *reinterpret_cast<const uint32_t *>(qrow + d0);
                            v1 = *reinterpret_cast<const uint32_t *>(qrow + d0 + 8);
                        }
                        qf[nt][ks][0] = v0; qf[nt][ks][1] = v1;
                    }
                } else {
                    // int8 KV: S^T = K Q^T runs on the int8 tensor-core path with K as it lies in the page.  An
                    // mma k-slot quad (4*tig .. 4*tig+3) is 4 consecutive d, so b0 = Q[h][32j + 4tig ..+3] and
                    // b1 = Q[h][32j + 16 + 4tig ..+3].  The per-channel K scale is folded into Q (..._Us2.h:2142);
                    // the K offset adds a per-head constant to every logit and cancels in the softmax
                    // (:2146-2155).  q*scale_k is brought to [1, 2) per head by a power of two (exact), then split
                    // into two int8 digits (split_digits: 2^-16 of the head's range, finer than the fp16 rounding
                    // of the non-quantised path); the factor returns to the logits in fp32 (hsc).
                    // FP8 (e4m3) KV: the same folding, but K is converted to fp16 (exact) and Q enters as
                    // fp16(q * scale_k * 2^n) in the k-slot permutation of the converted fragments: b0 = Q[h][16ks +
                    // 4tig, +1], b1 = Q[h][16ks + 4tig + 2, +3].  Scales are optional for FP8 (absent = 1).
                    const unsigned short *sk = (KVK == kKvF8 && p.aq_scale == nullptr) ? nullptr      // (int8: always present)
                                               : (sk_staged != nullptr && !p.aq_per_tensor) ? sk_staged
                                               : reinterpret_cast<const unsigned short *>(p.aq_scale) + (p.aq_per_tensor ? 0 : (size_t)n2 * kD);
                    auto qs4 = [&](int ks, float (&f)[4]) {
                        const int d0 = ks * 16 + tig * 4;
                        const uint2 qv = *reinterpret_cast<const uint2 *>(qrow + d0);
                        uint2 sv;
                        if (KVK == kKvF8 && sk == nullptr) sv.x = sv.y = (uint32_t)QTraits<QT>::from_float(1.f) * 0x10001u;
                        else if (p.aq_per_tensor) sv.x = sv.y = (uint32_t)sk[0] * 0x10001u;
                        else sv = *reinterpret_cast<const uint2 *>(sk + d0);
                        f[0] = QTraits<QT>::to_float(qv.x & 0xffff) * QTraits<QT>::to_float(sv.x & 0xffff);
                        f[1] = QTraits<QT>::to_float(qv.x >> 16) * QTraits<QT>::to_float(sv.x >> 16);
                        f[2] = QTraits<QT>::to_float(qv.y & 0xffff) * QTraits<QT>::to_float(sv.y & 0xffff);
                        f[3] = QTraits<QT>::to_float(qv.y >> 16) * QTraits<QT>::to_float(sv.y >> 16);
                    };
                    // (q * scale_k is evaluated once and kept: the accumulators' registers are free at a segment start,
                    // and this set-up is on the 8-bit stream's critical path -- every consumer warp runs it per segment)
                    float fq[8][4];
                    float amax = 0.f;
#pragma unroll
                    for (int ks = 0; ks < 8; ++ks) {
                        fq[ks][0] = fq[ks][1] = fq[ks][2] = fq[ks][3] = 0.f;
                        if (h < p.g) qs4(ks, fq[ks]);
                        amax = fmaxf(amax, fmaxf(fmaxf(fabsf(fq[ks][0]), fabsf(fq[ks][1])), fmaxf(fabsf(fq[ks][2]), fabsf(fq[ks][3]))));
                    }
                    amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 1));
                    amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 2));
                    const uint32_t ex = min(max((__float_as_uint(amax) >> 23) & 0xffu, 1u), 253u);
                    const float up = __uint_as_float((254u - ex) << 23) * (KVK == kKvI8 ? kQDigitRange : 1.f);   // int8: amax -> [63.5, 127); FP8: [1, 2)
                    const float down = __uint_as_float(ex << 23);                          // inverse of the power of two
#pragma unroll
                    for (int ks = 0; ks < 8; ++ks) {
                        uint32_t v1 = 0, v2 = 0;
                        {
                            const float (&f)[4] = fq[ks];
                            if (KVK == kKvI8) {
#pragma unroll
                                for (int e = 0; e < 4; ++e) {
                                    int d1, d2;
                                    split_digits(f[e] * up, d1, d2);
                                    v1 |= ((uint32_t)d1 & 0xffu) << (8 * e);
                                    v2 |= ((uint32_t)d2 & 0xffu) << (8 * e);
                                }
                            } else {
                                v1 = QTraits<MT>::pack2(f[0] * up, f[1] * up);
                                v2 = QTraits<MT>::pack2(f[2] * up, f[3] * up);
                            }
                        }
                        if (KVK == kKvI8) {
                            // ks = 2j: the b0 quad of k-step j, ks = 2j + 1: its b1 quad; digit 1 -> qf[2j], digit 2 -> qf[2j+1]
                            qf[nt][(ks & ~1)][ks & 1] = v1; qf[nt][(ks & ~1) + 1][ks & 1] = v2;
                        } else {
                            qf[nt][ks][0] = v1; qf[nt][ks][1] = v2;
                        }
                    }
                    // the softmax holds head nt*8 + 2*tig + e: fetch that head's factor
#pragma unroll
                    for (int e = 0; e < 2; ++e)
                        hsc[nt][e] = p.scale_log2 * (KVK == kKvI8 ? 1.f / (kQDigitRange * kQDigit2) : 1.f) *
                                     __shfl_sync(0xffffffffu, down, (tig * 2 + e) * 4);
                }
            }
            // ---- running state: O^T[d][head] fragments, m (log2 domain), l (thread-partial)
#pragma unroll
            for (int nt = 0; nt < NT; ++nt) {
#pragma unroll
                for (int mt = 0; mt < 8; ++mt)
#pragma unroll
                    for (int r = 0; r < 4; ++r) o[nt][mt][r] = 0.f;
                m_run[nt][0] = m_run[nt][1] = -INFINITY;
                l_run[nt][0] = l_run[nt][1] = 0.f;
            }

        }
        // ---- one page, or two pages of the same item at once (independent chains for the scheduler,
        // one max reduction and one rescale of O for both)
        if ((md.x >> 17) & 1) {
            int st[4]; uint32_t ph[4];
            st[0] = stage; ph[0] = phase;
#pragma unroll
            for (int i = 1; i < 4; ++i) {
                st[i] = st[i - 1] + 1; ph[i] = ph[i - 1];
                if (st[i] == nstages) { st[i] = 0; ph[i] ^= 1u; }
            }
            mbar_wait(&full[st[2]], ph[2]);
            const int4 md1 = metaS[st[2]];
            const unsigned metas[2] = {meta, (unsigned)md1.x & 0xffffu};
            const int tails[2] = {tail_page ? tail_cnt : p.bs, ((md1.x >> 16) & 1) ? (OPT ? (md1.w & 0xff) : md1.w) : p.bs};
            const int lps[2] = {OPT ? md.w >> 8 : 0, OPT ? md1.w >> 8 : 0};
            const int ks[2] = {st[0], st[2]}, vs[2] = {st[1], st[3]};
            const uint32_t vph[2] = {ph[1], ph[3]};
            cnt += 2;
            consume_pages<QT, KVK, NT, 2, OPT>(p, stage_u32, ks, vs, vph, metas, tails, lps, cur, smem + L.aux, qf, o, m_run, l_run, hsc, full, empty);
            stage = st[3] + 1; phase = ph[3];
            if (stage == nstages) { stage = 0; phase ^= 1u; }
        } else {
            int st1 = stage + 1; uint32_t ph1 = phase;
            if (st1 == nstages) { st1 = 0; ph1 ^= 1u; }
            const unsigned metas[1] = {meta};
            const int tails[1] = {tail_page ? tail_cnt : p.bs};
            const int lps[1] = {OPT ? md.w >> 8 : 0};
            const int ks[1] = {stage}, vs[1] = {st1};
            const uint32_t vph[1] = {ph1};
            cnt += 1;
            consume_pages<QT, KVK, NT, 1, OPT>(p, stage_u32, ks, vs, vph, metas, tails, lps, cur, smem + L.aux, qf, o, m_run, l_run, hsc, full, empty);
            stage = st1 + 1; phase = ph1;
            if (stage == nstages) { stage = 0; phase ^= 1u; }
        }
    }
    if (tid == 0)                                                    // post warps: done
        for (int w = 0; w < kPostWarps; ++w) post_message(mail, mfull, mempty, mi, mph, make_int4(-1, 0, 0, 0), make_int4(0, 0, 0, 0));
}

}  // namespace spa
