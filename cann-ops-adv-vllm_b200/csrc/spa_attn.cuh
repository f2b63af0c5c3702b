// Phase B (paged flash-decode over the union of the selected pages) and phase C (split combine).
//
// Work item = (b, kv head).  All g query heads of the GQA group are processed together, so every
// selected K/V page slice is fetched from HBM once per group (the compulsory traffic)
// even though each query head has its own page set: per-page head masks switch heads off.
// The flattened list of all items' union pages is cut into gridDim.x equal contiguous ranges
// (the scheme the reference authors prototyped in tests/test_block_lookup.py:22-48); a CTA
// streams its range through a TMA -> shared-memory ring and leaves one partial (O, m, l) per
// item it touched; phase C merges the partials of each item (flash-decode combine,
// sparse_paged_attention_split_Bbn2s2_Us2.h:2392-2495).
#pragma once
#include "spa_common.cuh"

namespace spa {

constexpr int kMergeHeadStride = 132;   // floats; +4 padding keeps the 4 lanes of a quad on distinct banks

struct AttnSmem {
    int stages, scratch, wml, meta, bars, prefix, total;
    __host__ __device__ static inline int align(int x, int a) { return (x + a - 1) / a * a; }
    __host__ __device__ AttnSmem(int nstages, int stage_bytes, int nt_tiles, int items) {
        int o = 0;
        stages = o;   o = align(o + nstages * stage_bytes, 1024);
        scratch = o;  o = align(o + 4 * nt_tiles * 8 * kMergeHeadStride * 4, 16);
        wml = o;      o = align(o + kConsumerWarps * nt_tiles * 8 * 2 * 4, 16);
        meta = o;     o = align(o + nstages * 4, 16);
        bars = o;     o = align(o + nstages * 2 * 8, 16);
        prefix = o;   o = align(o + (items + 2) * 4, 16);
        total = o;
    }
};

// head-dim index held by accumulator register r of m-tile mt (rows gid / gid+8 of the mma tile)
template <bool KVI8>
__device__ __forceinline__ int o_dim(int mt, int gid, int r) {
    return KVI8 ? mt * 16 + 2 * gid + (r >> 1) : mt * 16 + gid + (r >> 1) * 8;
}

__device__ __forceinline__ int upper_item(const int *prefix, int items, int pos) {
    // largest item with prefix[item] <= pos  (prefix is non-decreasing, prefix[items] = T > pos)
    int lo = 0, hi = items;          // invariant: prefix[lo] <= pos < prefix[hi]
    while (hi - lo > 1) {
        int mid = (lo + hi) >> 1;
        if (prefix[mid] <= pos) lo = mid; else hi = mid;
    }
    return lo;
}

// exclusive prefix sums of union_len into smem; prefix[items] = total
__device__ inline void build_prefix(const ChainParams &p, int *prefix) {
    const int tid = threadIdx.x, nt = blockDim.x;
    for (int i = tid; i < p.items; i += nt) prefix[i + 1] = p.union_len[i];
    if (tid == 0) prefix[0] = 0;
    __syncthreads();
    if (warp_id() == 0) {
        const int lane = lane_id();
        int base = 0;
        for (int i0 = 0; i0 < p.items; i0 += 32) {
            int i = i0 + lane;
            int v = (i < p.items) ? prefix[i + 1] : 0;
            int inc = v;
            for (int o = 1; o < 32; o <<= 1) {
                int t = __shfl_up_sync(0xffffffffu, inc, o);
                if (lane >= o) inc += t;
            }
            if (i < p.items) prefix[i + 1] = base + inc;
            base += __shfl_sync(0xffffffffu, inc, 31);
        }
    }
    __syncthreads();
}

template <typename QT, bool KVI8, int NT>
__device__ void attention_phase(const ChainParams &p, const CUtensorMap *tmK, const CUtensorMap *tmV,
                                unsigned char *smem) {
    constexpr int HP = NT * 8;
    AttnSmem L(p.nstages, p.stage_bytes, NT, p.items);
    unsigned char *stage_base = smem + L.stages;
    float *scratch = reinterpret_cast<float *>(smem + L.scratch);
    float *wml = reinterpret_cast<float *>(smem + L.wml);
    unsigned *metaS = reinterpret_cast<unsigned *>(smem + L.meta);
    uint64_t *full = reinterpret_cast<uint64_t *>(smem + L.bars);
    uint64_t *empty = full + p.nstages;
    int *prefix = reinterpret_cast<int *>(smem + L.prefix);

    const int tid = threadIdx.x, lane = lane_id(), warp = warp_id();
    const int nstages = p.nstages;
    const int G = gridDim.x, cta = blockIdx.x;

    build_prefix(p, prefix);
    const int T = prefix[p.items];
    const int start = (int)(((long long)cta * T) / G);
    const int end = (int)(((long long)(cta + 1) * T) / G);

    if (tid == 0) {
        for (int s = 0; s < nstages; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], kConsumerWarps); }
        fence_barrier_init();
        fence_proxy_async();
    }
    __syncthreads();
    if (start >= end) return;

    int stage = 0;
    uint32_t phase = 0;
    int pos = start;
    int item = upper_item(prefix, p.items, start);

    if (warp == kConsumerWarps) {
        // =============================== TMA producer ===============================
        if (lane == 0) { tma_prefetch_desc(tmK); tma_prefetch_desc(tmV); }
        while (pos < end) {
            const int seg_end = min(end, prefix[item + 1]);
            const int n2 = item % p.Nkv;
            const int *uphys = p.union_phys + (size_t)item * p.maxP - prefix[item];
            const unsigned *umeta = p.union_meta + (size_t)item * p.maxP - prefix[item];
            for (int base = pos; base < seg_end; base += 32) {
                const int idx = base + lane;
                int phys = 0; unsigned meta = 0;
                if (idx < seg_end) { phys = uphys[idx]; meta = umeta[idx]; }
                const int n = min(32, seg_end - base);
                for (int j = 0; j < n; ++j) {
                    const int ph = __shfl_sync(0xffffffffu, phys, j);
                    const unsigned mt = __shfl_sync(0xffffffffu, meta, j);
                    if (lane == 0) {
                        int c[4];
                        c[1] = (p.ax_head == 1) ? n2 : (p.ax_blk == 1 ? ph : 0);
                        c[2] = (p.ax_head == 2) ? n2 : (p.ax_blk == 2 ? ph : 0);
                        c[3] = (p.ax_head == 3) ? n2 : (p.ax_blk == 3 ? ph : 0);
#pragma unroll
                        for (int kv = 0; kv < 2; ++kv) {
                            mbar_wait(&empty[stage], phase ^ 1u);
                            metaS[stage] = mt;
                            mbar_expect_tx(&full[stage], (uint32_t)p.stage_bytes);
                            unsigned char *dst = stage_base + (size_t)stage * p.stage_bytes;
                            const CUtensorMap *tm = kv ? tmV : tmK;
                            if (KVI8) {
                                tma_load_4d(dst, tm, &full[stage], 0, c[1], c[2], c[3]);
                            } else {
                                tma_load_4d(dst, tm, &full[stage], 0, c[1], c[2], c[3]);
                                tma_load_4d(dst + p.stage_bytes / 2, tm, &full[stage], 64, c[1], c[2], c[3]);
                            }
                            if (++stage == nstages) { stage = 0; phase ^= 1u; }
                        }
                    }
                }
                __syncwarp();
            }
            pos = seg_end;
            ++item;
            while (item < p.items && prefix[item + 1] == prefix[item]) ++item;
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

    while (pos < end) {
        const int seg_end = min(end, prefix[item + 1]);
        const int b = item / p.Nkv, n2 = item % p.Nkv;
        const int tail_cnt = p.tail_count[item];

        // ---- Q fragments (B operand of S^T = K Q^T): b0 = Q[head][d0,d0+1], b1 = Q[head][d0+8,+9]
        uint32_t qf[NT][8][2];
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) {
            const int h = nt * 8 + gid;
            const unsigned short *qrow = qg + (size_t)b * p.q_sb + (size_t)(n2 * p.g + h) * p.q_sh;
#pragma unroll
            for (int ks = 0; ks < 8; ++ks) {
                // fp16/bf16 KV: k-slots of the mma map to d in natural order.  int8 KV: one
                // ldmatrix row-quad holds 4 consecutive d, so slot (2*tig+e) <-> d = 16ks+4tig+e and
                // slot (2*tig+8+e) <-> d = 16ks+4tig+2+e (same permutation on both operands).
                const int d0 = KVI8 ? ks * 16 + tig * 4 : ks * 16 + tig * 2;
                const int dstep = KVI8 ? 2 : 8;
                uint32_t v0 = 0, v1 = 0;
                if (h < p.g) {
                    v0 = *reinterpret_cast<const uint32_t *>(qrow + d0);
                    v1 = *reinterpret_cast<const uint32_t *>(qrow + d0 + dstep);
                    if (KVI8) {
                        // fold the per-channel K scale into Q (…_Us2.h:2142); the K offset adds a
                        // per-head constant to every logit and cancels in the softmax (:2146-2155)
                        const unsigned short *sk = reinterpret_cast<const unsigned short *>(p.aq_scale) +
                                                   (p.aq_per_tensor ? 0 : (size_t)n2 * kD);
                        float s0 = QTraits<QT>::to_float(sk[p.aq_per_tensor ? 0 : d0]);
                        float s1 = QTraits<QT>::to_float(sk[p.aq_per_tensor ? 0 : d0 + 1]);
                        float s8 = QTraits<QT>::to_float(sk[p.aq_per_tensor ? 0 : d0 + dstep]);
                        float s9 = QTraits<QT>::to_float(sk[p.aq_per_tensor ? 0 : d0 + dstep + 1]);
                        v0 = QTraits<QT>::pack2(QTraits<QT>::to_float(v0 & 0xffff) * s0,
                                                QTraits<QT>::to_float(v0 >> 16) * s1);
                        v1 = QTraits<QT>::pack2(QTraits<QT>::to_float(v1 & 0xffff) * s8,
                                                QTraits<QT>::to_float(v1 >> 16) * s9);
                    }
                }
                qf[nt][ks][0] = v0; qf[nt][ks][1] = v1;
            }
        }
        // ---- running state: O^T[d][head] fragments, m (log2 domain), l (thread-partial)
        float o[NT][8][4];
        float m_run[NT][2], l_run[NT][2];
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) {
#pragma unroll
            for (int mt = 0; mt < 8; ++mt)
#pragma unroll
                for (int r = 0; r < 4; ++r) o[nt][mt][r] = 0.f;
            m_run[nt][0] = m_run[nt][1] = -INFINITY;
            l_run[nt][0] = l_run[nt][1] = 0.f;
        }

        for (; pos < seg_end; ++pos) {
            // ------------------------------ K: S^T = K Q^T ------------------------------
            mbar_wait(&full[stage], phase);
            const unsigned meta = metaS[stage];
            float s[NT][4];
#pragma unroll
            for (int nt = 0; nt < NT; ++nt) s[nt][0] = s[nt][1] = s[nt][2] = s[nt][3] = 0.f;
            {
                const uint32_t kb = stage_u32 + (uint32_t)stage * p.stage_bytes;
                if (!KVI8) {
#pragma unroll
                    for (int ks = 0; ks < 8; ++ks) {
                        const int chunk = (ks & 3) * 2 + a_chk;
                        const uint32_t addr = kb + (ks >> 2) * (kBS * 128) + a_row * 128 +
                                              ((chunk ^ (a_row & 7)) << 4);
                        uint32_t a0, a1, a2, a3;
                        ldmatrix_x4(a0, a1, a2, a3, addr);
#pragma unroll
                        for (int nt = 0; nt < NT; ++nt)
                            mma_16816<QT>(s[nt], a0, a1, a2, a3, qf[nt][ks][0], qf[nt][ks][1]);
                    }
                } else {
                    // int8 page: [128 tokens][128 B]; one ldmatrix.x4 covers 16 tokens x 32 d
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const int chunk = j * 2 + a_chk;
                        const uint32_t addr = kb + a_row * 128 + ((chunk ^ (a_row & 7)) << 4);
                        uint32_t r0, r1, r2, r3;
                        ldmatrix_x4(r0, r1, r2, r3, addr);
#pragma unroll
                        for (int nt = 0; nt < NT; ++nt) {
                            mma_16816<QT>(s[nt], I8Pair<QT, 0, 1>::cvt(r0), I8Pair<QT, 0, 1>::cvt(r1),
                                          I8Pair<QT, 2, 3>::cvt(r0), I8Pair<QT, 2, 3>::cvt(r1),
                                          qf[nt][2 * j][0], qf[nt][2 * j][1]);
                            mma_16816<QT>(s[nt], I8Pair<QT, 0, 1>::cvt(r2), I8Pair<QT, 0, 1>::cvt(r3),
                                          I8Pair<QT, 2, 3>::cvt(r2), I8Pair<QT, 2, 3>::cvt(r3),
                                          qf[nt][2 * j + 1][0], qf[nt][2 * j + 1][1]);
                        }
                    }
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty[stage]);
            if (++stage == nstages) { stage = 0; phase ^= 1u; }

            // ------------------------------ online softmax ------------------------------
            uint32_t pb[NT][2];                      // P^T as B fragments (after transpose)
#pragma unroll
            for (int nt = 0; nt < NT; ++nt) {
                float pr[4];
                float alpha[2];
#pragma unroll
                for (int e = 0; e < 2; ++e) {
                    const int h = nt * 8 + tig * 2 + e;
                    const bool member = (meta >> h) & 1u;
                    const bool tail = (meta >> (16 + h)) & 1u;
                    const int ntok = member ? (tail ? tail_cnt : kBS) : 0;
                    float x0 = (t0 + gid < ntok) ? s[nt][e] * p.scale_log2 : -INFINITY;
                    float x1 = (t0 + gid + 8 < ntok) ? s[nt][2 + e] * p.scale_log2 : -INFINITY;
                    float mx = fmaxf(x0, x1);
                    mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 4));
                    mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 8));
                    mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 16));
                    const float m_new = fmaxf(m_run[nt][e], mx);
                    const float m_use = (m_new == -INFINITY) ? 0.f : m_new;
                    alpha[e] = fast_exp2(m_run[nt][e] - m_use);
                    const float p0 = fast_exp2(x0 - m_use);
                    const float p1 = fast_exp2(x1 - m_use);
                    l_run[nt][e] = l_run[nt][e] * alpha[e] + p0 + p1;
                    m_run[nt][e] = m_new;
                    pr[e] = p0; pr[2 + e] = p1;
                }
#pragma unroll
                for (int mt = 0; mt < 8; ++mt) {
                    o[nt][mt][0] *= alpha[0]; o[nt][mt][1] *= alpha[1];
                    o[nt][mt][2] *= alpha[0]; o[nt][mt][3] *= alpha[1];
                }
                // P rounded to the matmul dtype before PV (…_Us2.h:3062-3063); 8x8 transposes turn
                // the accumulator layout [tok][head] into the B-operand layout [k=tok][n=head]
                pb[nt][0] = movmatrix_trans(QTraits<QT>::pack2(pr[0], pr[1]));
                pb[nt][1] = movmatrix_trans(QTraits<QT>::pack2(pr[2], pr[3]));
            }

            // ------------------------------ V: O^T += V^T P^T ------------------------------
            mbar_wait(&full[stage], phase);
            {
                const uint32_t vb = stage_u32 + (uint32_t)stage * p.stage_bytes;
                if (!KVI8) {
#pragma unroll
                    for (int mt = 0; mt < 8; ++mt) {
                        const int chunk = (mt & 3) * 2 + v_chk;
                        const uint32_t addr = vb + (mt >> 2) * (kBS * 128) + v_row * 128 +
                                              ((chunk ^ (v_row & 7)) << 4);
                        uint32_t a0, a1, a2, a3;
                        ldmatrix_x4_trans(a0, a1, a2, a3, addr);
#pragma unroll
                        for (int nt = 0; nt < NT; ++nt)
                            mma_16816<QT>(o[nt][mt], a0, a1, a2, a3, pb[nt][0], pb[nt][1]);
                    }
                } else {
                    // int8 page, transposed 8x8 b16 loads: a register holds V[tok 2tig][d,d+1] and
                    // V[tok 2tig+1][d,d+1] with d = 16*chunk + 2*gid; m-tile row gid <-> d, row
                    // gid+8 <-> d+1 (see o_dim()).
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const int chunk = j * 2 + a_chk;
                        const uint32_t addr = vb + a_row * 128 + ((chunk ^ (a_row & 7)) << 4);
                        uint32_t r0, r1, r2, r3;
                        ldmatrix_x4_trans(r0, r1, r2, r3, addr);
#pragma unroll
                        for (int nt = 0; nt < NT; ++nt) {
                            mma_16816<QT>(o[nt][2 * j], I8Pair<QT, 0, 2>::cvt(r0), I8Pair<QT, 1, 3>::cvt(r0),
                                          I8Pair<QT, 0, 2>::cvt(r1), I8Pair<QT, 1, 3>::cvt(r1), pb[nt][0], pb[nt][1]);
                            mma_16816<QT>(o[nt][2 * j + 1], I8Pair<QT, 0, 2>::cvt(r2), I8Pair<QT, 1, 3>::cvt(r2),
                                          I8Pair<QT, 0, 2>::cvt(r3), I8Pair<QT, 1, 3>::cvt(r3), pb[nt][0], pb[nt][1]);
                        }
                    }
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty[stage]);
            if (++stage == nstages) { stage = 0; phase ^= 1u; }
        }

        // ---------------- segment flush: merge the 8 warps, write one partial ----------------
        // l: reduce the thread-partials over the 8 token rows held by different lanes
#pragma unroll
        for (int nt = 0; nt < NT; ++nt)
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                float l = l_run[nt][e];
                l += __shfl_xor_sync(0xffffffffu, l, 4);
                l += __shfl_xor_sync(0xffffffffu, l, 8);
                l += __shfl_xor_sync(0xffffffffu, l, 16);
                l_run[nt][e] = l;
                if (gid == 0) {
                    const int h = nt * 8 + tig * 2 + e;
                    wml[(warp * HP + h) * 2 + 0] = m_run[nt][e];
                    wml[(warp * HP + h) * 2 + 1] = l;
                }
            }
        named_bar_sync(1, kConsumerWarps * 32);
        float Mh[NT][2], Lh[NT][2];
#pragma unroll
        for (int nt = 0; nt < NT; ++nt)
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                const int h = nt * 8 + tig * 2 + e;
                float M = -INFINITY;
#pragma unroll
                for (int w = 0; w < kConsumerWarps; ++w) M = fmaxf(M, wml[(w * HP + h) * 2]);
                float Ls = 0.f;
#pragma unroll
                for (int w = 0; w < kConsumerWarps; ++w) {
                    const float mw = wml[(w * HP + h) * 2];
                    Ls += (mw == -INFINITY) ? 0.f : wml[(w * HP + h) * 2 + 1] * fast_exp2(mw - M);
                }
                const float sc = (m_run[nt][e] == -INFINITY) ? 0.f : fast_exp2(m_run[nt][e] - M);
#pragma unroll
                for (int mt = 0; mt < 8; ++mt) { o[nt][mt][e] *= sc; o[nt][mt][2 + e] *= sc; }
                Mh[nt][e] = M; Lh[nt][e] = Ls;
            }
        // pairwise tree over warps through 4 scratch buffers
#pragma unroll
        for (int half = kConsumerWarps / 2; half >= 1; half >>= 1) {
            if (warp >= half && warp < 2 * half) {
                float *buf = scratch + (size_t)(warp - half) * HP * kMergeHeadStride;
#pragma unroll
                for (int nt = 0; nt < NT; ++nt)
#pragma unroll
                    for (int mt = 0; mt < 8; ++mt)
#pragma unroll
                        for (int r = 0; r < 4; ++r) {
                            const int h = nt * 8 + tig * 2 + (r & 1);
                            const int d = o_dim<KVI8>(mt, gid, r);
                            buf[h * kMergeHeadStride + d] = o[nt][mt][r];
                        }
            }
            named_bar_sync(1, kConsumerWarps * 32);
            if (warp < half) {
                const float *buf = scratch + (size_t)warp * HP * kMergeHeadStride;
#pragma unroll
                for (int nt = 0; nt < NT; ++nt)
#pragma unroll
                    for (int mt = 0; mt < 8; ++mt)
#pragma unroll
                        for (int r = 0; r < 4; ++r) {
                            const int h = nt * 8 + tig * 2 + (r & 1);
                            const int d = o_dim<KVI8>(mt, gid, r);
                            o[nt][mt][r] += buf[h * kMergeHeadStride + d];
                        }
            }
            named_bar_sync(1, kConsumerWarps * 32);
        }
        if (warp == 0) {
            const size_t slot = (size_t)item + cta;
            float *po = p.part_o + slot * HP * kD;
            float *pml = p.part_ml + slot * HP * 2;
#pragma unroll
            for (int nt = 0; nt < NT; ++nt) {
#pragma unroll
                for (int mt = 0; mt < 8; ++mt)
#pragma unroll
                    for (int r = 0; r < 4; ++r) {
                        const int h = nt * 8 + tig * 2 + (r & 1);
                        const int d = o_dim<KVI8>(mt, gid, r);
                        po[h * kD + d] = o[nt][mt][r];
                    }
                if (gid == 0) {
#pragma unroll
                    for (int e = 0; e < 2; ++e) {
                        const int h = nt * 8 + tig * 2 + e;
                        pml[h * 2 + 0] = Mh[nt][e];
                        pml[h * 2 + 1] = Lh[nt][e];
                    }
                }
            }
        }
        ++item;
        while (item < p.items && prefix[item + 1] == prefix[item]) ++item;
    }
}

// ---- phase C: merge the partials of every (item, head) and write attention_out ---------------
template <typename QT, bool KVI8, int NT>
__device__ void combine_phase(const ChainParams &p, unsigned char *smem) {
    constexpr int HP = NT * 8;
    AttnSmem L(p.nstages, p.stage_bytes, NT, p.items);
    const int *prefix = reinterpret_cast<const int *>(smem + L.prefix);
    const int lane = lane_id();
    const int G = gridDim.x;
    const long long T = prefix[p.items];
    const int nwarps = blockDim.x / 32;
    const int units = p.items * p.g;
    for (int u = blockIdx.x * nwarps + warp_id(); u < units; u += G * nwarps) {
        const int item = u / p.g, h = u % p.g;
        const int b = item / p.Nkv, n2 = item % p.Nkv;
        const long long P0 = prefix[item], P1 = prefix[item + 1];
        float acc[4] = {0.f, 0.f, 0.f, 0.f};
        float Lsum = 0.f;
        if (P1 > P0) {
            const int c0 = (int)(((P0 + 1) * G + T - 1) / T) - 1;
            const int c1 = (int)((P1 * G + T - 1) / T) - 1;
            // CTAs whose page range is empty (T < gridDim.x) never wrote their slot: skip them
            float M = -INFINITY;
            for (int c = c0; c <= c1; ++c) {
                if ((long long)c * T / G == (long long)(c + 1) * T / G) continue;
                M = fmaxf(M, p.part_ml[((size_t)(item + c) * HP + h) * 2]);
            }
            for (int c = c0; c <= c1; ++c) {
                if ((long long)c * T / G == (long long)(c + 1) * T / G) continue;
                const size_t slot = (size_t)item + c;
                const float m = p.part_ml[(slot * HP + h) * 2];
                const float l = p.part_ml[(slot * HP + h) * 2 + 1];
                const float w = (m == -INFINITY) ? 0.f : fast_exp2(m - M);
                Lsum += l * w;
                const float4 v = *reinterpret_cast<const float4 *>(p.part_o + (slot * HP + h) * kD + lane * 4);
                acc[0] += v.x * w; acc[1] += v.y * w; acc[2] += v.z * w; acc[3] += v.w * w;
            }
        }
        const float inv = (Lsum > 0.f) ? 1.f / Lsum : 0.f;
        float r[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) r[e] = acc[e] * inv;
        if (KVI8 && Lsum > 0.f) {
            // V dequant applied once to the normalised sum: (O + offset_v) * scale_v (…_Us2.h:3655-3670)
            const unsigned short *sv = reinterpret_cast<const unsigned short *>(p.aq_scale) +
                                       (p.aq_per_tensor ? 1 : (size_t)(p.Nkv + n2) * kD);
            const unsigned short *ov = p.aq_offset ? reinterpret_cast<const unsigned short *>(p.aq_offset) +
                                       (p.aq_per_tensor ? 1 : (size_t)(p.Nkv + n2) * kD) : nullptr;
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const int d = p.aq_per_tensor ? 0 : lane * 4 + e;
                const float off = ov ? QTraits<QT>::to_float(ov[d]) : 0.f;
                r[e] = (r[e] + off) * QTraits<QT>::to_float(sv[d]);
            }
        }
        unsigned short *op = reinterpret_cast<unsigned short *>(p.out) + (size_t)b * p.out_sb +
                             (size_t)(n2 * p.g + h) * p.out_sh + lane * 4;
        uint2 pk;
        pk.x = QTraits<QT>::pack2(r[0], r[1]);
        pk.y = QTraits<QT>::pack2(r[2], r[3]);
        *reinterpret_cast<uint2 *>(op) = pk;
    }
}

}  // namespace spa
