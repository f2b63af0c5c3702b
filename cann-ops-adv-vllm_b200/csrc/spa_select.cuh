// Phase A of the chain: centroid scores, TopK, page selection, union-list construction.
//
// Work item = (b, kv head) with all g query heads of that GQA group, so the centroid table of the
// kv head is read once per item instead of once per query head (the reference re-streams it per
// (b, q head): compute_cent.cpp:171-174, cent_select.cpp:228-236).  An item is processed by a
// GROUP of 4 or 8 warps (two 4-warp groups work on different items concurrently when g <= 4 and
// there are more items than CTAs); groups synchronise with their own named barrier.
//
// Scores and TopK -- how bit-exact selection is obtained from tensor-core arithmetic:
//   1. approximate scores a[h][c] = q[h] . cent[c] on the tensor cores (mma.sync m16n8k16, fp16/bf16
//      operands, fp32 accumulate; M = 16 clusters, N = the group's heads, K = head dim);
//   2. |a - e| <= eps[h] for the exact score e, with eps = 2^-16 * ||q_h||_1 * max|cent| (the
//      products are exact in fp32; the bound covers 128 accumulation roundings with a 2^9 margin);
//   3. one 1024-bin histogram of order-preserving keys locates the bin holding the k-th largest a;
//      elements more than 2*eps above the bin are certainly selected, elements more than 2*eps
//      below it certainly not; only the few in between ("candidates") get their EXACT score
//      (float64 sum of exact products, order-independent) and are ranked by (score desc, index asc).
// The result equals the exact-arithmetic TopK of the oracle (oracle/sparse_chain.py:cent_scores,
// topk_indices), which is the tie-free centre of every fp32 evaluation order of the reference
// (compute_cent.cpp:277-343 VMLA + pair-add + BlockReduceSum; tests/utils.py:117 torch.matmul).
#pragma once
#include "spa_common.cuh"

namespace spa {

constexpr int kSelBins = 1024;            // histogram bins per head (power of two, multiple of 32)
constexpr int kXCap = kSelBins / 2;       // exact scores per warp kept in smem (aliases the histogram)
constexpr int kPG = 8;                    // logical pages per thread whose cluster id is held in registers
constexpr int kScoreTiles = 2;            // 16-cluster tiles loaded per batch by one warp

// flags
enum : int { kSelScores = 1, kSelWriteIndices = 2, kSelSelect = 4, kSelUnion = 8, kSelGather = 16 };

// ---- a group of warps working on one item ----------------------------------------------------
struct Grp {
    int gw;      // warps in the group (4 or 8)
    int gt;      // threads in the group
    int id;      // group index within the CTA
    int warp;    // warp index within the group
    int tid;     // thread index within the group
    int bar;     // named barrier id
};
__device__ __forceinline__ Grp make_group(int gw) {
    Grp G;
    G.gw = gw; G.gt = gw * 32;
    G.id = warp_id() / gw; G.warp = warp_id() % gw;
    G.tid = G.warp * 32 + lane_id();
    G.bar = 2 + G.id;
    return G;
}
__device__ __forceinline__ void grp_sync(const Grp &G) { named_bar_sync(G.bar, G.gt); }

// ---- shared-memory carve-up of ONE group (host + device) ---------------------------------------
struct SelSmem {
    int scs, bw;                         // score row stride (floats), bitmap words per head
    int sc, wreg, cand, sel, bitmap, pm, phys, srt, misc, total;
    __host__ __device__ static inline int al(int x) { return (x + 15) & ~15; }
    __host__ __device__ SelSmem(int g, int C, int k, int maxP, int gw, int flags) {
        const bool scores = (flags & kSelScores) != 0;
        scs = ((C + 3) & ~3) + 4;        // +4: the 4 lanes of a quad (heads 2t) hit distinct banks
        bw = (C + 31) / 32;
        int o = 0;
        sc = o;      o = al(o + (scores ? g * scs * 4 : 0));
        wreg = o;    o = al(o + (scores ? gw * kSelBins * 4 : 0));
        cand = o;    o = al(o + (scores ? gw * C * 2 : 0));
        sel = o;     o = al(o + ((scores && (flags & kSelWriteIndices)) ? gw * k * 4 : 0));
        bitmap = o;  o = al(o + (scores ? g * bw * 4 : 0));
        pm = o;      o = al(o + maxP * 4);
        phys = o;    o = al(o + maxP * 4);
        srt = o;     o = al(o + (scores ? 0 : g * k * 4));      // given indices (select_position)
        misc = o;    o = al(o + 16 * 4);
        total = (o + 127) & ~127;
    }
};

__device__ __forceinline__ float key_float(uint32_t key) {      // inverse of float_key()
    return __uint_as_float((key & 0x80000000u) ? (key & 0x7fffffffu) : ~key);
}

template <typename QT> __device__ __forceinline__ uint32_t absmax2(uint32_t m, uint32_t v);
template <> __device__ __forceinline__ uint32_t absmax2<__half>(uint32_t m, uint32_t v) {
    __half2 r = __hmax2(*reinterpret_cast<__half2 *>(&m), __habs2(*reinterpret_cast<__half2 *>(&v)));
    return *reinterpret_cast<uint32_t *>(&r);
}
template <> __device__ __forceinline__ uint32_t absmax2<__nv_bfloat16>(uint32_t m, uint32_t v) {
    __nv_bfloat162 r = __hmax2(*reinterpret_cast<__nv_bfloat162 *>(&m), __habs2(*reinterpret_cast<__nv_bfloat162 *>(&v)));
    return *reinterpret_cast<uint32_t *>(&r);
}

// ---- approximate scores on the tensor cores ----------------------------------------------------
// S^T[c][h] = cent[c][:] . q[h][:]  : A = 16 clusters x 16 d (row major), B = 16 d x 8 heads.
// The contraction index is order-free, so each thread of a quad owns a contiguous 64-byte quarter
// of a centroid row (4 x 128-bit loads, a quad covers the whole 256-byte row): k-slot (2t+e) of
// k-step s is d = 32t + 4s + e and slot (2t+8+e) is d = 32t + 4s + 2 + e, for A and B alike.
template <typename QT>
__device__ __forceinline__ float score_tiles(const ChainParams &p, int b, int n2, float *sc, int scs,
                                             const Grp &G) {
    const int lane = lane_id(), gid = lane >> 2, tig = lane & 3;
    const int g = p.g, C = p.C;
    const int ntl = (g + 7) >> 3;                      // head tiles (1 or 2)
    const unsigned short *q = reinterpret_cast<const unsigned short *>(p.q);
    uint32_t qf[2][8][2];
#pragma unroll
    for (int nt = 0; nt < 2; ++nt) {
        const int h = nt * 8 + gid;
        if (nt < ntl && h < g) {
            const uint2 *qp = reinterpret_cast<const uint2 *>(q + (size_t)b * p.q_sb + (size_t)(n2 * g + h) * p.q_sh + tig * 32);
#pragma unroll
            for (int s = 0; s < 8; ++s) { uint2 v = __ldg(qp + s); qf[nt][s][0] = v.x; qf[nt][s][1] = v.y; }
        } else {
#pragma unroll
            for (int s = 0; s < 8; ++s) { qf[nt][s][0] = 0u; qf[nt][s][1] = 0u; }
        }
    }
    const uint4 *cent = reinterpret_cast<const uint4 *>(p.cent) + (size_t)n2 * C * (kD / 8);
    const int ntiles = (C + 15) >> 4;
    uint32_t mx = 0u;
    for (int t0 = G.warp; t0 < ntiles; t0 += G.gw * kScoreTiles) {
        uint4 a[kScoreTiles][8];
#pragma unroll
        for (int u = 0; u < kScoreTiles; ++u) {
            const int mt = t0 + u * G.gw;
            const int r0 = mt * 16 + gid, r1 = r0 + 8;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                a[u][j] = (mt < ntiles && r0 < C) ? __ldg(cent + (size_t)r0 * 16 + tig * 4 + j) : make_uint4(0, 0, 0, 0);
                a[u][4 + j] = (mt < ntiles && r1 < C) ? __ldg(cent + (size_t)r1 * 16 + tig * 4 + j) : make_uint4(0, 0, 0, 0);
            }
        }
#pragma unroll
        for (int u = 0; u < kScoreTiles; ++u) {
            const int mt = t0 + u * G.gw;
            if (mt >= ntiles) break;
            float acc[2][4];
#pragma unroll
            for (int nt = 0; nt < 2; ++nt) acc[nt][0] = acc[nt][1] = acc[nt][2] = acc[nt][3] = 0.f;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const uint4 v0 = a[u][j], v1 = a[u][4 + j];
                mx = absmax2<QT>(mx, v0.x); mx = absmax2<QT>(mx, v0.y); mx = absmax2<QT>(mx, v0.z); mx = absmax2<QT>(mx, v0.w);
                mx = absmax2<QT>(mx, v1.x); mx = absmax2<QT>(mx, v1.y); mx = absmax2<QT>(mx, v1.z); mx = absmax2<QT>(mx, v1.w);
#pragma unroll
                for (int nt = 0; nt < 2; ++nt) {
                    if (nt < ntl) {
                        mma_16816<QT>(acc[nt], v0.x, v1.x, v0.y, v1.y, qf[nt][2 * j][0], qf[nt][2 * j][1]);
                        mma_16816<QT>(acc[nt], v0.z, v1.z, v0.w, v1.w, qf[nt][2 * j + 1][0], qf[nt][2 * j + 1][1]);
                    }
                }
            }
            const int r0 = mt * 16 + gid, r1 = r0 + 8;
#pragma unroll
            for (int nt = 0; nt < 2; ++nt) {
                const int h0 = nt * 8 + tig * 2;
                if (nt < ntl && h0 < g) {
                    if (r0 < C) sc[h0 * scs + r0] = acc[nt][0];
                    if (r1 < C) sc[h0 * scs + r1] = acc[nt][2];
                    if (h0 + 1 < g) {
                        if (r0 < C) sc[(h0 + 1) * scs + r0] = acc[nt][1];
                        if (r1 < C) sc[(h0 + 1) * scs + r1] = acc[nt][3];
                    }
                }
            }
        }
    }
    float m = fmaxf(QTraits<QT>::to_float((unsigned short)(mx & 0xffffu)), QTraits<QT>::to_float((unsigned short)(mx >> 16)));
#pragma unroll
    for (int o = 16; o; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    return m;
}

// exact partial dot of this lane's 4 head-dim elements (d = 4*lane .. 4*lane+3)
template <typename QT>
__device__ __forceinline__ double exact_partial(uint2 v, const double (&qd)[4]) {
    double s = QTraits<QT>::to_double((unsigned short)(v.x & 0xffffu)) * qd[0];
    s = fma(QTraits<QT>::to_double((unsigned short)(v.x >> 16)), qd[1], s);
    s = fma(QTraits<QT>::to_double((unsigned short)(v.y & 0xffffu)), qd[2], s);
    s = fma(QTraits<QT>::to_double((unsigned short)(v.y >> 16)), qd[3], s);
    return s;
}
__device__ __forceinline__ double warp_sum_double(double s) {
#pragma unroll
    for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    return s;
}

// exact scores of list[0..n) -> xs[0..n)   (one warp; rows come from L2, 4 in flight)
template <typename QT>
__device__ __forceinline__ void exact_scores(const uint2 *cent_rows, const double (&qd)[4], const int *list32,
                                             const unsigned short *list16, int n, double *xs) {
    const int lane = lane_id();
    for (int i0 = 0; i0 < n; i0 += 4) {
        uint2 v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int i = min(i0 + u, n - 1);
            const int c = list32 ? list32[i] : (int)list16[i];
            v[u] = __ldg(cent_rows + (size_t)c * (kD / 4) + lane);
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const double s = warp_sum_double(exact_partial<QT>(v[u], qd));
            if (lane == 0 && i0 + u < n) xs[i0 + u] = s;
        }
    }
    __syncwarp();
}

// ---- TopK of one head by one warp ------------------------------------------------------------------
// row[0..C): approximate scores.  Writes bitmap_h[0..bw) (membership of the k selected clusters) and,
// if sel != nullptr, sel[0..k) (unordered).  wreg: kSelBins words of warp-private smem.
template <typename QT>
__device__ void topk_head(const ChainParams &p, const unsigned short *qrow, const uint2 *cent_rows,
                          const float *row, float maxabs, unsigned *wreg, unsigned short *cand, int *sel,
                          unsigned *bitmap_h, double *gscratch, double (&qd)[4]) {
    const int lane = lane_id();
    const int C = p.C, k = p.k;
    const unsigned lt = (1u << lane) - 1u;
    // q_h as doubles (4 per lane) and its L1 norm
    float l1 = 0.f;
    {
        const uint2 qv = *reinterpret_cast<const uint2 *>(qrow + 4 * lane);
        const unsigned short e[4] = {(unsigned short)(qv.x & 0xffffu), (unsigned short)(qv.x >> 16),
                                     (unsigned short)(qv.y & 0xffffu), (unsigned short)(qv.y >> 16)};
#pragma unroll
        for (int i = 0; i < 4; ++i) { qd[i] = QTraits<QT>::to_double(e[i]); l1 += fabsf(QTraits<QT>::to_float(e[i])); }
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) l1 += __shfl_xor_sync(0xffffffffu, l1, o);
    const float band = ldexpf(l1 * maxabs, -15) * 1.0625f;        // 2 * eps, eps = 2^-16 ||q||_1 max|cent|

    // key range
    uint32_t lo = 0xffffffffu, hi = 0u;
    for (int c = lane; c < C; c += 32) {
        const uint32_t key = float_key(row[c]);
        lo = min(lo, key); hi = max(hi, key);
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) {
        lo = min(lo, __shfl_xor_sync(0xffffffffu, lo, o));
        hi = max(hi, __shfl_xor_sync(0xffffffffu, hi, o));
    }
    const int bits = 32 - __clz(hi - lo);                         // hi - lo < 2^bits
    const int shift = max(0, bits - 10);                          // 10 = log2(kSelBins)
    for (int i = lane; i < kSelBins; i += 32) wreg[i] = 0u;
    __syncwarp();
    for (int c = lane; c < C; c += 32) atomicAdd(&wreg[(float_key(row[c]) - lo) >> shift], 1u);
    __syncwarp();
    // bin holding the k-th largest: lane covers bins [32 lane, 32 lane + 32); suffix sums from the top
    int bstar, above;
    {
        const uint4 *hv = reinterpret_cast<const uint4 *>(wreg + lane * 32);
        int s = 0;
#pragma unroll
        for (int i = 0; i < 8; ++i) { const uint4 v = hv[i]; s += v.x + v.y + v.z + v.w; }
        int suf = s;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_down_sync(0xffffffffu, suf, o);
            if (lane + o < 32) suf += t;
        }
        const unsigned bal = __ballot_sync(0xffffffffu, suf >= k);
        const int ls = 31 - __clz(bal);                           // highest lane whose suffix reaches k
        int b_ = 0, cum = 0;
        if (lane == ls) {
            cum = suf - s;                                        // elements in lanes above
            for (b_ = lane * 32 + 31; b_ > lane * 32; --b_) {
                const int hb = wreg[b_];
                if (cum + hb >= k) break;
                cum += hb;
            }
        }
        bstar = __shfl_sync(0xffffffffu, b_, ls);
        above = __shfl_sync(0xffffffffu, cum, ls);
    }
    (void)above;
    __syncwarp();                                                 // histogram is dead from here on
    const unsigned long long lo_key = (unsigned long long)lo + ((unsigned long long)bstar << shift);
    const unsigned long long hi_key = lo_key + (1ull << shift);
    const float lo_val = key_float((uint32_t)lo_key);
    const float hi_val = (hi_key >= 0xFF800000ull) ? INFINITY : key_float((uint32_t)hi_key);

    // classify: certainly selected / candidate / certainly not
    int nsel = 0, ncand = 0;
    for (int c0 = 0; c0 < C; c0 += 32) {
        const int c = c0 + lane;
        const bool valid = c < C;
        const float a = valid ? row[c] : 0.f;
        const bool in = valid && (a - band >= hi_val);
        const bool cd = valid && !in && (a + band >= lo_val);
        const unsigned bi = __ballot_sync(0xffffffffu, in), bc = __ballot_sync(0xffffffffu, cd);
        if (lane == 0) bitmap_h[c0 >> 5] = bi;
        if (sel != nullptr && in) sel[nsel + __popc(bi & lt)] = c;
        if (cd) cand[ncand + __popc(bc & lt)] = (unsigned short)c;
        nsel += __popc(bi); ncand += __popc(bc);
    }
    const int need = k - nsel;
    __syncwarp();
    if (ncand == need) {
        for (int i = lane; i < ncand; i += 32) {
            const int c = cand[i];
            atomicOr(&bitmap_h[c >> 5], 1u << (c & 31));
            if (sel != nullptr) sel[nsel + i] = c;
        }
    } else {
        double *xs = (ncand <= kXCap) ? reinterpret_cast<double *>(wreg) : gscratch;
        exact_scores<QT>(cent_rows, qd, nullptr, cand, ncand, xs);
        for (int i = lane; i < ncand; i += 32) {
            const int e = cand[i];
            const double se = xs[i];
            int rank = 0;
            for (int j = 0; j < ncand; ++j) rank += precedes(xs[j], (int)cand[j], se, e) ? 1 : 0;
            if (rank < need) {
                atomicOr(&bitmap_h[e >> 5], 1u << (e & 31));
                if (sel != nullptr) sel[nsel + rank] = e;
            }
        }
    }
    __syncwarp();
}

// sel[0..k) -> out[0..k) in (exact score desc, index asc) order   (compute_cent op output)
template <typename QT>
__device__ void sort_selected(const ChainParams &p, const uint2 *cent_rows, const double (&qd)[4], const int *sel,
                              unsigned *wreg, double *gscratch, int *out) {
    const int k = p.k, lane = lane_id();
    double *xs = (k <= kXCap) ? reinterpret_cast<double *>(wreg) : gscratch;
    exact_scores<QT>(cent_rows, qd, sel, nullptr, k, xs);
    for (int i = lane; i < k; i += 32) {
        const int e = sel[i];
        const double se = xs[i];
        int rank = 0;
        for (int j = 0; j < k; ++j) rank += precedes(xs[j], sel[j], se, e) ? 1 : 0;
        out[rank] = e;
    }
    __syncwarp();
}

// ---- compaction of one bit-plane of pm[] into an ascending page list -------------------------
// One warp.  Writes out[0..min(count, cap)) = ascending i with (pm[i] >> bit) & 1, pads with the
// sentinel up to cap, clears the bit of entries beyond cap (so later consumers agree with the
// truncated list), returns the kept count.
__device__ inline int warp_compact_bit(unsigned *pm, int pl, int bit, int *out, int cap) {
    const int lane = lane_id();
    int base = 0;
    for (int i0 = 0; i0 < pl; i0 += 32) {
        int i = i0 + lane;
        bool on = (i < pl) && ((pm[i] >> bit) & 1u);
        unsigned bal = __ballot_sync(0xffffffffu, on);
        int pos = base + __popc(bal & ((1u << lane) - 1u));
        if (on) {
            if (pos < cap) { if (out) out[pos] = i; }
            else atomicAnd(&pm[i], ~((1u << bit) | (1u << (bit + 16))));
        }
        base += __popc(bal);
    }
    int cnt = min(base, cap);
    if (out) for (int j = cnt + lane; j < cap; j += 32) out[j] = kSentinel;
    return cnt;
}

// ---- union list of one item ------------------------------------------------------------------
// pm[i] = head mask | tail mask << 16 for logical page i.  Emits the ascending list of pages with
// a non-zero mask: physical block id (block_table value) + meta word, and the item's length.
// phys: block_table row, either staged in smem or the global row itself.
__device__ inline void warp_emit_union(const ChainParams &p, int item, const int *phys, const unsigned *pm, int pl) {
    const int lane = lane_id();
    int *uphys = p.union_phys + (size_t)item * p.maxP;
    unsigned *umeta = p.union_meta + (size_t)item * p.maxP;
    int base = 0;
    for (int i0 = 0; i0 < pl; i0 += 32) {
        int i = i0 + lane;
        unsigned m = (i < pl) ? pm[i] : 0u;
        bool on = (m & 0xffffu) != 0u;
        unsigned bal = __ballot_sync(0xffffffffu, on);
        if (on) {
            int pos = base + __popc(bal & ((1u << lane) - 1u));
            uphys[pos] = phys[i];
            umeta[pos] = m;
        }
        base += __popc(bal);
    }
    if (lane == 0) p.union_len[item] = base;
}

// ---- phase A for the fused op / cent_select / compute_cent / select_position ------------------
template <typename QT>
__device__ void select_item(const ChainParams &p, int item, int flags, unsigned char *smem, const Grp &G) {
    const int b = item / p.Nkv, n2 = item % p.Nkv;
    const int g = p.g, C = p.C, k = p.k;
    const int lane = lane_id();
    SelSmem L(g, C, k, p.maxP, G.gw, flags);
    float *sc = reinterpret_cast<float *>(smem + L.sc);
    unsigned *wreg = reinterpret_cast<unsigned *>(smem + L.wreg) + G.warp * kSelBins;
    unsigned short *cand = reinterpret_cast<unsigned short *>(smem + L.cand) + (size_t)G.warp * C;
    int *selw = (flags & kSelWriteIndices) ? reinterpret_cast<int *>(smem + L.sel) + G.warp * k : nullptr;
    unsigned *bitmap = reinterpret_cast<unsigned *>(smem + L.bitmap);
    unsigned *pm = reinterpret_cast<unsigned *>(smem + L.pm);
    int *physS = reinterpret_cast<int *>(smem + L.phys);
    int *srt = reinterpret_cast<int *>(smem + L.srt);
    float *misc = reinterpret_cast<float *>(smem + L.misc);
    const bool do_scores = (flags & kSelScores) != 0, do_select = (flags & kSelSelect) != 0;

    // ---- page-table gathers first: c[i] = block_ids[n2, block_table[b, i]] stays in registers
    //      until the selection is known (select_position.cpp:97-115 CopyIn/Gather)
    int seq = 0, pl = 0, rem = 0;
    bool has_tail = false;
    int creg[kPG];
    const int *bt = nullptr, *bid = nullptr;
    if (do_select) {
        seq = p.seq_len[b];
        pl = max(0, min((seq + p.bs - 1) / p.bs, p.maxP));
        rem = (pl > 0) ? seq - (pl - 1) * p.bs : 0;               // tokens on the last page
        has_tail = (pl > 0) && (rem < p.bs);
        bt = p.block_table + (size_t)b * p.bt_stride;
        bid = p.block_ids + (size_t)n2 * p.bid_stride;
        int btv[kPG];
#pragma unroll
        for (int j = 0; j < kPG; ++j) {
            const int i = G.tid + j * G.gt;
            btv[j] = (i < pl) ? __ldg(bt + i) : 0;
        }
#pragma unroll
        for (int j = 0; j < kPG; ++j) {
            const int i = G.tid + j * G.gt;
            creg[j] = 0;
            if (i < pl) {
                physS[i] = btv[j];
                creg[j] = __ldg(bid + max(0, min(btv[j], p.kvPageLen - 1)));   // clamp: utils.py:62
            }
        }
    }

    trace_mark(p, 8);
    if (do_scores) {
        const float wmax = score_tiles<QT>(p, b, n2, sc, L.scs, G);
        if (lane == 0) misc[G.warp] = wmax;
        trace_mark(p, 9);
        grp_sync(G);
        trace_mark(p, 10);
        float maxabs = 0.f;
        for (int w = 0; w < G.gw; ++w) maxabs = fmaxf(maxabs, misc[w]);
        const unsigned short *q = reinterpret_cast<const unsigned short *>(p.q);
        const uint2 *cent_rows = reinterpret_cast<const uint2 *>(p.cent) + (size_t)n2 * C * (kD / 4);
        double *gscratch = p.exact_scratch + ((size_t)blockIdx.x * kConsumerWarps + warp_id()) * C;
        for (int h = G.warp; h < g; h += G.gw) {
            const int hq = n2 * g + h;
            double qd[4];
            topk_head<QT>(p, q + (size_t)b * p.q_sb + (size_t)hq * p.q_sh, cent_rows, sc + h * L.scs, maxabs, wreg,
                          cand, selw, bitmap + h * L.bw, gscratch, qd);
            if (flags & kSelWriteIndices)
                sort_selected<QT>(p, cent_rows, qd, selw, wreg, gscratch, p.indices_out + ((size_t)b * p.Nq + hq) * k);
        }
    } else if (do_select) {
        // indices are given (select_position op): arbitrary values, duplicates allowed
        // (tests/select_position.py:45) -> direct comparison against the k entries per head.
        for (int i = G.tid; i < g * k; i += G.gt) {
            const int h = i / k;
            srt[i] = p.indices_in[(size_t)b * p.idx_sb + (size_t)(n2 * g + h) * p.idx_sh + (i % k)];
        }
    }
    if (!do_select) { grp_sync(G); return; }
    trace_mark(p, 11);
    grp_sync(G);
    trace_mark(p, 12);

    // ---- membership mask per logical page (select_position.cpp:122-183; tests/utils.py:57-80)
    auto page_mask = [&](int i, int c) {
        unsigned m = 0u;
        if (do_scores) {
            if (c >= 0 && c < C)
                for (int h = 0; h < g; ++h) m |= ((bitmap[h * L.bw + (c >> 5)] >> (c & 31)) & 1u) << h;
        } else {
            for (int h = 0; h < g; ++h) {
                bool hit = false;
                for (int j = 0; j < k; ++j) hit |= (srt[h * k + j] == c);
                m |= (hit ? 1u : 0u) << h;
            }
        }
        if (has_tail && i == pl - 1) m |= m << 16;
        pm[i] = m;
        if (flags & kSelGather)
            for (int h = 0; h < g; ++h) p.bt_gather[((size_t)b * p.Nq + n2 * g + h) * p.maxP + i] = c;
    };
#pragma unroll
    for (int j = 0; j < kPG; ++j) {
        const int i = G.tid + j * G.gt;
        if (i < pl) page_mask(i, creg[j]);
    }
    for (int i = G.tid + kPG * G.gt; i < pl; i += G.gt) {          // very long block tables
        const int ph = __ldg(bt + i);
        physS[i] = ph;
        page_mask(i, __ldg(bid + max(0, min(ph, p.kvPageLen - 1))));
    }
    if (flags & kSelGather)
        for (int i = pl + G.tid; i < p.maxP; i += G.gt)
            for (int h = 0; h < g; ++h) p.bt_gather[((size_t)b * p.Nq + n2 * g + h) * p.maxP + i] = 0;
    grp_sync(G);
    trace_mark(p, 13);
    // ---- per-head ascending compaction -> page_position / page_position_length
    for (int h = G.warp; h < g; h += G.gw) {
        const int hq = n2 * g + h;
        int *pp = p.page_position ? p.page_position + ((size_t)b * p.Nq + hq) * p.MP : nullptr;
        const int cnt = warp_compact_bit(pm, pl, h, pp, p.MP);
        if (p.page_position_length && lane < kPad)
            p.page_position_length[((size_t)b * p.Nq + hq) * kPad + lane] = cnt;
    }
    trace_mark(p, 14);
    if (flags & kSelUnion) {
        grp_sync(G);
        if (G.warp == 0) {
            warp_emit_union(p, item, physS, pm, pl);
            if (lane == 0) p.tail_count[item] = has_tail ? rem : p.bs;
        }
    }
    trace_mark(p, 15);
    grp_sync(G);
}

// ---- phase A' for the separate attention op: union list from a given block_position ----------
// block_position[b, n2 or hq, j]: logical page of sparse-sequence slot j, 0x7fffffff-terminated
// (sparse_paged_attention_split_Bbn2s2_Us2.h:258,277-279,293).  Slot j covers sparse tokens
// [j*bs, (j+1)*bs) clipped to actual_seq_lengths[b].  Whole CTA.
__device__ inline void union_from_block_position(const ChainParams &p, int item, unsigned char *smem) {
    const int b = item / p.Nkv, n2 = item % p.Nkv;
    const int g = p.g;
    const int tid = threadIdx.x, nt = blockDim.x;
    unsigned *pm = reinterpret_cast<unsigned *>(smem);
    for (int i = tid; i < p.maxP; i += nt) pm[i] = 0u;
    __syncthreads();
    long long actual = p.actual_seq[p.actual_n > 1 ? b : 0];
    if (actual < 0) actual = 0;
    long long nslots_ll = (actual + p.bs - 1) / p.bs;
    const int nslots = (nslots_ll < (long long)p.bp_len) ? (int)nslots_ll : p.bp_len;
    const int tail_slot = (actual % p.bs) ? (int)(actual / p.bs) : -1;
    const int nlists = (p.bp_heads == p.Nq) ? g : 1;
    for (int l = warp_id(); l < nlists; l += nt / 32) {
        const int row = (p.bp_heads == p.Nq) ? (n2 * g + l) : n2;
        const unsigned bits = (p.bp_heads == p.Nq) ? (1u << l) : ((1u << g) - 1u);
        const int *bp = p.bp_in + (size_t)b * p.bp_sb + (size_t)row * p.bp_sh;
        const int lane = lane_id();
        for (int j0 = 0; j0 < nslots; j0 += 32) {
            int j = j0 + lane;
            int v = (j < nslots) ? __ldg(bp + j) : kSentinel;
            bool ok = (j < nslots) && v != kSentinel && v >= 0 && v < p.maxP;
            unsigned bad = ~__ballot_sync(0xffffffffu, ok);
            int first_bad = bad ? (__ffs(bad) - 1) : 32;          // list ends at the first sentinel
            if (lane < first_bad) {
                unsigned m = bits | ((j == tail_slot) ? (bits << 16) : 0u);
                atomicOr(&pm[v], m);
            }
            if (first_bad < 32) break;
        }
    }
    __syncthreads();
    if (warp_id() == 0) {
        warp_emit_union(p, item, p.block_table + (size_t)b * p.bt_stride, pm, p.maxP);
        if (lane_id() == 0) p.tail_count[item] = (tail_slot >= 0) ? (int)(actual % p.bs) : p.bs;
    }
    __syncthreads();
}

// all items of phase A handled by this CTA's worker groups (warps 0..kConsumerWarps-1)
template <typename QT>
__device__ __forceinline__ void select_phase(const ChainParams &p, int flags, unsigned char *smem) {
    if (warp_id() >= kConsumerWarps) return;
    const Grp G = make_group(p.sel_gw);
    const int ngroups = kConsumerWarps / p.sel_gw;
    unsigned char *gs = smem + (size_t)G.id * p.sel_group_bytes;
    for (int item = blockIdx.x + G.id * gridDim.x; item < p.items; item += gridDim.x * ngroups)
        select_item<QT>(p, item, flags, gs, G);
}

// max_page_position_length[b, 0..7] = bs * max_h count   (cent_select.cpp:123-160)
__device__ inline void write_max_ppl(const ChainParams &p, int b) {
    if (warp_id() != 0) return;
    int m = 0;
    for (int h = lane_id(); h < p.Nq; h += 32)
        m = max(m, p.page_position_length[((size_t)b * p.Nq + h) * kPad]);
    for (int o = 16; o; o >>= 1) m = max(m, __shfl_xor_sync(0xffffffffu, m, o));
    if (lane_id() < kPad) p.max_ppl[(size_t)b * kPad + lane_id()] = (long long)m * p.bs;
}

}  // namespace spa
