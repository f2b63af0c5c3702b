// Phase A of the chain: centroid scores (exact), TopK, page selection, union-list construction.
//
// One CTA handles one work item (b, kv head) and all g query heads of that GQA group, so the
// centroid table of the kv head is read once per item instead of once per query head
// (the reference re-streams it per (b, q head): compute_cent.cpp:171-174, cent_select.cpp:228-236).
#pragma once
#include "spa_common.cuh"

namespace spa {

// ---- shared-memory carve-up for phase A (host + device) -------------------------------------
struct SelSmem {
    int qd, sc, hist, sel, srt, bitmap, pm, misc, total;
    __host__ __device__ static inline int align16(int x) { return (x + 15) & ~15; }
    __host__ __device__ SelSmem(int g, int C, int k, int hb, int nbins, int maxP) {
        int o = 0;
        qd = o;      o = align16(o + g * kD * 8);
        sc = o;      o = align16(o + hb * C * 8);
        hist = o;    o = align16(o + hb * nbins * 4);       // reused as boundary list (uint16 x 2*nbins)
        sel = o;     o = align16(o + hb * k * 4);
        srt = o;     o = align16(o + g * k * 4);            // also holds given indices (select_position)
        bitmap = o;  o = align16(o + g * ((C + 31) / 32) * 4);
        pm = o;      o = align16(o + maxP * 4);
        misc = o;    o = align16(o + 8 * kMaxGroup * 4);
        total = o;
    }
};

// misc slots (ints), each [kMaxGroup]
enum { kMiscMin = 0, kMiscMax = 1, kMiscSelCnt = 2, kMiscBndCnt = 3, kMiscBstar = 4, kMiscAbove = 5,
       kMiscShift = 6 };

// ---- scores: exact sum of exact products ----------------------------------------------------
// score[h][c] = sum_d q[h][d] * cent[c][d].  fp16/bf16 products are exact; the float64 running
// sum is exact for the value ranges in use, so the result is independent of summation order and
// equals the oracle's (oracle/sparse_chain.py:cent_scores).  Reference arithmetic: fp32
// accumulate with hardware-defined order (compute_cent.cpp:277-343, tests/utils.py:117).
template <typename QT, int HB>
__device__ __forceinline__ void score_rows(const ChainParams &p, int n2, int h0, int nh,
                                           const double *qd, double *sc) {
    const uint4 *cent = reinterpret_cast<const uint4 *>(p.cent) + (size_t)n2 * p.C * (kD / 8);
    for (int c = threadIdx.x; c < p.C; c += blockDim.x) {
        const uint4 *row = cent + (size_t)c * (kD / 8);
        double acc[HB];
#pragma unroll
        for (int h = 0; h < HB; ++h) acc[h] = 0.0;
#pragma unroll 4
        for (int j = 0; j < kD / 8; ++j) {
            uint4 v = __ldg(row + j);
            uint32_t w[4] = {v.x, v.y, v.z, v.w};
            double x[8];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                x[2 * e] = QTraits<QT>::to_double(static_cast<unsigned short>(w[e] & 0xffffu));
                x[2 * e + 1] = QTraits<QT>::to_double(static_cast<unsigned short>(w[e] >> 16));
            }
#pragma unroll
            for (int h = 0; h < HB; ++h) {
                if (h < nh) {
                    const double2 *qh = reinterpret_cast<const double2 *>(qd + (h0 + h) * kD + j * 8);
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        double2 qq = qh[e];
                        acc[h] = fma(x[2 * e], qq.x, acc[h]);
                        acc[h] = fma(x[2 * e + 1], qq.y, acc[h]);
                    }
                }
            }
        }
#pragma unroll
        for (int h = 0; h < HB; ++h)
            if (h < nh) sc[h * p.C + c] = acc[h];
    }
}

// ---- TopK over nh rows of C exact scores held in smem ---------------------------------------
// Output: sel[h][0..k) = the k selected cluster ids (unordered), and, if `sorted` != nullptr,
// sorted[h][0..k) in (score desc, index asc) order.  Algorithm: 32-bit order-preserving float keys
// -> one histogram pass locates the bin of the k-th largest; bins above it are taken wholesale,
// the boundary bin is resolved with exact float64 comparisons (ties -> lowest index).
__device__ inline void topk_rows(const ChainParams &p, int nh, const double *sc, unsigned *hist,
                                 int *sel, int *sorted, int *misc) {
    const int C = p.C, k = p.k, nbins = p.nbins;
    const int tid = threadIdx.x, nt = blockDim.x;
    int *kmin = misc + kMiscMin * kMaxGroup, *kmax = misc + kMiscMax * kMaxGroup;
    int *selcnt = misc + kMiscSelCnt * kMaxGroup, *bndcnt = misc + kMiscBndCnt * kMaxGroup;
    int *bstar = misc + kMiscBstar * kMaxGroup, *above = misc + kMiscAbove * kMaxGroup;
    int *shiftv = misc + kMiscShift * kMaxGroup;

    if (tid < nh) {
        kmin[tid] = (int)0xffffffffu; kmax[tid] = 0; selcnt[tid] = 0; bndcnt[tid] = 0;
    }
    for (int i = tid; i < nh * nbins; i += nt) hist[i] = 0;
    __syncthreads();
    // pass 1: per-row min / max key
    for (int h = 0; h < nh; ++h) {
        uint32_t lo = 0xffffffffu, hi = 0;
        for (int c = tid; c < C; c += nt) {
            uint32_t key = float_key((float)sc[h * C + c]);
            lo = min(lo, key); hi = max(hi, key);
        }
        for (int o = 16; o; o >>= 1) {
            lo = min(lo, __shfl_xor_sync(0xffffffffu, lo, o));
            hi = max(hi, __shfl_xor_sync(0xffffffffu, hi, o));
        }
        if (lane_id() == 0) {
            atomicMin(reinterpret_cast<unsigned *>(&kmin[h]), lo);
            atomicMax(reinterpret_cast<unsigned *>(&kmax[h]), hi);
        }
    }
    __syncthreads();
    if (tid < nh) {
        uint32_t range = (uint32_t)kmax[tid] - (uint32_t)kmin[tid];
        int bits = 32 - __clz(range);                 // range < 2^bits
        int lb = 31 - __clz(nbins);                   // log2(nbins)
        shiftv[tid] = max(0, bits - lb);
    }
    __syncthreads();
    // pass 2: histogram
    for (int h = 0; h < nh; ++h) {
        uint32_t lo = (uint32_t)kmin[h]; int sh = shiftv[h];
        for (int c = tid; c < C; c += nt) {
            uint32_t key = float_key((float)sc[h * C + c]);
            atomicAdd(&hist[h * nbins + ((key - lo) >> sh)], 1u);
        }
    }
    __syncthreads();
    // locate the bin holding the k-th largest: one warp per row, suffix sums from the top bin
    for (int h = warp_id(); h < nh; h += nt / 32) {
        const int lane = lane_id();
        const int per = nbins / 32;
        const unsigned *hrow = hist + h * nbins;
        int s = 0;
        for (int i = 0; i < per; ++i) s += hrow[lane * per + i];
        int suf = s;
        for (int o = 1; o < 32; o <<= 1) {
            int t = __shfl_down_sync(0xffffffffu, suf, o);
            if (lane + o < 32) suf += t;
        }
        unsigned bal = __ballot_sync(0xffffffffu, suf >= k);
        int ls = 31 - __clz(bal);                      // highest lane whose suffix reaches k
        if (lane == ls) {
            int cum = suf - s;                         // elements in lanes above
            int b = lane * per + per - 1;
            for (; b >= lane * per; --b) {
                int hb_ = hrow[b];
                if (cum + hb_ >= k) break;
                cum += hb_;
            }
            bstar[h] = b; above[h] = cum;
        }
    }
    __syncthreads();
    // collect: bins above -> selected; boundary bin -> candidate list (aliases the histogram)
    unsigned short *bnd = reinterpret_cast<unsigned short *>(hist);
    const int bcap = nbins * 2;                       // >= C by construction of nbins
    // histogram is dead from here on; but other rows' histograms share the region with this
    // row's candidate list, so finish reading bstar/above (done above) before writing.
    for (int h = 0; h < nh; ++h) {
        uint32_t lo = (uint32_t)kmin[h]; int sh = shiftv[h]; int bs_ = bstar[h];
        for (int c = tid; c < C; c += nt) {
            uint32_t key = float_key((float)sc[h * C + c]);
            int bin = (int)((key - lo) >> sh);
            if (bin > bs_) {
                int pos = atomicAdd(&selcnt[h], 1);
                sel[h * k + pos] = c;
            } else if (bin == bs_) {
                int pos = atomicAdd(&bndcnt[h], 1);
                bnd[h * bcap + pos] = (unsigned short)c;
            }
        }
    }
    __syncthreads();
    // boundary resolution by exact rank
    for (int h = 0; h < nh; ++h) {
        const int nb = bndcnt[h], need = k - above[h], base = above[h];
        for (int i = tid; i < nb; i += nt) {
            int e = bnd[h * bcap + i];
            double se = sc[h * C + e];
            int rank = 0;
            for (int j = 0; j < nb; ++j) {
                int f = bnd[h * bcap + j];
                rank += precedes(sc[h * C + f], f, se, e) ? 1 : 0;
            }
            if (rank < need) sel[h * k + base + rank] = e;
        }
    }
    __syncthreads();
    if (sorted != nullptr) {
        for (int i = tid; i < nh * k; i += nt) {
            int h = i / k;
            int e = sel[i];
            double se = sc[h * C + e];
            int rank = 0;
            for (int j = 0; j < k; ++j) {
                int f = sel[h * k + j];
                rank += precedes(sc[h * C + f], f, se, e) ? 1 : 0;
            }
            sorted[h * k + rank] = e;
        }
        __syncthreads();
    }
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
__device__ inline void warp_emit_union(const ChainParams &p, int item, int b, const unsigned *pm,
                                       int pl) {
    const int lane = lane_id();
    int *uphys = p.union_phys + (size_t)item * p.maxP;
    unsigned *umeta = p.union_meta + (size_t)item * p.maxP;
    const int *bt = p.block_table + (size_t)b * p.bt_stride;
    int base = 0;
    for (int i0 = 0; i0 < pl; i0 += 32) {
        int i = i0 + lane;
        unsigned m = (i < pl) ? pm[i] : 0u;
        bool on = (m & 0xffffu) != 0u;
        unsigned bal = __ballot_sync(0xffffffffu, on);
        if (on) {
            int pos = base + __popc(bal & ((1u << lane) - 1u));
            uphys[pos] = __ldg(bt + i);
            umeta[pos] = m;
        }
        base += __popc(bal);
    }
    if (lane == 0) p.union_len[item] = base;
}

// ---- phase A for the fused op / cent_select / compute_cent ----------------------------------
// flags
enum : int { kSelScores = 1, kSelWriteIndices = 2, kSelSelect = 4, kSelUnion = 8, kSelGather = 16 };

template <typename QT>
__device__ void select_item(const ChainParams &p, int item, int flags, unsigned char *smem) {
    const int b = item / p.Nkv, n2 = item % p.Nkv;
    const int g = p.g, C = p.C, k = p.k;
    const int tid = threadIdx.x, nt = blockDim.x;
    SelSmem L(g, C, k, p.hb, p.nbins, p.maxP);
    double *qd = reinterpret_cast<double *>(smem + L.qd);
    double *sc = reinterpret_cast<double *>(smem + L.sc);
    unsigned *hist = reinterpret_cast<unsigned *>(smem + L.hist);
    int *sel = reinterpret_cast<int *>(smem + L.sel);
    int *srt = reinterpret_cast<int *>(smem + L.srt);
    unsigned *bitmap = reinterpret_cast<unsigned *>(smem + L.bitmap);
    unsigned *pm = reinterpret_cast<unsigned *>(smem + L.pm);
    int *misc = reinterpret_cast<int *>(smem + L.misc);
    const int bw = (C + 31) / 32;

    if (flags & kSelScores) {
        const unsigned short *q = reinterpret_cast<const unsigned short *>(p.q);
        for (int i = tid; i < g * kD; i += nt) {
            int h = i / kD, d = i % kD;
            qd[i] = QTraits<QT>::to_double(q[(size_t)b * p.q_sb + (size_t)(n2 * g + h) * p.q_sh + d]);
        }
        for (int i = tid; i < g * bw; i += nt) bitmap[i] = 0u;
        __syncthreads();
        for (int h0 = 0; h0 < g; h0 += p.hb) {
            const int nh = min(p.hb, g - h0);
            if (p.hb >= 8)      score_rows<QT, 8>(p, n2, h0, nh, qd, sc);
            else if (p.hb >= 4) score_rows<QT, 4>(p, n2, h0, nh, qd, sc);
            else if (p.hb >= 2) score_rows<QT, 2>(p, n2, h0, nh, qd, sc);
            else                score_rows<QT, 1>(p, n2, h0, nh, qd, sc);
            __syncthreads();
            const bool want_sorted = (flags & kSelWriteIndices) != 0;
            topk_rows(p, nh, sc, hist, sel, want_sorted ? srt + h0 * k : nullptr, misc);
            // record the selection: bitmap for membership tests, indices for the op output
            for (int i = tid; i < nh * k; i += nt) {
                int h = h0 + i / k;
                int e = sel[i];
                atomicOr(&bitmap[h * bw + (e >> 5)], 1u << (e & 31));
                if (want_sorted)
                    p.indices_out[((size_t)b * p.Nq + n2 * g + h) * k + (i % k)] = srt[h0 * k + i];
            }
            __syncthreads();
        }
    }
    if (!(flags & kSelSelect)) return;

    // ---- select_position (tests/utils.py:4-84; select_position.cpp:122-183) ----
    const int seq = p.seq_len[b];
    int pl = (seq + p.bs - 1) / p.bs;
    pl = max(0, min(pl, p.maxP));
    const int *bt = p.block_table + (size_t)b * p.bt_stride;
    const int *bid = p.block_ids + (size_t)n2 * p.bid_stride;
    const int rem = (pl > 0) ? seq - (pl - 1) * p.bs : 0;          // tokens on the last page
    const bool has_tail = (pl > 0) && (rem < p.bs);
    if (!(flags & kSelScores)) {
        // indices are given (select_position op): arbitrary values, duplicates allowed
        // (tests/select_position.py:45) -> direct comparison against the k entries per head.
        for (int i = tid; i < g * k; i += nt) {
            int h = i / k;
            srt[i] = p.indices_in[(size_t)b * p.idx_sb + (size_t)(n2 * g + h) * p.idx_sh + (i % k)];
        }
        __syncthreads();
    }
    for (int i = tid; i < p.maxP; i += nt) {
        unsigned m = 0u;
        int c = 0;
        if (i < pl) {
            int phys = __ldg(bt + i);
            phys = max(0, min(phys, p.kvPageLen - 1));               // utils.py:62
            c = __ldg(bid + phys);
            if (flags & kSelScores) {
                if (c >= 0 && c < C)
                    for (int h = 0; h < g; ++h) m |= ((bitmap[h * bw + (c >> 5)] >> (c & 31)) & 1u) << h;
            } else {
                for (int h = 0; h < g; ++h) {
                    bool hit = false;
                    for (int j = 0; j < k; ++j) hit |= (srt[h * k + j] == c);
                    m |= (hit ? 1u : 0u) << h;
                }
            }
            if (has_tail && i == pl - 1) m |= m << 16;
        }
        pm[i] = m;
        if (flags & kSelGather) {
            for (int h = 0; h < g; ++h)
                p.bt_gather[((size_t)b * p.Nq + n2 * g + h) * p.maxP + i] = c;
        }
    }
    __syncthreads();
    // per-head ascending compaction -> page_position / page_position_length
    for (int h = warp_id(); h < g; h += nt / 32) {
        const int hq = n2 * g + h;
        int *pp = p.page_position ? p.page_position + ((size_t)b * p.Nq + hq) * p.MP : nullptr;
        int cnt = warp_compact_bit(pm, pl, h, pp, p.MP);
        if (p.page_position_length && lane_id() < kPad)
            p.page_position_length[((size_t)b * p.Nq + hq) * kPad + lane_id()] = cnt;
    }
    if (flags & kSelUnion) {
        __syncthreads();
        if (warp_id() == 0) {
            warp_emit_union(p, item, b, pm, pl);
            if (lane_id() == 0) p.tail_count[item] = has_tail ? rem : p.bs;
        }
    }
    __syncthreads();
}

// ---- phase A' for the separate attention op: union list from a given block_position ----------
// block_position[b, n2 or hq, j]: logical page of sparse-sequence slot j, 0x7fffffff-terminated
// (sparse_paged_attention_split_Bbn2s2_Us2.h:258,277-279,293).  Slot j covers sparse tokens
// [j*bs, (j+1)*bs) clipped to actual_seq_lengths[b].
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
        warp_emit_union(p, item, b, pm, p.maxP);
        if (lane_id() == 0) p.tail_count[item] = (tail_slot >= 0) ? (int)(actual % p.bs) : p.bs;
    }
    __syncthreads();
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
