// Phase A of the chain: centroid scores, TopK, page selection, union-list construction.
//
// Work item = (b, kv head) with all g query heads of that GQA group, so the centroid table of the
// kv head is read once per item instead of once per query head (the reference re-streams it per
// (b, q head): compute_cent.cpp:171-174, cent_select.cpp:228-236).  An item is processed by a
// GROUP of 4 or 8 warps (two 4-warp groups work on different items concurrently when g <= 4 and
// there are more items than CTAs); groups synchronise with their own named barrier.
//
// Scores and TopK -- how bit-exact selection is obtained from tensor-core arithmetic:
//   1. approximate scores a[h][c] = q[h] . cent[c] on the tensor cores: the kv head's centroid
//      table arrives as 128-row tiles by TMA (128B-swizzled, same shape as a K page), ldmatrix +
//      mma.sync m16n8k16 with fp32 accumulate (M = 16 clusters, N = the group's heads, K = head dim);
//   2. |a - e| <= eps[h] for the exact score e, with eps = 2^-16 * ||q_h||_1 * max|cent| (the
//      products are exact in fp32; the bound covers 128 truncating accumulation steps);
//   3. one 1024-bin histogram of order-preserving keys per head locates the bin holding the k-th
//      largest a; elements more than 2*eps above the bin are certainly selected, elements more
//      than 2*eps below it certainly not; the few in between ("candidates") are ranked by their
//      approximate score, and only those within 2*eps of the cut get their EXACT score (float64 sum
//      of exact products, order-independent), ranked by (score desc, index asc).
// The result equals the exact-arithmetic TopK of the oracle (oracle/sparse_chain.py:cent_scores,
// topk_indices), which is the tie-free centre of every fp32 evaluation order of the reference
// (compute_cent.cpp:277-343 VMLA + pair-add + BlockReduceSum; tests/utils.py:117 torch.matmul).
//
// The code is deliberately compact (little unrolling): it runs once per CTA from a cold
// instruction cache, where straight-line code costs more than loops.
#pragma once
#include "spa_common.cuh"

namespace spa {

constexpr int kSelBins = 1024;            // histogram bins per head (power of two, multiple of 32)
constexpr int kXCap = 256;                // exact scores per head kept in smem (more -> workspace)
constexpr int kCentTileRows = 128;        // centroid rows per TMA tile
constexpr int kCentTileBytes = kCentTileRows * kD * 2;

// flags (compile-time per kernel instantiation)
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

// per-head record in shared memory
struct HeadRec {
    unsigned kmin, kmax;                 // key range of the approximate scores
    int ncand;                           // candidates collected
    float l1;                            // ||q_h||_1
    float lo, inv, band;                 // histogram: bin(x) = floor(x * inv + lo) (lo = -min * inv); 2*eps
    int bstar;                           // bin of the k-th largest approximate score
    int total;                           // selected pages (compaction)
};

// ---- shared-memory carve-up of ONE group (host + device), offsets from the group base ---------
struct SelSmem {
    int scs, bw, nch;                    // score row stride (floats), bitmap words per head, 32-page chunks
    int sc, hist, cand, xs, sel, bitmap, pm, phys, cval, cnt, srt, hd, misc, total;
    __host__ __device__ static inline int al(int x) { return (x + 15) & ~15; }
    __host__ __device__ SelSmem(int g, int C, int k, int maxP, int flags) {
        const bool scores = (flags & kSelScores) != 0;
        scs = ((C + 3) & ~3) + 4;        // +4: the 4 lanes of a quad (heads 2t) hit distinct banks
        bw = (C + 31) / 32;
        nch = (maxP + 31) / 32;
        int o = 0;
        sc = o;      o = al(o + (scores ? g * scs * 4 : 0));
        hist = o;    o = al(o + (scores ? g * kSelBins * 4 : 0));
        cand = o;    o = al(o + (scores ? g * C * 2 : 0));
        xs = o;      o = al(o + (scores ? g * kXCap * 8 : 0));
        sel = o;     o = al(o + ((scores && (flags & kSelWriteIndices)) ? g * k * 4 : 0));
        bitmap = o;  o = al(o + (scores ? g * bw * 4 : 0));
        pm = o;      o = al(o + maxP * 4);
        phys = o;    o = al(o + maxP * 4);
        cval = o;    o = al(o + maxP * 4);
        cnt = o;     o = al(o + nch * (kMaxGroup + 1) * 4);
        srt = o;     o = al(o + (scores ? 0 : g * k * 4));      // given indices (select_position)
        hd = o;      o = al(o + kMaxGroup * (int)sizeof(HeadRec));
        misc = o;    o = al(o + 16 * 4);
        total = (o + 127) & ~127;
    }
};
// misc words: [0..7] per-warp max|cent|, [8] truncation flag, [9] union total, [10] queue base of the union

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

// ---- per-group TMA ring of centroid tiles -------------------------------------------------------
struct CentRing {
    unsigned char *base;                 // ns stages of kCentTileBytes (aliases the attention ring)
    uint64_t *full, *empty;              // [ns] each
    int ns;
};
__device__ __forceinline__ void issue_cent_tile(const CUtensorMap *tmC, const CentRing &R, int stage, int row0) {
    unsigned char *dst = R.base + (size_t)stage * kCentTileBytes;
    mbar_expect_tx(&R.full[stage], (uint32_t)kCentTileBytes);
    tma_load_2d(dst, tmC, &R.full[stage], 0, row0);
    tma_load_2d(dst + kCentTileBytes / 2, tmC, &R.full[stage], 64, row0);
}

// ---- approximate scores on the tensor cores ----------------------------------------------------
// S^T[c][h] = cent[c][:] . q[h][:]  per 128-row tile: A = 16 clusters x 16 d via ldmatrix from the
// swizzled tile, B = 16 d x 8 heads from registers.  Also gathers, per head, the key range of the
// scores and ||q_h||_1, and max|cent| (for the error band).
// Q fragments of the item's heads as mma B operands (b0 = Q[h][16ks+2t..], b1 = Q[h][16ks+2t+8..]) and
// ||q_h||_1 -> hd[h].l1.  Issued first thing in an item so the global-load latency overlaps the set-up.
template <typename QT>
__device__ __forceinline__ void load_q_frags(const ChainParams &p, int b, int n2, HeadRec *hd, const Grp &G,
                                             uint32_t (&qf)[2][8][2]) {
    const int lane = lane_id(), gid = lane >> 2, tig = lane & 3;
    const int g = p.g;
    const int ntl = (g + 7) >> 3;
    const unsigned short *q = reinterpret_cast<const unsigned short *>(p.q);
#pragma unroll
    for (int nt = 0; nt < 2; ++nt) {
        const int h = nt * 8 + gid;
        const bool on = nt < ntl && h < g;
        const unsigned short *qrow = q + (size_t)b * p.q_sb + (size_t)(n2 * g + (on ? h : 0)) * p.q_sh;
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) {
            uint32_t v0 = 0u, v1 = 0u;
            if (on) {
                v0 = *reinterpret_cast<const uint32_t *>(qrow + ks * 16 + tig * 2);
                v1 = *reinterpret_cast<const uint32_t *>(qrow + ks * 16 + tig * 2 + 8);
            }
            qf[nt][ks][0] = v0; qf[nt][ks][1] = v1;
        }
    }
}
template <typename QT>
__device__ __forceinline__ void q_l1_norms(const ChainParams &p, HeadRec *hd, const Grp &G, const uint32_t (&qf)[2][8][2]) {
    const int lane = lane_id(), gid = lane >> 2, tig = lane & 3;
    const int ntl = (p.g + 7) >> 3;
#pragma unroll
    for (int nt = 0; nt < 2; ++nt) {
        float l1 = 0.f;
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) {
            const uint32_t v0 = qf[nt][ks][0], v1 = qf[nt][ks][1];
            l1 += fabsf(QTraits<QT>::to_float((unsigned short)(v0 & 0xffffu))) + fabsf(QTraits<QT>::to_float((unsigned short)(v0 >> 16))) +
                  fabsf(QTraits<QT>::to_float((unsigned short)(v1 & 0xffffu))) + fabsf(QTraits<QT>::to_float((unsigned short)(v1 >> 16)));
        }
        l1 += __shfl_xor_sync(0xffffffffu, l1, 1);
        l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
        const int h = nt * 8 + gid;
        if (G.warp == 0 && tig == 0 && nt < ntl && h < p.g) hd[h].l1 = l1;
    }
}

template <typename QT>
__device__ __forceinline__ void score_item(const ChainParams &p, const CUtensorMap *tmC, int b, int n2, float *sc,
                                           int scs, HeadRec *hd, float *misc, const CentRing &R, unsigned &tcount,
                                           const Grp &G, const uint32_t (&qf)[2][8][2]) {
    const int lane = lane_id(), gid = lane >> 2, tig = lane & 3;
    const int g = p.g, C = p.C;
    const int ntl = (g + 7) >> 3;                      // head tiles (1 or 2)
    q_l1_norms<QT>(p, hd, G, qf);
    trace_clk(p, 18);
    const int ntiles = (C + kCentTileRows - 1) / kCentTileRows;
    const int sub = kConsumerWarps / G.gw;             // 16-row slabs per warp per tile
    unsigned kmn[2][2] = {{0xffffffffu, 0xffffffffu}, {0xffffffffu, 0xffffffffu}}, kmx[2][2] = {{0u, 0u}, {0u, 0u}};
    uint32_t mx4[4] = {0u, 0u, 0u, 0u};
    const uint32_t ring_u32 = smem_u32(R.base);
    const int a_chk = lane >> 4;
#pragma unroll 1
    for (int t = 0; t < ntiles; ++t) {
        const unsigned tc = tcount + (unsigned)t;
        const int s = (int)(tc % (unsigned)R.ns);
        const unsigned par = (tc / (unsigned)R.ns) & 1u;
        mbar_wait(&R.full[s], par);
#pragma unroll 1
        for (int mi = 0; mi < sub; ++mi) {
            const int t0 = (G.warp + mi * G.gw) * 16;
            const int a_row = t0 + (lane & 15);
            float acc[2][4], acd[2][4];            // even / odd k-steps: two independent mma chains
#pragma unroll
            for (int nt = 0; nt < 2; ++nt)
#pragma unroll
                for (int r = 0; r < 4; ++r) { acc[nt][r] = 0.f; acd[nt][r] = 0.f; }
            const uint32_t rowaddr = ring_u32 + (uint32_t)s * kCentTileBytes + a_row * 128;
#pragma unroll
            for (int ks = 0; ks < 8; ++ks) {
                const int chunk = (ks & 3) * 2 + a_chk;
                uint32_t a0, a1, a2, a3;
                ldmatrix_x4(a0, a1, a2, a3, rowaddr + (ks >> 2) * (kCentTileRows * 128) + ((chunk ^ (a_row & 7)) << 4));
                mx4[0] = absmax2<QT>(mx4[0], a0); mx4[1] = absmax2<QT>(mx4[1], a1);
                mx4[2] = absmax2<QT>(mx4[2], a2); mx4[3] = absmax2<QT>(mx4[3], a3);
                if (ks & 1) {
                    mma_16816<QT>(acd[0], a0, a1, a2, a3, qf[0][ks][0], qf[0][ks][1]);
                    if (ntl > 1) mma_16816<QT>(acd[1], a0, a1, a2, a3, qf[1][ks][0], qf[1][ks][1]);
                } else {
                    mma_16816<QT>(acc[0], a0, a1, a2, a3, qf[0][ks][0], qf[0][ks][1]);
                    if (ntl > 1) mma_16816<QT>(acc[1], a0, a1, a2, a3, qf[1][ks][0], qf[1][ks][1]);
                }
            }
#pragma unroll
            for (int nt = 0; nt < 2; ++nt)
#pragma unroll
                for (int r = 0; r < 4; ++r) acc[nt][r] += acd[nt][r];
            const int r0 = t * kCentTileRows + t0 + gid, r1 = r0 + 8;
#pragma unroll
            for (int nt = 0; nt < 2; ++nt) {
#pragma unroll
                for (int e = 0; e < 2; ++e) {
                    const int h = nt * 8 + tig * 2 + e;
                    if (nt < ntl && h < g) {
                        if (r0 < C) {
                            sc[h * scs + r0] = acc[nt][e];
                            const unsigned key = float_key(acc[nt][e]);
                            kmn[nt][e] = min(kmn[nt][e], key); kmx[nt][e] = max(kmx[nt][e], key);
                        }
                        if (r1 < C) {
                            sc[h * scs + r1] = acc[nt][2 + e];
                            const unsigned key = float_key(acc[nt][2 + e]);
                            kmn[nt][e] = min(kmn[nt][e], key); kmx[nt][e] = max(kmx[nt][e], key);
                        }
                    }
                }
            }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&R.empty[s]);
        if (G.tid == 0 && t + R.ns < ntiles) {
            mbar_wait(&R.empty[s], par);
            issue_cent_tile(tmC, R, s, n2 * C + (t + R.ns) * kCentTileRows);
        }
    }
    tcount += (unsigned)ntiles;
    trace_clk(p, 19);
    // key range per head: reduce over the 8 row-lanes, one shared atomic per (warp, head)
#pragma unroll
    for (int nt = 0; nt < 2; ++nt)
#pragma unroll
        for (int e = 0; e < 2; ++e) {
            unsigned lo = kmn[nt][e], hi = kmx[nt][e];
#pragma unroll
            for (int o = 4; o < 32; o <<= 1) {
                lo = min(lo, __shfl_xor_sync(0xffffffffu, lo, o));
                hi = max(hi, __shfl_xor_sync(0xffffffffu, hi, o));
            }
            const int h = nt * 8 + tig * 2 + e;
            if (gid == 0 && nt < ntl && h < g && lo <= hi) { atomicMin(&hd[h].kmin, lo); atomicMax(&hd[h].kmax, hi); }
        }
    const uint32_t mx = absmax2<QT>(absmax2<QT>(mx4[0], mx4[1]), absmax2<QT>(mx4[2], mx4[3]));
    float m = fmaxf(QTraits<QT>::to_float((unsigned short)(mx & 0xffffu)), QTraits<QT>::to_float((unsigned short)(mx >> 16)));
#pragma unroll
    for (int o = 16; o; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    if (lane == 0) misc[G.warp] = m;
}

// Centroid rows for the exact refinement: the shared-memory tiles when the whole table of the kv head
// is still resident in the group's TMA ring (C <= ns * 128 rows), else global memory (L2).
struct CentSrc {
    const uint2 *g;                      // [C][kD/4] rows of this kv head
    const unsigned char *ring;           // null: not resident
    int stage0, ns;                      // tile t of the item sits in stage (stage0 + t) % ns
    // this lane's 4 head-dim elements (d = 4*lane .. 4*lane+3) of row c
    __device__ __forceinline__ uint2 load(int c, int lane) const {
        if (ring == nullptr) return __ldg(g + (size_t)c * (kD / 4) + lane);
        const int t = c / kCentTileRows, r = c % kCentTileRows;
        const int byte = (lane & 15) * 8;                        // within the 128-byte half row (128B swizzle)
        const unsigned char *a = ring + (size_t)((stage0 + t) % ns) * kCentTileBytes + (lane >> 4) * (kCentTileBytes / 2) +
                                 r * 128 + ((((byte >> 4) ^ (r & 7)) << 4) | (byte & 15));
        return *reinterpret_cast<const uint2 *>(a);
    }
};

// exact partial dot of this lane's 4 head-dim elements (d = 4*lane .. 4*lane+3)
template <typename QT>
__device__ __forceinline__ double exact_partial(uint2 v, const double (&qd)[4]) {
    double s = QTraits<QT>::to_double((unsigned short)(v.x & 0xffffu)) * qd[0];
    s = fma(QTraits<QT>::to_double((unsigned short)(v.x >> 16)), qd[1], s);
    s = fma(QTraits<QT>::to_double((unsigned short)(v.y & 0xffffu)), qd[2], s);
    s = fma(QTraits<QT>::to_double((unsigned short)(v.y >> 16)), qd[3], s);
    return s;
}

// exact scores of list[0..n) -> xs[0..n)   (one warp; rows come from L2, 4 in flight)
template <typename QT>
static __device__ __noinline__ void exact_scores(const CentSrc cent_rows, const unsigned short *qrow, const int *list32,
                                          const unsigned short *list16, int n, double *xs) {
    const int lane = lane_id();
    double qd[4];
    {
        const uint2 qv = *reinterpret_cast<const uint2 *>(qrow + 4 * lane);
        qd[0] = QTraits<QT>::to_double((unsigned short)(qv.x & 0xffffu)); qd[1] = QTraits<QT>::to_double((unsigned short)(qv.x >> 16));
        qd[2] = QTraits<QT>::to_double((unsigned short)(qv.y & 0xffffu)); qd[3] = QTraits<QT>::to_double((unsigned short)(qv.y >> 16));
    }
#pragma unroll 1
    for (int i0 = 0; i0 < n; i0 += 4) {
        uint2 v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int i = min(i0 + u, n - 1);
            const int c = list32 ? list32[i] : (int)list16[i];
            v[u] = cent_rows.load(c, lane);
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            double s = exact_partial<QT>(v[u], qd);
#pragma unroll
            for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
            if (lane == 0 && i0 + u < n) xs[i0 + u] = s;
        }
    }
    __syncwarp();
}

// ---- TopK, step 1 (group): histogram of the order-preserving keys of every head ----------------
// Histogram bins are LINEAR in the score: bin(x) = floor((x - lo) * inv) with lo = the head's smallest score and
// inv = bins / range.  (Bins cut in the bit space of the float keys are very uneven when the scores straddle
// zero -- zero-mean centroids, fused_ifa_v5_case.py:55 -- and left ~50 candidates per head in the bin of the k-th
// score: c2 with randn centroids ran 12 us behind rand.)  sel_bin() is monotone non-decreasing in x, which is all
// the classification needs: bin(x) > bin(y) implies x > y.  Below the range: negative.
// (one FMA -- a single rounding, hence monotone -- and a floor conversion; c0 = -lo * inv)
__device__ __forceinline__ int sel_bin(float x, float c0, float inv) {
    return min(kSelBins - 1, __float2int_rd(fmaf(x, inv, c0)));
}
__device__ __forceinline__ void sel_bin_scale(unsigned kmin, unsigned kmax, float &c0, float &inv) {
    const float lo = key_float(kmin);
    const float range = key_float(kmax) - lo;
    inv = (range > 0.f && range < INFINITY) ? (float)kSelBins / range : 0.f;
    c0 = -lo * inv;
}
__device__ __forceinline__ void topk_hist(const ChainParams &p, const float *sc, int scs, unsigned *hist,
                                          const HeadRec *hd, const Grp &G) {
    // four heads per pass: four independent load -> key -> atomic chains per thread
#pragma unroll 1
    for (int h0 = 0; h0 < p.g; h0 += 4) {
        float lo[4], inv[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int h = min(h0 + u, p.g - 1);
            sel_bin_scale(hd[h].kmin, hd[h].kmax, lo[u], inv[u]);
        }
        const int nh = min(4, p.g - h0);
#pragma unroll 2
        for (int c = G.tid; c < p.C; c += G.gt) {
            float a[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) a[u] = sc[min(h0 + u, p.g - 1) * scs + c];
#pragma unroll
            for (int u = 0; u < 4; ++u)
                if (u < nh) atomicAdd(&hist[(h0 + u) * kSelBins + max(0, sel_bin(a[u], lo[u], inv[u]))], 1u);
        }
    }
}

// ---- TopK, step 2 (one warp per head): the bin holding the k-th largest score ------------------
__device__ __forceinline__ void topk_bin(const ChainParams &p, const unsigned *hh, HeadRec *r, float maxabs) {
    const int lane = lane_id(), k = p.k;
    // level 1: lane covers bins [32 lane, 32 lane + 32); suffix sums from the top
    int s = 0;
    {
        // rotated start: lanes of a quarter-warp hit 8 different 16-byte bank groups (stride is 128 B)
        const uint4 *hv = reinterpret_cast<const uint4 *>(hh + lane * 32);
#pragma unroll
        for (int i = 0; i < 8; ++i) { const uint4 v = hv[(i + lane) & 7]; s += v.x + v.y + v.z + v.w; }
    }
    int suf = s;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_down_sync(0xffffffffu, suf, o);
        if (lane + o < 32) suf += t;
    }
    const int ls = 31 - __clz(__ballot_sync(0xffffffffu, suf >= k));       // highest block whose suffix reaches k
    const int above_blk = __shfl_sync(0xffffffffu, suf - s, ls);           // elements in blocks above
    // level 2: lane = bin within block ls
    const int v = hh[ls * 32 + lane];
    int suf2 = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_down_sync(0xffffffffu, suf2, o);
        if (lane + o < 32) suf2 += t;
    }
    const int lb = 31 - __clz(__ballot_sync(0xffffffffu, above_blk + suf2 >= k));
    const int bstar = ls * 32 + lb;
    if (lane == 0) {
        sel_bin_scale(r->kmin, r->kmax, r->lo, r->inv);
        r->bstar = bstar;
        r->band = ldexpf(r->l1 * maxabs, -15) * 1.0625f;                   // 2 * eps
    }
}

// ---- TopK, step 3 (group): certainly selected -> bitmap, near the cut -> candidate list ----------
__device__ __forceinline__ void topk_classify(const ChainParams &p, const float *sc, int scs, HeadRec *hd,
                                              unsigned *bitmap, int bw, unsigned short *cand, const Grp &G) {
    const int lane = lane_id();
    const unsigned lt = (1u << lane) - 1u;
    // four heads per pass (independent chains).  A warp iteration covers one aligned 32-cluster chunk =
    // one bitmap word per head: ballots, no per-lane atomics.
#pragma unroll 1
    for (int h0 = 0; h0 < p.g; h0 += 4) {
        // sel_bin(x) > b*  <=>  x * inv + c0 >= b* + 1  (no bin above the last one);  sel_bin(x) >= b*  <=>  x * inv + c0 >= b*:
        // float compares instead of a floor conversion per element.  The band is folded into the constants.
        float inv[4], cin[4], ccd[4], tin[4], tcd[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int h = min(h0 + u, p.g - 1);
            inv[u] = hd[h].inv;
            cin[u] = fmaf(-hd[h].band, hd[h].inv, hd[h].lo);          // (a - band) * inv + c0, up to one rounding: the
            ccd[u] = fmaf(hd[h].band, hd[h].inv, hd[h].lo);           // band carries a 6 % safety factor for that
            tin[u] = (hd[h].bstar >= kSelBins - 1) ? INFINITY : (float)(hd[h].bstar + 1);
            tcd[u] = (float)hd[h].bstar;
        }
        const int nh = min(4, p.g - h0);
#pragma unroll 1
        for (int c0 = G.warp * 32; c0 < p.C; c0 += G.gt) {
            const int c = c0 + lane;
            unsigned bin[4], bcd[4];
            bool cd[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const float a = (c < p.C) ? sc[min(h0 + u, p.g - 1) * scs + c] : -INFINITY;
                // above every score of the k-th score's bin even after the error band: certainly selected;
                // able to reach that bin: candidate
                const bool in = (c < p.C) && fmaf(a, inv[u], cin[u]) >= tin[u];
                cd[u] = (c < p.C) && !in && fmaf(a, inv[u], ccd[u]) >= tcd[u];
                bin[u] = __ballot_sync(0xffffffffu, in);
                bcd[u] = __ballot_sync(0xffffffffu, cd[u]);
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                if (u >= nh) break;
                const int h = h0 + u;
                if (lane == 0) bitmap[h * bw + (c0 >> 5)] = bin[u];
                if (bcd[u]) {
                    int base = 0;
                    if (lane == 0) base = atomicAdd(&hd[h].ncand, __popc(bcd[u]));
                    base = __shfl_sync(0xffffffffu, base, 0);
                    if (cd[u]) cand[(size_t)h * p.C + base + __popc(bcd[u] & lt)] = (unsigned short)c;
                }
            }
        }
    }
}

// ---- TopK, step 4 (one warp per head): settle the candidates ------------------------------------
template <typename QT>
__device__ __forceinline__ void topk_resolve(const ChainParams &p, const unsigned short *qrow, const CentSrc cent_rows,
                                             const float *row, const HeadRec *r, unsigned short *cand,
                                             unsigned *bitmap_h, int bw, double *xs_smem, double *gscratch) {
    const int lane = lane_id();
    const unsigned lt = (1u << lane) - 1u;
    int nsel = 0;
    for (int w = lane; w < bw; w += 32) nsel += __popc(bitmap_h[w]);
#pragma unroll
    for (int o = 16; o; o >>= 1) nsel += __shfl_xor_sync(0xffffffffu, nsel, o);
    int need = p.k - nsel;
    int n = r->ncand;
    if (n != need && n <= 32) {
        // rank the candidates by approximate score; only those within the band of the cut stay open
        const float band = r->band;
        const int ci = (lane < n) ? (int)cand[lane] : 0x7fffffff;
        const float ai = (lane < n) ? row[ci] : -INFINITY;
        int rank = 0;
        for (int j = 0; j < n; ++j) {
            const float aj = __shfl_sync(0xffffffffu, ai, j);
            const int cj = __shfl_sync(0xffffffffu, ci, j);
            rank += (aj > ai || (aj == ai && cj < ci)) ? 1 : 0;
        }
        const unsigned bcut = __ballot_sync(0xffffffffu, lane < n && rank == need - 1);
        const float astar = __shfl_sync(0xffffffffu, ai, __ffs(bcut) - 1);
        const bool in = lane < n && (ai - band > astar);
        const bool amb = lane < n && !in && (ai + band >= astar);
        if (in) atomicOr(&bitmap_h[ci >> 5], 1u << (ci & 31));
        const unsigned bamb = __ballot_sync(0xffffffffu, amb);
        need -= __popc(__ballot_sync(0xffffffffu, in));
        n = __popc(bamb);
        __syncwarp();
        if (amb) cand[__popc(bamb & lt)] = (unsigned short)ci;
        __syncwarp();
    }
    if (n == need) {
        for (int i = lane; i < n; i += 32) {
            const int c = cand[i];
            atomicOr(&bitmap_h[c >> 5], 1u << (c & 31));
        }
    } else {
        double *xs = (n <= kXCap) ? xs_smem : gscratch;
        exact_scores<QT>(cent_rows, qrow, nullptr, cand, n, xs);
        for (int i = lane; i < n; i += 32) {
            const int e = cand[i];
            const double se = xs[i];
            int rank = 0;
            for (int j = 0; j < n; ++j) rank += precedes(xs[j], (int)cand[j], se, e) ? 1 : 0;
            if (rank < need) atomicOr(&bitmap_h[e >> 5], 1u << (e & 31));
        }
    }
    __syncwarp();
}

// compute_cent op output: the k selected clusters in (exact score desc, index asc) order  (one warp)
template <typename QT>
__device__ __forceinline__ void emit_sorted(const ChainParams &p, const unsigned short *qrow, const CentSrc cent_rows,
                                            const unsigned *bitmap_h, int bw, int *sel, double *xs_smem,
                                            double *gscratch, int *out) {
    const int lane = lane_id(), k = p.k;
    int base = 0;
    for (int w0 = 0; w0 < bw; w0 += 32) {
        unsigned word = (w0 + lane < bw) ? bitmap_h[w0 + lane] : 0u;
        int inc = __popc(word);
        const int mine = inc;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += t;
        }
        int pos = base + inc - mine;
        while (word) {
            const int bit = __ffs(word) - 1;
            sel[pos++] = (w0 + lane) * 32 + bit;
            word &= word - 1;
        }
        base += __shfl_sync(0xffffffffu, inc, 31);
    }
    __syncwarp();
    double *xs = (k <= kXCap) ? xs_smem : gscratch;
    exact_scores<QT>(cent_rows, qrow, sel, nullptr, k, xs);
    for (int i = lane; i < k; i += 32) {
        const int e = sel[i];
        const double se = xs[i];
        int rank = 0;
        for (int j = 0; j < k; ++j) rank += precedes(xs[j], sel[j], se, e) ? 1 : 0;
        out[rank] = e;
    }
    __syncwarp();
}

// ---- slow path (a head selected more pages than max_page_num): serial compaction ---------------
// One warp.  Writes out[0..min(count, cap)) = ascending i with (pm[i] >> bit) & 1, pads with the
// sentinel up to cap, clears the bit of entries beyond cap (so the union agrees with the
// truncated list), returns the kept count.
static __device__ __noinline__ int warp_compact_bit(unsigned *pm, int pl, int bit, int *out, int cap) {
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

// ---- publishing an item's union pages ------------------------------------------------------------
// The entries of an item sit at the fixed place flat[item * qstride + u), u < ulen, in ascending page order:
// writing them needs no reservation.  Every entry, record and chunk descriptor is a single 64-bit store that is
// non-zero once written, and a reader needs nothing else, so no ordering (fence) is required anywhere on this
// path: readers poll the word they need (spa_attn.cuh, take_entry / the producer's chunk loop).
//
// Who streams what is decided here, per item, by the publishing CTA alone (nothing global is known yet):
//   * the CTA keeps a PRIVATE PREFIX [0, S) of each of its own items -- together about static_pct % of a CTA's
//     fair share of the launch, estimated from this item -- and starts streaming it as soon as its own selection
//     is over, while other CTAs are still selecting (the reference separates the phases with SyncAll,
//     sparse_paged_fusion_attention.cpp:56-76);
//   * the rest [S, ulen) is cut into chunks of HALVING size (between 1/8 and 1/32 of a CTA's fair share -- larger when
//     an item is so long that its pslots - 1 chunks would not cover it otherwise): chunk j of item i is descriptor desc[j * items + i], and CTAs claim descriptors in index
//     order, so all the large chunks go first and the launch ends on the small ones (guided self-scheduling
//     without knowing the total).  No reservation: the slot is fixed;
//   * one release-add reports the item (high half of the kTail word) after an atomicMax of its chunk count
//     (low half): once every item has reported, a descriptor that is still zero is an empty slot and
//     rounds * items is the end of the queue;
//   * little work (own items adding up to <= small_pages pages): the item stays whole and private -- no
//     partial, no combine (five dependent L2 round trips after the item's last page).
// One thread.  m >= 0: also report the item's longest per-head page list (max_page_position_length).
static __device__ __noinline__ void publish_item(const ChainParams &p, PubState *pub, int item, int ulen, int tail_cnt, int m) {
    if (m >= 0) p.item_ppl[item] = m + 1;                             // picked up by a post warp (max_ppl_poll)
    int nch = 0;
    if (ulen > 0) {
        st_relaxed_u64(p.item_info + item, item_record(ulen, tail_cnt));
        const int G = (int)gridDim.x;
        const float fair = (float)p.items * (float)ulen / (float)G;    // a CTA's share, if every item were like this one
        const int n_mine = ((int)blockIdx.x < p.items) ? (p.items - 1 - (int)blockIdx.x) / G + 1 : 1;   // items this CTA selects
        const int used = pub->priv_used;
        int S = min(ulen, (int)(fair * (float)p.static_pct / (100.f * (float)n_mine) + 0.5f));
        if (used + ulen <= p.small_pages) S = ulen;
        if (pub->n_own >= kMaxOwn) S = 0;
        const int min_chunk = max(p.dyn_min, (int)(fair * (1.f / 32.f))), max_chunk = max(min_chunk, (int)(fair * (1.f / 8.f)));
        if (S > 0 && ulen - S < min_chunk) S = ulen;                   // no crumbs on either side
        if (S < p.dyn_min && S < ulen) S = 0;
        if (S > 0) {
            const int o = pub->n_own;
            pub->own_item[o] = item; pub->own_len[o] = S;
            pub->n_own = o + 1; pub->priv_used = used + S;
        }
        int at = S, left = ulen - S;
        while (left > 0) {
            int c = max(min_chunk, min((left * p.chunk_num) / p.chunk_den, max_chunk));
            const int slots_left = p.pslots - 1 - nch;                 // chunks this item may still be cut into
            c = max(c, (left + slots_left - 1) / slots_left);          // ... must cover what is left (few, long items)
            if (slots_left == 1 || left - c < min_chunk) c = left;
            st_relaxed_u64(p.desc + (size_t)nch * p.items + item, chunk_desc(item, at, c));
            at += c; left -= c; ++nch;
        }
    }
    if (nch > 0) atomicMax(p.ctrl + Ctrl::kTail, nch);
    asm volatile("red.release.gpu.global.add.s32 [%0], 1;" ::"l"(p.ctrl + Ctrl::kItemsDone) : "memory");
}
// A selection group hands its finished item (or, item < 0, "this group is done") to the publisher -- the TMA
// producer warp, idle during phase A -- so that the atomic's round trip stays off the selection's path.  One thread.
__device__ __forceinline__ void post_publish_request(PubState *pub, int gi, unsigned &seq, int item, int ulen, int tail_cnt, int m) {
    while (pub->ack_seq[gi] != seq) __nanosleep(20);                   // previous request taken (an item ago: never waits)
    pub->req[gi][0] = item; pub->req[gi][1] = ulen; pub->req[gi][2] = tail_cnt; pub->req[gi][3] = m;
    __threadfence_block();
    pub->req_seq[gi] = ++seq;
}
// Publisher loop of the TMA producer warp during phase A (lane 0).
__device__ __forceinline__ void publisher_loop(const ChainParams &p, PubState *pub, int ngroups) {
    unsigned seen[2] = {0u, 0u};
    int finished = 0;
    while (finished < ngroups) {
        bool any = false;
#pragma unroll
        for (int gi = 0; gi < 2; ++gi) {
            if (gi >= ngroups) break;
            const unsigned s = pub->req_seq[gi];
            if (s == seen[gi]) continue;
            __threadfence_block();
            const int item = pub->req[gi][0], ulen = pub->req[gi][1], tail = pub->req[gi][2], m = pub->req[gi][3];
            seen[gi] = s;
            pub->ack_seq[gi] = s;
            any = true;
            if (item < 0) ++finished; else publish_item(p, pub, item, ulen, tail, m);
        }
        if (!any) __nanosleep(64);
    }
}
// attention_out rows of an item without any page (seq 0 / nothing selected): zeros, as the combine of
// an empty sum gives.  nthreads threads, tid in [0, nthreads).
static __device__ __noinline__ void zero_item_output(const ChainParams &p, int b, int n2, int tid, int nthreads) {
    const bool i8out = p.pq_scale != nullptr;              // post-quant: int8 rows (InitAllZeroInt8Output, ..._Us2.h:996)
    for (int i = tid; i < p.g * (kD / 4); i += nthreads) {
        const int h = i / (kD / 4), d4 = i % (kD / 4);
        const size_t off = (size_t)b * p.out_sb + (size_t)(n2 * p.g + h) * p.out_sh + d4 * 4;
        if (i8out) { *reinterpret_cast<uint32_t *>(reinterpret_cast<unsigned char *>(p.out) + off) = 0u; continue; }
        *reinterpret_cast<uint2 *>(reinterpret_cast<unsigned short *>(p.out) + off) = make_uint2(0u, 0u);
        for (int r = 0; r < p.n_peer; ++r)                 // tensor parallel: the peers' gathered buffers too
            *reinterpret_cast<uint2 *>(reinterpret_cast<unsigned short *>(p.out_peer[r]) + off) = make_uint2(0u, 0u);
    }
    if (p.lse != nullptr)                                  // log of an empty sum
        for (int h = tid; h < p.g; h += nthreads) p.lse[(size_t)b * p.lse_sb + (size_t)(n2 * p.g + h) * p.lse_sh] = -INFINITY;
}
// Union list of one item (one warp, serial over chunks).  pm[i] = head mask | tail mask << 16 for
// logical page i; writes the ascending pages with a non-zero mask to the item's place in `flat`.  An entry has
// ONE tail flag: a page that is the tail page of some heads and a full page of others (possible only with
// per-head block_position lists) is listed twice, once per head subset.  Returns the number of entries.
static __device__ __noinline__ int warp_emit_union(const ChainParams &p, int item, const int *phys, const unsigned *pm, int pl,
                                            int tail_cnt) {
    const int lane = lane_id();
    const unsigned lt = (1u << lane) - 1u;
    unsigned long long *q = p.flat + (size_t)item * p.qstride;
    int base = 0;
    for (int i0 = 0; i0 < pl; i0 += 32) {
        const int i = i0 + lane;
        unsigned m = (i < pl) ? pm[i] : 0u;
        if (m != 0u && (unsigned)phys[i] >= (unsigned)p.num_blocks) m = 0u;      // block id outside the pool: skipped
        const unsigned mt = (m >> 16) & m & 0xffffu, mf = m & 0xffffu & ~mt;
        const unsigned bf = __ballot_sync(0xffffffffu, mf != 0u), bt = __ballot_sync(0xffffffffu, mt != 0u);
        const int id = (m != 0u) ? (p.opt ? i : phys[i]) : 0;                     // options build: the LOGICAL page travels
        if (mf) st_relaxed_u64(q + base + __popc(bf & lt), queue_entry(id, mf, false, item, p.bs));
        if (mt) st_relaxed_u64(q + base + __popc(bf) + __popc(bt & lt), queue_entry(id, mt, true, item, max(1, tail_cnt)));
        base += __popc(bf) + __popc(bt);
    }
    __syncwarp();
    return base;
}

// ---- phase A for one item: fused op / cent_select / compute_cent / select_position ------------
template <typename QT, int FLAGS>
__device__ void select_item(const ChainParams &p, const CUtensorMap *tmC, int item, unsigned char *smem,
                            const CentRing &R, unsigned &tcount, const Grp &G, PubState *pub, unsigned &pub_seq) {
    constexpr bool DS = (FLAGS & kSelScores) != 0, DP = (FLAGS & kSelSelect) != 0;
    constexpr int KG1 = kMaxGroup + 1;
    const int b = item / p.Nkv, n2 = item % p.Nkv;
    const int g = p.g, C = p.C, k = p.k, lane = lane_id();
    const int hq0 = n2 * g;
    const SelSmem L(g, C, k, p.maxP, FLAGS);
    float *sc = reinterpret_cast<float *>(smem + L.sc);
    unsigned *hist = reinterpret_cast<unsigned *>(smem + L.hist);
    unsigned short *cand = reinterpret_cast<unsigned short *>(smem + L.cand);
    double *xs = reinterpret_cast<double *>(smem + L.xs);
    int *selS = reinterpret_cast<int *>(smem + L.sel);
    unsigned *bitmap = reinterpret_cast<unsigned *>(smem + L.bitmap);
    unsigned *pm = reinterpret_cast<unsigned *>(smem + L.pm);
    int *physS = reinterpret_cast<int *>(smem + L.phys);
    int *cS = reinterpret_cast<int *>(smem + L.cval);
    int *cnt = reinterpret_cast<int *>(smem + L.cnt);
    int *srt = reinterpret_cast<int *>(smem + L.srt);
    HeadRec *hd = reinterpret_cast<HeadRec *>(smem + L.hd);
    float *misc = reinterpret_cast<float *>(smem + L.misc);
    int *miscI = reinterpret_cast<int *>(smem + L.misc);

    // ---- S0: start everything asynchronous, initialise the per-item state
    trace_clk(p, 16);
    int seq = 0, pl = 0, rem = 0;
    bool has_tail = false;
    const int *bid = nullptr;
    uint32_t qf[2][8][2];
    if (DS) load_q_frags<QT>(p, b, n2, hd, G, qf);                // in flight across the whole set-up
    if (DP) {
        seq = p.seq_len[b];                                       // first used after the block_table wait
        const int *bt = p.block_table + (size_t)b * p.bt_stride;
        bid = p.block_ids + (size_t)n2 * p.bid_stride;
        // whole block_table row -> smem (not just the pages of seq: no dependent load in front of the copy)
        for (int i = G.tid; i < p.maxP; i += G.gt) cp_async_4(&physS[i], bt + i);
    }
    if (DS) {
        if (G.tid == 0) {
            const int nt = min((C + kCentTileRows - 1) / kCentTileRows, R.ns);
            for (int t = 0; t < nt; ++t)
                issue_cent_tile(tmC, R, (int)((tcount + (unsigned)t) % (unsigned)R.ns), n2 * C + t * kCentTileRows);
        }
        if (G.tid < g) { hd[G.tid].kmin = 0xffffffffu; hd[G.tid].kmax = 0u; hd[G.tid].ncand = 0; }
        uint4 *hz = reinterpret_cast<uint4 *>(hist);
        for (int i = G.tid; i < g * (kSelBins / 4); i += G.gt) hz[i] = make_uint4(0u, 0u, 0u, 0u);
        for (int i = G.tid; i < g * L.bw; i += G.gt) bitmap[i] = 0u;
    }
    if (G.tid == 0) { miscI[8] = 0; miscI[9] = 0; miscI[10] = 0; }
    if (DP && p.page_position != nullptr) {
        // pad first: the real positions overwrite the head of each row later (ordered by the barriers)
        int *base = p.page_position + ((size_t)b * p.Nq + hq0) * p.MP;
        const int total = g * p.MP;
        if ((total & 3) == 0 && (reinterpret_cast<uintptr_t>(base) & 15) == 0) {
            int4 *b4 = reinterpret_cast<int4 *>(base);
            for (int i = G.tid; i < total / 4; i += G.gt) b4[i] = make_int4(kSentinel, kSentinel, kSentinel, kSentinel);
        } else {
            for (int i = G.tid; i < total; i += G.gt) base[i] = kSentinel;
        }
    }
    // c[i] = block_ids[n2, block_table[b, i]] (select_position.cpp:97-115), asynchronously into smem; the table
    // value is clamped as the golden does (tests/utils.py:62).  Needs the block_table row (a cold HBM round trip):
    // with scores to compute, that wait and the gather are issued AFTER the score loop -- the scores do not depend
    // on them, the row has long landed by then, and the gather hides under the TopK stages.  Page i is copied and
    // gathered by the same thread, so a thread only waits for its own copies.
    auto issue_gather = [&]() {
        cp_async_wait_all();
        pl = max(0, min((seq + p.bs - 1) / p.bs, p.maxP));
        rem = (pl > 0) ? seq - (pl - 1) * p.bs : 0;               // tokens on the last page
        has_tail = (pl > 0) && (rem < p.bs);
        for (int i = G.tid; i < pl; i += G.gt) cp_async_4(&cS[i], bid + max(0, min(physS[i], p.kvPageLen - 1)));
    };
    if (DP && !DS) issue_gather();
    grp_sync(G);
    trace_clk(p, 17);
    trace_mark(p, 8);

    const unsigned short *q = reinterpret_cast<const unsigned short *>(p.q);
    if (DS) {
        score_item<QT>(p, tmC, b, n2, sc, L.scs, hd, misc, R, tcount, G, qf);
        trace_clk(p, 20);
        if (DP) issue_gather();
        grp_sync(G);
        trace_clk(p, 21);
        trace_mark(p, 9);
        float maxabs = 0.f;
        for (int w = 0; w < G.gw; ++w) maxabs = fmaxf(maxabs, misc[w]);
        topk_hist(p, sc, L.scs, hist, hd, G);
        trace_clk(p, 22);
        grp_sync(G);
        trace_clk(p, 23);
        trace_mark(p, 10);
        for (int h = G.warp; h < g; h += G.gw) topk_bin(p, hist + h * kSelBins, &hd[h], maxabs);
        trace_clk(p, 24);
        grp_sync(G);
        trace_clk(p, 25);
        trace_mark(p, 11);
        topk_classify(p, sc, L.scs, hd, bitmap, L.bw, cand, G);
        trace_clk(p, 26);
        grp_sync(G);
        trace_clk(p, 27);
        trace_mark(p, 12);
        CentSrc cent_rows;
        cent_rows.g = reinterpret_cast<const uint2 *>(p.cent) + (size_t)n2 * C * (kD / 4);
        const int ntl_item = (C + kCentTileRows - 1) / kCentTileRows;
        cent_rows.ring = (ntl_item <= R.ns) ? R.base : nullptr;
        cent_rows.ns = R.ns;
        cent_rows.stage0 = (int)((tcount - (unsigned)ntl_item) % (unsigned)R.ns);   // score_item advanced tcount past this item
        double *gscratch = p.exact_scratch + ((size_t)blockIdx.x * kConsumerWarps + warp_id()) * C;
#pragma unroll 1
        for (int h = G.warp; h < g; h += G.gw) {
            const unsigned short *qrow = q + (size_t)b * p.q_sb + (size_t)(hq0 + h) * p.q_sh;
            topk_resolve<QT>(p, qrow, cent_rows, sc + h * L.scs, &hd[h], cand + (size_t)h * C, bitmap + h * L.bw, L.bw,
                             xs + h * kXCap, gscratch);
            if (FLAGS & kSelWriteIndices)
                emit_sorted<QT>(p, qrow, cent_rows, bitmap + h * L.bw, L.bw, selS + h * k, xs + h * kXCap, gscratch,
                                p.indices_out + ((size_t)b * p.Nq + hq0 + h) * k);
        }
        trace_clk(p, 28);
        trace_mark(p, 13);
    } else if (DP) {
        // indices are given (select_position op): arbitrary values, duplicates allowed
        // (tests/select_position.py:45) -> direct comparison against the k entries per head.
        for (int i = G.tid; i < g * k; i += G.gt)
            srt[i] = p.indices_in[(size_t)b * p.idx_sb + (size_t)(hq0 + i / k) * p.idx_sh + (i % k)];
    }
    if (!DP) { grp_sync(G); return; }
    cp_async_wait_all();
    grp_sync(G);

    // ---- membership mask per logical page (select_position.cpp:122-183; tests/utils.py:57-80)
#pragma unroll 1
    for (int i = G.tid; i < pl; i += G.gt) {
        const int c = cS[i];
        unsigned m = 0u;
        if (DS) {
            if (c >= 0 && c < C) {
                const unsigned *bw = bitmap + (c >> 5);
                const int sh = c & 31;
#pragma unroll 4
                for (int h = 0; h < g; ++h) m |= ((bw[h * L.bw] >> sh) & 1u) << h;
            }
        } else {
#pragma unroll 1
            for (int h = 0; h < g; ++h) {
                bool hit = false;
                for (int j = 0; j < k; ++j) hit |= (srt[h * k + j] == c);
                m |= (hit ? 1u : 0u) << h;
            }
        }
        if (p.sink_pages | p.recent_pages) {
            // anchor pages of the reference kernels (cent_select.cpp:350-362, fused ..._Us2.h:1930-1940): every
            // position whose block_table entry equals that of one of the first `sink` / last `recent` pages
            const int v = physS[i];
            bool forced = false;
            for (int j = 0; j < min(p.sink_pages, pl); ++j) forced |= (v == physS[j]);
            for (int j = max(pl - p.recent_pages, 0); j < pl; ++j) forced |= (v == physS[j]);
            if (forced) m |= (1u << g) - 1u;
        }
        if (has_tail && i == pl - 1) m |= m << 16;
        pm[i] = m;
        if (FLAGS & kSelGather)
            for (int h = 0; h < g; ++h) p.bt_gather[((size_t)b * p.Nq + hq0 + h) * p.maxP + i] = c;
    }
    if (FLAGS & kSelGather)
        for (int i = pl + G.tid; i < p.maxP; i += G.gt)
            for (int h = 0; h < g; ++h) p.bt_gather[((size_t)b * p.Nq + hq0 + h) * p.maxP + i] = 0;
    grp_sync(G);
    trace_clk(p, 29);
    trace_mark(p, 14);

    // ---- compaction: per-head ascending page lists + the group's union list, all warps in parallel.
    // pass 1: per 32-page chunk, the number of set bits of every head plane (and of the union)
    const int nch = (pl + 31) >> 5;
    const unsigned lt = (1u << lane) - 1u;
    // page budget per head: max_page_num, and pl / cap_den when the reference's workload clamp is on
    // (cent_select.cpp:218,410-414: the first pl/8 positions are kept)
    const int cap = (p.cap_den > 0) ? min(p.MP, pl / p.cap_den) : p.MP;
    const int tail_cnt = has_tail ? rem : p.bs;
#pragma unroll 1
    for (int ch = G.warp; ch < nch; ch += G.gw) {
        const int i = ch * 32 + lane;
        const unsigned m = (i < pl) ? pm[i] : 0u;
        // union (what is streamed): only pages whose block_table entry names a block of the pool -- a negative /
        // padding / out-of-range id is skipped, not read (include/spa_b200.h, "invalid block ids"); the per-head
        // position lists keep the golden's semantics and list the page all the same
        const bool inpool = !(FLAGS & kSelUnion) || (i < pl && (unsigned)physS[i] < (unsigned)p.num_blocks);
        int mine = 0;
#pragma unroll 1
        for (int h = 0; h <= g; ++h) {
            const bool on = (h < g) ? ((m >> h) & 1u) : ((m & 0xffffu) != 0u && inpool);
            const int c = __popc(__ballot_sync(0xffffffffu, on));
            if (lane == h) mine = c;
        }
        if (lane <= g) cnt[ch * KG1 + lane] = mine;
    }
    grp_sync(G);
    // exclusive prefix over chunks, one column (head / union) at a time; lanes = chunks
    {
#pragma unroll 1
        for (int col = G.warp; col <= g; col += G.gw) {
            int run = 0;
            for (int c0 = 0; c0 < nch; c0 += 32) {
                const int ch = c0 + lane;
                const int v = (ch < nch) ? cnt[ch * KG1 + col] : 0;
                int inc = v;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const int t = __shfl_up_sync(0xffffffffu, inc, o);
                    if (lane >= o) inc += t;
                }
                if (ch < nch) cnt[ch * KG1 + col] = run + inc - v;
                run += __shfl_sync(0xffffffffu, inc, 31);
            }
            if (lane == 0) {
                if (col < g) { hd[col].total = min(run, cap); if (run > cap) miscI[8] = 1; }
                else miscI[9] = run;
            }
        }
    }
    grp_sync(G);
    int *pp0 = p.page_position ? p.page_position + ((size_t)b * p.Nq + hq0) * p.MP : nullptr;
    if (miscI[8] == 0) {
        const int ulen = miscI[9];
        if ((FLAGS & kSelUnion) && G.tid == 0) {
            // the item goes to the publisher now: its chunks are on offer while pass 2 still writes the entries
            int m = 0;
            for (int h = 0; h < g; ++h) m = max(m, hd[h].total);
            post_publish_request(pub, G.id, pub_seq, item, ulen, tail_cnt, m);
        }
        // pass 2: scatter
        unsigned long long *qent = p.flat + (size_t)item * p.qstride;
#pragma unroll 1
        for (int ch = G.warp; ch < nch; ch += G.gw) {
            const int i = ch * 32 + lane;
            const unsigned m = (i < pl) ? pm[i] : 0u;
#pragma unroll 1
            for (int h = 0; h < g; ++h) {
                const bool on = (m >> h) & 1u;
                const unsigned bal = __ballot_sync(0xffffffffu, on);
                if (on && pp0) pp0[(size_t)h * p.MP + cnt[ch * KG1 + h] + __popc(bal & lt)] = i;
            }
            if (FLAGS & kSelUnion) {
                const bool on = (m & 0xffffu) != 0u && (unsigned)physS[i < pl ? i : 0] < (unsigned)p.num_blocks;
                const unsigned bal = __ballot_sync(0xffffffffu, on);
                if (on) st_relaxed_u64(qent + cnt[ch * KG1 + g] + __popc(bal & lt), queue_entry(p.opt ? i : physS[i], m, (m >> 16) != 0u, item, max(1, tail_cnt)));
            }
        }
        if (p.page_position_length)
            for (int i = G.tid; i < g * kPad; i += G.gt)
                p.page_position_length[((size_t)b * p.Nq + hq0) * kPad + i] = hd[i / kPad].total;
        trace_clk(p, 30);
        if ((FLAGS & kSelUnion) && ulen == 0) zero_item_output(p, b, n2, G.tid, G.gt);
    } else {
        // a head selected more than max_page_num pages: serial path that truncates consistently
        for (int h = G.warp; h < g; h += G.gw) {
            const int c = warp_compact_bit(pm, pl, h, pp0 ? pp0 + (size_t)h * p.MP : nullptr, cap);
            if (p.page_position_length && lane < kPad)
                p.page_position_length[((size_t)b * p.Nq + hq0 + h) * kPad + lane] = c;
            if (lane == 0) hd[h].total = c;
        }
        if (FLAGS & kSelUnion) {
            grp_sync(G);
            if (G.warp == 0) {
                const int ulen = warp_emit_union(p, item, physS, pm, pl, tail_cnt);
                if (ulen == 0) zero_item_output(p, b, n2, lane, 32);
                if (lane == 0) {
                    int m = 0;
                    for (int h = 0; h < g; ++h) m = max(m, hd[h].total);
                    post_publish_request(pub, G.id, pub_seq, item, ulen, tail_cnt, m);
                }
            }
        }
    }
    trace_clk(p, 31);
    trace_mark(p, 15);
    grp_sync(G);
}

// ---- phase A' for the separate attention op: union list from a given block_position ----------
// block_position[b, n2 or hq, j]: logical page of sparse-sequence slot j, 0x7fffffff-terminated
// (sparse_paged_attention_split_Bbn2s2_Us2.h:258,277-279,293).  Slot j covers sparse tokens
// [j*bs, (j+1)*bs) clipped to actual_seq_lengths[b].  Whole CTA.
__device__ inline void union_from_block_position(const ChainParams &p, int item, unsigned char *smem, PubState *pub) {
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
        const int tail_cnt = (tail_slot >= 0) ? (int)(actual % p.bs) : p.bs;
        const int ulen = warp_emit_union(p, item, p.block_table + (size_t)b * p.bt_stride, pm, p.maxP, tail_cnt);
        if (ulen == 0) zero_item_output(p, b, n2, lane_id(), 32);
        if (lane_id() == 0) publish_item(p, pub, item, ulen, tail_cnt, -1);
    }
    __syncthreads();
}

// ---- all items of phase A handled by this CTA's worker groups (warps 0..kConsumerWarps-1) --------
// smem: [ngroups x sel_ns centroid-tile stages][ngroups x group state]
template <typename QT, int FLAGS>
__device__ __forceinline__ void select_phase(const ChainParams &p, const CUtensorMap *tmC, unsigned char *smem, PubState *pub) {
    __shared__ uint64_t sel_bars[4][16];               // per group: full[0..8), empty[8..16)
    const int ngroups = kConsumerWarps / p.sel_gw;
    if ((FLAGS & kSelScores) && threadIdx.x == 0) {
        for (int gi = 0; gi < ngroups; ++gi)
            for (int s = 0; s < p.sel_ns; ++s) { mbar_init(&sel_bars[gi][s], 1); mbar_init(&sel_bars[gi][8 + s], p.sel_gw); }
        fence_barrier_init();
        fence_proxy_async();
        tma_prefetch_desc(tmC);
    }
    __syncthreads();
    if (warp_id() >= kConsumerWarps) {
        // fused chain: the TMA producer warp publishes the groups' items while they select (publish_item)
        if ((FLAGS & kSelUnion) && warp_id() == kConsumerWarps) {
            if (lane_id() == 0) publisher_loop(p, pub, ngroups);
            __syncwarp();
        }
        return;
    }
    const Grp G = make_group(p.sel_gw);
    CentRing R;
    R.base = smem + (size_t)G.id * p.sel_ns * kCentTileBytes;
    R.full = &sel_bars[G.id][0]; R.empty = &sel_bars[G.id][8]; R.ns = p.sel_ns;
    unsigned char *gs = smem + (size_t)ngroups * p.sel_ns * kCentTileBytes + (size_t)G.id * p.sel_group_bytes;
    unsigned tcount = 0, pub_seq = 0;
    for (int item = blockIdx.x + G.id * gridDim.x; item < p.items; item += gridDim.x * ngroups)
        select_item<QT, FLAGS>(p, tmC, item, gs, R, tcount, G, pub, pub_seq);
    if ((FLAGS & kSelUnion) && G.tid == 0) post_publish_request(pub, G.id, pub_seq, -1, 0, 0, 0);
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
