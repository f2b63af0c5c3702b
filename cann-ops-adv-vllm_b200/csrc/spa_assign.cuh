// Centroid maintenance: the step BEFORE the chain.  The chain consumes per-page cluster ids
// block_ids[Nkv, kvPageLen]; the reference never produces them (its tests draw them at random,
// tests/cent_select.py:49-50).  This kernel assigns pages to clusters: the keys of a page are pooled per KV
// head and the page goes to the centroid with the largest inner product -- the same metric ComputeCent
// ranks clusters by for a query (compute_cent.cpp:277-343).
//
// Defined arithmetic (the oracle restates it, oracle/sparse_chain.py: assign_page_clusters):
//   pooled[d] = (s0 + s1) + (s2 + s3), s_w = fp32 sum, in token order, of the valid tokens of the w-th quarter
//               of the page;
//   score[c]  = fp64 sum over d = 0..127, in order, of pooled[d] * cent[c, d]   (every product is exact);
//   block_ids[n2, page] = argmax_c score[c], ties -> lowest c.
// The kernel ranks with fp32 dot products and settles only the clusters within the rounding band of the best
// one in fp64, exactly as the TopK of the chain does (spa_select.cuh).
//
// Bound: HBM (one K page-head slice, bs*D*2 bytes, per unit; the centroid table is L2-resident).
#pragma once
#include "spa_common.cuh"

namespace spa {

struct AssignParams {
    const void *key; long long s_blk, s_row, s_head;     // K pool strides in elements
    const void *cent;                                     // [Nkv, C, D] contiguous
    const int *page_ids;                                  // [n] physical pages to assign
    const int *page_tokens;                               // [n] valid tokens per page, or null (= bs)
    int *block_ids; long long bid_stride;                 // [Nkv, kvPageLen] in/out
    float *page_mean;                                     // optional [n, Nkv, D] fp32 mean key of the page
    const void *aq_scale, *aq_offset;                     // int8 pool: K antiquant [Nkv, D] (or one value) in the centroid dtype
    int aq_per_tensor;
    int n, Nkv, C, bs, num_blocks, kvPageLen;
};

constexpr int kAssignThreads = 256;                     // 8 warps: warp w pools and resolves page w of the unit
constexpr int kAssignPages = 8;                          // pages of one KV head per unit = N of the score mma
constexpr int kAssignPad = 32;                           // 16-bit elements of row padding: 16-byte B loads of a quarter warp hit 32 distinct banks

// fp32 pooled keys enter the tensor cores as a sum of 16-bit pieces: 2 x fp16 (22 bits, after an exact 2^-7
// scaling that keeps 128-token sums of any fp16 keys inside the fp16 range) or 3 x bf16 (24 bits)
template <typename QT> struct AssignSplit;
template <> struct AssignSplit<__half> {
    static constexpr int NS = 2;
    static constexpr float kScale = 0.0078125f;          // 2^-7
    static constexpr float kRel = 4.76837158203125e-07f; // 2^-21: what the pieces miss, relative
    static constexpr float kAbs = 3.814697265625e-06f;   // 128 * 2^-25: fp16 subnormal floor, per unit of max|cent|
    static __device__ __forceinline__ uint32_t abs2(uint32_t v) { return v & 0x7fff7fffu; }
    static __device__ __forceinline__ uint32_t max2(uint32_t a, uint32_t b) {
        __half2 r = __hmax2(*reinterpret_cast<__half2 *>(&a), *reinterpret_cast<__half2 *>(&b));
        return *reinterpret_cast<uint32_t *>(&r);
    }
};
template <> struct AssignSplit<__nv_bfloat16> {
    static constexpr int NS = 3;
    static constexpr float kScale = 1.f;
    static constexpr float kRel = 1.1920928955078125e-07f;   // 2^-23
    static constexpr float kAbs = 0.f;
    static __device__ __forceinline__ uint32_t abs2(uint32_t v) { return v & 0x7fff7fffu; }
    static __device__ __forceinline__ uint32_t max2(uint32_t a, uint32_t b) {
        __nv_bfloat162 r = __hmax2(*reinterpret_cast<__nv_bfloat162 *>(&a), *reinterpret_cast<__nv_bfloat162 *>(&b));
        return *reinterpret_cast<uint32_t *>(&r);
    }
};

struct AssignSmem {
    int ps, pooled, l1, sc, total;
    __host__ __device__ AssignSmem(int C, int ns) {
        int o = 0;
        pooled = o; o += kAssignPages * kD * 4;                                  // fp32 pooled keys
        ps = o;     o += ns * kAssignPages * (kD + kAssignPad) * 2;              // their 16-bit pieces
        l1 = o;     o += kAssignPages * 2 * 4;                                   // per page: l1 of the scaled key, spare
        sc = o;     o += kAssignPages * C * 4;                                   // fp32 scores [page][cluster]
        total = o;
    }
};

// One unit = 8 listed pages x one KV head.  (1) warp w pools page w (HBM stream, 8-byte loads, 16 rows in
// flight per lane); (2) all warps: scores[cluster][page] = cent . pooled on the tensor cores (M = 16 clusters
// per tile, N = 8 pages, K = D), centroid fragments straight from L2 -- the table is read once per 8 pages;
// (3) warp w settles page w: maximum, clusters within the rounding band, exact fp64 scores, argmax.
template <typename QT, bool KVI8>
__global__ void __launch_bounds__(kAssignThreads, 4) spa_assign_kernel(const __grid_constant__ AssignParams p) {
    using SP = AssignSplit<QT>;
    constexpr int NS = SP::NS;
    constexpr int ROW = kD + kAssignPad;
    extern __shared__ __align__(16) unsigned char sm_raw[];
    const AssignSmem L(p.C, NS);
    float *pooled = reinterpret_cast<float *>(sm_raw + L.pooled);
    unsigned short *ps = reinterpret_cast<unsigned short *>(sm_raw + L.ps);
    float *l1s = reinterpret_cast<float *>(sm_raw + L.l1);
    float *sc = reinterpret_cast<float *>(sm_raw + L.sc);
    __shared__ float cmax_s[kAssignThreads / 32];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, gid = lane >> 2, tig = lane & 3;
    const unsigned short *kbase = reinterpret_cast<const unsigned short *>(p.key);
    const int q4 = p.bs >> 2;
    const int groups = (p.n + kAssignPages - 1) / kAssignPages;
    for (int u = blockIdx.x; u < groups * p.Nkv; u += gridDim.x) {
        const int grp = u / p.Nkv, n2 = u % p.Nkv;
        // ---- (1) pooling: this warp's page
        const int i = grp * kAssignPages + warp;
        int page = (i < p.n) ? p.page_ids[i] : -1;
        if (page >= p.num_blocks || page >= p.kvPageLen) page = -1;
        const int ntok = (page < 0) ? 0 : (p.page_tokens ? max(0, min(p.page_tokens[i], p.bs)) : p.bs);
        float quarter[4][4];
        if (KVI8) {
            // int8 pool: the integer sum of the page is exact (no order to define); antiquant once per channel:
            // pooled = (sum + ntok * offset) * scale in fp32 (x = (x_i8 + offset) * scale, ..._Us2.h:2142)
            const signed char *kp = reinterpret_cast<const signed char *>(p.key) + (size_t)max(page, 0) * p.s_blk +
                                    (size_t)n2 * p.s_head + lane * 4;
            int a0 = 0, a1 = 0, a2 = 0, a3 = 0;
            for (int t = 0; t < ntok; t += 16) {
                int v[16];
#pragma unroll
                for (int j = 0; j < 16; ++j)
                    v[j] = __ldcs(reinterpret_cast<const int *>(kp + (size_t)min(t + j, ntok - 1) * p.s_row));
#pragma unroll
                for (int j = 0; j < 16; ++j) {
                    if (t + j < ntok) {
                        a0 = __dp4a(v[j], 0x00000001, a0); a1 = __dp4a(v[j], 0x00000100, a1);
                        a2 = __dp4a(v[j], 0x00010000, a2); a3 = __dp4a(v[j], 0x01000000, a3);
                    }
                }
            }
            const int acc[4] = {a0, a1, a2, a3};
            const unsigned short *sv = reinterpret_cast<const unsigned short *>(p.aq_scale);
            const unsigned short *ov = reinterpret_cast<const unsigned short *>(p.aq_offset);
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const int d = p.aq_per_tensor ? 0 : n2 * kD + lane * 4 + e;
                const float off = ov ? QTraits<QT>::to_float(ov[d]) : 0.f;
                const float sc = QTraits<QT>::to_float(sv[d]);
                quarter[0][e] = __fmul_rn(__fadd_rn((float)acc[e], __fmul_rn((float)ntok, off)), sc);
                quarter[1][e] = quarter[2][e] = quarter[3][e] = 0.f;          // (x + 0) + (0 + 0) = x exactly
            }
        } else {
        {
            const unsigned short *kp = kbase + (size_t)max(page, 0) * p.s_blk + (size_t)n2 * p.s_head + lane * 4;
#pragma unroll
            for (int w = 0; w < 4; ++w) {
                float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
                const int t1 = min((w + 1) * q4, ntok);
                for (int t = w * q4; t < t1; t += 16) {
                    uint2 v[16];
#pragma unroll
                    for (int j = 0; j < 16; ++j)
                        v[j] = __ldcs(reinterpret_cast<const uint2 *>(kp + (size_t)min(t + j, t1 - 1) * p.s_row));   // all 16 in flight
#pragma unroll
                    for (int j = 0; j < 16; ++j) {
                        if (t + j < t1) {                     // token order, one fp32 add per element
                            s0 += QTraits<QT>::to_float((unsigned short)(v[j].x & 0xffffu));
                            s1 += QTraits<QT>::to_float((unsigned short)(v[j].x >> 16));
                            s2 += QTraits<QT>::to_float((unsigned short)(v[j].y & 0xffffu));
                            s3 += QTraits<QT>::to_float((unsigned short)(v[j].y >> 16));
                        }
                    }
                }
                quarter[w][0] = s0; quarter[w][1] = s1; quarter[w][2] = s2; quarter[w][3] = s3;
            }
        }
        }
        __syncthreads();                                      // the previous unit is done with the shared buffers
        {
            float l1 = 0.f;
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const float v = (quarter[0][e] + quarter[1][e]) + (quarter[2][e] + quarter[3][e]);
                const int d = lane * 4 + e;
                pooled[warp * kD + d] = v;
                if (p.page_mean && page >= 0) p.page_mean[((size_t)i * p.Nkv + n2) * kD + d] = ntok > 0 ? v / (float)ntok : 0.f;
                float r = v * SP::kScale;
                l1 += fabsf(r);
#pragma unroll
                for (int sidx = 0; sidx < NS; ++sidx) {
                    const unsigned short h = QTraits<QT>::from_float(r);
                    ps[(sidx * kAssignPages + warp) * ROW + d] = h;
                    r -= QTraits<QT>::to_float(h);            // exact: h is r rounded to fewer bits
                }
            }
#pragma unroll
            for (int o = 16; o; o >>= 1) l1 += __shfl_xor_sync(0xffffffffu, l1, o);
            if (lane == 0) l1s[warp * 2] = l1;
        }
        __syncthreads();
        // ---- (2) scores on the tensor cores: tile = 16 clusters x 8 pages, K = D in 8 steps
        uint32_t cm = 0u;
        const unsigned short *cbase = reinterpret_cast<const unsigned short *>(p.cent) + (size_t)n2 * p.C * kD;
        for (int c0 = warp * 16; c0 < p.C; c0 += (kAssignThreads / 32) * 16) {
            float acc[4] = {0.f, 0.f, 0.f, 0.f};
            const int r0 = min(c0 + gid, p.C - 1), r1 = min(c0 + gid + 8, p.C - 1);
            // The k order inside an mma is free as long as A and B agree.  Lane tig takes the 16 contiguous
            // bytes [32*kp + 8*tig, +8) of its two rows for each pair kp of k-steps: the whole tile is 8
            // independent 16-byte loads per lane (one L2 round trip), and the B pieces are read from the
            // same dims.
            const uint4 *a_lo = reinterpret_cast<const uint4 *>(cbase + (size_t)r0 * kD) + tig;
            const uint4 *a_hi = reinterpret_cast<const uint4 *>(cbase + (size_t)r1 * kD) + tig;
            uint4 alo[kD / 32], ahi[kD / 32];
#pragma unroll
            for (int kp = 0; kp < kD / 32; ++kp) { alo[kp] = __ldg(a_lo + kp * 4); ahi[kp] = __ldg(a_hi + kp * 4); }
#pragma unroll
            for (int kp = 0; kp < kD / 32; ++kp) {
                cm = SP::max2(cm, SP::max2(SP::max2(SP::abs2(alo[kp].x), SP::abs2(alo[kp].y)), SP::max2(SP::abs2(alo[kp].z), SP::abs2(alo[kp].w))));
                cm = SP::max2(cm, SP::max2(SP::max2(SP::abs2(ahi[kp].x), SP::abs2(ahi[kp].y)), SP::max2(SP::abs2(ahi[kp].z), SP::abs2(ahi[kp].w))));
#pragma unroll
                for (int sidx = 0; sidx < NS; ++sidx) {
                    const uint4 b = *reinterpret_cast<const uint4 *>(ps + (sidx * kAssignPages + gid) * ROW + kp * 32 + tig * 8);
                    mma_16816<QT>(acc, alo[kp].x, ahi[kp].x, alo[kp].y, ahi[kp].y, b.x, b.y);
                    mma_16816<QT>(acc, alo[kp].z, ahi[kp].z, alo[kp].w, ahi[kp].w, b.z, b.w);
                }
            }
            // accumulator: rows gid / gid + 8 (clusters), columns 2*tig, 2*tig + 1 (pages)
            if (c0 + gid < p.C) { sc[(2 * tig) * p.C + c0 + gid] = acc[0]; sc[(2 * tig + 1) * p.C + c0 + gid] = acc[1]; }
            if (c0 + gid + 8 < p.C) { sc[(2 * tig) * p.C + c0 + gid + 8] = acc[2]; sc[(2 * tig + 1) * p.C + c0 + gid + 8] = acc[3]; }
        }
        {
            float c = fmaxf(QTraits<QT>::to_float((unsigned short)(cm & 0xffffu)), QTraits<QT>::to_float((unsigned short)(cm >> 16)));
#pragma unroll
            for (int o = 16; o; o >>= 1) c = fmaxf(c, __shfl_xor_sync(0xffffffffu, c, o));
            if (lane == 0) cmax_s[warp] = c;
        }
        __syncthreads();
        // ---- (3) this warp's page: band around the maximum, exact scores inside it
        if (page >= 0) {
            float cmax = 0.f;
#pragma unroll
            for (int w = 0; w < kAssignThreads / 32; ++w) cmax = fmaxf(cmax, cmax_s[w]);
            const float *row = sc + warp * p.C;
            float smax = -INFINITY;
            for (int c = lane; c < p.C; c += 32) smax = fmaxf(smax, row[c]);
#pragma unroll
            for (int o = 16; o; o >>= 1) smax = fmaxf(smax, __shfl_xor_sync(0xffffffffu, smax, o));
            // |tensor-core score - exact| <= cmax * (l1 * (2^-16 accumulation + pieces) + subnormal floor)
            const float eps = 1.05f * cmax * (l1s[warp * 2] * (1.52587890625e-05f + SP::kRel) + SP::kAbs);
            const float cut = smax - 2.f * eps;
            double bv = -INFINITY;
            int bi = 0x7fffffff;
            const float *pv = pooled + warp * kD;
            for (int c = lane; c < p.C; c += 32) {
                if (!(row[c] >= cut)) continue;
                const uint4 *cr = reinterpret_cast<const uint4 *>(cbase + (size_t)c * kD);
                double acc = 0.0;
#pragma unroll 1
                for (int half = 0; half < 2; ++half) {            // the row in two batches of 8 independent loads
                    uint4 w[8];
#pragma unroll
                    for (int q = 0; q < 8; ++q) w[q] = __ldg(cr + half * 8 + q);
#pragma unroll
                    for (int q = 0; q < 8; ++q) {
                        const uint32_t ww[4] = {w[q].x, w[q].y, w[q].z, w[q].w};
#pragma unroll
                        for (int e = 0; e < 4; ++e) {              // d ascending: the defined summation order
                            const int d = half * 64 + q * 8 + e * 2;
                            acc += (double)pv[d] * QTraits<QT>::to_double((unsigned short)(ww[e] & 0xffffu));
                            acc += (double)pv[d + 1] * QTraits<QT>::to_double((unsigned short)(ww[e] >> 16));
                        }
                    }
                }
                if (acc > bv || (acc == bv && c < bi)) { bv = acc; bi = c; }
            }
#pragma unroll
            for (int o = 16; o; o >>= 1) {
                const double ov = __shfl_xor_sync(0xffffffffu, bv, o);
                const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
                if (ov > bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
            }
            if (lane == 0) p.block_ids[(size_t)n2 * p.bid_stride + page] = bi;
        }
    }
}

}  // namespace spa
