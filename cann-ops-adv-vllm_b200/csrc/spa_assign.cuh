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
    int n, Nkv, C, bs, num_blocks, kvPageLen;
};

constexpr int kAssignThreads = 128;

template <typename QT>
__global__ void __launch_bounds__(kAssignThreads, 8) spa_assign_kernel(const __grid_constant__ AssignParams p) {
    extern __shared__ float sm_scores[];                  // [C]
    __shared__ float pool[4][kD];
    __shared__ float pooled[kD];
    __shared__ float red[8];
    __shared__ double bestv[4];
    __shared__ int besti[4];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const unsigned short *kbase = reinterpret_cast<const unsigned short *>(p.key);
    const int q4 = p.bs >> 2;                             // tokens per warp
    for (int u = blockIdx.x; u < p.n * p.Nkv; u += gridDim.x) {
        const int i = u / p.Nkv, n2 = u % p.Nkv;
        const int page = p.page_ids[i];
        if (page < 0 || page >= p.num_blocks || page >= p.kvPageLen) continue;     // uniform per CTA
        const int ntok = p.page_tokens ? max(0, min(p.page_tokens[i], p.bs)) : p.bs;
        // ---- pooled key: warp w sums its quarter of the page in token order, lane = 4 head dims
        const unsigned short *kp = kbase + (size_t)page * p.s_blk + (size_t)n2 * p.s_head + lane * 4;
        float s[4] = {0.f, 0.f, 0.f, 0.f};
        const int t0 = warp * q4, t1 = min(t0 + q4, ntok);
        for (int t = t0; t < t1; t += 8) {
            uint2 v[8];
#pragma unroll
            for (int j = 0; j < 8; ++j)
                v[j] = (t + j < t1) ? __ldcs(reinterpret_cast<const uint2 *>(kp + (size_t)(t + j) * p.s_row)) : make_uint2(0u, 0u);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                if (t + j < t1) {
                    s[0] += QTraits<QT>::to_float((unsigned short)(v[j].x & 0xffffu));
                    s[1] += QTraits<QT>::to_float((unsigned short)(v[j].x >> 16));
                    s[2] += QTraits<QT>::to_float((unsigned short)(v[j].y & 0xffffu));
                    s[3] += QTraits<QT>::to_float((unsigned short)(v[j].y >> 16));
                }
            }
        }
        __syncthreads();                                   // previous unit done with pool/pooled/scores
#pragma unroll
        for (int e = 0; e < 4; ++e) pool[warp][lane * 4 + e] = s[e];
        __syncthreads();
        {
            const float v = (pool[0][tid] + pool[1][tid]) + (pool[2][tid] + pool[3][tid]);
            pooled[tid] = v;
            if (p.page_mean) p.page_mean[((size_t)i * p.Nkv + n2) * kD + tid] = ntok > 0 ? v / (float)ntok : 0.f;
            float a = fabsf(v);
#pragma unroll
            for (int o = 16; o; o >>= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
            if (lane == 0) red[warp] = a;
        }
        __syncthreads();
        const float l1 = (red[0] + red[1]) + (red[2] + red[3]);
        // ---- fp32 scores of every cluster, one cluster per thread and pass
        const uint4 *crow = reinterpret_cast<const uint4 *>(p.cent) + (size_t)n2 * p.C * (kD / 8);
        float smax = -INFINITY, cmax = 0.f;
        for (int c = tid; c < p.C; c += kAssignThreads) {
            float acc = 0.f;
#pragma unroll 4
            for (int d8 = 0; d8 < kD / 8; ++d8) {
                const uint4 w = __ldg(crow + (size_t)c * (kD / 8) + d8);
                const uint32_t ww[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const float c0 = QTraits<QT>::to_float((unsigned short)(ww[e] & 0xffffu));
                    const float c1 = QTraits<QT>::to_float((unsigned short)(ww[e] >> 16));
                    acc = fmaf(pooled[d8 * 8 + e * 2], c0, acc);
                    acc = fmaf(pooled[d8 * 8 + e * 2 + 1], c1, acc);
                    cmax = fmaxf(cmax, fmaxf(fabsf(c0), fabsf(c1)));
                }
            }
            sm_scores[c] = acc;
            smax = fmaxf(smax, acc);
        }
#pragma unroll
        for (int o = 16; o; o >>= 1) {
            smax = fmaxf(smax, __shfl_xor_sync(0xffffffffu, smax, o));
            cmax = fmaxf(cmax, __shfl_xor_sync(0xffffffffu, cmax, o));
        }
        if (lane == 0) { red[warp] = smax; red[4 + warp] = cmax; }
        __syncthreads();
        smax = fmaxf(fmaxf(red[0], red[1]), fmaxf(red[2], red[3]));
        cmax = fmaxf(fmaxf(red[4], red[5]), fmaxf(red[6], red[7]));
        // |fp32 score - exact| <= 128 * 2^-24 * sum|pooled*cent| <= 2^-17 * l1 * cmax; twice that as the margin
        const float eps = 1.01f * 1.52587890625e-05f * l1 * cmax;
        const float cut = smax - 2.f * eps;
        // ---- exact scores inside the band
        double bv = -INFINITY;
        int bi = 0x7fffffff;
        for (int c = tid; c < p.C; c += kAssignThreads) {
            if (!(sm_scores[c] >= cut)) continue;
            const unsigned short *cr = reinterpret_cast<const unsigned short *>(p.cent) + ((size_t)n2 * p.C + c) * kD;
            double acc = 0.0;
            for (int d = 0; d < kD; ++d) acc += (double)pooled[d] * QTraits<QT>::to_double(cr[d]);
            if (acc > bv || (acc == bv && c < bi)) { bv = acc; bi = c; }
        }
#pragma unroll
        for (int o = 16; o; o >>= 1) {
            const double ov = __shfl_xor_sync(0xffffffffu, bv, o);
            const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
            if (ov > bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
        }
        if (lane == 0) { bestv[warp] = bv; besti[warp] = bi; }
        __syncthreads();
        if (tid == 0) {
            for (int w = 1; w < 4; ++w)
                if (bestv[w] > bv || (bestv[w] == bv && besti[w] < bi)) { bv = bestv[w]; bi = besti[w]; }
            p.block_ids[(size_t)n2 * p.bid_stride + page] = bi;
        }
    }
}

}  // namespace spa
