/*
 * spa_b200.h -- C ABI of the B200-native centroid-sparse paged decode-attention chain.
 *
 * This is the drop-in boundary: it replaces the reference's aclnn two-phase C interface
 *   aclnn<Op>GetWorkspaceSize(tensors..., attrs..., outputs..., uint64_t *workspaceSize, aclOpExecutor **executor)
 *   aclnn<Op>(void *workspace, uint64_t workspaceSize, aclOpExecutor *executor, aclrtStream stream)
 * (reference: README.md:125-181, misc/aclnn_compute_cent.h, misc/aclnn_select_position.h,
 *  misc/aclnn_cent_select.h:27-51, src/transformer/sparse_paged_attention/ophost/aclnn_sparse_paged_attention.h,
 *  src/transformer/sparse_paged_fusion_attention/ophost/aclnn_sparse_paged_fusion_attention.h:60-81)
 * which the reference's PyTorch binding resolves with dlsym from libcust_opapi.so
 * (extension/csrc/pytorch_npu_helper.hpp:225-260).  Same shape here: plain C types only, no torch
 * types, no exceptions; phase 1 validates (where the reference's tiling checks live) and returns a
 * workspace size plus an executor ("plan"), phase 2 launches asynchronously on the given stream.
 *
 * Ownership (as in the reference, extension/csrc/add_custom.cpp:153,278-280,866-873): the caller
 * allocates every output and the workspace; kernels never allocate.  An executor stays valid (and
 * re-launchable, e.g. under CUDA-graph capture) as long as the tensors it was planned for live;
 * release it with spaDestroyExecutor.
 *
 * Optional tensors: pass NULL, or a descriptor with data == NULL / zero elements
 * (the reference maps empty at::Tensor to a null aclTensor, add_custom.cpp:32-35).
 */
#ifndef SPA_B200_H_
#define SPA_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SPA_API __attribute__((visibility("default")))

typedef enum {
    SPA_OK = 0,
    SPA_ERR_ARG = -1,          /* null pointer / malformed descriptor            */
    SPA_ERR_SHAPE = -2,        /* shapes inconsistent with each other or attrs   */
    SPA_ERR_DTYPE = -3,        /* unsupported dtype combination                  */
    SPA_ERR_UNSUPPORTED = -4,  /* legal in the reference op, not in this build   */
    SPA_ERR_CUDA = -5,         /* CUDA runtime / driver error                    */
    SPA_ERR_WORKSPACE = -6     /* workspace pointer null or too small            */
} spaStatus;

/* dtype codes (subset the reference binding converts, add_custom.cpp:43-70) */
typedef enum {
    SPA_F16 = 0, SPA_BF16 = 1, SPA_F32 = 2, SPA_I8 = 3, SPA_I32 = 4, SPA_I64 = 5,
    SPA_F8E4M3 = 6,            /* FP8 e4m3 KV pages (attention ops only; no reference analogue) */
    SPA_U8 = 7, SPA_BOOL = 8   /* attenMask (one byte per position; the reference takes bool / int8 / uint8)       */
} spaDtype;

#define SPA_MAX_DIMS 8
typedef struct {
    void *data;                    /* device pointer                                 */
    int32_t dtype;                 /* spaDtype                                       */
    int32_t ndim;
    int64_t sizes[SPA_MAX_DIMS];
    int64_t strides[SPA_MAX_DIMS]; /* in ELEMENTS                                    */
} spaTensor;

typedef struct spaExecutor spaExecutor;     /* replaces aclOpExecutor  */
typedef struct CUstream_st *spaStream;      /* == cudaStream_t         */

/* Thread-local text of the last error on this thread (replaces aclGetRecentErrMsg,
 * add_custom.cpp:263-265). */
SPA_API const char *spaGetLastError(void);
SPA_API int spaDestroyExecutor(spaExecutor *executor);
/* Library/ABI version, and the SM architecture the kernels were compiled for (100 => sm_100a). */
SPA_API int spaGetVersion(void);
SPA_API int spaGetCompiledArch(void);
/* Number of kernel launches an executor performs per spa<Op>() call (for bench accounting). */
SPA_API int spaExecutorNumLaunches(const spaExecutor *executor);
/* Profiling hook (the reference has device printf macros gated by V5_SPARSE_DEBUG_ENABLE,
 * sparse_paged_attention_split_Bbn2s2_Us2.h:28-36): when set, every CTA of the chain kernel stores
 * %globaltimer (ns) at its phase boundaries into deviceBuffer[grid][32] (uint64):
 * 0 start, 1 selection done, 3 page stream + combine done, 5 exit; 2 = last queue claim,
 * 4 = pages streamed, 6 = runs, 7 = producer done; 8..15 = stages of the CTA's last selection item;
 * 16..31 = SM clock probes inside the selection.  NULL switches it off (default).              */
SPA_API int spaExecutorGridSize(const spaExecutor *executor);
SPA_API int spaExecutorSetTrace(spaExecutor *executor, void *deviceBuffer);

/* The chain kernels keep a page queue and counters in the first part of the workspace; they must be
 * zero when a launch starts, and every launch leaves them zero again.  By default the library clears
 * them before each launch (one cudaMemsetAsync on the stream), because an aclnn-style workspace is
 * scratch the caller may reuse for anything between calls.  A caller that re-launches one executor on
 * the SAME workspace and never lets anything else write to it (a decode loop, a captured CUDA graph)
 * may set this flag: the workspace is then cleared once, on the first launch that sees it.       */
SPA_API int spaExecutorSetPersistentWorkspace(spaExecutor *executor, int enable);

/* Selection options the reference KERNELS hard-code but its golden (tests/utils.py) lacks -- off by default,
 * so the default behaviour is the golden's.  For CentSelect and SparsePagedFusionAttention.
 *   sinkPages, recentPages: every logical page whose block_table entry equals that of one of the first
 *       sinkPages / last recentPages pages of the sequence is selected for every head (reference: 1 and 2,
 *       cent_select.cpp:350-362; fused kernel sparse_paged_fusion_attention_split_Bbn2s2_Us2.h:1930-1940).
 *   maxFractionDen > 0: a head keeps at most ceil(seq/blockSize) / maxFractionDen pages, the lowest positions
 *       first (reference: 8, cent_select.cpp:218,410-414).  Entries of pagePosition beyond the kept length are
 *       padded with 0x7fffffff here (the reference leaves them; its consumers read only `length` entries). */
SPA_API int spaExecutorSetSelectOptions(spaExecutor *executor, int sinkPages, int recentPages, int maxFractionDen);

/* Optional softmax log-sum-exp output of the attention ops (the reference kernels' softmaxLse,
 * sparse_paged_attention_split_Bbn2s2_Us2.h:2601-2623): lse [B, Nq] (or [B, Nq, 1]) float32, caller-allocated,
 * lse[b,h] = ln(sum_i exp(scale * q.k_i)) over the tokens the row attended to; -inf when it attended to none.
 * Needed to merge this result with attention over other token ranges (chunked / prefix merging).
 * With int8 KV the kernel's logits differ from the true ones by a per-head constant the softmax cancels (K
 * offset, conversion bias); the log-sum-exp puts it back.  NULL switches it off.                           */
SPA_API int spaExecutorSetSoftmaxLse(spaExecutor *executor, const spaTensor *lse);

/* ---- ComputeCent: replaces aclnnComputeCentGetWorkspaceSize / aclnnComputeCent
 *      (misc/aclnn_compute_cent.h; kernel compute_cent/compute_cent.cpp:345).
 * query [B,Nq,D] (or [B,Nq,1,D]) f16/bf16, l1Cent [Nkv,C,D] same dtype,
 * indicesOut [B,Nq,k] int32 -- k is read from the OUTPUT's last dim exactly as the reference
 * tiling does (compute_cent_tiling.cpp:45-50).  Result: k largest q.cent scores per (b,head),
 * sorted descending, ties -> lowest index.                                                     */
SPA_API int spaComputeCentGetWorkspaceSize(const spaTensor *query, const spaTensor *l1Cent,
                                           const spaTensor *indicesOut,
                                           uint64_t *workspaceSize, spaExecutor **executor);
SPA_API int spaComputeCent(void *workspace, uint64_t workspaceSize, spaExecutor *executor,
                           spaStream stream);

/* ---- SelectPosition: replaces aclnnSelectPosition* (misc/aclnn_select_position.h; kernel
 *      select_position/select_position.cpp:242).  All int32.
 * blockIds [Nkv,kvPageLen], blockTable [maxB,maxP], seqLen [B], indices [B,Nq,k];
 * pagePositionOut [B,Nq,MP] (pad 0x7fffffff), pagePositionLengthOut [B,Nq,8],
 * blockTableGatherOut [B,Nq,maxP].  blockSize: tokens per page (reference constant 128,
 * select_position.cpp:7).                                                                      */
SPA_API int spaSelectPositionGetWorkspaceSize(const spaTensor *blockIds, const spaTensor *blockTable,
                                              const spaTensor *seqLen, const spaTensor *indices,
                                              int64_t blockSize,
                                              const spaTensor *pagePositionOut,
                                              const spaTensor *pagePositionLengthOut,
                                              const spaTensor *blockTableGatherOut,
                                              uint64_t *workspaceSize, spaExecutor **executor);
SPA_API int spaSelectPosition(void *workspace, uint64_t workspaceSize, spaExecutor *executor,
                              spaStream stream);

/* ---- CentSelect: replaces aclnnCentSelect* (misc/aclnn_cent_select.h:27-51; kernel
 *      cent_select/cent_select.cpp:538).  topK: reference constant 64 (cent_select_tiling.cpp:58).
 * maxPagePositionLengthOut [B,8] int64 = blockSize * max over heads of the count.
 * Batch is seqLen's (add_custom.cpp:988).                                                      */
SPA_API int spaCentSelectGetWorkspaceSize(const spaTensor *query, const spaTensor *l1Cent,
                                          const spaTensor *blockIds, const spaTensor *blockTable,
                                          const spaTensor *seqLen, int64_t topK, int64_t blockSize,
                                          const spaTensor *pagePositionOut,
                                          const spaTensor *pagePositionLengthOut,
                                          const spaTensor *maxPagePositionLengthOut,
                                          uint64_t *workspaceSize, spaExecutor **executor);
SPA_API int spaCentSelect(void *workspace, uint64_t workspaceSize, spaExecutor *executor,
                          spaStream stream);

/* ---- SparsePagedAttention: replaces aclnnSparsePagedAttention* (ophost/aclnn_sparse_paged_attention.h;
 *      kernel sparse_paged_attention/sparse_paged_attention.cpp:135).  Argument order follows the
 *      aclnn prototype: inputs, attrs, output.  key/value are single tensors (the reference passes
 *      1-element tensor lists, add_custom.py:70-75).
 * actualSeqLengths: DEVICE int64 [B] or [1] (the reference copies it to the host,
 *      add_custom.cpp:105-118; here it is read on device, no sync).
 * key / value dtype: the query dtype, int8 with antiquantScale (+ optional antiquantOffset) [2,Nkv,D] or [2] in the
 *      query dtype (x = (x_i8 + offset) * scale; row 0 = K, row 1 = V), or SPA_F8E4M3 (FP8 e4m3 pages -- no
 *      reference analogue; scale / offset optional with the same formula, absent = the stored numbers).
 * blockPosition [B, Nkv or Nq, MP] int32, 0x7fffffff-terminated.  A logical page listed more than once in a
 *      row is attended ONCE (the reference walks the slots and would weigh it twice; no reference test
 *      repeats a page).
 * Invalid block ids (both attention ops): a selected logical page whose blocktable entry is negative or
 *      >= the pool's block count (padding such as -1) is SKIPPED -- it contributes no tokens to the
 *      softmax; it is never read.  blockPosition / pagePositionLength of the fused op still list it, as the
 *      reference golden does (tests/utils.py:57-80 clamps the table value for the block_ids lookup only).
 *      At most 2^24 blocks per pool.
 * Optional inputs of the op definition (sparse_paged_attention_def.cpp:70-190; both attention ops, every KV dtype,
 * layout and block size; they select the "options build" of the kernel, csrc/spa_chain.cuh -- absent, the plain
 * build runs).  No reference test passes any of them (tests/ifa_v5_case.py:133-146):
 *   pseShift   [B or 1, Nq, 1, S] in the query dtype, S >= maxBlocksPerBatch * blockSize
 *              (sparse_paged_attention_tiling_check.cc:737-803): added to the scaled logits
 *              (sparse_paged_attention_split_Bbn2s2_Us2.h:2862-2874).
 *   attenMask  [B, (1, (1,)) S] bool / int8 / uint8, same S (tiling_check.cc:805-847): a non-zero byte SETS the
 *              logit to -1e12 after pseShift (:2876-2899, :757) -- finite, so a row whose tokens are all masked
 *              averages them.
 *     Both are indexed by the LOGICAL token position  page * blockSize + t  of the KV sequence.  The reference
 *     kernel, whose code is inherited from the dense op, indexes them by the position in the WALKED sequence
 *     (slot j of blockPosition -> positions j*blockSize ..): the same thing for a dense walk, but for a sparse walk
 *     it would make a mask or an ALiBi-style bias refer to compacted positions that differ from head to head;
 *     its own sources mark blockPosition support of such per-position inputs as unfinished (:1912 "TODO").
 *   quantScale2 (+ quantOffset2)  float32, or bfloat16 with a bfloat16 query; 1 element or [Nq, D]
 *              (sparse_paged_attention_tiling.cc:566-647): attentionOut must then be int8 and receives
 *              sat_s8(rint(o * scale2 + offset2)) (PostQuant, :3513-3600); rows that attended to nothing are 0.
 *              Not together with peer output (tensor parallelism).
 * dequantScale1, quantScale1, dequantScale2 (int8 QUERY: the sparse path takes f16 / bf16 queries only) and
 * kvPaddingSize (left padding, undefined for paged KV) must be absent (SPA_ERR_UNSUPPORTED otherwise).
 * Per-token antiquant (antiquant modes 1-5 of the tiling, sparse_paged_attention_tiling.cc:940-987) cannot be
 * requested through this op in the reference either: its definition has no antiquant_mode attribute and no
 * key/value_antiquant_scale inputs (sparse_paged_attention_def.cpp:200-310), ConvertContext leaves antiquantMode
 * null (tiling.cc:2035-2052, :2156), so the mode is always 0 = per channel / per tensor -- which is built.      */
SPA_API int spaSparsePagedAttentionGetWorkspaceSize(
    const spaTensor *query, const spaTensor *key, const spaTensor *value,
    const spaTensor *pseShift, const spaTensor *attenMask, const spaTensor *actualSeqLengths,
    const spaTensor *dequantScale1, const spaTensor *quantScale1, const spaTensor *dequantScale2,
    const spaTensor *quantScale2, const spaTensor *quantOffset2,
    const spaTensor *antiquantScale, const spaTensor *antiquantOffset,
    const spaTensor *blocktable, const spaTensor *kvPaddingSize, const spaTensor *blockPosition,
    int64_t numHeads, double scaleValue, const char *inputLayout, int64_t numKeyValueHeads,
    int64_t blockSize, int64_t innerPrecise,
    const spaTensor *attentionOut, uint64_t *workspaceSize, spaExecutor **executor);
SPA_API int spaSparsePagedAttention(void *workspace, uint64_t workspaceSize, spaExecutor *executor,
                                    spaStream stream);

/* ---- SparsePagedFusionAttention: replaces aclnnSparsePagedFusionAttention*
 *      (ophost/aclnn_sparse_paged_fusion_attention.h:60-81; kernel
 *      sparse_paged_fusion_attention/sparse_paged_fusion_attention.cpp:156).  Argument order as in
 *      the aclnn prototype: inputs, attrs, the three aux outputs, attentionOut.
 * One launch: CentSelect (writes blockPosition [B,Nq,MP], pagePositionLength [B,Nq,8],
 * maxPagePositionLength [B,8] in place) then attention over the selected pages.
 * Exactness of the selection: scores are defined by EXACT arithmetic (fp16/bf16 products summed without
 * rounding; ties -> lowest index), which float64 accumulation delivers while, per (head, cluster) dot
 * product, the largest and the smallest non-zero |q_d * cent_d| lie within 2^40 of each other (always true
 * for fp16; for bf16 it excludes tables that mix magnitudes like 1e-30 and 1e+3 in one row).  Outside that
 * range two scores closer than float64 rounding of the largest term may be ordered either way.
 * topK: reference tiling constant (cent_select: 64).  actualSeqLengths is accepted and ignored
 * (the reference kernel substitutes the on-device maxPagePositionLength,
 * sparse_paged_fusion_attention.cpp:68).                                                        */
SPA_API int spaSparsePagedFusionAttentionGetWorkspaceSize(
    const spaTensor *query, const spaTensor *key, const spaTensor *value,
    const spaTensor *pseShift, const spaTensor *attenMask, const spaTensor *actualSeqLengths,
    const spaTensor *dequantScale1, const spaTensor *quantScale1, const spaTensor *dequantScale2,
    const spaTensor *quantScale2, const spaTensor *quantOffset2,
    const spaTensor *antiquantScale, const spaTensor *antiquantOffset,
    const spaTensor *blocktable, const spaTensor *kvPaddingSize,
    const spaTensor *l1Cent, const spaTensor *blockIds, const spaTensor *totalSeqLen,
    int64_t numHeads, double scaleValue, const char *inputLayout, int64_t numKeyValueHeads,
    int64_t blockSize, int64_t innerPrecise, int64_t topK,
    const spaTensor *blockPosition, const spaTensor *pagePositionLength,
    const spaTensor *maxPagePositionLength,
    const spaTensor *attentionOut, uint64_t *workspaceSize, spaExecutor **executor);
SPA_API int spaSparsePagedFusionAttention(void *workspace, uint64_t workspaceSize,
                                          spaExecutor *executor, spaStream stream);

/* ---- SparsePagedFusionAttentionStatic: allocation-free, single-call variant of the fused op.  It has NO
 *      counterpart in the reference; it was added for callers that
 *      cannot call cudaMalloc (code under CUDA-graph capture, tight serving loops, harnesses that forbid it).
 * Same tensors and attributes as spaSparsePagedFusionAttentionGetWorkspaceSize (same validation, same
 * kernel, default select options, no softmax LSE), launched at once on `stream`: the caller passes inputs
 * and outputs only.  The workspace is a file-scope `static __device__` array of SPA_STATIC_SCRATCH_BYTES
 * inside the library -- every GPU that loads the library holds that much device memory until the process
 * exits -- and the plan lives on the stack: the call performs no cudaMalloc / cudaFree and no host or
 * library allocation.  A problem whose workspace exceeds the array (B*Nkv work items x (page entries +
 * 9 or 16 fp32 partials) + exact-score scratch: BASELINE configs c2 / c3 / c4 need 13.7 / 25.1 / 31.1 MB) is
 * refused with SPA_ERR_WORKSPACE and a message naming both sizes; use the two-phase entry points then.
 * All calls on a device share the one array, which keeps the previous call's contents: the words that
 * must start at zero are cleared on `stream` in front of every launch, everything else is written before
 * it is read.  Calls that OVERLAP on the GPU (two streams, two host threads, a graph replay next to an
 * eager call) race on the array -- serialise them on one stream.                                      */
#define SPA_STATIC_SCRATCH_BYTES (32u << 20)
SPA_API uint64_t spaStaticScratchBytes(void);
SPA_API int spaSparsePagedFusionAttentionStatic(
    const spaTensor *query, const spaTensor *key, const spaTensor *value,
    const spaTensor *pseShift, const spaTensor *attenMask, const spaTensor *actualSeqLengths,
    const spaTensor *dequantScale1, const spaTensor *quantScale1, const spaTensor *dequantScale2,
    const spaTensor *quantScale2, const spaTensor *quantOffset2,
    const spaTensor *antiquantScale, const spaTensor *antiquantOffset,
    const spaTensor *blocktable, const spaTensor *kvPaddingSize,
    const spaTensor *l1Cent, const spaTensor *blockIds, const spaTensor *totalSeqLen,
    int64_t numHeads, double scaleValue, const char *inputLayout, int64_t numKeyValueHeads,
    int64_t blockSize, int64_t innerPrecise, int64_t topK,
    const spaTensor *blockPosition, const spaTensor *pagePositionLength,
    const spaTensor *maxPagePositionLength,
    const spaTensor *attentionOut, spaStream stream);

/* ---- AssignPageClusters: centroid maintenance, the step BEFORE the chain.  The reference
 *      consumes blockIds [Nkv, kvPageLen] but never produces them (tests/cent_select.py:49-50 draws them at
 *      random), so there is no aclnn interface to cite; the two-phase shape is the same.
 * For every listed physical page and every KV head: pool the page's keys (fp32 sum of the valid tokens) and
 * store the index of the centroid with the largest inner product -- the metric ComputeCent ranks clusters by
 * (compute_cent.cpp:277-343) -- into blockIds[n2, page]; ties -> lowest index.  Run it when a page fills up
 * (or for every page after the centroids changed).
 * key [blocks,bs,Nkv,D] / [blocks,bs,Nkv*D] / [blocks,Nkv,bs,D] in the dtype of l1Cent (f16/bf16), or int8 with
 * the attention ops' antiquantScale / antiquantOffset ([2,Nkv,D] or [2] in the l1Cent dtype; the K row is used:
 * pooled = (integer sum + tokens * offset) * scale), l1Cent [Nkv,C,D], pageIds [n] int32, pageTokens [n] int32
 * or absent (= blockSize),
 * blockIds [Nkv,kvPageLen] int32 in/out, pageMeanOut optional [n,Nkv,D] float32 (mean key of each page, the
 * input of a k-means centroid update).  Entries of pageIds outside the pool are skipped.                  */
SPA_API int spaAssignPageClustersGetWorkspaceSize(const spaTensor *key, const spaTensor *l1Cent,
                                                  const spaTensor *pageIds, const spaTensor *pageTokens,
                                                  const spaTensor *antiquantScale, const spaTensor *antiquantOffset,
                                                  int64_t blockSize, const spaTensor *blockIds,
                                                  const spaTensor *pageMeanOut,
                                                  uint64_t *workspaceSize, spaExecutor **executor);
SPA_API int spaAssignPageClusters(void *workspace, uint64_t workspaceSize, spaExecutor *executor,
                                  spaStream stream);

/* ---- KV-head tensor parallelism: the output all-gather fused into the launch (the reference
 *      has no distributed code on this path, so there is no interface to cite).  One process per GPU; rank r
 *      owns a contiguous range of KV heads and computes attentionOut[:, its query heads, :].  With peer output
 *      set, every output row is stored into the gathered buffer [B, Nq_total, D] of EVERY rank by NVLink peer
 *      stores from the combine epilogue, and the launch ends with a flag exchange (release/acquire at system
 *      scope): when the launch completes on a rank's stream, that rank's gathered buffer holds the rows of all
 *      ranks.  No NCCL call, no extra launch.
 *
 * spaPeerBufferCreate   cudaMalloc + zero-fill + CUDA IPC handle (64 bytes) to send to the other processes.
 * spaPeerBufferOpen     map a peer's buffer into this process (cudaIpcOpenMemHandle, peer access enabled).
 * spaExecutorSetPeerOutput
 *     gatheredOut[r], r in [0, world): rank r's image of the gathered buffer as mapped in THIS process
 *     (gatheredOut[rank] is the local one).  The executor's attentionOut must be a view into the local
 *     image; the same byte offset and strides are used in every peer image.
 *     signals[r]: rank r's signal block (>= 64 bytes, zero when first used, never written by the caller);
 *     word [i] = launches flagged by rank i, word [8] = own launch count, word [9] != 0: an exchange timed out
 *     (a peer did not launch within 2 s), words [10..11] = uint64 total ns this rank spent in the exchanges.
 *     Every rank must launch the same number of times.  A caller that re-launches before its PEERS have
 *     consumed their gathered buffer must alternate between two buffers (two executors): a rank cannot
 *     finish launch t before every peer has finished launch t, so buffer t%2 is never overwritten while a
 *     peer still reads it.   world <= 1 or gatheredOut == NULL switches peer output off.                   */
/* Test hook: occupy `numSMs` SMs (one CTA each, nearly all of the SM's shared memory) for `nanoseconds` on
 * `stream`.  Used by the tests to run the chain kernel while part of the device is unavailable: the chain
 * kernel is launched cooperatively (all its CTAs must be co-resident), so it starts when the SMs are free
 * again instead of spinning on a partial grid.  No reference counterpart.                                  */
SPA_API int spaDebugHoldSMs(int numSMs, uint64_t nanoseconds, spaStream stream);

#define SPA_IPC_HANDLE_BYTES 64
#define SPA_MAX_PEERS 8
SPA_API int spaPeerBufferCreate(uint64_t bytes, void **devPtr, unsigned char *ipcHandle);
SPA_API int spaPeerBufferOpen(const unsigned char *ipcHandle, void **devPtr);
SPA_API int spaPeerBufferClose(void *devPtr);
SPA_API int spaPeerBufferDestroy(void *devPtr);
SPA_API int spaExecutorSetPeerOutput(spaExecutor *executor, int rank, int world,
                                     void *const *gatheredOut, void *const *signals);

#ifdef __cplusplus
}
#endif
#endif /* SPA_B200_H_ */
