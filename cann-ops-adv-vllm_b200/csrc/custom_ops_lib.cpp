// PyTorch binding of the B200 sparse decode-attention chain: module `custom_ops_lib` and the
// `myops` operator library, drop-in for the reference's extension/csrc/{add_custom.cpp,
// registration.cpp, pytorch_npu_helper.hpp}.  It does what EXEC_NPU_CMD does there
// (pytorch_npu_helper.hpp:721-800): convert at::Tensor -> C descriptors, call
// <Op>GetWorkspaceSize, allocate the workspace from the caching allocator, launch on the current
// stream -- but against the C ABI in include/spa_b200.h and CUDA streams.  No compute lives here.
#include <ATen/cuda/CUDAContext.h>
#include <c10/cuda/CUDAGuard.h>
#include <torch/extension.h>
#include <torch/library.h>

#include <cstdlib>
#include <cstring>
#include <list>
#include <string>
#include <tuple>
#include <vector>

#include "../../include/spa_b200.h"

namespace {

// Reference constants (add_custom.cpp:940,958,987; select_position.cpp:7); keyword-overridable
// from Python, see custom_ops/add_custom.py.
constexpr int64_t kDefaultK = 64;
constexpr int64_t kDefaultMaxPageNum = 256;
constexpr int64_t kDefaultBlockSize = 128;
constexpr int64_t kPad = 8;

int to_spa_dtype(at::ScalarType t) {
    switch (t) {
        case at::kHalf: return SPA_F16;
        case at::kBFloat16: return SPA_BF16;
        case at::kFloat: return SPA_F32;
        case at::kChar: return SPA_I8;
        case at::kFloat8_e4m3fn: return SPA_F8E4M3;
        case at::kInt: return SPA_I32;
        case at::kLong: return SPA_I64;
        case at::kByte: return SPA_U8;             // attention_mask
        case at::kBool: return SPA_BOOL;
        default: TORCH_CHECK(false, "Unsupported tensor dtype");   // add_custom.cpp:69
    }
}

// undefined or zero-element tensor => absent (add_custom.cpp:32-35)
struct Desc {
    spaTensor t;
    bool present;
    const spaTensor *ptr() const { return present ? &t : nullptr; }
};

Desc desc(const at::Tensor &x, const char *name, bool keep_empty = false) {
    Desc d{};
    d.present = false;
    if (!x.defined()) return d;
    if (x.numel() == 0 && !keep_empty) return d;
    TORCH_CHECK(x.is_cuda(), name, ": expected a CUDA tensor");
    TORCH_CHECK(x.dim() <= SPA_MAX_DIMS, name, ": too many dims");
    d.t.data = x.numel() ? x.data_ptr() : nullptr;
    d.t.dtype = to_spa_dtype(x.scalar_type());
    d.t.ndim = (int32_t)x.dim();
    for (int i = 0; i < x.dim(); ++i) { d.t.sizes[i] = x.size(i); d.t.strides[i] = x.stride(i); }
    d.present = true;
    return d;
}

void check(int rc, const char *what) {
    TORCH_CHECK(rc == SPA_OK, what, " failed (", rc, "): ", spaGetLastError());
}

struct ExecGuard {
    spaExecutor *e = nullptr;
    ~ExecGuard() { if (e) spaDestroyExecutor(e); }
};

// ---- executor cache of the fused op (per thread; entries own an executor and its private workspace) ----
struct CachedExec { std::vector<int64_t> key; spaExecutor *exec; uint64_t ws; at::Tensor workspace; };
std::list<CachedExec> &plan_cache() {
    static thread_local std::list<CachedExec> *c = new std::list<CachedExec>();   // never destroyed: no CUDA calls at exit
    return *c;
}
bool plan_cache_enabled() {
    // off while any scheduling knob is set from the environment (they are read when an executor is planned)
    for (const char *v : {"SPA_SMALL_PAGES", "SPA_STATIC_PCT", "SPA_DYN_RUNS", "SPA_DYN_MIN", "SPA_SEL_GW", "SPA_PAIR", "SPA_COOP", "SPA_AUX_STAGE"})
        if (getenv(v)) return false;
    const char *s = getenv("SPA_PLAN_CACHE");
    return !(s && s[0] == '0');
}
void key_add(std::vector<int64_t> &k, const Desc &d) {
    if (!d.present) { k.push_back(-1); return; }
    k.push_back((int64_t)(uintptr_t)d.t.data); k.push_back(d.t.dtype); k.push_back(d.t.ndim);
    for (int i = 0; i < d.t.ndim; ++i) { k.push_back(d.t.sizes[i]); k.push_back(d.t.strides[i]); }
}
CachedExec *plan_cache_find(const std::vector<int64_t> &key) {
    auto &c = plan_cache();
    for (auto it = c.begin(); it != c.end(); ++it)
        if (it->key == key) { c.splice(c.begin(), c, it); return &c.front(); }
    return nullptr;
}
void plan_cache_insert(std::vector<int64_t> &&key, spaExecutor *e, uint64_t ws, const at::Tensor &workspace) {
    auto &c = plan_cache();
    c.push_front(CachedExec{std::move(key), e, ws, workspace});
    while (c.size() > 16) { spaDestroyExecutor(c.back().exec); c.pop_back(); }
}

void set_lse(spaExecutor *e, const c10::optional<at::Tensor> &lse, const char *what) {
    if (!lse.has_value() || !lse->defined() || lse->numel() == 0) return;
    Desc d = desc(*lse, "softmax_lse");
    check(spaExecutorSetSoftmaxLse(e, d.ptr()), what);
}

at::Tensor alloc_ws(uint64_t bytes, const at::Tensor &like) {
    return at::empty({(int64_t)std::max<uint64_t>(bytes, 256)}, like.options().dtype(at::kByte));
}

at::Tensor seq_on_device(const at::Tensor &t, const at::Tensor &like) {
    // the reference copies actual_seq_lengths to the HOST (add_custom.cpp:105-118); here the kernel
    // reads it on device, so a host tensor is uploaded instead.
    if (!t.defined() || t.numel() == 0) return t;
    at::Tensor r = t.to(at::kLong);
    if (!r.is_cuda()) r = r.to(like.device(), /*non_blocking=*/true);
    return r.contiguous();
}

// ------------------------------------------------------------------------------------------
at::Tensor compute_cent_k(const at::Tensor &query, const at::Tensor &l1_cent, int64_t k) {
    TORCH_CHECK(query.is_cuda(), "compute_cent: the CUDA extension only runs on CUDA tensors");
    c10::cuda::CUDAGuard guard(query.device());
    at::Tensor result = at::empty({query.size(0), query.size(1), k}, query.options().dtype(at::kInt));
    Desc q = desc(query, "query"), c = desc(l1_cent, "l1_cent"), r = desc(result, "indices", true);
    uint64_t ws = 0; ExecGuard g;
    check(spaComputeCentGetWorkspaceSize(q.ptr(), c.ptr(), r.ptr(), &ws, &g.e), "compute_cent");
    at::Tensor wsp = alloc_ws(ws, query);
    check(spaComputeCent(wsp.data_ptr(), ws, g.e, (spaStream)at::cuda::getCurrentCUDAStream().stream()), "compute_cent");
    return result;
}
at::Tensor compute_cent(const at::Tensor &query, const at::Tensor &l1_cent) {
    return compute_cent_k(query, l1_cent, kDefaultK);
}

std::tuple<at::Tensor, at::Tensor, at::Tensor> select_position_ex(
    const at::Tensor &block_ids, const at::Tensor &block_table, const at::Tensor &seq_len,
    const at::Tensor &indices, int64_t max_page_num, int64_t block_size) {
    TORCH_CHECK(block_ids.is_cuda(), "select_position: the CUDA extension only runs on CUDA tensors");
    c10::cuda::CUDAGuard guard(block_ids.device());
    const int64_t B = indices.size(0), Nq = indices.size(1), maxP = block_table.size(1);
    TORCH_CHECK(seq_len.dim() == 1 && seq_len.size(0) <= B, "select_position: seq_len must be [B]");
    const int64_t Bs = seq_len.size(0);
    auto opt = block_ids.options();
    at::Tensor pp = at::empty({B, Nq, max_page_num}, opt);
    at::Tensor ppl = at::empty({B, Nq, kPad}, opt);
    at::Tensor gather = at::empty({B, Nq, maxP}, opt);
    at::Tensor pp_v = pp.narrow(0, 0, Bs), ppl_v = ppl.narrow(0, 0, Bs), ga_v = gather.narrow(0, 0, Bs);
    Desc a = desc(block_ids, "block_ids"), b = desc(block_table, "block_table"), c = desc(seq_len, "seq_len"),
         d = desc(indices, "indices"), o1 = desc(pp_v, "page_position"), o2 = desc(ppl_v, "page_position_length"),
         o3 = desc(ga_v, "block_table_gather");
    uint64_t ws = 0; ExecGuard g;
    check(spaSelectPositionGetWorkspaceSize(a.ptr(), b.ptr(), c.ptr(), d.ptr(), block_size, o1.ptr(), o2.ptr(),
                                            o3.ptr(), &ws, &g.e), "select_position");
    at::Tensor wsp = alloc_ws(ws, block_ids);
    check(spaSelectPosition(wsp.data_ptr(), ws, g.e, (spaStream)at::cuda::getCurrentCUDAStream().stream()), "select_position");
    return std::make_tuple(pp, ppl, gather);
}
std::tuple<at::Tensor, at::Tensor, at::Tensor> select_position(const at::Tensor &block_ids,
                                                               const at::Tensor &block_table,
                                                               const at::Tensor &seq_len,
                                                               const at::Tensor &indices) {
    return select_position_ex(block_ids, block_table, seq_len, indices, kDefaultMaxPageNum, kDefaultBlockSize);
}

std::tuple<at::Tensor, at::Tensor, at::Tensor> cent_select_ex(
    const at::Tensor &query, const at::Tensor &l1_cent, const at::Tensor &block_ids,
    const at::Tensor &block_table, const at::Tensor &seq_len, int64_t k, int64_t max_page_num,
    int64_t block_size, int64_t sink_pages, int64_t recent_pages, int64_t max_fraction_den) {
    TORCH_CHECK(query.is_cuda(), "cent_select: the CUDA extension only runs on CUDA tensors");
    c10::cuda::CUDAGuard guard(query.device());
    const int64_t B = seq_len.size(0), Nq = query.size(1);          // add_custom.cpp:988
    auto opt = block_ids.options();
    at::Tensor pp = at::empty({B, Nq, max_page_num}, opt);
    at::Tensor ppl = at::empty({B, Nq, kPad}, opt);
    at::Tensor mppl = at::empty({B, kPad}, opt.dtype(at::kLong));
    Desc q = desc(query, "query"), c = desc(l1_cent, "l1_cent"), a = desc(block_ids, "block_ids"),
         b = desc(block_table, "block_table"), s = desc(seq_len, "seq_len"), o1 = desc(pp, "page_position", true),
         o2 = desc(ppl, "page_position_length", true), o3 = desc(mppl, "max_page_position_length", true);
    if (B == 0) return std::make_tuple(pp, ppl, mppl);
    uint64_t ws = 0; ExecGuard g;
    check(spaCentSelectGetWorkspaceSize(q.ptr(), c.ptr(), a.ptr(), b.ptr(), s.ptr(), k, block_size, o1.ptr(),
                                        o2.ptr(), o3.ptr(), &ws, &g.e), "cent_select");
    if (sink_pages | recent_pages | max_fraction_den)
        check(spaExecutorSetSelectOptions(g.e, (int)sink_pages, (int)recent_pages, (int)max_fraction_den), "cent_select");
    at::Tensor wsp = alloc_ws(ws, query);
    check(spaCentSelect(wsp.data_ptr(), ws, g.e, (spaStream)at::cuda::getCurrentCUDAStream().stream()), "cent_select");
    return std::make_tuple(pp, ppl, mppl);
}
std::tuple<at::Tensor, at::Tensor, at::Tensor> cent_select(const at::Tensor &query, const at::Tensor &l1_cent,
                                                           const at::Tensor &block_ids,
                                                           const at::Tensor &block_table,
                                                           const at::Tensor &seq_len) {
    return cent_select_ex(query, l1_cent, block_ids, block_table, seq_len, kDefaultK, kDefaultMaxPageNum,
                          kDefaultBlockSize, 0, 0, 0);
}

// centroid maintenance (no reference counterpart): updates block_ids in place, optionally returns page means
at::Tensor assign_page_clusters(const at::Tensor &key, const at::Tensor &l1_cent, const at::Tensor &page_ids,
                                const c10::optional<at::Tensor> &page_tokens, at::Tensor block_ids, int64_t block_size,
                                bool return_mean, const c10::optional<at::Tensor> &antiquant_scale,
                                const c10::optional<at::Tensor> &antiquant_offset) {
    TORCH_CHECK(key.is_cuda(), "assign_page_clusters: the CUDA extension only runs on CUDA tensors");
    c10::cuda::CUDAGuard guard(key.device());
    at::Tensor mean;
    if (return_mean) mean = at::empty({page_ids.numel(), l1_cent.size(0), l1_cent.size(2)}, key.options().dtype(at::kFloat));
    Desc k = desc(key, "key"), c = desc(l1_cent, "l1_cent"), pi = desc(page_ids, "page_ids"),
         pt = desc(page_tokens.has_value() ? *page_tokens : at::Tensor(), "page_tokens"),
         bi = desc(block_ids, "block_ids"), pm = desc(mean, "page_mean"),
         as = desc(antiquant_scale.has_value() ? *antiquant_scale : at::Tensor(), "antiquant_scale"),
         ao = desc(antiquant_offset.has_value() ? *antiquant_offset : at::Tensor(), "antiquant_offset");
    uint64_t ws = 0; ExecGuard g;
    check(spaAssignPageClustersGetWorkspaceSize(k.ptr(), c.ptr(), pi.ptr(), pt.ptr(), as.ptr(), ao.ptr(), block_size, bi.ptr(),
                                                pm.ptr(), &ws, &g.e),
          "assign_page_clusters");
    check(spaAssignPageClusters(nullptr, 0, g.e, (spaStream)at::cuda::getCurrentCUDAStream().stream()), "assign_page_clusters");
    return return_mean ? mean : block_ids;
}

const at::Tensor &single(const std::vector<at::Tensor> &l, const char *name) {
    TORCH_CHECK(l.size() == 1, name, ": expected a 1-element tensor list (add_custom.py:70-75)");
    return l[0];
}

at::Tensor sparse_paged_attention_ex(
    const at::Tensor &query, const std::vector<at::Tensor> &key_list, const std::vector<at::Tensor> &value_list,
    const at::Tensor &pse_shift, const at::Tensor &attention_mask, const at::Tensor &actual_seq_lengths,
    const at::Tensor &dequant_scale1, const at::Tensor &quant_scale1, const at::Tensor &dequant_scale2,
    const at::Tensor &quant_scale2, const at::Tensor &quant_offset2, const at::Tensor &antiquant_scale,
    const at::Tensor &antiquant_offset, const at::Tensor &blocktable, const at::Tensor &kv_padding_size,
    const at::Tensor &blockposition, int64_t num_heads, double scale_value, std::string input_layout,
    int64_t num_key_value_heads, int64_t block_size, int64_t inner_precise,
    const c10::optional<at::Tensor> &softmax_lse) {
    TORCH_CHECK(query.is_cuda(), "sparse_paged_attention: the CUDA extension only runs on CUDA tensors");
    c10::cuda::CUDAGuard guard(query.device());
    const at::Tensor &key = single(key_list, "key_list"), &value = single(value_list, "value_list");
    // add_custom.cpp:559-560 allocates the query's dtype; with quant_scale2 the op's output is int8
    // (sparse_paged_attention_tiling.cc:570-602), which the reference binding cannot express -- here it can
    at::Tensor out = at::empty(query.sizes(), quant_scale2.defined() && quant_scale2.numel() ? query.options().dtype(at::kChar) : query.options());
    at::Tensor asl = seq_on_device(actual_seq_lengths, query);
    Desc q = desc(query, "query"), k = desc(key, "key"), v = desc(value, "value"), pse = desc(pse_shift, "pse_shift"),
         am = desc(attention_mask, "attention_mask"), sl = desc(asl, "actual_seq_lengths"),
         d1 = desc(dequant_scale1, "dequant_scale1"), q1 = desc(quant_scale1, "quant_scale1"),
         d2 = desc(dequant_scale2, "dequant_scale2"), q2 = desc(quant_scale2, "quant_scale2"),
         qo2 = desc(quant_offset2, "quant_offset2"), as = desc(antiquant_scale, "antiquant_scale"),
         ao = desc(antiquant_offset, "antiquant_offset"), bt = desc(blocktable, "blocktable"),
         kp = desc(kv_padding_size, "kv_padding_size"), bp = desc(blockposition, "blockposition"),
         o = desc(out, "attention_out", true);
    if (query.numel() == 0) return out;
    uint64_t ws = 0; ExecGuard g;
    check(spaSparsePagedAttentionGetWorkspaceSize(q.ptr(), k.ptr(), v.ptr(), pse.ptr(), am.ptr(), sl.ptr(), d1.ptr(),
                                                  q1.ptr(), d2.ptr(), q2.ptr(), qo2.ptr(), as.ptr(), ao.ptr(), bt.ptr(),
                                                  kp.ptr(), bp.ptr(), num_heads, scale_value, input_layout.c_str(),
                                                  num_key_value_heads, block_size, inner_precise, o.ptr(), &ws, &g.e),
          "sparse_paged_attention");
    set_lse(g.e, softmax_lse, "sparse_paged_attention");
    at::Tensor wsp = alloc_ws(ws, query);
    check(spaSparsePagedAttention(wsp.data_ptr(), ws, g.e, (spaStream)at::cuda::getCurrentCUDAStream().stream()),
          "sparse_paged_attention");
    return out;
}
at::Tensor sparse_paged_attention(
    const at::Tensor &query, const std::vector<at::Tensor> &key_list, const std::vector<at::Tensor> &value_list,
    const at::Tensor &pse_shift, const at::Tensor &attention_mask, const at::Tensor &actual_seq_lengths,
    const at::Tensor &dequant_scale1, const at::Tensor &quant_scale1, const at::Tensor &dequant_scale2,
    const at::Tensor &quant_scale2, const at::Tensor &quant_offset2, const at::Tensor &antiquant_scale,
    const at::Tensor &antiquant_offset, const at::Tensor &blocktable, const at::Tensor &kv_padding_size,
    const at::Tensor &blockposition, int64_t num_heads, double scale_value, std::string input_layout,
    int64_t num_key_value_heads, int64_t block_size, int64_t inner_precise) {
    return sparse_paged_attention_ex(query, key_list, value_list, pse_shift, attention_mask, actual_seq_lengths,
                                     dequant_scale1, quant_scale1, dequant_scale2, quant_scale2, quant_offset2,
                                     antiquant_scale, antiquant_offset, blocktable, kv_padding_size, blockposition,
                                     num_heads, scale_value, input_layout, num_key_value_heads, block_size,
                                     inner_precise, c10::nullopt);
}

// ---- fused op: plan object so callers can keep a launch-ready executor (CUDA graphs, serving) --
struct FusionPlan {
    spaExecutor *exec = nullptr;
    uint64_t ws_bytes = 0;
    at::Tensor workspace;
    std::vector<at::Tensor> keep;     // keeps the planned tensors alive
    ~FusionPlan() { if (exec) spaDestroyExecutor(exec); }
};

std::shared_ptr<FusionPlan> make_fusion_plan(
    const at::Tensor &query, const at::Tensor &key, const at::Tensor &value, const at::Tensor &antiquant_scale,
    const at::Tensor &antiquant_offset, const at::Tensor &blocktable, const at::Tensor &l1_cent,
    const at::Tensor &block_ids, const at::Tensor &total_seq_len, const at::Tensor &block_position,
    const at::Tensor &page_position_length, const at::Tensor &max_page_position_length,
    const at::Tensor &attention_out, int64_t num_heads, double scale_value, const std::string &input_layout,
    int64_t num_key_value_heads, int64_t block_size, int64_t inner_precise, int64_t k, int64_t sink_pages,
    int64_t recent_pages, int64_t max_fraction_den, const c10::optional<at::Tensor> &softmax_lse,
    const c10::optional<at::Tensor> &pse_shift, const c10::optional<at::Tensor> &attention_mask,
    const c10::optional<at::Tensor> &quant_scale2, const c10::optional<at::Tensor> &quant_offset2) {
    TORCH_CHECK(query.is_cuda(), "sparse_paged_fusion_attention: the CUDA extension only runs on CUDA tensors");
    c10::cuda::CUDAGuard guard(query.device());
    auto plan = std::make_shared<FusionPlan>();
    auto opt = [](const c10::optional<at::Tensor> &t) { return t.has_value() ? *t : at::Tensor(); };
    Desc pse = desc(opt(pse_shift), "pse_shift"), am = desc(opt(attention_mask), "attention_mask"),
         q2 = desc(opt(quant_scale2), "quant_scale2"), qo2 = desc(opt(quant_offset2), "quant_offset2");
    Desc q = desc(query, "query"), kk = desc(key, "key"), v = desc(value, "value"),
         as = desc(antiquant_scale, "antiquant_scale"), ao = desc(antiquant_offset, "antiquant_offset"),
         bt = desc(blocktable, "blocktable"), lc = desc(l1_cent, "l1_cent"), bi = desc(block_ids, "block_ids"),
         ts = desc(total_seq_len, "total_seq_len"), bp = desc(block_position, "block_position", true),
         pl = desc(page_position_length, "page_position_length", true),
         mp = desc(max_page_position_length, "max_page_position_length", true),
         o = desc(attention_out, "attention_out", true);
    check(spaSparsePagedFusionAttentionGetWorkspaceSize(
              q.ptr(), kk.ptr(), v.ptr(), pse.ptr(), am.ptr(), nullptr, nullptr, nullptr, nullptr, q2.ptr(), qo2.ptr(),
              as.ptr(), ao.ptr(), bt.ptr(), nullptr, lc.ptr(), bi.ptr(), ts.ptr(), num_heads, scale_value,
              input_layout.c_str(), num_key_value_heads, block_size, inner_precise, k, bp.ptr(), pl.ptr(), mp.ptr(),
              o.ptr(), &plan->ws_bytes, &plan->exec),
          "sparse_paged_fusion_attention");
    if (sink_pages | recent_pages | max_fraction_den)
        check(spaExecutorSetSelectOptions(plan->exec, (int)sink_pages, (int)recent_pages, (int)max_fraction_den),
              "sparse_paged_fusion_attention");
    set_lse(plan->exec, softmax_lse, "sparse_paged_fusion_attention");
    if (softmax_lse.has_value() && softmax_lse->defined()) plan->keep.push_back(*softmax_lse);
    for (const auto *t : {&pse_shift, &attention_mask, &quant_scale2, &quant_offset2})
        if (t->has_value() && (*t)->defined()) plan->keep.push_back(**t);
    plan->workspace = alloc_ws(plan->ws_bytes, query);
    // the plan owns its workspace and nothing else writes to it: the queue words need clearing only once
    check(spaExecutorSetPersistentWorkspace(plan->exec, 1), "spaExecutorSetPersistentWorkspace");
    for (const at::Tensor &t : std::vector<at::Tensor>{query, key, value, antiquant_scale, antiquant_offset, blocktable, l1_cent, block_ids,
                  total_seq_len, block_position, page_position_length, max_page_position_length, attention_out})
        plan->keep.push_back(t);
    return plan;
}

void launch_fusion_plan(const std::shared_ptr<FusionPlan> &plan) {
    c10::cuda::CUDAGuard guard(plan->workspace.device());
    check(spaSparsePagedFusionAttention(plan->workspace.data_ptr(), plan->ws_bytes, plan->exec,
                                        (spaStream)at::cuda::getCurrentCUDAStream().stream()),
          "sparse_paged_fusion_attention");
}

std::tuple<at::Tensor, at::Tensor, at::Tensor> sparse_paged_fusion_attention_ex(
    const at::Tensor &query, const std::vector<at::Tensor> &key_list, const std::vector<at::Tensor> &value_list,
    const at::Tensor &pse_shift, const at::Tensor &attention_mask, const at::Tensor &actual_seq_lengths,
    const at::Tensor &dequant_scale1, const at::Tensor &quant_scale1, const at::Tensor &dequant_scale2,
    const at::Tensor &quant_scale2, const at::Tensor &quant_offset2, const at::Tensor &antiquant_scale,
    const at::Tensor &antiquant_offset, const at::Tensor &blocktable, const at::Tensor &kv_padding_size,
    const at::Tensor &l1_cent, const at::Tensor &block_ids, const at::Tensor &total_seq_len,
    const at::Tensor &block_position, const at::Tensor &page_position_length,
    const at::Tensor &max_page_position_length, int64_t num_heads, double scale_value, std::string input_layout,
    int64_t num_key_value_heads, int64_t block_size, int64_t inner_precise, int64_t k, int64_t sink_pages,
    int64_t recent_pages, int64_t max_fraction_den, const c10::optional<at::Tensor> &softmax_lse) {
    TORCH_CHECK(query.is_cuda(), "sparse_paged_fusion_attention: the CUDA extension only runs on CUDA tensors");
    c10::cuda::CUDAGuard guard(query.device());
    const at::Tensor &key = single(key_list, "key_list"), &value = single(value_list, "value_list");
    (void)actual_seq_lengths;    // ignored by the fused kernel (sparse_paged_fusion_attention.cpp:68)
    // add_custom.cpp:800-801; int8 with quant_scale2 (see sparse_paged_attention_ex)
    at::Tensor out = at::empty(query.sizes(), quant_scale2.defined() && quant_scale2.numel() ? query.options().dtype(at::kChar) : query.options());
    Desc q = desc(query, "query"), kk = desc(key, "key"), v = desc(value, "value"), pse = desc(pse_shift, "pse_shift"),
         am = desc(attention_mask, "attention_mask"), d1 = desc(dequant_scale1, "dequant_scale1"),
         q1 = desc(quant_scale1, "quant_scale1"), d2 = desc(dequant_scale2, "dequant_scale2"),
         q2 = desc(quant_scale2, "quant_scale2"), qo2 = desc(quant_offset2, "quant_offset2"),
         as = desc(antiquant_scale, "antiquant_scale"), ao = desc(antiquant_offset, "antiquant_offset"),
         bt = desc(blocktable, "blocktable"), kp = desc(kv_padding_size, "kv_padding_size"),
         lc = desc(l1_cent, "l1_cent"), bi = desc(block_ids, "block_ids"), ts = desc(total_seq_len, "total_seq_len"),
         bp = desc(block_position, "block_position", true), pl = desc(page_position_length, "page_position_length", true),
         mp = desc(max_page_position_length, "max_page_position_length", true), o = desc(out, "attention_out", true);
    if (total_seq_len.numel() == 0) return std::make_tuple(out, block_position, max_page_position_length);
    spaStream stream = (spaStream)at::cuda::getCurrentCUDAStream().stream();
    // Executor cache (aclnn keeps repeatable executors the same way, add_custom.cpp:847): a decode loop calls
    // this op with the same tensors' addresses and shapes step after step; the planned executor and its
    // private, already-clean workspace are then reused -- no validation, descriptor encoding, allocation or
    // workspace clear on the step's critical path.
    std::vector<int64_t> ckey;
    const bool use_cache = plan_cache_enabled();
    if (use_cache) {
        ckey.reserve(256);
        for (const Desc *d : {&q, &kk, &v, &pse, &am, &d1, &q1, &d2, &q2, &qo2, &as, &ao, &bt, &kp, &lc, &bi, &ts, &bp, &pl, &mp, &o})
            key_add(ckey, *d);
        Desc ls = desc(softmax_lse.has_value() ? *softmax_lse : at::Tensor(), "softmax_lse");
        key_add(ckey, ls);
        int64_t sbits; std::memcpy(&sbits, &scale_value, sizeof(sbits));
        for (int64_t a : {num_heads, sbits, num_key_value_heads, block_size, inner_precise, k, sink_pages, recent_pages,
                          max_fraction_den, (int64_t)(uintptr_t)stream, (int64_t)query.get_device()})
            ckey.push_back(a);
        for (char ch : input_layout) ckey.push_back(ch);
        if (CachedExec *hit = plan_cache_find(ckey)) {
            check(spaSparsePagedFusionAttention(hit->workspace.data_ptr(), hit->ws, hit->exec, stream),
                  "sparse_paged_fusion_attention");
            return std::make_tuple(out, block_position, max_page_position_length);
        }
    }
    uint64_t ws = 0; ExecGuard g;
    // C-level order: inputs, attrs, the three aux outputs, attention_out (add_custom.cpp:836-844)
    check(spaSparsePagedFusionAttentionGetWorkspaceSize(
              q.ptr(), kk.ptr(), v.ptr(), pse.ptr(), am.ptr(), nullptr, d1.ptr(), q1.ptr(), d2.ptr(), q2.ptr(), qo2.ptr(),
              as.ptr(), ao.ptr(), bt.ptr(), kp.ptr(), lc.ptr(), bi.ptr(), ts.ptr(), num_heads, scale_value,
              input_layout.c_str(), num_key_value_heads, block_size, inner_precise, k, bp.ptr(), pl.ptr(), mp.ptr(),
              o.ptr(), &ws, &g.e),
          "sparse_paged_fusion_attention");
    if (sink_pages | recent_pages | max_fraction_den)
        check(spaExecutorSetSelectOptions(g.e, (int)sink_pages, (int)recent_pages, (int)max_fraction_den),
              "sparse_paged_fusion_attention");
    set_lse(g.e, softmax_lse, "sparse_paged_fusion_attention");
    at::Tensor wsp = alloc_ws(ws, query);
    if (use_cache) check(spaExecutorSetPersistentWorkspace(g.e, 1), "sparse_paged_fusion_attention");
    check(spaSparsePagedFusionAttention(wsp.data_ptr(), ws, g.e, stream), "sparse_paged_fusion_attention");
    if (use_cache) { plan_cache_insert(std::move(ckey), g.e, ws, wsp); g.e = nullptr; }
    return std::make_tuple(out, block_position, max_page_position_length);       // add_custom.cpp:931
}

std::tuple<at::Tensor, at::Tensor, at::Tensor> sparse_paged_fusion_attention(
    const at::Tensor &query, const std::vector<at::Tensor> &key_list, const std::vector<at::Tensor> &value_list,
    const at::Tensor &pse_shift, const at::Tensor &attention_mask, const at::Tensor &actual_seq_lengths,
    const at::Tensor &dequant_scale1, const at::Tensor &quant_scale1, const at::Tensor &dequant_scale2,
    const at::Tensor &quant_scale2, const at::Tensor &quant_offset2, const at::Tensor &antiquant_scale,
    const at::Tensor &antiquant_offset, const at::Tensor &blocktable, const at::Tensor &kv_padding_size,
    const at::Tensor &l1_cent, const at::Tensor &block_ids, const at::Tensor &total_seq_len,
    const at::Tensor &block_position, const at::Tensor &page_position_length,
    const at::Tensor &max_page_position_length, int64_t num_heads, double scale_value, std::string input_layout,
    int64_t num_key_value_heads, int64_t block_size, int64_t inner_precise) {
    return sparse_paged_fusion_attention_ex(query, key_list, value_list, pse_shift, attention_mask,
                                            actual_seq_lengths, dequant_scale1, quant_scale1, dequant_scale2,
                                            quant_scale2, quant_offset2, antiquant_scale, antiquant_offset, blocktable,
                                            kv_padding_size, l1_cent, block_ids, total_seq_len, block_position,
                                            page_position_length, max_page_position_length, num_heads, scale_value,
                                            input_layout, num_key_value_heads, block_size, inner_precise, kDefaultK,
                                            0, 0, 0, c10::nullopt);
}

at::Tensor incre_flash_attention_v4(
    const at::Tensor &query, const std::vector<at::Tensor> &key_list, const std::vector<at::Tensor> &value_list,
    const at::Tensor &pse_shift, const at::Tensor &attention_mask, const at::Tensor &actual_seq_lengths,
    const at::Tensor &dequant_scale1, const at::Tensor &quant_scale1, const at::Tensor &dequant_scale2,
    const at::Tensor &quant_scale2, const at::Tensor &quant_offset2, const at::Tensor &antiquant_scale,
    const at::Tensor &antiquant_offset, const at::Tensor &blocktable, const at::Tensor &kv_padding_size,
    int64_t num_heads, double scale_value, std::string input_layout, int64_t num_key_value_heads, int64_t block_size,
    int64_t inner_precise) {
    // The reference routes this entry to the STOCK CANN IncreFlashAttention op (its dense baseline,
    // add_custom.cpp:120-318).  Here the dense paged decode is the sparse kernel over EVERY page: an identity
    // block_position per KV head, clipped by actual_seq_lengths.  Only the paged layout with a block table
    // is served (what the sparse tests compare against); the stock op's other options are refused by the
    // attention op's validation.
    TORCH_CHECK(query.is_cuda(), "incre_flash_attention_v4: the CUDA extension only runs on CUDA tensors");
    TORCH_CHECK(blocktable.defined() && blocktable.dim() == 2 && blocktable.numel() > 0,
                "incre_flash_attention_v4: only the paged KV layout (blocktable [B, maxBlocks]) is supported");
    c10::cuda::CUDAGuard guard(query.device());
    const int64_t B = query.size(0), maxP = blocktable.size(1);
    if (num_key_value_heads <= 0) num_key_value_heads = num_heads;      // the reference's default: 0 = MHA
    at::Tensor ident = at::arange(maxP, blocktable.options().dtype(at::kInt))
                           .view({1, 1, maxP}).expand({B, num_key_value_heads, maxP}).contiguous();
    return sparse_paged_attention(query, key_list, value_list, pse_shift, attention_mask, actual_seq_lengths,
                                  dequant_scale1, quant_scale1, dequant_scale2, quant_scale2, quant_offset2,
                                  antiquant_scale, antiquant_offset, blocktable, kv_padding_size, ident, num_heads,
                                  scale_value, input_layout, num_key_value_heads, block_size, inner_precise);
}

// ---- Meta (shape inference) impls: output shapes/dtypes of the CUDA impls above, no data.  Mirrors the
// role of op_plugin/python/meta/_meta_registrations.py:38-93 (FakeTensor / torch.compile tracing).
at::Tensor meta_attention(
    const at::Tensor &query, const std::vector<at::Tensor> &, const std::vector<at::Tensor> &, const at::Tensor &,
    const at::Tensor &, const at::Tensor &, const at::Tensor &, const at::Tensor &, const at::Tensor &,
    const at::Tensor &quant_scale2, const at::Tensor &, const at::Tensor &, const at::Tensor &, const at::Tensor &,
    const at::Tensor &, const at::Tensor &, int64_t, double, std::string, int64_t, int64_t, int64_t) {
    return at::empty(query.sizes(), quant_scale2.numel() ? query.options().dtype(at::kChar) : query.options());
}
at::Tensor meta_attention_v4(
    const at::Tensor &query, const std::vector<at::Tensor> &, const std::vector<at::Tensor> &, const at::Tensor &,
    const at::Tensor &, const at::Tensor &, const at::Tensor &, const at::Tensor &, const at::Tensor &,
    const at::Tensor &, const at::Tensor &, const at::Tensor &, const at::Tensor &, const at::Tensor &,
    const at::Tensor &, int64_t, double, std::string, int64_t, int64_t, int64_t) {
    return at::empty(query.sizes(), query.options());
}
at::Tensor meta_compute_cent(const at::Tensor &query, const at::Tensor &) {
    return at::empty({query.size(0), query.size(1), kDefaultK}, query.options().dtype(at::kInt));
}
std::tuple<at::Tensor, at::Tensor, at::Tensor> meta_select_position(const at::Tensor &block_ids,
                                                                    const at::Tensor &block_table,
                                                                    const at::Tensor &, const at::Tensor &indices) {
    auto opt = block_ids.options().dtype(at::kInt);
    const int64_t B = indices.size(0), Nq = indices.size(1);
    return std::make_tuple(at::empty({B, Nq, kDefaultMaxPageNum}, opt), at::empty({B, Nq, kPad}, opt),
                           at::empty({B, Nq, block_table.size(1)}, opt));
}
std::tuple<at::Tensor, at::Tensor, at::Tensor> meta_cent_select(const at::Tensor &query, const at::Tensor &,
                                                                const at::Tensor &block_ids, const at::Tensor &,
                                                                const at::Tensor &seq_len) {
    auto opt = block_ids.options().dtype(at::kInt);
    const int64_t B = seq_len.size(0), Nq = query.size(1);
    return std::make_tuple(at::empty({B, Nq, kDefaultMaxPageNum}, opt), at::empty({B, Nq, kPad}, opt),
                           at::empty({B, kPad}, opt.dtype(at::kLong)));
}
std::tuple<at::Tensor, at::Tensor, at::Tensor> meta_fusion(
    const at::Tensor &query, const std::vector<at::Tensor> &, const std::vector<at::Tensor> &, const at::Tensor &,
    const at::Tensor &, const at::Tensor &, const at::Tensor &, const at::Tensor &, const at::Tensor &,
    const at::Tensor &quant_scale2, const at::Tensor &, const at::Tensor &, const at::Tensor &, const at::Tensor &,
    const at::Tensor &, const at::Tensor &, const at::Tensor &, const at::Tensor &, const at::Tensor &block_position,
    const at::Tensor &, const at::Tensor &max_page_position_length, int64_t, double, std::string, int64_t, int64_t,
    int64_t) {
    // fresh tensors of the caller-allocated shapes (a functional schema must not return its inputs)
    return std::make_tuple(at::empty(query.sizes(), quant_scale2.numel() ? query.options().dtype(at::kChar) : query.options()),
                           at::empty(block_position.sizes(), block_position.options()),
                           at::empty(max_page_position_length.sizes(), max_page_position_length.options()));
}

}  // namespace

// Schema strings identical to the reference's (extension/csrc/registration.cpp:43-57).
TORCH_LIBRARY(myops, m) {
    m.def("incre_flash_attention_v4(Tensor query, Tensor[] key_list, Tensor[] value_list, Tensor pse_shift, Tensor attention_mask, Tensor actual_seq_lengths, Tensor dequant_scale1, Tensor quant_scale1, Tensor dequant_scale2, Tensor quant_scale2, Tensor quant_offset2, Tensor antiquant_scale, Tensor antiquant_offset, Tensor blocktable, Tensor kv_padding_size, int num_heads, float scale_value, str input_layout, int num_key_value_heads, int block_size, int inner_precise) -> Tensor");
    m.def("sparse_paged_attention(Tensor query, Tensor[] key_list, Tensor[] value_list, Tensor pse_shift, Tensor attention_mask,Tensor actual_seq_lengths, Tensor dequant_scale1, Tensor quant_scale1, Tensor dequant_scale2,Tensor quant_scale2, Tensor quant_offset2, Tensor antiquant_scale, Tensor antiquant_offset, Tensor blocktable,Tensor kv_padding_size, Tensor blockposition, int num_heads, float scale_value, str input_layout,int num_key_value_heads, int block_size, int inner_precise) ->Tensor");
    m.def("compute_cent(Tensor query, Tensor l1_cent) -> Tensor");
    m.def("select_position(Tensor block_ids, Tensor block_table, Tensor seq_len, Tensor indices) -> (Tensor, Tensor, Tensor)");
    m.def("cent_select(Tensor query, Tensor l1_cent, Tensor block_ids, Tensor block_table, Tensor seq_len) -> (Tensor, Tensor, Tensor)");
    m.def("sparse_paged_fusion_attention("
          "Tensor query, Tensor[] key_list, Tensor[] value_list, Tensor pse_shift, Tensor attention_mask, Tensor actual_seq_lengths, "
          "Tensor dequant_scale1, Tensor quant_scale1, Tensor dequant_scale2, Tensor quant_scale2, Tensor quant_offset2, "
          "Tensor antiquant_scale, Tensor antiquant_offset, Tensor blocktable, Tensor kv_padding_size, "
          "Tensor l1_cent, Tensor block_ids, Tensor total_seq_len, "
          "Tensor block_position, Tensor page_position_length, Tensor max_page_position_length, "
          "int num_heads, float scale_value, str input_layout, int num_key_value_heads, int block_size, int inner_precise) -> (Tensor, Tensor, Tensor)");
}

// The reference registers for PrivateUse1 (NPU), registration.cpp:60-68; here the CUDA key.
TORCH_LIBRARY_IMPL(myops, CUDA, m) {
    m.impl("incre_flash_attention_v4", &incre_flash_attention_v4);
    m.impl("sparse_paged_attention", &sparse_paged_attention);
    m.impl("compute_cent", &compute_cent);
    m.impl("select_position", &select_position);
    m.impl("cent_select", &cent_select);
    m.impl("sparse_paged_fusion_attention", &sparse_paged_fusion_attention);
}

TORCH_LIBRARY_IMPL(myops, Meta, m) {
    m.impl("incre_flash_attention_v4", &meta_attention_v4);
    m.impl("sparse_paged_attention", &meta_attention);
    m.impl("compute_cent", &meta_compute_cent);
    m.impl("select_position", &meta_select_position);
    m.impl("cent_select", &meta_cent_select);
    m.impl("sparse_paged_fusion_attention", &meta_fusion);
}

PYBIND11_MODULE(custom_ops_lib, m) {
    // the six reference entry points (registration.cpp:71-79)
    m.def("incre_flash_attention_v4", &incre_flash_attention_v4, "FlashAttention V4 (dense paged decode on the sparse kernel)");
    m.def("sparse_paged_attention", &sparse_paged_attention, "Sparse Paged Attention");
    m.def("compute_cent", &compute_cent, "Compute Cent");
    m.def("select_position", &select_position, "Select Position");
    m.def("cent_select", &cent_select, "Cent Select");
    m.def("sparse_paged_fusion_attention", &sparse_paged_fusion_attention, "Fused Sparse Paged Fusion Attention");
    // keyword-extended variants used by custom_ops/add_custom.py (defaults == reference constants)
    m.def("compute_cent_k", &compute_cent_k);
    m.def("select_position_ex", &select_position_ex);
    m.def("cent_select_ex", &cent_select_ex);
    m.def("sparse_paged_fusion_attention_ex", &sparse_paged_fusion_attention_ex);
    m.def("sparse_paged_attention_ex", &sparse_paged_attention_ex);
    m.def("assign_page_clusters", &assign_page_clusters);
    pybind11::class_<FusionPlan, std::shared_ptr<FusionPlan>>(m, "FusionPlan")
        .def_property_readonly("workspace_bytes", [](const FusionPlan &p) { return p.ws_bytes; })
        .def_property_readonly("num_launches", [](const FusionPlan &p) { return spaExecutorNumLaunches(p.exec); })
        .def_property_readonly("grid_size", [](const FusionPlan &p) { return spaExecutorGridSize(p.exec); })
        .def("set_peer_output", [](FusionPlan &p, int64_t rank, int64_t world, const std::vector<int64_t> &gathered,
                                   const std::vector<int64_t> &signals) {
            TORCH_CHECK((int64_t)gathered.size() == world && (int64_t)signals.size() == world, "set_peer_output: need one pointer per rank");
            std::vector<void *> g(world), s(world);
            for (int64_t i = 0; i < world; ++i) { g[i] = (void *)(uintptr_t)gathered[i]; s[i] = (void *)(uintptr_t)signals[i]; }
            check(spaExecutorSetPeerOutput(p.exec, (int)rank, (int)world, g.data(), s.data()), "spaExecutorSetPeerOutput");
        })
        .def("set_trace", [](FusionPlan &p, const at::Tensor &buf) {
            if (!buf.defined() || buf.numel() == 0) { spaExecutorSetTrace(p.exec, nullptr); return; }
            TORCH_CHECK(buf.is_cuda() && buf.scalar_type() == at::kLong && buf.is_contiguous() &&
                        buf.numel() >= (int64_t)spaExecutorGridSize(p.exec) * 32, "trace: need int64 CUDA [grid,32]");
            p.keep.push_back(buf);
            spaExecutorSetTrace(p.exec, buf.data_ptr());
        });
    // ---- KV-head tensor parallelism: peer buffers for the fused output all-gather (include/spa_b200.h)
    m.def("peer_buffer_create", [](int64_t bytes) {
        void *ptr = nullptr;
        unsigned char h[SPA_IPC_HANDLE_BYTES];
        check(spaPeerBufferCreate((uint64_t)bytes, &ptr, h), "spaPeerBufferCreate");
        return std::make_pair((int64_t)(uintptr_t)ptr, pybind11::bytes((const char *)h, SPA_IPC_HANDLE_BYTES));
    });
    m.def("peer_buffer_open", [](const std::string &handle) {
        TORCH_CHECK(handle.size() == SPA_IPC_HANDLE_BYTES, "peer_buffer_open: need a 64-byte IPC handle");
        void *ptr = nullptr;
        check(spaPeerBufferOpen((const unsigned char *)handle.data(), &ptr), "spaPeerBufferOpen");
        return (int64_t)(uintptr_t)ptr;
    });
    m.def("peer_buffer_close", [](int64_t ptr) { check(spaPeerBufferClose((void *)(uintptr_t)ptr), "spaPeerBufferClose"); });
    m.def("peer_buffer_destroy", [](int64_t ptr) { check(spaPeerBufferDestroy((void *)(uintptr_t)ptr), "spaPeerBufferDestroy"); });
    m.def("peer_buffer_tensor", [](int64_t ptr, int64_t bytes, int64_t device) {
        // uint8 view of a peer buffer (local or mapped); the caller keeps the buffer alive
        return at::from_blob((void *)(uintptr_t)ptr, {bytes}, at::TensorOptions().dtype(at::kByte).device(at::kCUDA, (c10::DeviceIndex)device));
    });
    m.def("make_fusion_plan", &make_fusion_plan);
    m.def("launch_fusion_plan", &launch_fusion_plan);
    m.def("abi_version", []() { return spaGetVersion(); });
    m.def("compiled_arch", []() { return spaGetCompiledArch(); });
}
