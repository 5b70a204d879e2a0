// Thin torch C++ extension over the C ABI of libpa_b200.so (include/pa_b200.h).
//
// north_star: "Host code stays Python/PyTorch and calls hand-written sm_100a CUDA through a thin torch C++
// extension (C-ABI)".  This file is that extension: every function takes torch tensors, validates them the
// way the Python mirror of the reference API documents (same exception types: ValueError / TypeError /
// IndexError / RuntimeError), builds the strided PaKvCache descriptor, picks up the current device and
// stream, and calls the SAME extern "C" entry point a ctypes / cgo / JNI binding would call.  Nothing is
// computed here and no kernel lives here; libpa_b200.so stays free of torch.  What it buys is host time:
// the ctypes path costs ~20 us of Python per call (descriptor building, argument marshalling, stream and
// device lookups), this one ~3 us - the difference is most of an eager engine step at toy model sizes.
#include <torch/extension.h>
#include <torch/csrc/Exceptions.h>

#include <c10/cuda/CUDAGuard.h>
#include <c10/cuda/CUDAStream.h>

#include "../../include/pa_b200.h"

namespace {

using at::Tensor;
using c10::optional;

// Messages are assembled with std::string appends only: the stream-formatting path of c10::str (std::ostringstream
// instantiated in THIS translation unit) crashed on the GPU boxes for any message that held a number, while the
// same binary worked in the build container; nothing here is worth a second C++ runtime dependency.
inline void app(std::string& s, const char* v) { s += v ? v : "?"; }
inline void app(std::string& s, const std::string& v) { s += v; }
inline void app(std::string& s, int64_t v) { s += std::to_string(v); }
inline void app(std::string& s, int v) { s += std::to_string(v); }
inline void app(std::string& s, size_t v) { s += std::to_string(v); }
inline void app(std::string& s, c10::ScalarType v) { s += c10::toString(v); }
inline void app(std::string& s, const c10::Device& v) { s += v.str(); }
template <typename... A>
std::string cat(const A&... a) {
  std::string s;
  (app(s, a), ...);
  return s;
}

void need_cuda(const Tensor& t, const char* name) {
  TORCH_CHECK(t.is_cuda(), cat(name, " is on ", t.device(),
              ": the B200 paged-attention path has no CPU fallback; allocate the tensor on a CUDA device"));
}
void need_cuda_or_pinned(const Tensor& t, const char* name) {
  TORCH_CHECK(t.is_cuda() || t.is_pinned(), cat(name,
              " is pageable host memory: the B200 paged-attention path has no CPU fallback; allocate it on a CUDA "
              "device (or pin it for zero-copy access)"));
}
void status(int rc, const char* what) {
  if (rc == PA_OK) return;
  const char* text = pa_status_string(rc);
  const std::string msg = std::string(what) + " failed: " + (text ? text : "?") + " (status " + std::to_string(rc) + ")";
  TORCH_CHECK(false, msg);
}

// Either one tensor kv_cache[2, NB, Hkv, BS, D] ('hbsd', 01_paged_attention/paged_attention.py:26-30) or a pair
// (k_cache, v_cache)[NB, BS, H, D] ('bshd', 05_mini_vllm/mini_vllm.py:544-552); described by strides, never copied.
PaKvCache describe(const Tensor& a, const optional<Tensor>& b) {
  PaKvCache c;
  int64_t nb, h, bs, d, sb, sh, ss, sd;
  if (b.has_value()) {
    const Tensor& v = *b;
    need_cuda(a, "k_cache");
    need_cuda(v, "v_cache");
    TORCH_CHECK_VALUE(a.dim() == 4 && v.sizes() == a.sizes() && v.strides() == a.strides() && v.scalar_type() == a.scalar_type(), cat("'bshd' cache must be a (k_cache, v_cache) pair of identical [NB, BS, H, D] tensors"));
    nb = a.size(0); bs = a.size(1); h = a.size(2); d = a.size(3);
    sb = a.stride(0); ss = a.stride(1); sh = a.stride(2); sd = a.stride(3);
    c.k = a.data_ptr();
    c.v = v.data_ptr();
  } else {
    need_cuda(a, "kv_cache");
    TORCH_CHECK_VALUE(a.dim() == 5 && a.size(0) == 2, cat("'hbsd' cache must be kv_cache[2, NB, Hkv, BS, D]"));
    nb = a.size(1); h = a.size(2); bs = a.size(3); d = a.size(4);
    sb = a.stride(1); sh = a.stride(2); ss = a.stride(3); sd = a.stride(4);
    c.k = a.data_ptr();
    c.v = static_cast<uint8_t*>(a.data_ptr()) + a.stride(0) * a.element_size();
  }
  TORCH_CHECK_VALUE(sd == 1 || d <= 1, cat("the head_dim axis of the KV cache must be contiguous"));
  TORCH_CHECK_TYPE(a.scalar_type() == at::kBFloat16 || a.scalar_type() == at::kFloat, cat("unsupported KV cache dtype ", a.scalar_type(), " (bf16 / fp32 only)"));
  c.stride_block = sb; c.stride_head = sh; c.stride_slot = ss;
  c.num_blocks = (int32_t)nb; c.num_kv_heads = (int32_t)h; c.block_size = (int32_t)bs; c.head_dim = (int32_t)d;
  c.dtype = a.scalar_type() == at::kFloat ? PA_F32 : PA_BF16;
  c.reserved = 0;
  return c;
}

void* stream_of(const Tensor& t) { return c10::cuda::getCurrentCUDAStream(t.device().index()).stream(); }

// ---- KV slot write (pa_reshape_and_cache) ---------------------------------------------------------------
void reshape_and_cache(Tensor key, Tensor value, const Tensor& ca, const optional<Tensor>& cb, const Tensor& slot_mapping) {
  const PaKvCache c = describe(ca, cb);
  need_cuda(key, "key");
  need_cuda(value, "value");
  need_cuda(slot_mapping, "slot_mapping");
  TORCH_CHECK_VALUE(key.dim() == 3 && key.sizes() == value.sizes(), cat("key / value must both be [T, Hkv, D]"));
  TORCH_CHECK_VALUE(key.size(1) == c.num_kv_heads && key.size(2) == c.head_dim, cat("key is [T, ", key.size(1), ", ", key.size(2),
                    "] but the cache holds ", c.num_kv_heads, " heads of dim ", c.head_dim));
  TORCH_CHECK_TYPE(key.scalar_type() == ca.scalar_type() && value.scalar_type() == ca.scalar_type(), cat("key/value dtype ",
                   key.scalar_type(), "/", value.scalar_type(), " != cache dtype ", ca.scalar_type()));
  TORCH_CHECK_VALUE(slot_mapping.numel() == key.size(0), cat("slot_mapping must have one entry per token"));
  if (key.stride(2) != 1 && key.size(2) > 1) key = key.contiguous();
  if (value.strides() != key.strides()) { value = value.contiguous(); key = key.contiguous(); }
  const Tensor slots = slot_mapping.scalar_type() == at::kLong && slot_mapping.is_contiguous()
                           ? slot_mapping : slot_mapping.to(at::kLong).contiguous();
  int rc = PA_OK;
  {
    c10::cuda::CUDAGuard guard(ca.device());
    rc = pa_reshape_and_cache(key.data_ptr(), value.data_ptr(), key.stride(0), key.stride(1), &c,
                                slots.data_ptr<int64_t>(), (int32_t)key.size(0), stream_of(ca));
  }
  status(rc, "pa_reshape_and_cache");
}

// ---- fused RoPE + QKV split + KV slot write (pa_rope_and_cache) ---------------------------------------------
Tensor rope_and_cache(Tensor qkv, const Tensor& positions, Tensor cos_cached, Tensor sin_cached, const Tensor& ca,
                      const optional<Tensor>& cb, const Tensor& slot_mapping, int64_t num_q_heads, optional<Tensor> q_out) {
  const PaKvCache c = describe(ca, cb);
  need_cuda(qkv, "qkv"); need_cuda(positions, "positions"); need_cuda(cos_cached, "cos_cached");
  need_cuda(sin_cached, "sin_cached"); need_cuda(slot_mapping, "slot_mapping");
  const int64_t d = c.head_dim, hkv = c.num_kv_heads;
  TORCH_CHECK_VALUE(qkv.dim() == 2 && qkv.size(1) == (num_q_heads + 2 * hkv) * d, cat("qkv must be [T, ", (num_q_heads + 2 * hkv) * d, "]"));
  TORCH_CHECK_TYPE(qkv.scalar_type() == cos_cached.scalar_type() && qkv.scalar_type() == sin_cached.scalar_type() &&
                   qkv.scalar_type() == ca.scalar_type(), cat("qkv, the cos / sin tables and the cache must share one dtype"));
  TORCH_CHECK_VALUE(cos_cached.sizes() == sin_cached.sizes() && cos_cached.dim() == 2 && cos_cached.size(1) == d, cat("cos_cached / sin_cached must be [max_len, D]"));
  const int64_t t = qkv.size(0);
  TORCH_CHECK_VALUE(positions.numel() == t && slot_mapping.numel() == t, cat("positions and slot_mapping must have one entry per token"));
  if (qkv.stride(1) != 1) qkv = qkv.contiguous();
  cos_cached = cos_cached.contiguous();
  sin_cached = sin_cached.contiguous();
  Tensor out = q_out.has_value() ? *q_out : at::empty({t, num_q_heads, d}, qkv.options());
  const Tensor pos = positions.scalar_type() == at::kLong && positions.is_contiguous() ? positions : positions.to(at::kLong).contiguous();
  const Tensor slots = slot_mapping.scalar_type() == at::kLong && slot_mapping.is_contiguous()
                           ? slot_mapping : slot_mapping.to(at::kLong).contiguous();
  int rc = PA_OK;
  {
    c10::cuda::CUDAGuard guard(ca.device());
    rc = pa_rope_and_cache(qkv.data_ptr(), qkv.stride(0), (int32_t)num_q_heads, cos_cached.data_ptr(), sin_cached.data_ptr(),
                             cos_cached.stride(0), (int32_t)cos_cached.size(0), pos.data_ptr<int64_t>(), slots.data_ptr<int64_t>(),
                             (int32_t)t, out.data_ptr(), out.stride(0), out.stride(1), &c, stream_of(ca));
  }
  status(rc, "pa_rope_and_cache");
  return out;
}

// ---- batched decode (pa_decode / pa_decode_gather) --------------------------------------------------------------
Tensor decode(Tensor query, const Tensor& ca, const optional<Tensor>& cb, Tensor block_tables, Tensor seq_lens,
              optional<Tensor> out_opt, const Tensor& workspace, int64_t peers_ptr) {
  const PaKvCache c = describe(ca, cb);
  need_cuda_or_pinned(query, "query");
  need_cuda(block_tables, "block_tables");
  need_cuda(seq_lens, "seq_lens");
  TORCH_CHECK_VALUE(query.dim() == 3, cat("query must be [B, Hq, D]"));
  const int64_t b = query.size(0), hq = query.size(1), d = query.size(2);
  TORCH_CHECK_VALUE(d == c.head_dim, cat("query head_dim ", d, " != cache head_dim ", c.head_dim));
  TORCH_CHECK_VALUE(hq % c.num_kv_heads == 0, cat("num_q_heads ", hq, " is not a multiple of num_kv_heads ", c.num_kv_heads));
  TORCH_CHECK_TYPE(query.scalar_type() == ca.scalar_type(), cat("query dtype ", query.scalar_type(), " != cache dtype ", ca.scalar_type()));
  TORCH_CHECK_TYPE(block_tables.scalar_type() == at::kInt && seq_lens.scalar_type() == at::kInt, cat("block_tables and seq_lens must be int32 device tensors"));
  TORCH_CHECK_VALUE(block_tables.dim() == 2 && block_tables.size(0) == b && seq_lens.numel() == b, cat("block_tables must be [B, max_blocks] and seq_lens [B]"));
  if (block_tables.size(1) > 1 && block_tables.stride(1) != 1) block_tables = block_tables.contiguous();
  if (query.stride(2) != 1 && d > 1) query = query.contiguous();
  seq_lens = seq_lens.contiguous();
  Tensor out;
  if (out_opt.has_value()) {
    out = *out_opt;
    need_cuda_or_pinned(out, "out");
    TORCH_CHECK_VALUE(out.dim() == 3 && out.size(0) == b && out.size(1) == hq && out.size(2) == d &&
                      out.scalar_type() == query.scalar_type() && (out.stride(2) == 1 || d <= 1), cat("out must be [B, Hq, D] with the query's dtype and a contiguous last axis"));
  } else {
    out = at::empty({b, hq, d}, ca.options());
  }
  const size_t need = pa_decode_workspace_bytes((int32_t)b, (int32_t)hq, c.num_kv_heads, (int32_t)d);
  const size_t have = (size_t)workspace.numel() * workspace.element_size();
  TORCH_CHECK_VALUE(workspace.is_cuda() && have >= need, cat("workspace must be a CUDA tensor of at least ", need, " bytes"));
  int rc = PA_OK;
  {
    c10::cuda::CUDAGuard guard(ca.device());
    if (peers_ptr == 0)
      rc = pa_decode(query.data_ptr(), query.stride(0), query.stride(1), out.data_ptr(), out.stride(0), out.stride(1), &c,
                       block_tables.data_ptr<int32_t>(), (int32_t)block_tables.stride(0), seq_lens.data_ptr<int32_t>(), (int32_t)b,
                       (int32_t)hq, workspace.data_ptr(), have, stream_of(ca));
    else
      rc = pa_decode_gather(query.data_ptr(), query.stride(0), query.stride(1), out.data_ptr(), out.stride(0), out.stride(1), &c,
                              block_tables.data_ptr<int32_t>(), (int32_t)block_tables.stride(0), seq_lens.data_ptr<int32_t>(),
                              (int32_t)b, (int32_t)hq, workspace.data_ptr(), have,
                              reinterpret_cast<const PaPeerGather*>(peers_ptr), stream_of(ca));
  }
  status(rc, peers_ptr == 0 ? "pa_decode" : "pa_decode_gather");
  return out;
}

int64_t decode_workspace_bytes(int64_t b, int64_t hq, int64_t hkv, int64_t d) {
  return (int64_t)pa_decode_workspace_bytes((int32_t)b, (int32_t)hq, (int32_t)hkv, (int32_t)d);
}

// ---- prefill ---------------------------------------------------------------------------------------------
bool prefill_tc_available(const Tensor& ca, const optional<Tensor>& cb, int64_t num_q_heads) {
  const PaKvCache c = describe(ca, cb);
  return pa_prefill_tc_available(&c, (int32_t)num_q_heads) != 0;
}

// Returns the status: PA_ERR_UNSUPPORTED (-2) means "not a tensor-core shape" and the caller loops per sequence.
int64_t prefill_varlen(Tensor query, const Tensor& ca, const optional<Tensor>& cb, const Tensor& block_tables,
                       const Tensor& cu_q_lens, const Tensor& ctx_lens, int64_t max_q_len, const Tensor& out) {
  const PaKvCache c = describe(ca, cb);
  need_cuda(query, "query"); need_cuda(block_tables, "block_tables"); need_cuda(cu_q_lens, "cu_q_lens"); need_cuda(ctx_lens, "ctx_lens");
  TORCH_CHECK_VALUE(query.dim() == 3 && query.size(2) == c.head_dim && query.scalar_type() == ca.scalar_type(), cat("query must be [T, Hq, D] in the cache dtype"));
  TORCH_CHECK_VALUE(block_tables.dim() == 2 && block_tables.scalar_type() == at::kInt && block_tables.stride(1) == 1, cat("block_tables must be int32 [B, max_blocks] with a contiguous last axis"));
  TORCH_CHECK_VALUE(cu_q_lens.scalar_type() == at::kInt && ctx_lens.scalar_type() == at::kInt, cat("cu_q_lens and ctx_lens must be int32"));
  const int64_t b = block_tables.size(0);
  TORCH_CHECK_VALUE(cu_q_lens.numel() == b + 1 && ctx_lens.numel() == b, cat("cu_q_lens must have B + 1 entries and ctx_lens B"));
  TORCH_CHECK_VALUE(query.size(1) % c.num_kv_heads == 0, cat("num_q_heads must be a multiple of num_kv_heads"));
  TORCH_CHECK_VALUE(query.stride(2) == 1 || query.size(2) <= 1, cat("query must have a contiguous last axis"));
  if (query.size(0) == 0 || b == 0) return PA_OK;
  const Tensor cu = cu_q_lens.contiguous(), cl = ctx_lens.contiguous();
  int rc = PA_OK;
  {
    c10::cuda::CUDAGuard guard(query.device());
    rc = pa_prefill_varlen(query.data_ptr(), query.stride(0), query.stride(1), out.data_ptr(), out.stride(0), out.stride(1), &c,
                                     block_tables.data_ptr<int32_t>(), (int32_t)block_tables.stride(0), (int32_t)block_tables.size(1),
                                     cu.data_ptr<int32_t>(), cl.data_ptr<int32_t>(), (int32_t)b, (int32_t)max_q_len,
                                     (int32_t)query.size(1), stream_of(query));
  }
  if (rc != PA_ERR_UNSUPPORTED) status(rc, "pa_prefill_varlen");
  return rc;
}

void prefill_chunk(const Tensor& query, const Tensor& ca, const optional<Tensor>& cb, const Tensor& block_table, int64_t start_pos,
                   const Tensor& out, const Tensor& workspace) {
  const PaKvCache c = describe(ca, cb);
  need_cuda(query, "query");
  need_cuda(block_table, "block_table");
  TORCH_CHECK_VALUE(query.dim() == 3 && query.size(2) == c.head_dim && query.scalar_type() == ca.scalar_type(), cat("query must be [T, Hq, D] in the cache dtype"));
  TORCH_CHECK_VALUE(query.size(1) % c.num_kv_heads == 0, cat("num_q_heads must be a multiple of num_kv_heads"));
  TORCH_CHECK_VALUE((query.stride(2) == 1 || query.size(2) <= 1) && block_table.scalar_type() == at::kInt && block_table.is_contiguous(), cat("query needs a contiguous last axis and block_table must be contiguous int32"));
  const int64_t t = query.size(0);
  const int64_t need_blocks = (start_pos + t + c.block_size - 1) / c.block_size;
  // what the reference's block_table[i // BS] raises (paged_attention.py:119-123)
  TORCH_CHECK_INDEX(block_table.numel() >= need_blocks, cat("block_table has ", block_table.numel(), " entries, positions up to ",
                    start_pos + t, " need ", need_blocks));
  int rc = PA_OK;
  {
    c10::cuda::CUDAGuard guard(query.device());
    rc = pa_prefill(query.data_ptr(), query.stride(0), query.stride(1), out.data_ptr(), out.stride(0), out.stride(1), &c,
                      block_table.data_ptr<int32_t>(), (int32_t)start_pos, (int32_t)t, (int32_t)query.size(1), workspace.data_ptr(),
                      (size_t)workspace.numel() * workspace.element_size(), stream_of(query));
  }
  status(rc, "pa_prefill");
}

// ---- persistent block tables (pa_block_table_step) -----------------------------------------------------------------
void block_table_step(const Tensor& table, const optional<Tensor>& updates, const optional<Tensor>& num_updates,
                      const optional<Tensor>& row_of, const optional<Tensor>& ctx_lens, const optional<Tensor>& step_table,
                      int64_t block_size) {
  need_cuda(table, "table");
  TORCH_CHECK_VALUE(table.scalar_type() == at::kInt && table.dim() == 2 && table.stride(1) == 1, cat("table must be int32 [rows, W] with a contiguous last axis"));
  const int32_t* upd = nullptr;
  const int32_t* nupd = nullptr;
  int32_t max_upd = 0;
  if (updates.has_value() && updates->numel()) {
    need_cuda(*updates, "updates");
    TORCH_CHECK(num_updates.has_value(), cat("num_updates is required with updates"));
    need_cuda(*num_updates, "num_updates");
    TORCH_CHECK_VALUE(updates->scalar_type() == at::kInt && num_updates->scalar_type() == at::kInt && updates->is_contiguous(), cat("updates must be contiguous int32 [cap, 3] and num_updates an int32 device tensor"));
    max_upd = (int32_t)(updates->numel() / 3);
    upd = updates->data_ptr<int32_t>();
    nupd = num_updates->data_ptr<int32_t>();
  }
  const int32_t* rows = nullptr;
  const int32_t* lens = nullptr;
  int32_t* st = nullptr;
  int32_t n_gather = 0, st_stride = 0;
  if (row_of.has_value() && row_of->numel()) {
    TORCH_CHECK(ctx_lens.has_value() && step_table.has_value(), cat("ctx_lens and step_table are required with row_of"));
    need_cuda(*row_of, "row_of"); need_cuda(*ctx_lens, "ctx_lens"); need_cuda(*step_table, "step_table");
    TORCH_CHECK_VALUE(row_of->scalar_type() == at::kInt && ctx_lens->scalar_type() == at::kInt && step_table->scalar_type() == at::kInt, cat("row_of, ctx_lens and step_table must be int32"));
    n_gather = (int32_t)row_of->numel();
    TORCH_CHECK_VALUE(ctx_lens->numel() == n_gather && step_table->dim() == 2 && step_table->size(0) >= n_gather &&
                      step_table->stride(1) == 1 && row_of->is_contiguous() && ctx_lens->is_contiguous(), cat("row_of / ctx_lens must be contiguous [n] and step_table [>= n, W]"));
    rows = row_of->data_ptr<int32_t>();
    lens = ctx_lens->data_ptr<int32_t>();
    st = step_table->data_ptr<int32_t>();
    st_stride = (int32_t)step_table->stride(0);
  }
  int rc = PA_OK;
  {
    c10::cuda::CUDAGuard guard(table.device());
    rc = pa_block_table_step(table.data_ptr<int32_t>(), (int32_t)table.stride(0), (int32_t)table.size(0), upd, nupd, max_upd, rows, lens,
                               n_gather, (int32_t)block_size, st, st_stride, stream_of(table));
  }
  status(rc, "pa_block_table_step");
}

}  // namespace

PYBIND11_MODULE(TORCH_EXTENSION_NAME, m) {
  m.doc() = "thin torch C++ extension over libpa_b200.so (include/pa_b200.h)";
  // every function goes through torch::wrap_pybind_function: c10 errors become the matching Python exception
  // (ValueError / TypeError / IndexError / RuntimeError) WITHOUT the C++ backtrace - pybind11's default translator
  // calls what(), whose symbolised backtrace crashed the interpreter once a CUDA context was up
  m.def("abi_version", []() { return pa_abi_version(); });
  // test hook: raise each exception kind the shim uses (tests/test_c_abi.py checks the Python types)
  m.def("_raise", torch::wrap_pybind_function([](int64_t kind, int64_t n) {
    if (kind == 0) TORCH_CHECK(false, cat("runtime ", n));
    if (kind == 1) TORCH_CHECK_VALUE(false, cat("value ", n));
    if (kind == 2) TORCH_CHECK_TYPE(false, cat("type ", n));
    if (kind == 3) TORCH_CHECK_INDEX(false, cat("index ", n, " of ", n + 1));
    if (kind == 4) status(-3, "pa_test");
  }));
  m.def("reshape_and_cache", torch::wrap_pybind_function(reshape_and_cache));
  m.def("rope_and_cache", torch::wrap_pybind_function(rope_and_cache));
  m.def("decode", torch::wrap_pybind_function(decode));
  m.def("decode_workspace_bytes", torch::wrap_pybind_function(decode_workspace_bytes));
  m.def("prefill_tc_available", torch::wrap_pybind_function(prefill_tc_available));
  m.def("prefill_varlen", torch::wrap_pybind_function(prefill_varlen));
  m.def("prefill_chunk", torch::wrap_pybind_function(prefill_chunk));
  m.def("block_table_step", torch::wrap_pybind_function(block_table_step));
}
