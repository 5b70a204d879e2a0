// pa_decode_static: pa_decode for callers that cannot allocate device memory (include/pa_b200.h).
//
// No counterpart in the reference.  Some callers may not call
// cudaMalloc at all - code under CUDA-graph capture, tight serving loops, harnesses that forbid it - and
// pa_decode wants a caller-owned workspace (counter header + split-K partials).  This entry point is pa_decode
// with that workspace taken from ONE file-scope device array of PA_DECODE_STATIC_BYTES: every GPU whose context
// loads this module holds the array until the process exits.  Nothing here (nor anything pa_decode reaches:
// planning, tensor-map encoding, the kernel launches) calls cudaMalloc / cudaFree or an allocating library.
//
// State of the array between calls.  It is zero when the module is loaded (static storage) and keeps the last
// call's contents afterwards.  What must start at zero is the unit ticket and the exited-CTA counter (the first
// two words): the kernel puts them back itself when its last CTA exits, and this entry point clears them again,
// stream-ordered, before every launch - eight bytes, so a call never depends on how the previous one ended.  What
// must NOT be cleared is the launch epoch word behind them: it advances by one per launch, and a partial word an
// earlier call left in the array never validates because it carries that call's epoch.  Every call on a device
// SHARES the array: two calls that can overlap on the GPU (different
// streams, or two host threads launching on unordered streams) race on the counters and on each other's
// partials - order them on one stream or with events.  Concurrent use of pa_decode with a caller-owned
// workspace is unaffected.
#include <atomic>

#include <cuda_runtime.h>

#include "../../include/pa_b200.h"

namespace {

// The scratch itself.  In host code this name is only the host-side shadow of the symbol, NOT a device address:
// the address is asked for once per device with cudaGetSymbolAddress (below) and that pointer is what the
// kernels get.
static __device__ __align__(256) unsigned char g_decode_static_ws[PA_DECODE_STATIC_BYTES];

constexpr int kMaxDevices = 64;
std::atomic<void*> g_ws_ptr[kMaxDevices];

// Device address of the array on the CURRENT device, looked up once per device with that device current.
int static_workspace(void** ptr) {
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess) { cudaGetLastError(); return PA_ERR_NO_DEVICE; }
  if (dev < 0 || dev >= kMaxDevices) return PA_ERR_UNSUPPORTED;
  void* p = g_ws_ptr[dev].load(std::memory_order_acquire);
  if (!p) {
    const cudaError_t e = cudaGetSymbolAddress(&p, g_decode_static_ws);
    if (e != cudaSuccess || !p) { cudaGetLastError(); return PA_ERR_CUDA; }
    g_ws_ptr[dev].store(p, std::memory_order_release);     // two racing first calls store the same address
  }
  *ptr = p;
  return PA_OK;
}

}  // namespace

extern "C" int pa_decode_static(const void* q, int64_t q_stride_b, int64_t q_stride_h, void* out, int64_t o_stride_b,
                                int64_t o_stride_h, const PaKvCache* cache, const int32_t* block_tables,
                                int32_t bt_stride, const int32_t* seq_lens, int32_t batch, int32_t num_q_heads,
                                void* stream) {
  if (batch < 0 || !cache) return PA_ERR_BAD_ARG;
  if (batch == 0) return PA_OK;
  if (cache->num_kv_heads <= 0 || cache->head_dim <= 0 || num_q_heads <= 0) return PA_ERR_BAD_ARG;
  // the same sizing rule a caller of pa_decode follows; a shape that needs more than the array is refused
  // before anything is launched
  const size_t need = pa_decode_workspace_bytes(batch, num_q_heads, cache->num_kv_heads, cache->head_dim);
  if (need > (size_t)PA_DECODE_STATIC_BYTES) return PA_ERR_STATIC_LIMIT;
  void* ws = nullptr;
  const int rc = static_workspace(&ws);
  if (rc != PA_OK) return rc;
  // the two counters that must be zero on entry (not the epoch word that follows them); a memset node, no allocation
  if (cudaMemsetAsync(ws, 0, 8, static_cast<cudaStream_t>(stream)) != cudaSuccess) { cudaGetLastError(); return PA_ERR_CUDA; }
  return pa_decode(q, q_stride_b, q_stride_h, out, o_stride_b, o_stride_h, cache, block_tables, bt_stride, seq_lens,
                   batch, num_q_heads, ws, (size_t)PA_DECODE_STATIC_BYTES, stream);
}
