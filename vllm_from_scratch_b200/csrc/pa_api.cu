// C ABI of libpa_b200.so (include/pa_b200.h): argument validation, split-K planning
// and kernel selection.  Nothing here touches torch; the Python host
// (vllm_from_scratch_b200/paged_attention.py) passes raw device pointers and the
// current stream.
#include <atomic>
#include <math.h>
#include <mutex>
#include <string.h>
#include <vector>

#include "../../include/pa_b200.h"
#include "../../include/pa_b200_debug.h"
#include "pa_common.cuh"

namespace {

std::atomic<int> g_force_path{0};
std::atomic<int> g_units_per_sm{16};    // fast path: smallest split-K unit = total / (units_per_sm * SMs)
std::atomic<int> g_guided_k{3};         // fast path: largest unit = remaining / (guided_k * SMs)
std::atomic<int> g_stages{0};           // fast path: per-warp ring depth override (0 = as deep as fits)
std::atomic<int> g_claim_limit{32};     // fast path: claim-ahead bound in blocks per CTA
std::atomic<long long*> g_dbg{nullptr};

int sm_count() {
  static std::atomic<int> cached[64];
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) { cudaGetLastError(); return -1; }
  if (dev >= 0 && dev < 64 && cached[dev].load() > 0) return cached[dev].load();
  int n = 0;
  if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) { cudaGetLastError(); return -1; }
  if (dev >= 0 && dev < 64) cached[dev].store(n);
  return n;
}

size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

struct WsLayout {
  size_t plan_off, ml_off, o_off, total;
};

// The workspace starts with a fixed-size counter header that must be zero when pa_decode is
// entered and is zero again when it returns (pa_decode_workspace_init zeroes it once).
constexpr size_t kHeaderBytes = 1 << 20;

// rows = partial rows per slot, slots = partial slots.  `flagged`: the TMA kernel's self-validating partial
// rows - (m, l) and the head_dim values as 16-byte words {value, flag, value, flag} - in the O area alone.
WsLayout ws_layout(int batch, int64_t slots, int rows, int head_dim, bool flagged) {
  WsLayout w;
  w.plan_off = kHeaderBytes;
  w.ml_off = w.plan_off + align_up((size_t)(batch + 3) * 4, 256);
  w.o_off = w.ml_off + align_up(flagged ? 0 : (size_t)slots * rows * sizeof(float2), 256);
  w.total = w.o_off + align_up(flagged ? 2 * (size_t)slots * rows * (head_dim / 2 + 1) * 16   // units' partials + level-2
                                       : (size_t)slots * rows * head_dim * sizeof(float), 256);
  return w;
}

// Per-kernel device timing for the roofline report (pa_profile_begin / pa_profile_end): three
// events per pa_decode call - before the split-K kernel, between it and the reduce, after.
struct ProfState {
  std::mutex mu;
  std::vector<cudaEvent_t> ev;     // 3 per call
  int cap = 0, used = 0;
  bool on = false;
} g_prof;

// The workspace is sized for the largest units_per_sm the tuning hook accepts, so changing the
// knob never invalidates a workspace the caller already holds.
constexpr int kMaxUnitsPerSm = 64;

// fast path partial slots: one per work unit
int64_t tma_slots(int batch, int num_kv_heads, int sms) {
  return pa_decode_tma_max_units(batch, num_kv_heads, kMaxUnitsPerSm * sms);
}

bool tma_header_fits(int batch, int num_kv_heads) {
  const int64_t segs = (int64_t)batch * num_kv_heads;
  return (size_t)(pa_decode_tma_header_ints() + segs) * 4 <= kHeaderBytes;
}

__global__ void pa_iota_lens_kernel(int32_t* lens, int32_t first, int32_t n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) lens[i] = first + i;
}

}  // namespace

extern "C" int pa_abi_version(void) { return PA_B200_ABI_VERSION; }

extern "C" const char* pa_status_string(int status) {
  switch (status) {
    case PA_OK: return "ok";
    case PA_ERR_BAD_ARG: return "bad argument";
    case PA_ERR_UNSUPPORTED: return "unsupported shape or dtype";
    case PA_ERR_MISALIGNED: return "misaligned pointer or stride";
    case PA_ERR_WORKSPACE: return "workspace too small";
    case PA_ERR_CUDA: return "CUDA call failed";
    case PA_ERR_NO_DEVICE: return "no sm_100 device is current";
    case PA_ERR_STATIC_LIMIT: return "shape needs more scratch than the static array of pa_decode_static holds";
    default: return "unknown status";
  }
}

extern "C" int pa_decode_force_path(int path) { return g_force_path.exchange(path); }

extern "C" int pa_decode_set_tuning(int units_per_sm, int guided_k, int stages, int claim_limit) {
  if (units_per_sm < 1 || units_per_sm > kMaxUnitsPerSm || guided_k < 1 || stages < 0 || claim_limit < 1)
    return PA_ERR_BAD_ARG;
  g_claim_limit.store(claim_limit);
  g_units_per_sm.store(units_per_sm);
  g_guided_k.store(guided_k);
  g_stages.store(stages);
  return PA_OK;
}

extern "C" int pa_decode_set_pdl(int on) { return pa_decode_tma_set_pdl(on); }

extern "C" int pa_decode_debug_times(void* device_buffer) {
  g_dbg.store(static_cast<long long*>(device_buffer));
  return PA_OK;
}

extern "C" int pa_decode_num_workers(void) {
  const int sms = sm_count();
  if (sms <= 0) return PA_ERR_NO_DEVICE;
  return sms;                                   // one persistent CTA per SM
}

extern "C" size_t pa_decode_workspace_bytes(int32_t batch, int32_t num_q_heads, int32_t num_kv_heads,
                                            int32_t head_dim) {
  if (batch <= 0 || num_q_heads <= 0 || num_kv_heads <= 0 || head_dim <= 0) return kHeaderBytes;
  int sms = sm_count();
  if (sms <= 0) sms = 148;                       // sizing without a device: assume a B200
  const int g = (num_q_heads + num_kv_heads - 1) / num_kv_heads;
  const int nt = (g + 7) / 8;
  int ctas;
  const int64_t slots_tma = tma_slots(batch, num_kv_heads, sms);
  const int64_t slots_gen = (int64_t)batch * num_kv_heads * nt + pa_decode_generic_workers(sms, &ctas);
  const size_t a = ws_layout(batch, slots_tma, g, head_dim, true).total;
  const size_t b = ws_layout(batch, slots_gen, 8, head_dim, false).total;
  return a > b ? a : b;
}

extern "C" int pa_decode_workspace_init(void* workspace, size_t workspace_bytes, void* stream) {
  if (!workspace || workspace_bytes < kHeaderBytes) return PA_ERR_WORKSPACE;
  // all of it: the header must be zero, and the partial area must not hold a stale word that looks flagged
  return cudaMemsetAsync(workspace, 0, workspace_bytes, static_cast<cudaStream_t>(stream)) == cudaSuccess ? PA_OK
                                                                                                        : PA_ERR_CUDA;
}

static int decode_impl(const void* q, int64_t q_stride_b, int64_t q_stride_h, void* out, int64_t o_stride_b,
                       int64_t o_stride_h, const PaKvCache* cache, const int32_t* block_tables,
                       int32_t bt_stride, const int32_t* seq_lens, int32_t batch, int32_t num_q_heads,
                       void* workspace, size_t workspace_bytes, const PaPeerGather* peers, void* stream) {
  if (batch < 0 || !cache) return PA_ERR_BAD_ARG;
  if (batch == 0) return PA_OK;
  if (!q || !out || !cache->k || !cache->v || !block_tables || !seq_lens || !workspace) return PA_ERR_BAD_ARG;
  if (cache->dtype != PA_BF16 && cache->dtype != PA_F32) return PA_ERR_UNSUPPORTED;
  const int Hkv = cache->num_kv_heads, D = cache->head_dim, BS = cache->block_size;
  if (Hkv <= 0 || D <= 0 || BS <= 0 || num_q_heads <= 0 || bt_stride < 0) return PA_ERR_BAD_ARG;
  if (num_q_heads % Hkv != 0) return PA_ERR_BAD_ARG;
  if (D > 256) return PA_ERR_UNSUPPORTED;
  if (batch > PA_MAX_BATCH_SMEM) return PA_ERR_UNSUPPORTED;
  if (reinterpret_cast<uintptr_t>(workspace) & 15) return PA_ERR_MISALIGNED;
  const int sms = sm_count();
  if (sms <= 0) return PA_ERR_NO_DEVICE;

  PaDecodeParams p;
  p.num_remote = 0;
  p.local_sig = nullptr;
  for (int r = 0; r < 7; ++r) { p.remote_out[r] = nullptr; p.remote_sig[r] = nullptr; }
  p.ll_epoch = nullptr;
  if (peers) {
    const bool ll = peers->ll_epoch != nullptr;            // LL form: flags travel with the data, no counters
    if (peers->num_remote < 0 || peers->num_remote > 7 || (!ll && !peers->local_sig)) return PA_ERR_BAD_ARG;
    for (int r = 0; r < peers->num_remote; ++r) {
      if (!peers->remote_out[r] || (!ll && !peers->remote_sig[r])) return PA_ERR_BAD_ARG;
      if (reinterpret_cast<uintptr_t>(peers->remote_out[r]) & (ll ? 7 : 3)) return PA_ERR_MISALIGNED;
      p.remote_out[r] = peers->remote_out[r];
      p.remote_sig[r] = ll ? nullptr : peers->remote_sig[r];
    }
    p.num_remote = peers->num_remote;
    p.local_sig = ll ? nullptr : peers->local_sig;
    p.ll_epoch = peers->ll_epoch;
    // the LL word of an output word is addressed by the word's offset in a CONTIGUOUS [B, Hq, D] slot
    if (ll && (o_stride_h != cache->head_dim || o_stride_b != (int64_t)num_q_heads * cache->head_dim))
      return PA_ERR_BAD_ARG;
  }
  p.q = q; p.out = out; p.k = cache->k; p.v = cache->v;
  p.block_tables = block_tables; p.seq_lens = seq_lens;
  p.q_stride_b = q_stride_b; p.q_stride_h = q_stride_h;
  p.o_stride_b = o_stride_b; p.o_stride_h = o_stride_h;
  p.kv_stride_block = cache->stride_block; p.kv_stride_head = cache->stride_head; p.kv_stride_slot = cache->stride_slot;
  p.bt_stride = bt_stride;
  p.batch = batch; p.num_q_heads = num_q_heads; p.num_kv_heads = Hkv; p.head_dim = D; p.block_size = BS;
  p.num_blocks = cache->num_blocks;
  p.group = num_q_heads / Hkv;
  p.pos_head = 0;
  p.scale_log2e = (float)((1.0 / sqrt((double)D)) * 1.4426950408889634);

  const bool tma = g_force_path.load() == 0 && pa_decode_tma_supported(p, cache->dtype) &&
                   tma_header_fits(batch, Hkv);
  if (peers && !tma) return PA_ERR_UNSUPPORTED;          // the fused gather lives in the fast-path kernel only
  int ctas = 0;
  bool stream_path = false;
  int64_t slots;
  int rows;
  p.num_stages = g_stages.load();
  p.guided_k = g_guided_k.load();
  p.claim_limit = g_claim_limit.load();
  p.target_units = g_units_per_sm.load() * sms;
  p.dbg = g_dbg.load();
  if (tma) {
    p.tile_q = p.group;
    p.q_tiles = 1;
    p.num_workers = sms;
    slots = tma_slots(batch, Hkv, sms);
    rows = p.group;
  } else {
    p.tile_q = 8;
    p.q_tiles = (p.group + 7) / 8;
    // 16-byte aligned rows: the cp.async streaming kernel; anything else: the plain generic kernel
    stream_path = g_force_path.load() != 2 && pa_decode_stream_supported(p, cache->dtype);
    p.num_workers = stream_path ? pa_decode_stream_workers(p, cache->dtype, sms, &ctas) : pa_decode_generic_workers(sms, &ctas);
    slots = (int64_t)batch * Hkv * p.q_tiles + p.num_workers;
    rows = p.tile_q;
  }
  const WsLayout w = ws_layout(batch, slots, rows, D, tma);
  if (workspace_bytes < w.total) return PA_ERR_WORKSPACE;
  uint8_t* ws = static_cast<uint8_t*>(workspace);
  p.counters = reinterpret_cast<int32_t*>(ws);
  p.plan = reinterpret_cast<int32_t*>(ws + w.plan_off);
  p.part_ml = reinterpret_cast<float2*>(ws + w.ml_off);
  p.part_o = reinterpret_cast<float*>(ws + w.o_off);
  p.part_o2 = tma ? p.part_o + (size_t)slots * rows * (D / 2 + 1) * 4 : nullptr;

  cudaStream_t st = static_cast<cudaStream_t>(stream);
  cudaEvent_t* pe = nullptr;
  if (g_prof.on) {
    std::lock_guard<std::mutex> lock(g_prof.mu);
    if (g_prof.on && g_prof.used < g_prof.cap) pe = &g_prof.ev[3 * g_prof.used++];
  }
  if (pe) cudaEventRecord(pe[0], st);
  int rc = tma ? pa_launch_decode_tma(p, sms, st)
               : stream_path ? pa_launch_decode_stream(p, cache->dtype, ctas, st)
                             : pa_launch_decode_generic(p, cache->dtype, ctas, st);
  if (rc) return rc;
  if (pe) cudaEventRecord(pe[1], st);
  if (!tma) rc = pa_launch_reduce(p, cache->dtype, st);      // the TMA kernel merges its own partials
  if (pe) cudaEventRecord(pe[2], st);
  return rc;
}

extern "C" int pa_decode(const void* q, int64_t q_stride_b, int64_t q_stride_h, void* out, int64_t o_stride_b,
                         int64_t o_stride_h, const PaKvCache* cache, const int32_t* block_tables,
                         int32_t bt_stride, const int32_t* seq_lens, int32_t batch, int32_t num_q_heads,
                         void* workspace, size_t workspace_bytes, void* stream) {
  return decode_impl(q, q_stride_b, q_stride_h, out, o_stride_b, o_stride_h, cache, block_tables, bt_stride, seq_lens,
                     batch, num_q_heads, workspace, workspace_bytes, nullptr, stream);
}

extern "C" int pa_decode_gather(const void* q, int64_t q_stride_b, int64_t q_stride_h, void* out, int64_t o_stride_b,
                                int64_t o_stride_h, const PaKvCache* cache, const int32_t* block_tables,
                                int32_t bt_stride, const int32_t* seq_lens, int32_t batch, int32_t num_q_heads,
                                void* workspace, size_t workspace_bytes, const PaPeerGather* peers, void* stream) {
  if (!peers) return PA_ERR_BAD_ARG;
  return decode_impl(q, q_stride_b, q_stride_h, out, o_stride_b, o_stride_h, cache, block_tables, bt_stride, seq_lens,
                     batch, num_q_heads, workspace, workspace_bytes, peers, stream);
}

// ---- peer memory plumbing for the fused gather ------------------------------------------
namespace {
// Spins until every rank's arrival counter has reached `expected` (counters are monotonic: the
// caller passes step * segments_per_rank).  Bounded: after `timeout_ns` it records the failure in
// status[0] and returns instead of hanging the GPU.
// `epoch` (device counter, may be null): when given, the step number lives on the device - the
// kernel waits for (epoch + 1) * segments and then advances epoch - so the launch has no
// per-step host argument and can sit in a CUDA graph.
__global__ void pa_peer_wait_kernel(const uint32_t* sig, int world, uint32_t expected, unsigned long long timeout_ns,
                                    int32_t* status, uint32_t* epoch, uint32_t segments) {
  const int r = threadIdx.x;
  if (epoch) {
    const uint32_t e = *reinterpret_cast<volatile uint32_t*>(epoch);
    expected = (e + 1u) * segments;
    __syncwarp();
    if (r == 0) *epoch = e + 1u;
  }
  if (r >= world) return;
  if (status && *reinterpret_cast<volatile int32_t*>(status) != 0) return;   // a peer already failed: do not wait again
  unsigned long long t0, t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t0));
  for (;;) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(sig + r) : "memory");
    if ((int32_t)(v - expected) >= 0) break;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    if (t - t0 > timeout_ns) {
      if (status) atomicExch(status, 1 + r);
      break;
    }
    __nanosleep(64);
  }
}
}  // namespace

extern "C" int pa_peer_wait(const uint32_t* local_sig, int32_t world, uint32_t expected, int64_t timeout_us,
                            int32_t* status, void* stream) {
  if (!local_sig || world < 1 || world > 8 || timeout_us <= 0) return PA_ERR_BAD_ARG;
  pa_peer_wait_kernel<<<1, 32, 0, static_cast<cudaStream_t>(stream)>>>(
      local_sig, world, expected, (unsigned long long)timeout_us * 1000ull, status, nullptr, 0u);
  return cudaGetLastError() == cudaSuccess ? PA_OK : PA_ERR_CUDA;
}

extern "C" int pa_peer_wait_next(const uint32_t* local_sig, int32_t world, uint32_t* epoch, uint32_t segments,
                                 int64_t timeout_us, int32_t* status, void* stream) {
  if (!local_sig || !epoch || world < 1 || world > 8 || timeout_us <= 0) return PA_ERR_BAD_ARG;
  pa_peer_wait_kernel<<<1, 32, 0, static_cast<cudaStream_t>(stream)>>>(
      local_sig, world, 0u, (unsigned long long)timeout_us * 1000ull, status, epoch, segments);
  return cudaGetLastError() == cudaSuccess ? PA_OK : PA_ERR_CUDA;
}

namespace {
// Receiver side of the LL gather: every 8-byte word {data, flag} that a remote rank stored into this
// rank's LL area is polled until its flag equals this step's number, and its data word is written
// into the gather set.  The last block to finish advances the step number.  Bounded like the wait.
__global__ void __launch_bounds__(128)
pa_peer_unpack_kernel(const uint2* ll, uint32_t* gathered, int world, int rank, int words, uint32_t* epoch,
                      uint32_t* done, unsigned long long timeout_ns, int32_t* status) {
  const uint32_t flag = *reinterpret_cast<volatile uint32_t*>(epoch) + 1u;
  const int total = world * words;
  bool dead = status && *reinterpret_cast<volatile int32_t*>(status) != 0;
  unsigned long long t0;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t0));
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int r = i / words;
    if (r == rank) continue;                               // the local slot was written in place by the decode kernel
    uint2 v;
    for (unsigned spin = 0;; ++spin) {
      asm volatile("ld.volatile.global.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "l"(ll + i) : "memory");
      if (v.y == flag || dead) break;
      if ((spin & 255u) == 255u) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
        if (t - t0 > timeout_ns) {
          if (status) atomicExch(status, 1 + r);
          dead = true;
        }
      }
    }
    gathered[i] = v.x;
  }
  // The kernel is launched with programmatic stream serialisation right behind pa_decode_gather (which lets its
  // dependents go at its very start): these small CTAs (128 threads, no shared memory) fit next to a decode CTA, so
  // the polling above overlaps the local decode instead of starting after it.  Nothing above depends on the local
  // decode (its own slot is skipped); everything below does - the step number may only advance once the local
  // kernel, which stamps its outgoing words with it, has finished, and the consumers behind this kernel on the
  // stream need the local slot complete as well.
  asm volatile("griddepcontrol.wait;" ::: "memory");
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    if (atomicAdd(done, 1u) == gridDim.x - 1) {
      *done = 0u;
      *epoch = flag;
    }
  }
}
}  // namespace

extern "C" int pa_peer_unpack(const void* ll_local, void* gathered, int32_t world, int32_t rank, int32_t words_per_slot,
                              uint32_t* epoch, uint32_t* done, int64_t timeout_us, int32_t* status, void* stream) {
  if (!ll_local || !gathered || !epoch || !done || world < 1 || world > 8 || rank < 0 || rank >= world ||
      words_per_slot <= 0 || timeout_us <= 0)
    return PA_ERR_BAD_ARG;
  if ((reinterpret_cast<uintptr_t>(ll_local) & 7) || (reinterpret_cast<uintptr_t>(gathered) & 3)) return PA_ERR_MISALIGNED;
  const int64_t total = (int64_t)world * words_per_slot;
  if (total > 0x7fffffff) return PA_ERR_UNSUPPORTED;
  int blocks = (int)((total + 128 * 8 - 1) / (128 * 8));   // ~8 words per thread
  if (blocks > 128) blocks = 128;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(blocks);
  cfg.blockDim = dim3(128);
  cfg.stream = static_cast<cudaStream_t>(stream);
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, pa_peer_unpack_kernel, static_cast<const uint2*>(ll_local), static_cast<uint32_t*>(gathered),
                            (int)world, (int)rank, (int)words_per_slot, epoch, done,
                            (unsigned long long)timeout_us * 1000ull, status) == cudaSuccess ? PA_OK : PA_ERR_CUDA;
}

extern "C" int pa_peer_alloc(size_t bytes, void** ptr, void* ipc_handle_64) {
  if (!ptr || !ipc_handle_64 || bytes == 0) return PA_ERR_BAD_ARG;
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  void* d = nullptr;
  if (cudaMalloc(&d, bytes) != cudaSuccess) return PA_ERR_CUDA;
  cudaIpcMemHandle_t h;
  if (cudaMemset(d, 0, bytes) != cudaSuccess || cudaDeviceSynchronize() != cudaSuccess ||
      cudaIpcGetMemHandle(&h, d) != cudaSuccess) {
    cudaFree(d);
    return PA_ERR_CUDA;
  }
  memcpy(ipc_handle_64, &h, 64);
  *ptr = d;
  return PA_OK;
}

extern "C" int pa_peer_open(const void* ipc_handle_64, void** ptr) {
  if (!ipc_handle_64 || !ptr) return PA_ERR_BAD_ARG;
  cudaIpcMemHandle_t h;
  memcpy(&h, ipc_handle_64, 64);
  return cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess) == cudaSuccess ? PA_OK : PA_ERR_CUDA;
}

extern "C" int pa_peer_close(void* ptr) {
  return ptr && cudaIpcCloseMemHandle(ptr) == cudaSuccess ? PA_OK : PA_ERR_CUDA;
}

extern "C" int pa_peer_free(void* ptr) {
  return ptr && cudaFree(ptr) == cudaSuccess ? PA_OK : PA_ERR_CUDA;
}

extern "C" int pa_profile_begin(int32_t max_calls) {
  if (max_calls <= 0) return PA_ERR_BAD_ARG;
  std::lock_guard<std::mutex> lock(g_prof.mu);
  for (cudaEvent_t e : g_prof.ev) cudaEventDestroy(e);
  g_prof.ev.assign((size_t)3 * max_calls, nullptr);
  for (auto& e : g_prof.ev)
    if (cudaEventCreate(&e) != cudaSuccess) return PA_ERR_CUDA;
  g_prof.cap = max_calls;
  g_prof.used = 0;
  g_prof.on = true;
  return PA_OK;
}

extern "C" int pa_profile_end(float* decode_ms, float* reduce_ms, int32_t cap) {
  std::lock_guard<std::mutex> lock(g_prof.mu);
  g_prof.on = false;
  int n = g_prof.used < cap ? g_prof.used : cap;
  int rc = n;
  for (int i = 0; i < n; ++i) {
    if (cudaEventSynchronize(g_prof.ev[3 * i + 2]) != cudaSuccess ||
        cudaEventElapsedTime(decode_ms + i, g_prof.ev[3 * i], g_prof.ev[3 * i + 1]) != cudaSuccess ||
        cudaEventElapsedTime(reduce_ms + i, g_prof.ev[3 * i + 1], g_prof.ev[3 * i + 2]) != cudaSuccess) {
      rc = PA_ERR_CUDA;
      break;
    }
  }
  for (cudaEvent_t e : g_prof.ev) cudaEventDestroy(e);
  g_prof.ev.clear();
  g_prof.cap = g_prof.used = 0;
  return rc;
}

// ---- prefill = one decode "sequence" per query token over the shared block table --------
namespace {
constexpr int kPrefillChunk = 4096;     // query tokens per decode launch (< PA_MAX_BATCH_SMEM)
}

extern "C" size_t pa_prefill_workspace_bytes(int32_t num_tokens, int32_t num_q_heads, int32_t num_kv_heads,
                                             int32_t head_dim) {
  const int chunk = num_tokens < kPrefillChunk ? (num_tokens > 0 ? num_tokens : 1) : kPrefillChunk;
  return pa_decode_workspace_bytes(chunk, num_q_heads, num_kv_heads, head_dim) + align_up((size_t)chunk * 4, 256);
}

namespace {
std::atomic<int> g_prefill_force{0};

int fill_prefill_params(PaPrefillParams& p, const void* q, int64_t q_stride_t, int64_t q_stride_h, void* out,
                        int64_t o_stride_t, int64_t o_stride_h, const PaKvCache* cache, int32_t num_q_heads) {
  memset(&p, 0, sizeof(p));
  if (!cache || cache->num_kv_heads <= 0 || num_q_heads <= 0 || num_q_heads % cache->num_kv_heads) return PA_ERR_BAD_ARG;
  p.q = q; p.out = out; p.k = cache->k; p.v = cache->v;
  p.q_stride_t = q_stride_t; p.q_stride_h = q_stride_h; p.o_stride_t = o_stride_t; p.o_stride_h = o_stride_h;
  p.kv_stride_block = cache->stride_block; p.kv_stride_head = cache->stride_head; p.kv_stride_slot = cache->stride_slot;
  p.num_q_heads = num_q_heads; p.num_kv_heads = cache->num_kv_heads; p.head_dim = cache->head_dim;
  p.num_blocks = cache->num_blocks;
  p.group = num_q_heads / cache->num_kv_heads;
  p.scale_log2e = (float)((1.0 / sqrt((double)cache->head_dim)) * 1.4426950408889634);
  return PA_OK;
}
}  // namespace

extern "C" int pa_prefill_force_path(int path) { return g_prefill_force.exchange(path); }

extern "C" int pa_prefill_tc_available(const PaKvCache* cache, int32_t num_q_heads) {
  PaPrefillParams p;
  if (!cache || !cache->k || !cache->v) return 0;
  if (fill_prefill_params(p, cache->k, 0, 0, cache->k, 0, 0, cache, num_q_heads) != PA_OK) return 0;
  p.q_stride_t = p.o_stride_t = (int64_t)num_q_heads * cache->head_dim;
  p.q_stride_h = p.o_stride_h = cache->head_dim;
  p.batch = 1; p.max_q_len = 1;
  return pa_prefill_tc_supported(p, cache->dtype, cache->block_size) ? 1 : 0;
}

extern "C" int pa_prefill_varlen(const void* q, int64_t q_stride_t, int64_t q_stride_h, void* out, int64_t o_stride_t,
                                 int64_t o_stride_h, const PaKvCache* cache, const int32_t* block_tables,
                                 int32_t bt_stride, int32_t max_blocks, const int32_t* cu_q_lens,
                                 const int32_t* ctx_lens, int32_t batch, int32_t max_q_len, int32_t num_q_heads,
                                 void* stream) {
  if (batch < 0 || max_q_len < 0 || max_blocks < 0 || !cache) return PA_ERR_BAD_ARG;
  if (batch == 0 || max_q_len == 0) return PA_OK;
  if (!q || !out || !cache->k || !cache->v || !block_tables || !cu_q_lens || !ctx_lens || bt_stride < 0)
    return PA_ERR_BAD_ARG;
  if (sm_count() <= 0) return PA_ERR_NO_DEVICE;
  PaPrefillParams p;
  int rc = fill_prefill_params(p, q, q_stride_t, q_stride_h, out, o_stride_t, o_stride_h, cache, num_q_heads);
  if (rc) return rc;
  p.block_tables = block_tables; p.bt_stride = bt_stride; p.max_blocks = max_blocks;
  p.cu_q_lens = cu_q_lens; p.ctx_lens = ctx_lens;
  p.batch = batch; p.max_q_len = max_q_len;
  if (!pa_prefill_tc_supported(p, cache->dtype, cache->block_size)) return PA_ERR_UNSUPPORTED;
  return g_prefill_force.load() == 2 ? pa_launch_prefill_tc(p, static_cast<cudaStream_t>(stream))
                                     : pa_launch_prefill_tc2(p, static_cast<cudaStream_t>(stream));
}

extern "C" int pa_prefill(const void* q, int64_t q_stride_t, int64_t q_stride_h, void* out, int64_t o_stride_t,
                          int64_t o_stride_h, const PaKvCache* cache, const int32_t* block_table,
                          int32_t start_pos, int32_t num_tokens, int32_t num_q_heads, void* workspace,
                          size_t workspace_bytes, void* stream) {
  if (num_tokens < 0 || start_pos < 0 || !cache) return PA_ERR_BAD_ARG;
  if (num_tokens == 0) return PA_OK;
  if (!workspace || (reinterpret_cast<uintptr_t>(workspace) & 15)) return PA_ERR_BAD_ARG;
  if (cache->dtype != PA_BF16 && cache->dtype != PA_F32) return PA_ERR_UNSUPPORTED;
  if (workspace_bytes < pa_prefill_workspace_bytes(num_tokens, num_q_heads, cache->num_kv_heads, cache->head_dim))
    return PA_ERR_WORKSPACE;
  if (g_prefill_force.load() != 1 && q && out && block_table && cache->k && cache->v) {
    PaPrefillParams tp;
    if (fill_prefill_params(tp, q, q_stride_t, q_stride_h, out, o_stride_t, o_stride_h, cache, num_q_heads) == PA_OK) {
      tp.block_tables = block_table; tp.bt_stride = 0;
      tp.batch = 1; tp.num_tokens = num_tokens; tp.max_q_len = num_tokens;
      tp.ctx_len_single = start_pos + num_tokens;
      if (pa_prefill_tc_supported(tp, cache->dtype, cache->block_size)) {
        if (sm_count() <= 0) return PA_ERR_NO_DEVICE;
        return g_prefill_force.load() == 2 ? pa_launch_prefill_tc(tp, static_cast<cudaStream_t>(stream))
                                           : pa_launch_prefill_tc2(tp, static_cast<cudaStream_t>(stream));
      }
    }
  }
  const int es = cache->dtype == PA_F32 ? 4 : 2;
  const int chunk_cap = num_tokens < kPrefillChunk ? num_tokens : kPrefillChunk;
  // the decode workspace (with its zeroed counter header) sits at offset 0; the lens array follows it
  const size_t dws_bytes = pa_decode_workspace_bytes(chunk_cap, num_q_heads, cache->num_kv_heads, cache->head_dim);
  uint8_t* dws = static_cast<uint8_t*>(workspace);
  int32_t* lens = reinterpret_cast<int32_t*>(dws + dws_bytes);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  for (int t0 = 0; t0 < num_tokens; t0 += kPrefillChunk) {
    const int n = num_tokens - t0 < kPrefillChunk ? num_tokens - t0 : kPrefillChunk;
    pa_iota_lens_kernel<<<(n + 255) / 256, 256, 0, st>>>(lens, start_pos + t0 + 1, n);
    if (cudaGetLastError() != cudaSuccess) return PA_ERR_CUDA;
    const int rc = pa_decode(static_cast<const uint8_t*>(q) + (int64_t)t0 * q_stride_t * es, q_stride_t, q_stride_h,
                             static_cast<uint8_t*>(out) + (int64_t)t0 * o_stride_t * es, o_stride_t, o_stride_h,
                             cache, block_table, /*bt_stride=*/0, lens, n, num_q_heads, dws, dws_bytes, stream);
    if (rc) return rc;
  }
  return PA_OK;
}
