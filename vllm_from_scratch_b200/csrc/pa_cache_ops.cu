// KV-cache data movement: slot write (reshape-and-cache), block copy, block swap.
// All three are pure byte copies - bit-exact by construction - and HBM- (or PCIe-) bound.
//
//  * pa_reshape_and_cache : the slot-mapping scatter of new K/V rows.  Replaces the Python
//    write loops 01_paged_attention/paged_attention.py:48-51 and :119-123 (layout
//    [2, NB, Hkv, BS, D]) and the index_put 05_mini_vllm/mini_vllm.py:424-427 (layout
//    [NB, BS, H, D]); the cache layout is described by strides, so one kernel serves both.
//  * pa_copy_blocks / pa_swap_blocks : no reference code exists;
//    definitional dst[b_dst] = src[b_src] over every plane of a pool.
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/pa_b200.h"

namespace {

__device__ __forceinline__ uint4 ld_stream(const uint4* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}

struct RacParams {
  const uint8_t* key;
  const uint8_t* value;
  uint8_t* k_cache;
  uint8_t* v_cache;
  const int64_t* slots;
  int64_t src_stride_token, src_stride_head;           // bytes
  int64_t stride_block, stride_head, stride_slot;      // bytes
  int32_t num_tokens, num_heads, block_size;
  int32_t row_units;                                   // units (16 B or esize) per [D] row
  int64_t num_slots;                                   // NB * BS, for the bounds check
};

// One work item = one unit of one (token, head) row, for K and for V.  UNIT_T = uint4 when every
// address is 16-byte aligned (the normal case), else the element type.  Each thread keeps UNROLL
// independent loads in flight (2 * UNROLL with K and V).
template <typename UNIT_T, int UNROLL>
__global__ void __launch_bounds__(256)
pa_reshape_and_cache_kernel(const RacParams p) {
  const int64_t per_token = (int64_t)p.num_heads * p.row_units;
  const int64_t total = (int64_t)p.num_tokens * per_token;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t base = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; base < total; base += stride * UNROLL) {
    UNIT_T kv[UNROLL], vv[UNROLL];
    int64_t dst[UNROLL];
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) {
      const int64_t i = base + u * stride;
      dst[u] = -1;
      if (i < total) {
        const int64_t t = i / per_token;
        const int r = (int)(i - t * per_token);
        const int h = r / p.row_units, c = r - h * p.row_units;
        const int64_t slot = __ldg(p.slots + t);
        if (slot >= 0 && slot < p.num_slots) {             // negative slot = padding token
          const int64_t blk = slot / p.block_size, off = slot - blk * p.block_size;
          const int64_t src = t * p.src_stride_token + h * p.src_stride_head + (int64_t)c * sizeof(UNIT_T);
          dst[u] = blk * p.stride_block + h * p.stride_head + off * p.stride_slot + (int64_t)c * sizeof(UNIT_T);
          kv[u] = *reinterpret_cast<const UNIT_T*>(p.key + src);
          vv[u] = *reinterpret_cast<const UNIT_T*>(p.value + src);
        }
      }
    }
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) {
      if (dst[u] >= 0) {
        *reinterpret_cast<UNIT_T*>(p.k_cache + dst[u]) = kv[u];
        *reinterpret_cast<UNIT_T*>(p.v_cache + dst[u]) = vv[u];
      }
    }
  }
}

// grid.x = chunks per block, grid.y = pair, grid.z = plane.  One CTA copies kCopyChunk bytes.
constexpr int kCopyThreads = 256;
constexpr int kCopyUnroll = 4;
constexpr int kCopyChunk = kCopyThreads * kCopyUnroll * 16;      // 16 KiB

__global__ void __launch_bounds__(kCopyThreads)
pa_copy_blocks_kernel(uint8_t* base, int64_t plane_stride, int64_t block_stride, int64_t block_bytes,
                      const int64_t* __restrict__ pairs) {
  const int64_t src_b = __ldg(pairs + 2 * blockIdx.y), dst_b = __ldg(pairs + 2 * blockIdx.y + 1);
  if (src_b == dst_b) return;
  const uint8_t* src = base + blockIdx.z * plane_stride + src_b * block_stride;
  uint8_t* dst = base + blockIdx.z * plane_stride + dst_b * block_stride;
  const int64_t lo = (int64_t)blockIdx.x * kCopyChunk;
  uint4 v[kCopyUnroll];
#pragma unroll
  for (int u = 0; u < kCopyUnroll; ++u) {
    const int64_t o = lo + ((int64_t)u * kCopyThreads + threadIdx.x) * 16;
    if (o < block_bytes) v[u] = ld_stream(reinterpret_cast<const uint4*>(src + o));
  }
#pragma unroll
  for (int u = 0; u < kCopyUnroll; ++u) {
    const int64_t o = lo + ((int64_t)u * kCopyThreads + threadIdx.x) * 16;
    if (o < block_bytes) *reinterpret_cast<uint4*>(dst + o) = v[u];
  }
}

// Gather / scatter between a pool and a dense staging area laid out [plane][pair][block_bytes]:
// GATHER: staging[z][y] = pool[z][pairs[y].src];  scatter: pool[z][pairs[y].dst] = staging[z][y].
template <bool GATHER>
__global__ void __launch_bounds__(kCopyThreads)
pa_stage_blocks_kernel(uint8_t* pool, int64_t plane_stride, int64_t block_stride, uint8_t* staging,
                       int64_t block_bytes, const int64_t* __restrict__ pairs) {
  const int64_t b = __ldg(pairs + 2 * blockIdx.y + (GATHER ? 0 : 1));
  uint8_t* in_pool = pool + blockIdx.z * plane_stride + b * block_stride;
  uint8_t* in_stage = staging + ((int64_t)blockIdx.z * gridDim.y + blockIdx.y) * block_bytes;
  const uint8_t* src = GATHER ? in_pool : in_stage;
  uint8_t* dst = GATHER ? in_stage : in_pool;
  const int64_t lo = (int64_t)blockIdx.x * kCopyChunk;
  uint4 v[kCopyUnroll];
#pragma unroll
  for (int u = 0; u < kCopyUnroll; ++u) {
    const int64_t o = lo + ((int64_t)u * kCopyThreads + threadIdx.x) * 16;
    if (o < block_bytes) v[u] = ld_stream(reinterpret_cast<const uint4*>(src + o));
  }
#pragma unroll
  for (int u = 0; u < kCopyUnroll; ++u) {
    const int64_t o = lo + ((int64_t)u * kCopyThreads + threadIdx.x) * 16;
    if (o < block_bytes) *reinterpret_cast<uint4*>(dst + o) = v[u];
  }
}

__global__ void pa_slot_mapping_kernel(const int32_t* __restrict__ bt, int32_t start_pos, int32_t n, int32_t bs,
                                       int64_t* __restrict__ slots) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int pos = start_pos + i;
  slots[i] = (int64_t)__ldg(bt + pos / bs) * bs + pos % bs;
}

// ---- persistent block tables -----------------------------------------
// The reference rebuilds every request's block table from Python lists each step
// (05_mini_vllm/mini_vllm.py:729-739 builds the step's tensors, :195-207 appends block ids).  Here the
// tables live on the device across steps ([rows, stride] int32, one row per live request); a step
// ships only the scheduler's delta - (row, column, block id) triples - and then gathers the rows of
// the requests it scheduled, in scheduling order, into the dense [batch, stride] table the attention
// kernels read.  The update count is read from DEVICE memory so both launches can sit in a CUDA graph.
__global__ void pa_table_update_kernel(int32_t* __restrict__ table, int32_t stride, int32_t rows,
                                       const int32_t* __restrict__ updates, const int32_t* __restrict__ num_updates,
                                       int32_t max_updates) {
  const int n = min(max(__ldg(num_updates), 0), max_updates);
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const int r = __ldg(updates + 3 * i), c = __ldg(updates + 3 * i + 1), v = __ldg(updates + 3 * i + 2);
    if ((unsigned)r < (unsigned)rows && (unsigned)c < (unsigned)stride) table[(int64_t)r * stride + c] = v;
  }
}

// grid.x = gathered row j: step_table[j, 0 : n) = table[row_of[j], 0 : n), n = ceil(ctx_lens[j] / BS)
// entries (all a kernel may dereference for that request); row_of[j] < 0 (padding) leaves the row alone.
__global__ void __launch_bounds__(128)
pa_table_gather_kernel(const int32_t* __restrict__ table, int32_t stride, int32_t rows,
                       const int32_t* __restrict__ row_of, const int32_t* __restrict__ ctx_lens, int32_t block_size,
                       int32_t* __restrict__ step_table, int32_t step_stride) {
  const int j = blockIdx.x;
  const int r = __ldg(row_of + j);
  if ((unsigned)r >= (unsigned)rows) return;
  const int len = max(__ldg(ctx_lens + j), 0);
  const int n = min(min((len + block_size - 1) / block_size, stride), step_stride);
  const int32_t* src = table + (int64_t)r * stride;
  int32_t* dst = step_table + (int64_t)j * step_stride;
  for (int c = threadIdx.x; c < n; c += blockDim.x) dst[c] = src[c];
}

bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

}  // namespace

extern "C" int pa_slot_mapping(const int32_t* block_table, int32_t start_pos, int32_t num_tokens,
                               int32_t block_size, int64_t* slots, void* stream) {
  if (num_tokens < 0 || start_pos < 0 || block_size <= 0) return PA_ERR_BAD_ARG;
  if (num_tokens == 0) return PA_OK;
  if (!block_table || !slots) return PA_ERR_BAD_ARG;
  pa_slot_mapping_kernel<<<(num_tokens + 255) / 256, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      block_table, start_pos, num_tokens, block_size, slots);
  return cudaGetLastError() == cudaSuccess ? PA_OK : PA_ERR_CUDA;
}

extern "C" int pa_block_table_step(int32_t* table, int32_t table_stride, int32_t num_rows,
                                   const int32_t* updates, const int32_t* num_updates, int32_t max_updates,
                                   const int32_t* row_of, const int32_t* ctx_lens, int32_t num_gather,
                                   int32_t block_size, int32_t* step_table, int32_t step_stride, void* stream) {
  if (table_stride <= 0 || num_rows <= 0 || max_updates < 0 || num_gather < 0 || block_size <= 0) return PA_ERR_BAD_ARG;
  if (!table) return PA_ERR_BAD_ARG;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (max_updates > 0) {
    if (!updates || !num_updates) return PA_ERR_BAD_ARG;
    int ctas = (max_updates + 255) / 256;
    if (ctas > 64) ctas = 64;
    pa_table_update_kernel<<<ctas, 256, 0, st>>>(table, table_stride, num_rows, updates, num_updates, max_updates);
  }
  if (num_gather > 0) {
    if (!row_of || !ctx_lens || !step_table || step_stride <= 0) return PA_ERR_BAD_ARG;
    pa_table_gather_kernel<<<num_gather, 128, 0, st>>>(table, table_stride, num_rows, row_of, ctx_lens, block_size,
                                                       step_table, step_stride);
  }
  return cudaGetLastError() == cudaSuccess ? PA_OK : PA_ERR_CUDA;
}

extern "C" int pa_reshape_and_cache(const void* key, const void* value, int64_t kv_stride_token,
                                    int64_t kv_stride_head, const PaKvCache* cache,
                                    const int64_t* slot_mapping, int32_t num_tokens, void* stream) {
  if (!cache || num_tokens < 0) return PA_ERR_BAD_ARG;
  if (num_tokens == 0) return PA_OK;
  if (!key || !value || !cache->k || !cache->v || !slot_mapping) return PA_ERR_BAD_ARG;
  if (cache->dtype != PA_BF16 && cache->dtype != PA_F32) return PA_ERR_UNSUPPORTED;
  if (cache->head_dim <= 0 || cache->num_kv_heads <= 0 || cache->block_size <= 0 || cache->num_blocks <= 0)
    return PA_ERR_BAD_ARG;
  const int es = cache->dtype == PA_F32 ? 4 : 2;
  RacParams p;
  p.key = static_cast<const uint8_t*>(key);
  p.value = static_cast<const uint8_t*>(value);
  p.k_cache = static_cast<uint8_t*>(cache->k);
  p.v_cache = static_cast<uint8_t*>(cache->v);
  p.slots = slot_mapping;
  p.src_stride_token = kv_stride_token * es;
  p.src_stride_head = kv_stride_head * es;
  p.stride_block = cache->stride_block * es;
  p.stride_head = cache->stride_head * es;
  p.stride_slot = cache->stride_slot * es;
  p.num_tokens = num_tokens;
  p.num_heads = cache->num_kv_heads;
  p.block_size = cache->block_size;
  p.num_slots = (int64_t)cache->num_blocks * cache->block_size;
  const int64_t row_bytes = (int64_t)cache->head_dim * es;
  const bool vec = aligned16(key) && aligned16(value) && aligned16(cache->k) && aligned16(cache->v) &&
                   row_bytes % 16 == 0 &&
                   ((p.src_stride_token | p.src_stride_head | p.stride_block | p.stride_head | p.stride_slot) & 15) == 0;
  p.row_units = (int32_t)(vec ? row_bytes / 16 : cache->head_dim);
  const int64_t total = (int64_t)num_tokens * p.num_heads * p.row_units;
#ifndef PA_RAC_UNROLL
#define PA_RAC_UNROLL 4     // 8 with streaming loads measured 22.4 us against 15.9 us at T = 8192 (fewer resident warps)
#endif
  constexpr int UNROLL = PA_RAC_UNROLL;      // 2 x UNROLL 16-byte loads in flight per thread
  int64_t ctas = (total + 256 * UNROLL - 1) / (256 * UNROLL);
  if (ctas > 148 * 16) ctas = 148 * 16;                    // grid-stride beyond 16 CTAs per SM
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (vec)
    pa_reshape_and_cache_kernel<uint4, UNROLL><<<(int)ctas, 256, 0, st>>>(p);
  else if (es == 4)
    pa_reshape_and_cache_kernel<uint32_t, UNROLL><<<(int)ctas, 256, 0, st>>>(p);
  else
    pa_reshape_and_cache_kernel<uint16_t, UNROLL><<<(int)ctas, 256, 0, st>>>(p);
  return cudaGetLastError() == cudaSuccess ? PA_OK : PA_ERR_CUDA;
}

extern "C" int pa_copy_blocks(void* base, int64_t plane_stride_bytes, int32_t num_planes,
                              int64_t block_stride_bytes, int64_t block_bytes, const int64_t* pairs,
                              int32_t num_pairs, void* stream) {
  if (num_pairs < 0 || num_planes < 0 || block_bytes < 0) return PA_ERR_BAD_ARG;
  if (num_pairs == 0 || num_planes == 0 || block_bytes == 0) return PA_OK;
  if (!base || !pairs) return PA_ERR_BAD_ARG;
  if (!aligned16(base) || ((plane_stride_bytes | block_stride_bytes | block_bytes) & 15)) return PA_ERR_MISALIGNED;
  if (num_pairs > 65535 || num_planes > 65535) return PA_ERR_UNSUPPORTED;
  dim3 grid((unsigned)((block_bytes + kCopyChunk - 1) / kCopyChunk), (unsigned)num_pairs, (unsigned)num_planes);
  pa_copy_blocks_kernel<<<grid, kCopyThreads, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<uint8_t*>(base), plane_stride_bytes, block_stride_bytes, block_bytes, pairs);
  return cudaGetLastError() == cudaSuccess ? PA_OK : PA_ERR_CUDA;
}

extern "C" int pa_swap_blocks(void* dst_base, int64_t dst_plane_stride_bytes, int64_t dst_block_stride_bytes,
                              const void* src_base, int64_t src_plane_stride_bytes, int64_t src_block_stride_bytes,
                              int32_t num_planes, int64_t block_bytes, const int64_t* pairs_host,
                              int32_t num_pairs, int32_t direction, void* stream) {
  if (num_pairs < 0 || num_planes < 0 || block_bytes < 0) return PA_ERR_BAD_ARG;
  if (num_pairs == 0 || num_planes == 0 || block_bytes == 0) return PA_OK;
  if (!dst_base || !src_base || !pairs_host) return PA_ERR_BAD_ARG;
  if (direction != 0 && direction != 1) return PA_ERR_BAD_ARG;
  const cudaMemcpyKind kind = direction == 0 ? cudaMemcpyDeviceToHost : cudaMemcpyHostToDevice;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  uint8_t* dst = static_cast<uint8_t*>(dst_base);
  const uint8_t* src = static_cast<const uint8_t*>(src_base);
  // Blocks that are consecutive on both sides (and dense: stride == bytes) travel as one run;
  // all planes of a run go in ONE strided 2-D copy when the pitches allow it.
  const bool dense = dst_block_stride_bytes == block_bytes && src_block_stride_bytes == block_bytes;
  const int64_t kMaxPitch = 0x7fffffff;   // cudaMemcpy2D pitch limit (cudaDevAttrMaxPitch)
  int i = 0;
  while (i < num_pairs) {
    int j = i + 1;
    if (dense)
      while (j < num_pairs && pairs_host[2 * j] == pairs_host[2 * (j - 1)] + 1 &&
             pairs_host[2 * j + 1] == pairs_host[2 * (j - 1) + 1] + 1)
        ++j;
    const int64_t width = (int64_t)(j - i) * block_bytes;
    const uint8_t* s0 = src + pairs_host[2 * i] * src_block_stride_bytes;
    uint8_t* d0 = dst + pairs_host[2 * i + 1] * dst_block_stride_bytes;
    cudaError_t e = cudaSuccess;
    // wide runs go as one 1-D copy per plane (a 2-row 2-D copy of 32 MiB rows measured 66 % of the
    // 1-D rate); narrow runs over many planes as ONE 2-D copy (fewer driver calls)
    if (num_planes > 1 && width < (256 << 10) && dst_plane_stride_bytes <= kMaxPitch && src_plane_stride_bytes <= kMaxPitch &&
        dst_plane_stride_bytes >= width && src_plane_stride_bytes >= width) {
      e = cudaMemcpy2DAsync(d0, (size_t)dst_plane_stride_bytes, s0, (size_t)src_plane_stride_bytes, (size_t)width,
                            (size_t)num_planes, kind, st);
    } else {
      for (int pl = 0; pl < num_planes && e == cudaSuccess; ++pl)
        e = cudaMemcpyAsync(d0 + pl * dst_plane_stride_bytes, s0 + pl * src_plane_stride_bytes, (size_t)width, kind, st);
    }
    if (e != cudaSuccess) return PA_ERR_CUDA;
    i = j;
  }
  return PA_OK;
}

extern "C" int pa_swap_blocks_staged(void* dev_base, int64_t dev_plane_stride_bytes, int64_t dev_block_stride_bytes,
                                     void* host_base, int64_t host_plane_stride_bytes, int64_t host_block_stride_bytes,
                                     int32_t num_planes, int64_t block_bytes, const int64_t* pairs_host,
                                     const int64_t* pairs_dev, int32_t num_pairs, int32_t direction,
                                     void* staging, int64_t staging_bytes, void* stream) {
  if (num_pairs < 0 || num_planes < 0 || block_bytes < 0) return PA_ERR_BAD_ARG;
  if (num_pairs == 0 || num_planes == 0 || block_bytes == 0) return PA_OK;
  if (!dev_base || !host_base || !pairs_host || !pairs_dev || !staging) return PA_ERR_BAD_ARG;
  if (direction != 0 && direction != 1) return PA_ERR_BAD_ARG;
  if (host_block_stride_bytes != block_bytes) return PA_ERR_UNSUPPORTED;      // host blocks must be dense
  if (staging_bytes < (int64_t)num_planes * num_pairs * block_bytes) return PA_ERR_BAD_ARG;
  if (!aligned16(dev_base) || !aligned16(staging) ||
      ((dev_plane_stride_bytes | dev_block_stride_bytes | block_bytes) & 15)) return PA_ERR_MISALIGNED;
  if (num_pairs > 65535 || num_planes > 65535) return PA_ERR_UNSUPPORTED;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  uint8_t* stg = static_cast<uint8_t*>(staging);
  uint8_t* host = static_cast<uint8_t*>(host_base);
  const dim3 grid((unsigned)((block_bytes + kCopyChunk - 1) / kCopyChunk), (unsigned)num_pairs, (unsigned)num_planes);
  const int hcol = direction == 0 ? 1 : 0;                 // which column of a pair is the host block
  if (direction == 0)
    pa_stage_blocks_kernel<true><<<grid, kCopyThreads, 0, st>>>(static_cast<uint8_t*>(dev_base), dev_plane_stride_bytes,
                                                                dev_block_stride_bytes, stg, block_bytes, pairs_dev);
  // one 1-D pinned copy per (plane, run of consecutive host blocks)
  int i = 0;
  while (i < num_pairs) {
    int j = i + 1;
    while (j < num_pairs && pairs_host[2 * j + hcol] == pairs_host[2 * (j - 1) + hcol] + 1) ++j;
    const size_t width = (size_t)(j - i) * block_bytes;
    for (int pl = 0; pl < num_planes; ++pl) {
      uint8_t* h = host + pl * host_plane_stride_bytes + pairs_host[2 * i + hcol] * block_bytes;
      uint8_t* s = stg + ((int64_t)pl * num_pairs + i) * block_bytes;
      const cudaError_t e = direction == 0 ? cudaMemcpyAsync(h, s, width, cudaMemcpyDeviceToHost, st)
                                           : cudaMemcpyAsync(s, h, width, cudaMemcpyHostToDevice, st);
      if (e != cudaSuccess) return PA_ERR_CUDA;
    }
    i = j;
  }
  if (direction == 1)
    pa_stage_blocks_kernel<false><<<grid, kCopyThreads, 0, st>>>(static_cast<uint8_t*>(dev_base), dev_plane_stride_bytes,
                                                                 dev_block_stride_bytes, stg, block_bytes, pairs_dev);
  return cudaGetLastError() == cudaSuccess ? PA_OK : PA_ERR_CUDA;
}
