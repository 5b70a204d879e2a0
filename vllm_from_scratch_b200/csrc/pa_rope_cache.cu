// Fused RoPE + QKV split + KV slot write: one pass from the QKV projection's output to rotated Q
// and to rotated K / plain V sitting in their cache slots.
//
// Replaces steps 1b-3 of the engine's attention (05_mini_vllm/mini_vllm.py:410-427):
//     q, k, v = qkv.chunk(3); cos, sin = rope(positions)                       (:410-417, :372-375)
//     q = q * cos + rotate_half(q) * sin ; k = k * cos + rotate_half(k) * sin   (:418-419, :355-357)
//     k_cache.view(-1, H, D)[slot_mapping] = k ; v_cache...[slot_mapping] = v   (:424-427)
// which in torch is 3 chunk views, 2 table gathers, 8 elementwise kernels and 2 index_puts, each
// a full pass over the activations.  Here every element of qkv is read once and written once.
//
// Bit-exactness: torch evaluates the formula op by op in the tensor dtype - a product is rounded
// before the sum.  The kernel does the same: no FMA contraction (__fmul_rn / __fadd_rn), and for
// bf16 every product and the sum are rounded to bf16 exactly where torch rounds them.  cos / sin
// come from the module's own cached tables (RotaryEmbedding.cos_cached / sin_cached, :366-370),
// indexed by position, so table construction stays the reference's.
//
// qkv row layout: [ Hq * D | Hkv * D | Hkv * D ] (GQA-ready; the engine has Hq == Hkv).
// HBM-bound: algorithmic bytes = T * (Hq + 2 Hkv) * D * esize read + the same written + tables.
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/pa_b200.h"

namespace {

struct RopeParams {
  const uint8_t* qkv;
  const uint8_t* cos_t;
  const uint8_t* sin_t;
  const int64_t* positions;
  const int64_t* slots;
  uint8_t* q_out;
  uint8_t* k_cache;
  uint8_t* v_cache;
  int64_t qkv_stride_t;                                // bytes
  int64_t table_stride;                                // bytes per position row
  int64_t q_stride_t, q_stride_h;                      // bytes
  int64_t stride_block, stride_head, stride_slot;      // bytes
  int64_t num_slots;
  int32_t num_tokens, hq, hkv, head_dim, block_size, max_pos;
};

template <typename T> struct Vec;
template <> struct Vec<float> { static constexpr int N = 4; };
template <> struct Vec<__nv_bfloat16> { static constexpr int N = 8; };

__device__ __forceinline__ float rnd(float x, float) { return x; }
__device__ __forceinline__ float rnd(float x, __nv_bfloat16) { return __bfloat162float(__float2bfloat16_rn(x)); }
__device__ __forceinline__ float ldf(const float* p) { return *p; }
__device__ __forceinline__ float ldf(const __nv_bfloat16* p) { return __bfloat162float(*p); }
__device__ __forceinline__ void stf(float* p, float x) { *p = x; }
__device__ __forceinline__ void stf(__nv_bfloat16* p, float x) { *p = __float2bfloat16_rn(x); }

// out_lo = x1 * cos_lo + (-x2) * sin_lo ; out_hi = x2 * cos_hi + x1 * sin_hi, rounded like torch
template <typename T>
__device__ __forceinline__ void rotate(float x1, float x2, float cl, float sl, float ch, float sh, float& lo, float& hi) {
  const T tag{};
  lo = __fadd_rn(rnd(__fmul_rn(x1, cl), tag), rnd(__fmul_rn(-x2, sl), tag));
  hi = __fadd_rn(rnd(__fmul_rn(x2, ch), tag), rnd(__fmul_rn(x1, sh), tag));
}

// One work item = W consecutive elements of the LOW half of one (token, head slot) row together
// with their partners in the high half.  W = 16 bytes worth when everything is 16-byte aligned,
// else 1.  Head slots [0, Hq) are Q, [Hq, Hq + Hkv) K, the rest V.
template <typename T, int W>
__global__ void __launch_bounds__(256)
pa_rope_and_cache_kernel(const RopeParams p) {
  const int half = p.head_dim / 2;
  const int items_per_row = half / W;
  const int slots_per_token = p.hq + 2 * p.hkv;
  const int64_t total = (int64_t)p.num_tokens * slots_per_token * items_per_row;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % items_per_row) * W;
    const int64_t r = i / items_per_row;
    const int hs = (int)(r % slots_per_token);
    const int64_t t = r / slots_per_token;
    const T* src = reinterpret_cast<const T*>(p.qkv + t * p.qkv_stride_t) + (int64_t)hs * p.head_dim;
    T* dst;
    if (hs < p.hq) {
      dst = reinterpret_cast<T*>(p.q_out + t * p.q_stride_t + (int64_t)hs * p.q_stride_h);
    } else {
      const int64_t slot = __ldg(p.slots + t);
      if (slot < 0 || slot >= p.num_slots) continue;             // padding token: no cache write
      const int64_t blk = slot / p.block_size, off = slot - blk * p.block_size;
      const bool is_k = hs < p.hq + p.hkv;
      const int h = is_k ? hs - p.hq : hs - p.hq - p.hkv;
      dst = reinterpret_cast<T*>((is_k ? p.k_cache : p.v_cache) + blk * p.stride_block + h * p.stride_head +
                                 off * p.stride_slot);
    }
    T lo[W], hi[W];
    if (W > 1) {
      *reinterpret_cast<uint4*>(lo) = *reinterpret_cast<const uint4*>(src + c);
      *reinterpret_cast<uint4*>(hi) = *reinterpret_cast<const uint4*>(src + half + c);
    } else {
      lo[0] = src[c];
      hi[0] = src[half + c];
    }
    if (hs < p.hq + p.hkv) {                                     // Q and K rotate, V passes through
      int64_t pos = __ldg(p.positions + t);
      pos = pos < 0 ? 0 : (pos >= p.max_pos ? p.max_pos - 1 : pos);
      const T* ct = reinterpret_cast<const T*>(p.cos_t + pos * p.table_stride);
      const T* st = reinterpret_cast<const T*>(p.sin_t + pos * p.table_stride);
#pragma unroll
      for (int j = 0; j < W; ++j) {
        float a, b;
        rotate<T>(ldf(lo + j), ldf(hi + j), ldf(ct + c + j), ldf(st + c + j), ldf(ct + half + c + j),
                  ldf(st + half + c + j), a, b);
        stf(lo + j, a);
        stf(hi + j, b);
      }
    }
    if (W > 1) {
      *reinterpret_cast<uint4*>(dst + c) = *reinterpret_cast<uint4*>(lo);
      *reinterpret_cast<uint4*>(dst + half + c) = *reinterpret_cast<uint4*>(hi);
    } else {
      dst[c] = lo[0];
      dst[half + c] = hi[0];
    }
  }
}

template <typename T>
int launch(const RopeParams& p, bool vec, cudaStream_t st) {
  const int w = vec ? Vec<T>::N : 1;
  const int64_t total = (int64_t)p.num_tokens * (p.hq + 2 * p.hkv) * (p.head_dim / 2 / w);
  int64_t blocks = (total + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  if (blocks < 1) blocks = 1;
  if (vec) pa_rope_and_cache_kernel<T, Vec<T>::N><<<(int)blocks, 256, 0, st>>>(p);
  else pa_rope_and_cache_kernel<T, 1><<<(int)blocks, 256, 0, st>>>(p);
  return cudaGetLastError() == cudaSuccess ? PA_OK : PA_ERR_CUDA;
}

}  // namespace

extern "C" int pa_rope_and_cache(const void* qkv, int64_t qkv_stride_t, int32_t num_q_heads, const void* cos_table,
                                 const void* sin_table, int64_t table_stride, int32_t max_positions,
                                 const int64_t* positions, const int64_t* slot_mapping, int32_t num_tokens,
                                 void* q_out, int64_t q_stride_t, int64_t q_stride_h, const PaKvCache* cache,
                                 void* stream) {
  if (num_tokens < 0 || !cache) return PA_ERR_BAD_ARG;
  if (num_tokens == 0) return PA_OK;
  if (!qkv || !cos_table || !sin_table || !positions || !slot_mapping || !q_out || !cache->k || !cache->v)
    return PA_ERR_BAD_ARG;
  if (cache->dtype != PA_BF16 && cache->dtype != PA_F32) return PA_ERR_UNSUPPORTED;
  const int D = cache->head_dim;
  if (D <= 0 || (D & 1) || num_q_heads <= 0 || cache->num_kv_heads <= 0 || cache->block_size <= 0 || max_positions <= 0)
    return PA_ERR_BAD_ARG;
  const int es = cache->dtype == PA_F32 ? 4 : 2;
  RopeParams p;
  p.qkv = static_cast<const uint8_t*>(qkv);
  p.cos_t = static_cast<const uint8_t*>(cos_table);
  p.sin_t = static_cast<const uint8_t*>(sin_table);
  p.positions = positions; p.slots = slot_mapping;
  p.q_out = static_cast<uint8_t*>(q_out);
  p.k_cache = static_cast<uint8_t*>(cache->k); p.v_cache = static_cast<uint8_t*>(cache->v);
  p.qkv_stride_t = qkv_stride_t * es; p.table_stride = table_stride * es;
  p.q_stride_t = q_stride_t * es; p.q_stride_h = q_stride_h * es;
  p.stride_block = cache->stride_block * es; p.stride_head = cache->stride_head * es; p.stride_slot = cache->stride_slot * es;
  p.num_slots = (int64_t)cache->num_blocks * cache->block_size;
  p.num_tokens = num_tokens; p.hq = num_q_heads; p.hkv = cache->num_kv_heads; p.head_dim = D;
  p.block_size = cache->block_size; p.max_pos = max_positions;
  const int vecn = 16 / es;
  const uintptr_t ptrs = reinterpret_cast<uintptr_t>(qkv) | reinterpret_cast<uintptr_t>(q_out) |
                         reinterpret_cast<uintptr_t>(cache->k) | reinterpret_cast<uintptr_t>(cache->v);
  const int64_t strides = p.qkv_stride_t | p.q_stride_t | p.q_stride_h | p.stride_block | p.stride_head | p.stride_slot;
  const bool vec = ((D / 2) % vecn == 0) && !(ptrs & 15) && !(strides & 15) && !(((int64_t)D * es) & 15);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  return cache->dtype == PA_F32 ? launch<float>(p, vec, st) : launch<__nv_bfloat16>(p, vec, st);
}
