// Shared device-side definitions for the paged-attention kernels (sm_100a).
//
// Work decomposition of the GENERIC path ("stream-K over KV blocks"; the bf16 fast path,
// pa_decode_tma.cu, uses dynamic work units instead and documents its own):
//   The decode batch is flattened into one linear space of KV blocks,
//       L = cum[s] * HS + hseg * nblk(s) + b
//   (s = sequence, hseg = head segment of that sequence, b = logical block,
//    cum[s] = sum of nblk over earlier sequences, HS = head segments per
//    sequence).  The space is cut into `num_workers` equal contiguous ranges of
//   Q blocks; one persistent WARP ("worker") streams each range.  A range may
//   start / end inside a (sequence, head) segment, so a worker emits one partial
//   (unnormalised O, running max m, running sum l) per segment it touches, into
//   partial slot `segment_id + worker_id` (strictly increasing along the linear
//   order, hence unique; at most num_segments + num_workers slots).  Segments
//   that lie inside a single worker are normalised and written directly.  The
//   reduce kernel merges the partials of split segments.
//   This keeps every worker's byte count equal for ragged batches, which is what
//   an HBM-bound kernel needs to finish all SMs at the same time.
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#define PA_MIN_BLOCKS_PER_WORKER 4
#define PA_MAX_BATCH_SMEM 8191   // cum[] lives in shared memory: (B + 1) * 4 bytes

struct PaDecodeParams {
  const void* q;
  void* out;
  const void* k;
  const void* v;
  const int32_t* block_tables;
  const int32_t* seq_lens;
  int64_t q_stride_b, q_stride_h;
  int64_t o_stride_b, o_stride_h;
  int64_t kv_stride_block, kv_stride_head, kv_stride_slot;
  int32_t bt_stride;
  int32_t batch, num_q_heads, num_kv_heads, head_dim, block_size;
  int32_t num_blocks;     // NB: extent of the block axis (TMA zero-fills block ids outside [0, NB))
  int32_t group;          // g = Hq / Hkv
  int32_t tile_q;         // q heads per head segment (partial rows per slot)
  int32_t q_tiles;        // ceil(g / tile_q); HS = Hkv * q_tiles
  int32_t num_workers;
  int32_t pos_head;       // TMA path: which of tensor-map dims 3 / 4 is the head axis (the other is the block)
  int32_t num_stages;     // TMA path: ring depth (K + V tile pairs) per CTA
  int32_t target_units;   // TMA path: split-K work units aimed at per launch (units_per_sm * SMs)
  int32_t guided_k;       // TMA path: guided self-scheduling divisor (unit ~ remaining / (guided_k * CTAs))
  int32_t claim_limit;    // TMA path: published-but-unconsumed blocks per CTA above which no new unit is claimed
  float scale_log2e;      // (1 / sqrt(D)) * log2(e)
  // workspace
  float* part_o;          // [slots][tile_q][D]
  float* part_o2;         // TMA path: second partial area of the same size (level-2 partials of two-level folds)
  float2* part_ml;        // [slots][tile_q]  (m in log2 domain, l)
  int32_t* plan;          // [0] = Q, [1] = total linear blocks, [2 .. 2 + B] = cum[0..B]  (generic path)
  int32_t* counters;      // zero-initialised counter header, left zero by every launch (TMA path)
  long long* dbg;         // optional per-CTA timestamps (pa_decode_debug_times), else null
  // fused output gather over NVLink peer memory (pa_decode_gather; TMA path only): every final
  // output row is also stored into this rank's slot of each remote rank's gather buffer, and each
  // completed (sequence, KV head) segment is counted into this rank's arrival counter on every rank
  int32_t num_remote;     // 0 = plain pa_decode
  void* remote_out[7];
  uint32_t* remote_sig[7];
  uint32_t* local_sig;    // null = no signalling
  // LL ("flag in data") form of the same gather: when ll_epoch is set, remote_out[r] are 8-byte-word
  // buffers - every 4-byte output word travels as {data, step number} in ONE 8-byte store - and nothing
  // is fenced or counted on the sender; the receiver's unpack kernel polls the step numbers
  const uint32_t* ll_epoch;   // device word holding the number of completed steps; this step's flag = *ll_epoch + 1
};

struct PaCursor {
  int s;      // sequence
  int h;      // head segment within the sequence
  int b;      // logical block within the sequence
  int n;      // nblk(s)
};

__host__ __device__ __forceinline__ int pa_ceil_div(int a, int b) { return (a + b - 1) / b; }

// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) is a PER-DEVICE attribute: one process may drive
// several GPUs (paged_attention._DeviceOf), so "set once" is tracked per device ordinal.
#include <atomic>
struct PaPerDeviceFlag {
  std::atomic<bool> set[64];
  // returns the device ordinal if the attribute still has to be set on the current device, -2 if it
  // is already set, -1 if the device cannot be determined (then: set it every time)
  int need() {
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess) { cudaGetLastError(); return -1; }
    if (dev < 0 || dev >= 64) return -1;
    return set[dev].load(std::memory_order_acquire) ? -2 : dev;
  }
  void done(int dev) { if (dev >= 0 && dev < 64) set[dev].store(true, std::memory_order_release); }
};

// CTA-wide exclusive scan of nblk(s) into s_cum[0..B]; every thread of the CTA must call.
// scratch: >= 33 ints of shared memory.
__device__ __forceinline__ void pa_build_cum(const int32_t* __restrict__ seq_lens, int batch,
                                             int block_size, int* s_cum, int* scratch) {
  const int tid = threadIdx.x, nthr = blockDim.x;
  const int lane = tid & 31, warp = tid >> 5, nwarps = (nthr + 31) >> 5;
  const int per = pa_ceil_div(batch, nthr);
  const int lo = min(tid * per, batch), hi = min(lo + per, batch);
  int sum = 0;
  for (int i = lo; i < hi; ++i) sum += pa_ceil_div(max(__ldg(seq_lens + i), 0), block_size);
  int incl = sum;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    int t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += t;
  }
  if (lane == 31) scratch[warp] = incl;
  __syncthreads();
  if (warp == 0) {
    int w = lane < nwarps ? scratch[lane] : 0;
    int wi = w;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      int t = __shfl_up_sync(0xffffffffu, wi, o);
      if (lane >= o) wi += t;
    }
    scratch[lane] = wi - w;    // exclusive prefix of warp sums
    if (lane == 31) scratch[32] = wi;
  }
  __syncthreads();
  int run = scratch[warp] + incl - sum;
  for (int i = lo; i < hi; ++i) {
    s_cum[i] = run;
    run += pa_ceil_div(max(__ldg(seq_lens + i), 0), block_size);
  }
  if (tid == 0) s_cum[batch] = scratch[32];
  __syncthreads();
}

// Map a linear block index to (s, h, b).  s_cum is in shared memory.
__device__ __forceinline__ PaCursor pa_locate(const int* s_cum, int batch, int hs, int lin) {
  const int key = lin / hs;
  int lo = 0, hi = batch;          // find largest s with s_cum[s] <= key and s_cum[s + 1] > key
  while (hi - lo > 1) {
    int mid = (lo + hi) >> 1;
    if (s_cum[mid] <= key) lo = mid; else hi = mid;
  }
  // lo is the LAST index with s_cum[lo] <= key, so s_cum[lo + 1] > key: the sequence is non-empty.
  PaCursor c;
  c.s = lo;
  c.n = s_cum[lo + 1] - s_cum[lo];
  const int r = lin - s_cum[lo] * hs;
  c.h = r / c.n;
  c.b = r - c.h * c.n;
  return c;
}

__device__ __forceinline__ void pa_advance(PaCursor& c, const int* s_cum, int batch, int hs) {
  if (++c.b < c.n) return;
  c.b = 0;
  if (++c.h < hs) return;
  c.h = 0;
  do {
    ++c.s;
    if (c.s >= batch) { c.n = 1; return; }     // past the end; never dereferenced
    c.n = s_cum[c.s + 1] - s_cum[c.s];
  } while (c.n == 0);
}

__device__ __forceinline__ float pa_to_float(float x) { return x; }
__device__ __forceinline__ float pa_to_float(__nv_bfloat16 x) { return __bfloat162float(x); }
__device__ __forceinline__ void pa_store(float* p, float x) { *p = x; }
__device__ __forceinline__ void pa_store(__nv_bfloat16* p, float x) { *p = __float2bfloat16_rn(x); }

// host-side launchers implemented by the individual translation units
struct PaKvCache;
int pa_launch_decode_tma(const PaDecodeParams& p, int num_sms, cudaStream_t stream);
int pa_launch_decode_generic(const PaDecodeParams& p, int dtype, int ctas, cudaStream_t stream);
int pa_launch_reduce(const PaDecodeParams& p, int dtype, cudaStream_t stream);
bool pa_decode_tma_supported(const PaDecodeParams& p, int dtype);
int pa_decode_tma_header_ints();
int pa_decode_tma_set_pdl(int on);
int64_t pa_decode_tma_max_units(int batch, int num_kv_heads, int target_units);
int pa_decode_generic_workers(int num_sms, int* ctas);
int pa_decode_stream_workers(const PaDecodeParams& p, int dtype, int num_sms, int* ctas);
bool pa_decode_stream_supported(const PaDecodeParams& p, int dtype);
int pa_launch_decode_stream(const PaDecodeParams& p, int dtype, int ctas, cudaStream_t stream);

// ---- tensor-core (tcgen05) paged prefill, pa_prefill_tc.cu --------------------------------
// A varlen batch of query chunks: sequence s owns query tokens [cu_q_lens[s], cu_q_lens[s + 1])
// at positions [ctx_lens[s] - q_len, ctx_lens[s]) and attends causally to positions [0, pos]
// read through its block table (the chunk's own K / V already written by reshape-and-cache).
// cu_q_lens == nullptr: ONE sequence of num_tokens query tokens whose context is ctx_len_single.
struct PaPrefillParams {
  const void* q;
  void* out;
  const void* k;
  const void* v;
  const int32_t* block_tables;
  const int32_t* cu_q_lens;
  const int32_t* ctx_lens;
  int64_t q_stride_t, q_stride_h;
  int64_t o_stride_t, o_stride_h;
  int64_t kv_stride_block, kv_stride_head, kv_stride_slot;
  int32_t bt_stride;
  int32_t batch, num_tokens, ctx_len_single, max_q_len;
  int32_t num_q_heads, num_kv_heads, head_dim, group;
  int32_t num_blocks;     // NB: extent of the block axis of the tensor maps
  int32_t max_blocks;     // block-table entries readable per sequence (0 = unknown: trust ctx_lens)
  int32_t pos_head;       // which of tensor-map dims 3 / 4 is the head axis
  float scale_log2e;
  int32_t debug;          // bring-up knob, 0 = normal
  long long* trace;       // PA_PREFILL_TRACE=1: per-tile clock64 stamps of CTA (0,0,0), else null
  long long* cta_times;   // pa_prefill_debug_times: 8 words per CTA (smid, globaltimer stamps, KV tiles), else null
};
bool pa_prefill_tc_supported(const PaPrefillParams& p, int dtype, int block_size);
int pa_launch_prefill_tc(const PaPrefillParams& p, cudaStream_t stream);    // one 128-row tile per CTA
int pa_launch_prefill_tc2(const PaPrefillParams& p, cudaStream_t stream);   // two tiles per CTA, ping-pong
