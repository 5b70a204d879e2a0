// Cross-partition reduce of the split-K decode (second launch of pa_decode).
//
// The decode kernels (pa_decode_tma.cu / pa_decode_generic.cu) cut the flattened
// (sequence, head segment, KV block) space into equal worker ranges; a segment that
// straddles a range boundary leaves one partial per worker it touches:
//     part_o [slot][row][D]  unnormalised sum_t exp2(s_t - m) * V_t     (fp32)
//     part_ml[slot][row]     (m = running max in the log2 domain, l = sum_t exp2(s_t - m))
// with slot = segment_id + worker_id.  This kernel recomputes which workers touched each
// segment from the plan the decode kernel left in the workspace (per-worker range length
// and the exclusive scan of blocks per sequence), merges their partials with the usual
// log-sum-exp weights, normalises and writes out[s, head, :] in the cache dtype.
// Segments that lie inside one worker were already written by the decode kernel and are
// skipped; sequences with seq_len == 0 get zeros (defined extension, SURVEY.md 8b - the
// reference raises in torch.stack([]), 01_paged_attention/paged_attention.py:93).
//
// One warp per (segment, row); lanes stride over head_dim.  The work is tiny next to the
// KV stream (partials are ~1 % of the KV bytes), so the only goal here is low latency.
#include "pa_common.cuh"

namespace {

constexpr int kWarpsR = 4;

template <typename T>
__global__ void __launch_bounds__(kWarpsR * 32)
pa_reduce_kernel(const PaDecodeParams p) {
  const int lane = threadIdx.x & 31;
  const int hs = p.num_kv_heads * p.q_tiles;
  const int unit = blockIdx.x * kWarpsR + (threadIdx.x >> 5);     // (segment, row)
  const int seg = unit / p.tile_q, row = unit - seg * p.tile_q;
  if (seg >= p.batch * hs) return;
  const int s = seg / hs, h = seg - s * hs;
  const int kvh = h / p.q_tiles, qt = h - kvh * p.q_tiles;
  if (row >= min(p.tile_q, p.group - qt * p.tile_q)) return;      // padding row of the tile
  const int head = kvh * p.group + qt * p.tile_q + row;
  const int D = p.head_dim;
  T* op = static_cast<T*>(p.out) + (int64_t)s * p.o_stride_b + (int64_t)head * p.o_stride_h;

  const int per = p.plan[0];
  const int c0 = p.plan[2 + s], n = p.plan[3 + s] - c0;
  if (n == 0) {
    for (int d = lane; d < D; d += 32) pa_store(op + d, 0.f);
    return;
  }
  const int lin0 = c0 * hs + h * n;
  const int w0 = lin0 / per, w1 = (lin0 + n - 1) / per;
  if (w0 == w1) return;                                           // written by the decode kernel

  float m = -INFINITY;
  for (int w = w0; w <= w1; ++w) m = fmaxf(m, p.part_ml[(int64_t)(seg + w) * p.tile_q + row].x);
  float l = 0.f;
  for (int w = w0; w <= w1; ++w) {
    const float2 ml = p.part_ml[(int64_t)(seg + w) * p.tile_q + row];
    l += ml.y * exp2f(ml.x - m);
  }
  const float inv = 1.f / l;
  for (int d = lane; d < D; d += 32) {
    float o = 0.f;
    for (int w = w0; w <= w1; ++w) {
      const int64_t r = (int64_t)(seg + w) * p.tile_q + row;
      o = fmaf(p.part_o[r * D + d], exp2f(p.part_ml[r].x - m), o);
    }
    pa_store(op + d, o * inv);
  }
}

}  // namespace

int pa_launch_reduce(const PaDecodeParams& p, int dtype, cudaStream_t stream) {
  const int64_t units = (int64_t)p.batch * p.num_kv_heads * p.q_tiles * p.tile_q;
  if (units == 0) return 0;
  const int ctas = (int)((units + kWarpsR - 1) / kWarpsR);
  if (dtype == 1)
    pa_reduce_kernel<float><<<ctas, kWarpsR * 32, 0, stream>>>(p);
  else
    pa_reduce_kernel<__nv_bfloat16><<<ctas, kWarpsR * 32, 0, stream>>>(p);
  return cudaGetLastError() == cudaSuccess ? 0 : -5;
}
