// Paged-attention decode, generic CUDA-core path: fp32 caches (the reference's default,
// 01_paged_attention/paged_attention.py:19, parity <= 1e-5) and every bf16 shape the TMA
// path does not take (block_size != 16, head_dim not 64/128, odd strides).
//
// Same work decomposition and partial format as the fast path (pa_common.cuh): a worker is
// a warp that streams a contiguous range of KV blocks.  Head segments are (KV head, tile of
// up to 8 query heads of its group).  Per chunk of 32 tokens:
//   QK^T : lane <-> token.  The lane walks its K row with 16-byte loads and dots it with the
//          query tile held in shared memory (broadcast reads) - fp32 FMA, no tensor cores,
//          so an fp32 cache is read and multiplied at full precision.
//   softmax: online, exp2 domain, max over the chunk by warp shuffles.
//   PV   : lane <-> head_dim column.  For each token the weight and the row offset are
//          broadcast by shuffle and the V row is read coalesced.
#include "pa_common.cuh"

namespace {

constexpr int kTQ = 8;          // query heads per head segment
constexpr int kWarpsG = 4;      // workers per CTA
constexpr int kCtasPerSm = 4;

template <typename T> struct VecOf;
template <> struct VecOf<float> { static constexpr int N = 4; };
template <> struct VecOf<__nv_bfloat16> { static constexpr int N = 8; };

template <typename T>
__device__ __forceinline__ void load_row_vec(const T* ptr, float (&f)[VecOf<T>::N]);
template <>
__device__ __forceinline__ void load_row_vec<float>(const float* ptr, float (&f)[4]) {
  const float4 v = __ldg(reinterpret_cast<const float4*>(ptr));
  f[0] = v.x; f[1] = v.y; f[2] = v.z; f[3] = v.w;
}
template <>
__device__ __forceinline__ void load_row_vec<__nv_bfloat16>(const __nv_bfloat16* ptr, float (&f)[8]) {
  const uint4 v = __ldg(reinterpret_cast<const uint4*>(ptr));
  const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    f[2 * i] = __uint_as_float(w[i] << 16);
    f[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
  }
}

// T: cache / q / out element type.  DPL: head_dim columns per lane (D <= 32 * DPL).
// VEC: 16-byte K-row loads are legal (alignment + D % N == 0).
template <typename T, int DPL, bool VEC>
__global__ void __launch_bounds__(kWarpsG * 32)
pa_decode_generic_kernel(const PaDecodeParams p) {
  extern __shared__ uint8_t smem_raw[];
  // q tiles: [kWarpsG][kTQ][D] float | scratch[40] | cum[B + 1]
  const int D = p.head_dim;
  float* s_q = reinterpret_cast<float*>(smem_raw);
  int* scratch = reinterpret_cast<int*>(s_q + kWarpsG * kTQ * D);
  int* s_cum = scratch + 40;

  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int BS = p.block_size;
  pa_build_cum(p.seq_lens, p.batch, BS, s_cum, scratch);

  const int hs = p.num_kv_heads * p.q_tiles;
  const int total = s_cum[p.batch] * hs;
  const int per = max(pa_ceil_div(total, p.num_workers), PA_MIN_BLOCKS_PER_WORKER);
  if (blockIdx.x == 0) {
    for (int i = threadIdx.x; i <= p.batch; i += blockDim.x) p.plan[2 + i] = s_cum[i];
    if (threadIdx.x == 0) { p.plan[0] = per; p.plan[1] = total; }
  }
  const int worker = warp * gridDim.x + blockIdx.x;
  const int lin0 = worker * per;
  const int count = min(per, total - lin0);
  if (count <= 0) return;

  const T* kbase = static_cast<const T*>(p.k);
  const T* vbase = static_cast<const T*>(p.v);
  const T* qbase = static_cast<const T*>(p.q);
  T* obase = static_cast<T*>(p.out);
  float* my_q = s_q + warp * kTQ * D;
  const int g = p.group;

  PaCursor cc = pa_locate(s_cum, p.batch, hs, lin0);
  int done = 0;
  while (done < count) {
    // ---- one segment piece: blocks [cc.b, b_end) of (cc.s, cc.h) ------------------------
    const int nb_here = min(cc.n - cc.b, count - done);
    const int b_end = cc.b + nb_here;
    const int kvh = cc.h / p.q_tiles, qt = cc.h - kvh * p.q_tiles;
    const int gq = min(kTQ, g - qt * kTQ);                 // live query heads in this tile
    const int hq0 = kvh * g + qt * kTQ;
    const int seq_len = __ldg(p.seq_lens + cc.s);
    const int tok_end = min(b_end * BS, seq_len);
    const int32_t* bt = p.block_tables + (int64_t)cc.s * p.bt_stride;

    __syncwarp();
    for (int i = lane; i < kTQ * D; i += 32) {
      const int j = i / D, d = i - j * D;
      my_q[i] = j < gq ? pa_to_float(qbase[(int64_t)cc.s * p.q_stride_b + (int64_t)(hq0 + j) * p.q_stride_h + d]) : 0.f;
    }
    __syncwarp();

    float m_run[kTQ], l_run[kTQ], acc[kTQ][DPL];
#pragma unroll
    for (int j = 0; j < kTQ; ++j) {
      m_run[j] = -INFINITY;
      l_run[j] = 0.f;
#pragma unroll
      for (int i = 0; i < DPL; ++i) acc[j][i] = 0.f;
    }

    for (int t0 = cc.b * BS; t0 < tok_end; t0 += 32) {
      const int t = t0 + lane;
      const bool valid = t < tok_end;
      int64_t off = 0;
      if (valid) {
        const int blk = t / BS;
        off = (int64_t)__ldg(bt + blk) * p.kv_stride_block + (int64_t)kvh * p.kv_stride_head +
              (int64_t)(t - blk * BS) * p.kv_stride_slot;
      }
      // ---- scores: lane's token against the query tile -------------------------------
      float sc[kTQ];
#pragma unroll
      for (int j = 0; j < kTQ; ++j) sc[j] = 0.f;
      if (valid) {
        const T* kr = kbase + off;
        if (VEC) {
          constexpr int N = VecOf<T>::N;
          for (int d = 0; d < D; d += N) {
            float kf[N];
            load_row_vec<T>(kr + d, kf);
#pragma unroll
            for (int j = 0; j < kTQ; ++j) {
              if (j < gq) {
#pragma unroll
                for (int e = 0; e < N; e += 4) {
                  const float4 qv = *reinterpret_cast<const float4*>(my_q + j * D + d + e);
                  sc[j] = fmaf(kf[e], qv.x, sc[j]);
                  sc[j] = fmaf(kf[e + 1], qv.y, sc[j]);
                  sc[j] = fmaf(kf[e + 2], qv.z, sc[j]);
                  sc[j] = fmaf(kf[e + 3], qv.w, sc[j]);
                }
              }
            }
          }
        } else {
          for (int d = 0; d < D; ++d) {
            const float kf = pa_to_float(kr[d]);
#pragma unroll
            for (int j = 0; j < kTQ; ++j)
              if (j < gq) sc[j] = fmaf(kf, my_q[j * D + d], sc[j]);
          }
        }
      }
      // ---- online softmax ------------------------------------------------------------------
      float pw[kTQ];
#pragma unroll
      for (int j = 0; j < kTQ; ++j) {
        if (j < gq) {
          const float s = valid ? sc[j] * p.scale_log2e : -INFINITY;
          float mx = s;
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
          const float m_new = fmaxf(m_run[j], mx);       // finite: lane 0's token is valid
          const float alpha = exp2f(m_run[j] - m_new);
          m_run[j] = m_new;
          pw[j] = exp2f(s - m_new);
          l_run[j] = l_run[j] * alpha + pw[j];            // per-lane partial sum
#pragma unroll
          for (int i = 0; i < DPL; ++i) acc[j][i] *= alpha;
        } else {
          pw[j] = 0.f;
        }
      }
      // ---- PV: lane <-> column ---------------------------------------------------------------
      const int ntok = min(32, tok_end - t0);
      const int off_lo = (int)(off & 0xffffffff), off_hi = (int)(off >> 32);
      for (int tt = 0; tt < ntok; ++tt) {
        const int64_t o = ((int64_t)__shfl_sync(0xffffffffu, off_hi, tt) << 32) |
                          (uint32_t)__shfl_sync(0xffffffffu, off_lo, tt);
        const T* vr = vbase + o;
        float vv[DPL];
#pragma unroll
        for (int i = 0; i < DPL; ++i) {
          const int d = lane + 32 * i;
          vv[i] = d < D ? pa_to_float(vr[d]) : 0.f;
        }
#pragma unroll
        for (int j = 0; j < kTQ; ++j) {
          if (j < gq) {
            const float pj = __shfl_sync(0xffffffffu, pw[j], tt);
#pragma unroll
            for (int i = 0; i < DPL; ++i) acc[j][i] = fmaf(pj, vv[i], acc[j][i]);
          }
        }
      }
    }

    // ---- emit: final result if the whole segment is ours, else a partial ------------------
    const int seg_lin0 = lin0 + done - cc.b;
    const bool whole = (seg_lin0 / per) == ((seg_lin0 + cc.n - 1) / per);
    const int slot = cc.s * hs + cc.h + worker;
#pragma unroll
    for (int j = 0; j < kTQ; ++j) {
      if (j < gq) {
        float l = l_run[j];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) l += __shfl_xor_sync(0xffffffffu, l, o);
        if (whole) {
          const float inv = 1.f / l;
          T* op = obase + (int64_t)cc.s * p.o_stride_b + (int64_t)(hq0 + j) * p.o_stride_h;
#pragma unroll
          for (int i = 0; i < DPL; ++i) {
            const int d = lane + 32 * i;
            if (d < D) pa_store(op + d, acc[j][i] * inv);
          }
        } else {
          float* po = p.part_o + ((int64_t)slot * kTQ + j) * D;
#pragma unroll
          for (int i = 0; i < DPL; ++i) {
            const int d = lane + 32 * i;
            if (d < D) po[d] = acc[j][i];
          }
          if (lane == 0) p.part_ml[(int64_t)slot * kTQ + j] = make_float2(m_run[j], l);
        }
      }
    }

    done += nb_here;
    // jump the cursor to the end of this piece, then step to the next segment
    cc.b = b_end - 1;
    pa_advance(cc, s_cum, p.batch, hs);
  }
}

template <typename T, int DPL>
int launch_t(const PaDecodeParams& p, bool vec, int ctas, cudaStream_t stream) {
  const size_t smem = (size_t)kWarpsG * kTQ * p.head_dim * 4 + 40 * 4 + (size_t)(p.batch + 1) * 4;
  if (smem > 200 * 1024) return -2;
  auto kv = pa_decode_generic_kernel<T, DPL, true>;
  auto ks = pa_decode_generic_kernel<T, DPL, false>;
  auto kern = vec ? kv : ks;
  if (smem > 48 * 1024 &&
      cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess)
    return -5;
  kern<<<ctas, kWarpsG * 32, smem, stream>>>(p);
  return cudaGetLastError() == cudaSuccess ? 0 : -5;
}

template <typename T>
int launch_d(const PaDecodeParams& p, bool vec, int ctas, cudaStream_t stream) {
  const int D = p.head_dim;
  if (D <= 32) return launch_t<T, 1>(p, vec, ctas, stream);
  if (D <= 64) return launch_t<T, 2>(p, vec, ctas, stream);
  if (D <= 128) return launch_t<T, 4>(p, vec, ctas, stream);
  if (D <= 256) return launch_t<T, 8>(p, vec, ctas, stream);
  return -2;
}

}  // namespace

int pa_decode_generic_workers(int num_sms, int* ctas) {
  *ctas = num_sms * kCtasPerSm;
  return num_sms * kCtasPerSm * kWarpsG;
}

int pa_launch_decode_generic(const PaDecodeParams& p, int dtype, int ctas, cudaStream_t stream) {
  const int esize = dtype == 1 ? 4 : 2;
  const int n = 16 / esize;
  const bool vec = (p.head_dim % n == 0) && (p.head_dim % 4 == 0) &&
                   ((reinterpret_cast<uintptr_t>(p.k) & 15) == 0) &&
                   (((p.kv_stride_block | p.kv_stride_head | p.kv_stride_slot) % n) == 0);
  if (dtype == 1) return launch_d<float>(p, vec, ctas, stream);
  return launch_d<__nv_bfloat16>(p, vec, ctas, stream);
}
