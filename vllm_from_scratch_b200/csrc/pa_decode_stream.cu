// Paged-attention decode, generic STREAMING path: fp32 caches (the reference's default dtype,
// 01_paged_attention/paged_attention.py:19, parity <= 1e-5) and every bf16 shape the TMA path does not
// take, as long as K / V rows are 16-byte aligned (anything else stays on pa_decode_generic.cu).
//
// Round 1's generic kernel read K rows straight from global memory, one lane per token: a warp kept only
// a handful of 512-byte row loads in flight and streamed an fp32 C2 cache at 1.8 TB/s (0.27 of the copy
// peak).  This kernel keeps the same work decomposition and partial format (pa_common.cuh: a worker is a
// warp that streams a contiguous range of the flattened (sequence, head segment, KV block) space, leaves one
// partial per segment piece, pa_reduce.cu merges) but every warp stages its K / V rows through a private
// cp.async ring in shared memory - 8 warps x R stages x <= 8 KiB, ~190 KiB in flight per SM, the same depth
// the TMA kernel runs at - and does the math out of shared memory:
//   QK^T : 8 lanes per token, 4 tokens at a time.  A lane dots the 16-byte chunks q8, q8 + 8, ... of its
//          token's K row with the query tile (fp32 in shared memory, broadcast reads) for up to 8 query
//          heads; three xor-shuffles finish the dot.  Every LDS.128 phase is one token's 128 contiguous
//          bytes: conflict-free without padding.
//   softmax: online, exp2 domain, once per stage (<= 8 tokens); the weights go through a 256-byte
//          shared-memory tile so the PV phase reads them as broadcasts.
//   PV   : lane <-> 16-byte column chunk of the V row, fp32 FMA.  m, l are lane-uniform.
// fp32 FMA throughout (no tensor cores): an fp32 cache is read and multiplied at full precision.
#include <type_traits>

#include "pa_common.cuh"

namespace {

constexpr int kTQ = 8;              // query heads per head segment (same tile as pa_decode_generic.cu / pa_reduce.cu)
constexpr int kWarpsS = 8;          // workers per CTA (16 where the registers allow: GQ = 4, one chunk per lane); one persistent CTA per SM
constexpr int kMaxTS = 8;           // tokens per stage, at most
constexpr int kSmemBudgetS = 227 * 1024;

template <typename T> struct Chunk;
template <> struct Chunk<float> {
  static constexpr int N = 4;
  static __device__ __forceinline__ void load(uint32_t addr, float (&f)[4]) {
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(f[0]), "=f"(f[1]), "=f"(f[2]), "=f"(f[3]) : "r"(addr));
  }
};
template <> struct Chunk<__nv_bfloat16> {
  static constexpr int N = 8;
  static __device__ __forceinline__ void load(uint32_t addr, float (&f)[8]) {
    uint32_t w[4];
    asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]) : "r"(addr));
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      f[2 * i] = __uint_as_float(w[i] << 16);
      f[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
    }
  }
};

struct StreamLayout {
  uint32_t q, w, scratch, cum, ring, total;
  int ts, stages, stage_bytes;
};
// Shared-memory carve-up, identical on host and device.  stages < 2: the shape does not fit.
__host__ __device__ inline StreamLayout stream_layout(int head_dim, int esize, int batch, int warps, int gq) {
  StreamLayout L;
  const int row = head_dim * esize;
  int ts = kMaxTS;
  while (ts > 1 && 2 * ts * row > 65536 / warps) ts >>= 1;      // 8 KiB stages at 8 warps, 4 KiB at 16
  L.ts = ts;
  L.stage_bytes = 2 * ts * row;
  uint32_t o = 0;
  L.q = o;        o += (uint32_t)warps * gq * head_dim * 4;
  L.w = o;        o += (uint32_t)warps * kMaxTS * gq * 4;
  L.scratch = o;  o += 40 * 4;
  L.cum = o;      o += (uint32_t)(batch + 1) * 4;
  o = (o + 15u) & ~15u;
  L.ring = o;
  const int fit = ((int)kSmemBudgetS - (int)o - 16) / (warps * L.stage_bytes);
  L.stages = fit > 4 ? 4 : fit;
  L.total = o + 16 + (uint32_t)(L.stages > 0 ? L.stages : 0) * warps * L.stage_bytes;
  return L;
}

__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_wait(int pending) {     // at most `pending` groups still in flight
  switch (pending) {
    case 0: asm volatile("cp.async.wait_group 0;" ::: "memory"); break;
    case 1: asm volatile("cp.async.wait_group 1;" ::: "memory"); break;
    case 2: asm volatile("cp.async.wait_group 2;" ::: "memory"); break;
    default: asm volatile("cp.async.wait_group 3;" ::: "memory"); break;
  }
}

// T: cache / q / out element type.  CPL: 16-byte chunks of a V row per lane (head_dim <= 32 * CPL * N).
// GQ: query heads computed per head segment (4 when the GQA group is <= 4, else the full tile of 8): the
// kernel is bound by instruction issue at 8 warps per SM, so nothing is spent on predicated-off heads.
// WARPS: 8, or 16 with half-size stages (4 tokens: one token group) - the loop is bound by instruction issue and
// fixed-latency stalls, and four warps per scheduler hide what two cannot; 16 x 32 threads leave 128 registers each.
template <typename T, int CPL, int GQ, int WARPS>
__global__ void __launch_bounds__(WARPS * 32, 1)
pa_decode_stream_kernel(const PaDecodeParams p) {
  constexpr int N = Chunk<T>::N;
  extern __shared__ __align__(16) uint8_t smem_s[];
  const int D = p.head_dim, BS = p.block_size;
  const StreamLayout L = stream_layout(D, (int)sizeof(T), p.batch, WARPS, GQ);
  const int TS = L.ts, R = L.stages;
  const int row_bytes = D * (int)sizeof(T);
  const int cpr = row_bytes / 16;                      // 16-byte chunks per row
  float* s_q = reinterpret_cast<float*>(smem_s + L.q);
  int* scratch = reinterpret_cast<int*>(smem_s + L.scratch);
  int* s_cum = reinterpret_cast<int*>(smem_s + L.cum);

  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
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
  float* my_q = s_q + warp * GQ * D;
  float* my_w = reinterpret_cast<float*>(smem_s + L.w) + warp * kMaxTS * GQ;
  const uint32_t ring_base = (static_cast<uint32_t>(__cvta_generic_to_shared(smem_s + L.ring)) + 15u) & ~15u;
  const uint32_t my_ring = ring_base + (uint32_t)warp * R * L.stage_bytes;
  const uint32_t my_w_s = static_cast<uint32_t>(__cvta_generic_to_shared(my_w));
  const uint32_t my_q_s = static_cast<uint32_t>(__cvta_generic_to_shared(my_q));
  const int g = p.group;
  const int tg = lane >> 3, q8 = lane & 7;              // QK^T: token of the group of 4, eighth of the row
  const int r_first = lane / cpr, c_first = lane % cpr, r_step = 32 / cpr, c_step = 32 % cpr;   // cp.async chunk walk

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
    const int64_t head_off = (int64_t)kvh * p.kv_stride_head;

    __syncwarp();
    for (int i = lane; i < GQ * D; i += 32) {            // rows past the live heads are zeros: their scores stay 0
      const int j = i / D, d = i - j * D;
      my_q[i] = j < gq ? pa_to_float(qbase[(int64_t)cc.s * p.q_stride_b + (int64_t)(hq0 + j) * p.q_stride_h + d]) : 0.f;
    }
    __syncwarp();

    float m_run[GQ], l_run[GQ], acc[GQ][CPL * N];
#pragma unroll
    for (int j = 0; j < GQ; ++j) {
      m_run[j] = -INFINITY;
      l_run[j] = 0.f;
#pragma unroll
      for (int i = 0; i < CPL * N; ++i) acc[j][i] = 0.f;
    }

    // stage = a run of <= TS consecutive tokens inside one block: K rows, then V rows.  The producer cursor
    // (pt: token, pblk / poff: its block and offset in the block) and the consumer cursor advance by the same rule.
    auto issue = [&](int blk, int off, int n, int st) {
      const int64_t base = (int64_t)__ldg(bt + blk) * p.kv_stride_block + head_off + (int64_t)off * p.kv_stride_slot;
      const uint32_t dst0 = my_ring + (uint32_t)st * L.stage_bytes;
      // chunk i = lane, lane + 32, ... of the 2 n rows (K rows then V rows), cpr chunks per row: (row, chunk) advance
      // without a division
      int r = r_first, c = c_first;
      while (r < 2 * n) {
        const bool is_v = r >= n;
        const int tt = is_v ? r - n : r;
        const T* src = (is_v ? vbase : kbase) + base + (int64_t)tt * p.kv_stride_slot + c * N;
        cp_async16(dst0 + (uint32_t)((is_v ? TS + tt : tt) * row_bytes + c * 16), src);
        r += r_step;
        c += c_step;
        if (c >= cpr) { c -= cpr; ++r; }
      }
    };

    int pt = cc.b * BS, pblk = cc.b, poff = 0, ct = pt, coff = 0, ps = 0, cs = 0;
    auto produce = [&]() {
      if (pt < tok_end) {
        const int n = min(min(TS, BS - poff), tok_end - pt);
        issue(pblk, poff, n, ps);
        pt += n;
        poff += n;
        if (poff == BS) { poff = 0; ++pblk; }
        ps = ps + 1 == R ? 0 : ps + 1;
      }
      cp_commit();
    };
    for (int i = 0; i < R - 1; ++i) produce();
    while (ct < tok_end) {
      produce();
      cp_wait(R - 1);
      __syncwarp();
      const int n = min(min(TS, BS - coff), tok_end - ct);
      const uint32_t kst = my_ring + (uint32_t)cs * L.stage_bytes;
      const uint32_t vst = kst + (uint32_t)TS * row_bytes;

      // ---- scores of the stage's tokens: lane's tokens tg (group 0) and 4 + tg (group 1) ------------
      float sc[2][GQ];
#pragma unroll
      for (int j = 0; j < GQ; ++j) sc[0][j] = sc[1][j] = 0.f;
      const bool v0 = tg < n, v1 = 4 + tg < n;
      auto qk = [&](auto ng_c) {
        constexpr int NG = decltype(ng_c)::value;              // token groups of the stage: 2 (8 tokens) or 1 (<= 4)
        for (int c = q8; c < cpr; c += 8) {
          float k0[N], k1[N];
#pragma unroll
          for (int e = 0; e < N; ++e) k0[e] = k1[e] = 0.f;
          if (v0) Chunk<T>::load(kst + (uint32_t)(tg * row_bytes + c * 16), k0);
          if (NG == 2 && v1) Chunk<T>::load(kst + (uint32_t)((4 + tg) * row_bytes + c * 16), k1);
#pragma unroll
          for (int j = 0; j < GQ; ++j) {
#pragma unroll
            for (int e = 0; e < N; e += 4) {
              float qv[4];
              Chunk<float>::load(my_q_s + (uint32_t)((j * D + c * N + e) * 4), qv);
#pragma unroll
              for (int x = 0; x < 4; ++x) {
                sc[0][j] = fmaf(k0[e + x], qv[x], sc[0][j]);
                if (NG == 2) sc[1][j] = fmaf(k1[e + x], qv[x], sc[1][j]);
              }
            }
          }
        }
#pragma unroll
        for (int gi = 0; gi < NG; ++gi) {
#pragma unroll
          for (int j = 0; j < GQ; ++j) {
            float v = sc[gi][j];
            v += __shfl_xor_sync(0xffffffffu, v, 1);
            v += __shfl_xor_sync(0xffffffffu, v, 2);
            v += __shfl_xor_sync(0xffffffffu, v, 4);
            sc[gi][j] = v;
          }
        }
      };
      if (TS > 4) qk(std::integral_constant<int, 2>{}); else qk(std::integral_constant<int, 1>{});
      // ---- online softmax over the stage -------------------------------------------------------
#pragma unroll
      for (int j = 0; j < GQ; ++j) {
        const float s0 = v0 ? sc[0][j] * p.scale_log2e : -INFINITY;
        const float s1 = v1 ? sc[1][j] * p.scale_log2e : -INFINITY;
        float mx = fmaxf(s0, s1);
        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 8));
        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 16));
        const float m_new = fmaxf(m_run[j], mx);         // finite: the stage's first token is valid
        const float alpha = exp2f(m_run[j] - m_new);
        m_run[j] = m_new;
        const float p0 = exp2f(s0 - m_new), p1 = exp2f(s1 - m_new);
        float ps_ = p0 + p1;                               // every lane of a token's group holds the same value
        ps_ += __shfl_xor_sync(0xffffffffu, ps_, 8);
        ps_ += __shfl_xor_sync(0xffffffffu, ps_, 16);
        l_run[j] = l_run[j] * alpha + ps_;
#pragma unroll
        for (int i = 0; i < CPL * N; ++i) acc[j][i] *= alpha;
        if (q8 == 0) {
          my_w[tg * GQ + j] = p0;
          my_w[(4 + tg) * GQ + j] = p1;
        }
      }
      __syncwarp();
      // ---- PV: lane <-> 16-byte column chunk ---------------------------------------------------
      for (int t = 0; t < n; ++t) {
        float w[GQ];
#pragma unroll
        for (int j0 = 0; j0 < GQ; j0 += 4) {
          float w4[4];
          Chunk<float>::load(my_w_s + (uint32_t)((t * GQ + j0) * 4), w4);
#pragma unroll
          for (int j = 0; j < 4; ++j) w[j0 + j] = w4[j];
        }
#pragma unroll
        for (int ci = 0; ci < CPL; ++ci) {
          const int c = lane + 32 * ci;
          if (c < cpr) {
            float vf[N];
            Chunk<T>::load(vst + (uint32_t)(t * row_bytes + c * 16), vf);
#pragma unroll
            for (int j = 0; j < GQ; ++j) {
#pragma unroll
              for (int e = 0; e < N; ++e) acc[j][ci * N + e] = fmaf(w[j], vf[e], acc[j][ci * N + e]);
            }
          }
        }
      }
      __syncwarp();                                      // the stage (and the weight tile) may be overwritten
      ct += n;
      coff += n;
      if (coff == BS) coff = 0;
      cs = cs + 1 == R ? 0 : cs + 1;
    }
    cp_wait(0);

    // ---- emit: final result if the whole segment is ours, else a partial ------------------
    const int seg_lin0 = lin0 + done - cc.b;
    const bool whole = (seg_lin0 / per) == ((seg_lin0 + cc.n - 1) / per);
    const int slot = cc.s * hs + cc.h + worker;
#pragma unroll
    for (int j = 0; j < GQ; ++j) {
      if (j < gq) {
        const float l = l_run[j];
        if (whole) {
          const float inv = 1.f / l;
          T* op = obase + (int64_t)cc.s * p.o_stride_b + (int64_t)(hq0 + j) * p.o_stride_h;
#pragma unroll
          for (int ci = 0; ci < CPL; ++ci) {
            const int c = lane + 32 * ci;
            if (c < cpr) {
#pragma unroll
              for (int e = 0; e < N; ++e) pa_store(op + c * N + e, acc[j][ci * N + e] * inv);
            }
          }
        } else {
          float* po = p.part_o + ((int64_t)slot * kTQ + j) * D;
#pragma unroll
          for (int ci = 0; ci < CPL; ++ci) {
            const int c = lane + 32 * ci;
            if (c < cpr) {
#pragma unroll
              for (int e = 0; e < N; ++e) po[c * N + e] = acc[j][ci * N + e];
            }
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

template <typename T, int CPL, int GQ, int WARPS>
int launch_s(const PaDecodeParams& p, int ctas, cudaStream_t stream) {
  const StreamLayout L = stream_layout(p.head_dim, (int)sizeof(T), p.batch, WARPS, GQ);
  auto kern = pa_decode_stream_kernel<T, CPL, GQ, WARPS>;
  static PaPerDeviceFlag attr;
  const int dev = attr.need();
  if (dev != -2) {
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBudgetS) != cudaSuccess) return -5;
    attr.done(dev);
  }
  kern<<<ctas, WARPS * 32, L.total, stream>>>(p);
  return cudaGetLastError() == cudaSuccess ? 0 : -5;
}

}  // namespace

// 16 warps where 128 registers per thread are enough (GQ = 4, one 16-byte chunk per lane) and a ring of >= 2
// half-size stages fits; otherwise 8
static int stream_warps(const PaDecodeParams& p, int dtype) {
  const int esize = dtype == 1 ? 4 : 2;
  // ... and only while a half-size stage still holds 8 tokens (rows <= 256 bytes: bf16 D <= 128, fp32 D <= 64): with
  // 4-token stages the per-stage softmax and pipeline overhead eats the gain (fp32 D 128: 305 us against 285 us at
  // 8 warps; bf16 D 128: 243 us against 300 us)
  const StreamLayout L16 = stream_layout(p.head_dim, esize, p.batch, 16, 4);
  const int wide = (p.group <= 4 && p.head_dim * esize / 16 <= 32 && L16.stages >= 2 && L16.ts == kMaxTS) ? 16 : kWarpsS;
#ifdef PA_STREAM_WARPS
  return PA_STREAM_WARPS == 16 ? wide : kWarpsS;
#else
  return wide;
#endif
}

int pa_decode_stream_workers(const PaDecodeParams& p, int dtype, int num_sms, int* ctas) {
  *ctas = num_sms;
  return num_sms * stream_warps(p, dtype);
}

// 16-byte aligned rows of whole 16-byte chunks, at most 32 * CPL chunks per row, and a ring of >= 2 stages
bool pa_decode_stream_supported(const PaDecodeParams& p, int dtype) {
  const int esize = dtype == 1 ? 4 : 2, n = 16 / esize;
  if (p.head_dim % n) return false;
  if ((reinterpret_cast<uintptr_t>(p.k) | reinterpret_cast<uintptr_t>(p.v)) & 15) return false;
  if ((p.kv_stride_block | p.kv_stride_head | p.kv_stride_slot) % n) return false;
  if (p.head_dim * esize / 16 > 64) return false;
  if (p.batch > PA_MAX_BATCH_SMEM) return false;
  return stream_layout(p.head_dim, esize, p.batch, kWarpsS, p.group <= 4 ? 4 : 8).stages >= 2;
}

int pa_launch_decode_stream(const PaDecodeParams& p, int dtype, int ctas, cudaStream_t stream) {
  const int esize = dtype == 1 ? 4 : 2;
  const int cpr = p.head_dim * esize / 16;
  const bool g4 = p.group <= 4;               // every head segment has <= 4 live query heads
  const bool w16 = stream_warps(p, dtype) == 16;
  if (dtype == 1) {
    if (cpr <= 32) return g4 ? (w16 ? launch_s<float, 1, 4, 16>(p, ctas, stream) : launch_s<float, 1, 4, 8>(p, ctas, stream))
                             : launch_s<float, 1, 8, 8>(p, ctas, stream);
    return g4 ? launch_s<float, 2, 4, 8>(p, ctas, stream) : launch_s<float, 2, 8, 8>(p, ctas, stream);
  }
  if (cpr <= 32) return g4 ? (w16 ? launch_s<__nv_bfloat16, 1, 4, 16>(p, ctas, stream) : launch_s<__nv_bfloat16, 1, 4, 8>(p, ctas, stream))
                           : launch_s<__nv_bfloat16, 1, 8, 8>(p, ctas, stream);
  return g4 ? launch_s<__nv_bfloat16, 2, 4, 8>(p, ctas, stream) : launch_s<__nv_bfloat16, 2, 8, 8>(p, ctas, stream);
}
