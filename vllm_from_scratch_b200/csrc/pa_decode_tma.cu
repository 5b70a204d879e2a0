// Paged-attention decode, bf16, B200 fast path.
//
// Replaces the per-sequence Python loop of the reference
// (01_paged_attention/paged_attention.py:80-101) for bf16 caches with
// block_size 16 and head_dim 64 / 128.
//
// Shape of the kernel (HBM-bound streaming gather, ~4 flop/byte):
//  * persistent grid: one CTA per SM, NWARPS independent warps ("workers") per CTA;
//    every worker streams an equal, contiguous range of the flattened
//    (sequence, KV head, block) space (see pa_common.cuh);
//  * each worker owns a ring of STAGES shared-memory stages; one stage = the K tile
//    and the V tile of one physical block for one KV head ([16 tokens x D] bf16
//    each).  Lane 0 looks the physical block up in the block table and issues the
//    TMA tensor loads (cp.async.bulk.tensor, 128-byte swizzle) that complete on the
//    stage's mbarrier; the worker refills a stage as soon as it has consumed it, so
//    NWARPS * STAGES * 8 KiB stay in flight per SM;
//  * math on legacy tensor cores (mma.sync m16n8k16 bf16, fp32 accumulate) in the
//    TRANSPOSED formulation so the GQA query heads sit in the n = 8 dimension:
//        S^T[16 tok x 8 heads] = K[16 x D] . Q^T[D x 8]
//        O^T[D x 8 heads]     += V^T[D x 16 tok] . P^T[16 tok x 8 heads]
//    A operands come from shared memory with ldmatrix (K) / ldmatrix.trans (V),
//    conflict-free under the TMA swizzle; P^T moves from the S^T accumulator layout
//    to the B-operand layout with movmatrix.  tcgen05 needs M >= 64 and is not a fit
//    for a 4-row GQA group; the kernel is bound by HBM, not by the tensor pipe;
//  * online softmax in fp32 (exp2 domain), warp-shuffle reductions over tokens.
#include <cuda.h>
#include <mutex>
#include <string.h>

#include "pa_common.cuh"

namespace {

constexpr int kBS = 16;          // tokens per block on this path
constexpr int kWarps = 8;        // workers per CTA
constexpr int kStages = 3;       // ring depth per worker

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t done;
  do {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.b32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(bar), "r"(parity)
        : "memory");
  } while (!done);
}
__device__ __forceinline__ void tma_load_4d(uint32_t dst, const CUtensorMap* map, uint32_t bar,
                                            int c0, int c1, int c2, int c3, uint64_t policy) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint"
      " [%0], [%1, {%3, %4, %5, %6}], [%2], %7;"
      ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "l"(policy)
      : "memory");
}
__device__ __forceinline__ void ldsm_x4(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0, %1, %2, %3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr) : "memory");
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0, %1, %2, %3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr) : "memory");
}
__device__ __forceinline__ uint32_t movm_t(uint32_t a) {
  uint32_t d;
  asm volatile("movmatrix.sync.aligned.m8n8.trans.b16 %0, %1;" : "=r"(d) : "r"(a));
  return d;
}
__device__ __forceinline__ void mma_bf16(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, "
      "{%0, %1, %2, %3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
  __nv_bfloat162 t = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&t);
}
__device__ __forceinline__ float bf16lo(uint32_t u) { return __uint_as_float(u << 16); }
__device__ __forceinline__ float bf16hi(uint32_t u) { return __uint_as_float(u & 0xffff0000u); }

// HD: head_dim (64 / 128).  NT: n-tiles of 8 query heads per KV head (g <= 8 * NT).
template <int HD, int NT>
__global__ void __launch_bounds__(kWarps * 32, 1)
pa_decode_tma_kernel(const __grid_constant__ CUtensorMap tm_k, const __grid_constant__ CUtensorMap tm_v,
                     const PaDecodeParams p) {
  constexpr int HALVES = HD / 64;             // 128-byte column halves of a row
  constexpr int HALF_BYTES = kBS * 128;       // one [16 x 64] bf16 box
  constexpr int TILE_BYTES = HALVES * HALF_BYTES;
  constexpr int STAGE_BYTES = 2 * TILE_BYTES; // K tile + V tile
  constexpr int KSTEPS = HD / 16;

  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
  // [kWarps][kStages][STAGE_BYTES] | mbar[kWarps][kStages] (8 B each) | scratch[33] | cum[B + 1]
  const uint32_t ring_bytes = kWarps * kStages * STAGE_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_gen + ring_bytes);
  int* scratch = reinterpret_cast<int*>(bars + kWarps * kStages);
  int* s_cum = scratch + 40;

  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;

  if (threadIdx.x == 0) {
#pragma unroll
    for (int i = 0; i < kWarps * kStages; ++i) mbar_init(smem_u32(bars + i), 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tm_k) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tm_v) : "memory");
  }
  pa_build_cum(p.seq_lens, p.batch, kBS, s_cum, scratch);   // ends with __syncthreads()

  const int hs = p.num_kv_heads;              // one head segment per KV head (q_tiles == 1)
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

  uint64_t policy;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(policy));

  const uint32_t my_ring = smem_base + warp * (kStages * STAGE_BYTES);
  const uint32_t my_bars = smem_u32(bars + warp * kStages);

  PaCursor pc = pa_locate(s_cum, p.batch, hs, lin0);
  PaCursor cc = pc;

  auto produce = [&](int it) {
    if (lane == 0) {
      const int stage = it % kStages;
      const int phys = __ldg(p.block_tables + (int64_t)pc.s * p.bt_stride + pc.b);
      const uint32_t bar = my_bars + stage * 8;
      const uint32_t dst = my_ring + stage * STAGE_BYTES;
      const int c1 = p.pos_head == 1 ? pc.h : (p.pos_block == 1 ? phys : 0);
      const int c2 = p.pos_head == 2 ? pc.h : (p.pos_block == 2 ? phys : 0);
      const int c3 = p.pos_head == 3 ? pc.h : (p.pos_block == 3 ? phys : 0);
      mbar_expect_tx(bar, STAGE_BYTES);
#pragma unroll
      for (int hf = 0; hf < HALVES; ++hf) {
        tma_load_4d(dst + hf * HALF_BYTES, &tm_k, bar, hf * 64, c1, c2, c3, policy);
        tma_load_4d(dst + TILE_BYTES + hf * HALF_BYTES, &tm_v, bar, hf * 64, c1, c2, c3, policy);
      }
    }
    pa_advance(pc, s_cum, p.batch, hs);
  };

  const int prologue = min(kStages, count);
  for (int it = 0; it < prologue; ++it) produce(it);

  // per-lane constant pieces of the ldmatrix addresses
  const int k_tok = (lane & 7) + ((lane >> 3) & 1) * 8;       // K: matrices 0/1 = tokens 0-7 / 8-15
  const int k_csel = lane >> 4;                               //    matrices 2/3 = next 8 dims
  const int v_tok = (lane & 7) + ((lane >> 4) & 1) * 8;       // V^T: matrices 2/3 = tokens 8-15
  const int v_csel = (lane >> 3) & 1;                         //      matrices 1/3 = next 8 dims
  const uint32_t k_row = k_tok * 128, v_row = v_tok * 128;
  const int k_x = k_tok & 7, v_x = v_tok & 7;

  const __nv_bfloat16* qbase = static_cast<const __nv_bfloat16*>(p.q);
  __nv_bfloat16* obase = static_cast<__nv_bfloat16*>(p.out);
  const int g = p.group;

  uint32_t qf[NT][KSTEPS][2];
  float acc[KSTEPS][NT][4];
  float m_run[NT][2], l_run[NT][2];
  bool seg_start = true;
  int seg_lin0 = 0;

  for (int it = 0; it < count; ++it) {
    const int stage = it % kStages;
    if (seg_start) {
      seg_start = false;
      seg_lin0 = lin0 + it - cc.b;        // linear index of block 0 of this segment
      // Q^T B-fragments: b0 = Q[head = lane/4][dim = ks*16 + (lane%4)*2 + {0,1}], b1 = dim + 8
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) {
        const int hq = nt * 8 + (lane >> 2);
        const __nv_bfloat16* qp = qbase + (int64_t)cc.s * p.q_stride_b + (int64_t)(cc.h * g + hq) * p.q_stride_h +
                                  (lane & 3) * 2;
#pragma unroll
        for (int ks = 0; ks < KSTEPS; ++ks) {
          uint32_t b0 = 0, b1 = 0;
          if (hq < g) {
            b0 = __ldg(reinterpret_cast<const uint32_t*>(qp + ks * 16));
            b1 = __ldg(reinterpret_cast<const uint32_t*>(qp + ks * 16 + 8));
          }
          qf[nt][ks][0] = b0;
          qf[nt][ks][1] = b1;
        }
        m_run[nt][0] = m_run[nt][1] = -INFINITY;
        l_run[nt][0] = l_run[nt][1] = 0.f;
      }
#pragma unroll
      for (int mt = 0; mt < KSTEPS; ++mt)
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) acc[mt][nt][0] = acc[mt][nt][1] = acc[mt][nt][2] = acc[mt][nt][3] = 0.f;
    }

    mbar_wait(my_bars + stage * 8, (it / kStages) & 1);
    const uint32_t kt = my_ring + stage * STAGE_BYTES;
    const uint32_t vt = kt + TILE_BYTES;

    // ---- S^T = K . Q^T ----------------------------------------------------------
    float sc[NT][2][4];
#pragma unroll
    for (int nt = 0; nt < NT; ++nt)
#pragma unroll
      for (int c = 0; c < 2; ++c) sc[nt][c][0] = sc[nt][c][1] = sc[nt][c][2] = sc[nt][c][3] = 0.f;
#pragma unroll
    for (int ks = 0; ks < KSTEPS; ++ks) {
      uint32_t a[4];
      const int chunk = (ks & 3) * 2 + k_csel;
      ldsm_x4(a, kt + (ks >> 2) * HALF_BYTES + k_row + ((chunk ^ k_x) << 4));
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) mma_bf16(sc[nt][ks & 1], a, qf[nt][ks][0], qf[nt][ks][1]);
    }

    // ---- online softmax over this block's 16 tokens ----------------------------------
    const int seq_len = __ldg(p.seq_lens + cc.s);
    const int valid = seq_len - cc.b * kBS;          // >= 1; < 16 only in the last block
    const int t0 = lane >> 2;
    uint32_t pb[NT][2];
    bool need_rescale = false;
    float alpha[NT][2];
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
      float s0[2], s1[2];
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        s0[j] = (sc[nt][0][j] + sc[nt][1][j]) * p.scale_log2e;           // token t0,     head 2*(lane%4)+j
        s1[j] = (sc[nt][0][j + 2] + sc[nt][1][j + 2]) * p.scale_log2e;   // token t0 + 8
        if (t0 >= valid) s0[j] = -INFINITY;
        if (t0 + 8 >= valid) s1[j] = -INFINITY;
        float mx = fmaxf(s0[j], s1[j]);
        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 4));
        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 8));
        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 16));
        const float m_new = fmaxf(m_run[nt][j], mx);     // finite: token 0 of the block is always valid
        alpha[nt][j] = exp2f(m_run[nt][j] - m_new);      // exp2(-inf) = 0 on the first block
        need_rescale |= (m_new != m_run[nt][j]);
        m_run[nt][j] = m_new;
        s0[j] = exp2f(s0[j] - m_new);
        s1[j] = exp2f(s1[j] - m_new);
      }
      const uint32_t p0 = pack_bf16(s0[0], s0[1]);       // token t0:     heads (2c, 2c+1)
      const uint32_t p1 = pack_bf16(s1[0], s1[1]);       // token t0 + 8
      // the normaliser sums the SAME bf16-rounded weights the PV product uses
      l_run[nt][0] = l_run[nt][0] * alpha[nt][0] + bf16lo(p0) + bf16lo(p1);
      l_run[nt][1] = l_run[nt][1] * alpha[nt][1] + bf16hi(p0) + bf16hi(p1);
      pb[nt][0] = movm_t(p0);                            // -> B fragment: k = tokens 0-7,  n = head
      pb[nt][1] = movm_t(p1);                            //                k = tokens 8-15
    }
    if (__any_sync(0xffffffffu, need_rescale)) {
#pragma unroll
      for (int mt = 0; mt < KSTEPS; ++mt)
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) {
          acc[mt][nt][0] *= alpha[nt][0];
          acc[mt][nt][1] *= alpha[nt][1];
          acc[mt][nt][2] *= alpha[nt][0];
          acc[mt][nt][3] *= alpha[nt][1];
        }
    }

    // A V row of a token past seq_len may hold anything (Inf/NaN bit patterns in a recycled
    // block); 0 * NaN would poison the product, so clear those rows in shared memory.
    if (valid < kBS) {
      for (int i = lane; i < (kBS - valid) * (HD / 8); i += 32) {
        const int row = valid + i / (HD / 8), c = i % (HD / 8);
        const uint32_t a = vt + (c >> 3) * HALF_BYTES + row * 128 + (((c & 7) ^ (row & 7)) << 4);
        asm volatile("st.shared.v4.b32 [%0], {%1, %1, %1, %1};" ::"r"(a), "r"(0) : "memory");
      }
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      __syncwarp();
    }

    // ---- O^T += V^T . P^T -------------------------------------------------------
#pragma unroll
    for (int mt = 0; mt < KSTEPS; ++mt) {
      uint32_t a[4];
      const int chunk = (mt & 3) * 2 + v_csel;
      ldsm_x4_t(a, vt + (mt >> 2) * HALF_BYTES + v_row + ((chunk ^ v_x) << 4));
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) mma_bf16(acc[mt][nt], a, pb[nt][0], pb[nt][1]);
    }

    // stage consumed by every lane -> refill it
    __syncwarp();
    if (it + kStages < count) produce(it + kStages);

    // ---- end of this (sequence, head) segment, or of the worker's range -------------
    const bool seg_end = (cc.b + 1 == cc.n);
    if (seg_end || it + 1 == count) {
      // segment [seg_lin0, seg_lin0 + n) lies inside one worker <=> write the final result
      const bool whole = (seg_lin0 / per) == ((seg_lin0 + cc.n - 1) / per);
      const int slot = cc.s * hs + cc.h + worker;
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) {
        float l[2];
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          l[j] = l_run[nt][j];
          l[j] += __shfl_xor_sync(0xffffffffu, l[j], 4);
          l[j] += __shfl_xor_sync(0xffffffffu, l[j], 8);
          l[j] += __shfl_xor_sync(0xffffffffu, l[j], 16);
        }
        const int hq = nt * 8 + (lane & 3) * 2;      // this lane's two heads: hq, hq + 1
        if (whole) {
          const float inv0 = 1.f / l[0], inv1 = 1.f / l[1];
          __nv_bfloat16* o0 = obase + (int64_t)cc.s * p.o_stride_b + (int64_t)(cc.h * g + hq) * p.o_stride_h;
          __nv_bfloat16* o1 = o0 + p.o_stride_h;
#pragma unroll
          for (int mt = 0; mt < KSTEPS; ++mt) {
            const int d = mt * 16 + (lane >> 2);
            if (hq < g) {
              o0[d] = __float2bfloat16_rn(acc[mt][nt][0] * inv0);
              o0[d + 8] = __float2bfloat16_rn(acc[mt][nt][2] * inv0);
            }
            if (hq + 1 < g) {
              o1[d] = __float2bfloat16_rn(acc[mt][nt][1] * inv1);
              o1[d + 8] = __float2bfloat16_rn(acc[mt][nt][3] * inv1);
            }
          }
        } else {
          float* po = p.part_o + ((int64_t)slot * p.tile_q + hq) * HD;
#pragma unroll
          for (int mt = 0; mt < KSTEPS; ++mt) {
            const int d = mt * 16 + (lane >> 2);
            if (hq < g) { po[d] = acc[mt][nt][0]; po[d + 8] = acc[mt][nt][2]; }
            if (hq + 1 < g) { po[HD + d] = acc[mt][nt][1]; po[HD + d + 8] = acc[mt][nt][3]; }
          }
          if ((lane >> 2) == 0) {
            if (hq < g) p.part_ml[(int64_t)slot * p.tile_q + hq] = make_float2(m_run[nt][0], l[0]);
            if (hq + 1 < g) p.part_ml[(int64_t)slot * p.tile_q + hq + 1] = make_float2(m_run[nt][1], l[1]);
          }
        }
      }
      seg_start = true;
    }
    pa_advance(cc, s_cum, p.batch, hs);
  }
}

// ------------------------------------------------------------------------------------
// host side: tensor maps
// ------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* sym = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(sym);
  });
  return fn;
}

struct MapKey {
  const void* base;
  int64_t sb, sh, ss;
  int32_t nb, nh, hd;
  bool operator==(const MapKey& o) const { return memcmp(this, &o, sizeof(MapKey)) == 0; }
};
struct MapEntry {
  MapKey key;
  CUtensorMap map;
  int pos_head, pos_block;
  bool used;
};
constexpr int kMapCache = 64;
MapEntry g_maps[kMapCache];
int g_map_next = 0;
std::mutex g_map_mu;

// Builds (or fetches) the 4-D tensor map of one K or V plane: dim0 = head_dim columns
// (box 64 = 128 bytes, SWIZZLE_128B), dims 1..3 = {slot, head, block} ordered by ascending
// stride; the box covers the 16 slots of one (block, head).
int get_map(const void* base, const PaDecodeParams& p, CUtensorMap* out, int* pos_head, int* pos_block) {
  MapKey key;
  memset(&key, 0, sizeof(key));
  key.base = base;
  key.sb = p.kv_stride_block; key.sh = p.kv_stride_head; key.ss = p.kv_stride_slot;
  key.nb = 0; key.nh = p.num_kv_heads; key.hd = p.head_dim;
  std::lock_guard<std::mutex> lock(g_map_mu);
  for (int i = 0; i < kMapCache; ++i)
    if (g_maps[i].used && g_maps[i].key == key) {
      *out = g_maps[i].map; *pos_head = g_maps[i].pos_head; *pos_block = g_maps[i].pos_block;
      return 0;
    }
  EncodeTiledFn enc = get_encode_fn();
  if (!enc) return -5;
  struct Dim { int64_t stride; uint64_t size; uint32_t box; int what; };   // what: 0 slot, 1 head, 2 block
  // The block extent is deliberately the largest the strides allow: block ids are validated by
  // the caller's allocator, and TMA zero-fills anything out of range instead of faulting.
  Dim d[3] = {{p.kv_stride_slot, (uint64_t)kBS, (uint32_t)kBS, 0},
              {p.kv_stride_head, (uint64_t)p.num_kv_heads, 1u, 1},
              {p.kv_stride_block, (uint64_t)1 << 31, 1u, 2}};
  for (int i = 0; i < 3; ++i)
    for (int j = i + 1; j < 3; ++j)
      if (d[j].stride < d[i].stride) { Dim t = d[i]; d[i] = d[j]; d[j] = t; }
  cuuint64_t gdim[4] = {(cuuint64_t)p.head_dim, d[0].size, d[1].size, d[2].size};
  cuuint64_t gstr[3] = {(cuuint64_t)d[0].stride * 2, (cuuint64_t)d[1].stride * 2, (cuuint64_t)d[2].stride * 2};
  cuuint32_t box[4] = {64, d[0].box, d[1].box, d[2].box};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUtensorMap m;
  CUresult r = enc(&m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), gdim, gstr, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return -5;
  int ph = 0, pb = 0;
  for (int i = 0; i < 3; ++i) {
    if (d[i].what == 1) ph = i + 1;
    if (d[i].what == 2) pb = i + 1;
  }
  MapEntry& e = g_maps[g_map_next];
  g_map_next = (g_map_next + 1) % kMapCache;
  e.key = key; e.map = m; e.pos_head = ph; e.pos_block = pb; e.used = true;
  *out = m; *pos_head = ph; *pos_block = pb;
  return 0;
}

template <int HD, int NT>
int launch(const PaDecodeParams& p, const CUtensorMap& mk, const CUtensorMap& mv, int num_sms, cudaStream_t stream) {
  constexpr int STAGE_BYTES = 2 * (HD / 64) * kBS * 128;
  const size_t smem = 1024 + (size_t)kWarps * kStages * STAGE_BYTES + kWarps * kStages * 8 + 40 * 4 +
                      (size_t)(p.batch + 1) * 4 + 16;
  auto kern = pa_decode_tma_kernel<HD, NT>;
  static bool attr_set = false;
  if (!attr_set) {
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024) != cudaSuccess) return -5;
    attr_set = true;
  }
  if (smem > 227 * 1024) return -2;
  kern<<<num_sms, kWarps * 32, smem, stream>>>(mk, mv, p);
  return cudaGetLastError() == cudaSuccess ? 0 : -5;
}

}  // namespace

int pa_decode_tma_workers(int num_sms) { return num_sms * kWarps; }

bool pa_decode_tma_supported(const PaDecodeParams& p, int dtype) {
  if (dtype != 0) return false;                                   // bf16 only
  if (p.block_size != kBS) return false;
  if (p.head_dim != 64 && p.head_dim != 128) return false;
  if (p.group > 16) return false;
  if (p.batch > PA_MAX_BATCH_SMEM) return false;
  if ((reinterpret_cast<uintptr_t>(p.k) | reinterpret_cast<uintptr_t>(p.v)) & 15) return false;
  if ((p.kv_stride_block | p.kv_stride_head | p.kv_stride_slot) & 7) return false;   // 16-byte strides
  if (p.kv_stride_block <= 0 || p.kv_stride_head <= 0 || p.kv_stride_slot <= 0) return false;
  if ((reinterpret_cast<uintptr_t>(p.q) & 3) || ((p.q_stride_b | p.q_stride_h) & 1)) return false;
  return true;
}

int pa_launch_decode_tma(const PaDecodeParams& p_in, int num_sms, cudaStream_t stream) {
  PaDecodeParams p = p_in;
  CUtensorMap mk, mv;
  int ph, pb;
  int rc = get_map(p.k, p, &mk, &ph, &pb);
  if (rc) return rc;
  rc = get_map(p.v, p, &mv, &ph, &pb);
  if (rc) return rc;
  p.pos_head = ph;
  p.pos_block = pb;
  const int nt = (p.group + 7) / 8;
  if (p.head_dim == 128) return nt == 1 ? launch<128, 1>(p, mk, mv, num_sms, stream) : launch<128, 2>(p, mk, mv, num_sms, stream);
  return nt == 1 ? launch<64, 1>(p, mk, mv, num_sms, stream) : launch<64, 2>(p, mk, mv, num_sms, stream);
}
