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

// ------------------------------------------------------------------------------------
// launch plan (uniform, recomputed by every thread from the block prefix sums)
// ------------------------------------------------------------------------------------
// The flattened (sequence, KV head, block) space [0, total) is split in two regions:
//   static  [0, S)      : worker w owns [w * per_s, (w + 1) * per_s) - one long contiguous range,
//                         so it leaves at most one partial per segment it touches;
//   pool    [S, total)  : cut into NC chunks of CH blocks that the workers claim through an atomic
//                         counter once their static range is done.  SMs do not all get the same
//                         share of the memory system (ncu: per-SM active cycles spread 75 % .. 95 % of
//                         the kernel with a purely static split); the pool lets the fast ones absorb
//                         the difference so that every SM streams until the end.
// Partials of pool chunks are merged in two levels (groups of kGroup consecutive chunks, then the
// segment) so that no merge reads more than a few dozen partials.
struct Plan {
  int total, per_s, S, CH, NC;
};

constexpr int kGroup = 32;          // pool chunks per first-level merge group
constexpr int kChunkMin = 4;        // smallest pool chunk, in blocks (32 KiB of K+V at D = 128)
constexpr int kPoolDiv = 4;         // pool = 1 / kPoolDiv of the work
constexpr int kHeaderInts = 16;     // counter header: [0] pool ticket, [1] exited workers

__device__ __forceinline__ Plan make_plan(int total, int W, int nc_max, int pool_div, int chunk_min) {
  Plan pl;
  pl.total = total;
  if (pool_div > 0 && total >= W * 16) {
    pl.per_s = (total - total / pool_div) / W;
    pl.S = pl.per_s * W;
  } else {
    pl.per_s = max(pa_ceil_div(total, W), PA_MIN_BLOCKS_PER_WORKER);
    pl.S = total;
  }
  const int pool = total - pl.S;
  pl.CH = max(chunk_min, pa_ceil_div(pool, nc_max));
  pl.NC = pa_ceil_div(pool, pl.CH);
  return pl;
}

// Which ranges touch the segment [a, b) of the linear space.
struct SegInfo {
  int w0, ns;        // static workers w0 .. w0 + ns - 1
  int c0, c1;        // pool chunks c0 .. c1 (valid when ng > 0)
  int g0, ng;        // chunk groups g0 .. g0 + ng - 1
  int expected;      // arrivals at the segment counter = ns + ng
};

__device__ __forceinline__ SegInfo seg_info(const Plan& pl, int a, int b) {
  SegInfo si;
  si.w0 = 0; si.ns = 0; si.c0 = 0; si.c1 = -1; si.g0 = 0; si.ng = 0;
  if (a < pl.S) {
    si.w0 = a / pl.per_s;
    si.ns = (min(b, pl.S) - 1) / pl.per_s - si.w0 + 1;
  }
  if (b > pl.S) {
    si.c0 = (max(a, pl.S) - pl.S) / pl.CH;
    si.c1 = (b - 1 - pl.S) / pl.CH;
    si.g0 = si.c0 / kGroup;
    si.ng = si.c1 / kGroup - si.g0 + 1;
  }
  si.expected = si.ns + si.ng;
  return si;
}

__device__ __forceinline__ float4 ldcg4(const float* p) { return __ldcg(reinterpret_cast<const float4*>(p)); }

// Merge the partials in slots [lo1, lo1 + n1) and [lo2, lo2 + n2) of one segment (every row of the
// GQA group) with the usual log-sum-exp weights.  Either normalises and writes out[s, head, :], or
// writes the merged (unnormalised) partial into `dst_slot`.  Called by a whole warp; lanes own 4
// consecutive head_dim columns.  Partials were written by other SMs: read them through L2 (ld.cg).
template <int HD>
__device__ __noinline__ void merge_partials(const PaDecodeParams& p, int lo1, int n1, int lo2, int n2,
                                               bool to_out, int dst_slot, int s, int kvh, int lane) {
  constexpr int RC = 4;                                   // rows merged together
  constexpr int SU = 8;                                   // sources in flight per round (x RC loads)
  const int g = p.group, n = n1 + n2;
  const bool col_ok = lane * 4 < HD;
  auto slot_of = [&](int i) { return i < n1 ? lo1 + i : lo2 + (i - n1); };
  for (int r0 = 0; r0 < g; r0 += RC) {
    // pass 1: running max per row over all sources (lane <-> source)
    float m[RC];
#pragma unroll
    for (int r = 0; r < RC; ++r) m[r] = -INFINITY;
    for (int i = lane; i < n; i += 32) {
      const int64_t base = (int64_t)slot_of(i) * g + r0;
#pragma unroll
      for (int r = 0; r < RC; ++r)
        if (r0 + r < g) m[r] = fmaxf(m[r], __ldcg(&p.part_ml[base + r]).x);
    }
#pragma unroll
    for (int r = 0; r < RC; ++r)
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) m[r] = fmaxf(m[r], __shfl_xor_sync(0xffffffffu, m[r], o));
    float l[RC];
    float4 acc[RC];
#pragma unroll
    for (int r = 0; r < RC; ++r) { l[r] = 0.f; acc[r] = make_float4(0.f, 0.f, 0.f, 0.f); }
    for (int i0 = 0; i0 < n; i0 += 32) {
      float w_mine[RC];                                   // lane j: weight of source i0 + j
#pragma unroll
      for (int r = 0; r < RC; ++r) w_mine[r] = 0.f;
      if (i0 + lane < n) {
        const int64_t base = (int64_t)slot_of(i0 + lane) * g + r0;
#pragma unroll
        for (int r = 0; r < RC; ++r)
          if (r0 + r < g) {
            const float2 ml = __ldcg(&p.part_ml[base + r]);
            w_mine[r] = exp2f(ml.x - m[r]);
            l[r] += ml.y * w_mine[r];
          }
      }
      const int cnt = min(32, n - i0);
      for (int j0 = 0; j0 < cnt; j0 += SU) {
        float4 v[SU][RC];
#pragma unroll
        for (int u = 0; u < SU; ++u) {
          const bool live = j0 + u < cnt && col_ok;
          const int64_t base = live ? ((int64_t)slot_of(i0 + j0 + u) * g + r0) * HD + lane * 4 : 0;
#pragma unroll
          for (int r = 0; r < RC; ++r) {
            v[u][r] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (live && r0 + r < g) v[u][r] = ldcg4(p.part_o + base + r * HD);
          }
        }
#pragma unroll
        for (int u = 0; u < SU; ++u)
#pragma unroll
          for (int r = 0; r < RC; ++r) {
            const float w = __shfl_sync(0xffffffffu, w_mine[r], (j0 + u) & 31);   // 0 past the end
            acc[r].x = fmaf(w, v[u][r].x, acc[r].x); acc[r].y = fmaf(w, v[u][r].y, acc[r].y);
            acc[r].z = fmaf(w, v[u][r].z, acc[r].z); acc[r].w = fmaf(w, v[u][r].w, acc[r].w);
          }
      }
    }
#pragma unroll
    for (int r = 0; r < RC; ++r) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) l[r] += __shfl_xor_sync(0xffffffffu, l[r], o);
      if (r0 + r >= g) continue;
      const int row = r0 + r;
      if (to_out) {
        const float inv = 1.f / l[r];
        __nv_bfloat16* op = static_cast<__nv_bfloat16*>(p.out) + (int64_t)s * p.o_stride_b +
                            (int64_t)(kvh * g + row) * p.o_stride_h + lane * 4;
        if (col_ok) {
          if ((reinterpret_cast<uintptr_t>(op) & 7) == 0) {
            uint2 pk;
            pk.x = pack_bf16(acc[r].x * inv, acc[r].y * inv);
            pk.y = pack_bf16(acc[r].z * inv, acc[r].w * inv);
            *reinterpret_cast<uint2*>(op) = pk;
          } else {
            op[0] = __float2bfloat16_rn(acc[r].x * inv); op[1] = __float2bfloat16_rn(acc[r].y * inv);
            op[2] = __float2bfloat16_rn(acc[r].z * inv); op[3] = __float2bfloat16_rn(acc[r].w * inv);
          }
        }
      } else {
        if (col_ok) *reinterpret_cast<float4*>(p.part_o + ((int64_t)dst_slot * g + row) * HD + lane * 4) = acc[r];
        if (lane == 0) p.part_ml[(int64_t)dst_slot * g + row] = make_float2(m[r], l[r]);
      }
    }
  }
}

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
  constexpr uint32_t FULL = 0xffffffffu;

  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
  // [kWarps][kStages][STAGE_BYTES] | mbar[kWarps][kStages] | meta[kWarps][kStages] (2 x int4) | scratch[40] | cum[B + 1]
  const uint32_t ring_bytes = kWarps * kStages * STAGE_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_gen + ring_bytes);
  int4* meta = reinterpret_cast<int4*>(bars + kWarps * kStages);
  int* scratch = reinterpret_cast<int*>(meta + 2 * kWarps * kStages);
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

  const int g = p.group;
  __nv_bfloat16* obase = static_cast<__nv_bfloat16*>(p.out);

  // sequences with no tokens: defined result is zeros (the reference raises, paged_attention.py:93)
  for (int s = blockIdx.x * blockDim.x + threadIdx.x; s < p.batch; s += gridDim.x * blockDim.x)
    if (s_cum[s + 1] == s_cum[s])
      for (int h = 0; h < p.num_q_heads; ++h)
        for (int d = 0; d < HD; ++d) obase[(int64_t)s * p.o_stride_b + (int64_t)h * p.o_stride_h + d] = __float2bfloat16_rn(0.f);

  const int hs = p.num_kv_heads;              // one head segment per KV head (q_tiles == 1)
  const int W = p.num_workers;
  const Plan pl = make_plan(s_cum[p.batch] * hs, W, p.nc_max, p.pool_div, p.chunk_min);
  const int segs = p.batch * hs;
  const int slot2_base = segs + W + p.nc_max;                    // group partials live after the range partials
  int* const ticket = p.counters;                                // [0] pool ticket, [1] exited workers
  int* const grp_cnt = p.counters + kHeaderInts;                 // [seg + group]
  int* const seg_cnt = grp_cnt + segs + pa_ceil_div(p.nc_max, kGroup);   // [seg]

  const int worker = warp * gridDim.x + blockIdx.x;
  unsigned long long dbg_t0 = 0, dbg_t1 = 0;
  int dbg_static = 0;
  if (p.dbg) asm volatile("mov.u64 %0, %globaltimer;" : "=l"(dbg_t0));
  const uint32_t my_ring = smem_base + warp * (kStages * STAGE_BYTES);
  const uint32_t my_bars = smem_u32(bars + warp * kStages);
  int4* my_meta = meta + 2 * warp * kStages;

  uint64_t policy;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(policy));

  // ---- producer: walks this worker's ranges (static range, then claimed pool chunks) ----------------
  int p_uid = worker;                                            // range id: worker, or W + chunk
  int p_left = min(max(pl.S - worker * pl.per_s, 0), pl.per_s);  // blocks left in the current range
  bool p_first = true;
  PaCursor pc;
  pc.s = pc.h = pc.b = 0; pc.n = 1;
  if (p_left > 0) pc = pa_locate(s_cum, p.batch, hs, worker * pl.per_s);
  bool pool_open = pl.NC > 0;
  int nxt_raw = 0;                                               // lane 0: chunk claimed one range ahead
  if (pool_open && lane == 0) nxt_raw = atomicAdd(ticket, 1);
  dbg_static = p_left > 0 ? p_left - 1 : -1;                    // consumer iteration that ends the static range
  int produced = 0;
  int bt_reg = 0, bt_s = -1, bt_base = 0;                        // 32 block-table entries cached across the lanes

  auto produce = [&]() {
    if (p_left == 0) {
      if (!pool_open) return;
      const int c = __shfl_sync(FULL, nxt_raw, 0);
      if (c >= pl.NC) { pool_open = false; return; }
      if (lane == 0) nxt_raw = atomicAdd(ticket, 1);
      const int lin = pl.S + c * pl.CH;
      p_uid = W + c;
      p_left = min(pl.CH, pl.total - lin);
      pc = pa_locate(s_cum, p.batch, hs, lin);
      p_first = true;
    }
    if (bt_s != pc.s || pc.b < bt_base || pc.b >= bt_base + 32) {
      bt_s = pc.s;
      bt_base = pc.b;
      bt_reg = (bt_base + lane < pc.n) ? __ldg(p.block_tables + (int64_t)pc.s * p.bt_stride + bt_base + lane) : 0;
    }
    const int phys = __shfl_sync(FULL, bt_reg, pc.b - bt_base);
    const bool last_blk = pc.b + 1 == pc.n;
    const int first = (p_first || pc.b == 0) ? 1 : 0;
    const int last = (p_left == 1 || last_blk) ? 1 : 0;
    if (lane == 0) {
      const int stage = produced % kStages;
      const int valid = last_blk ? __ldg(p.seq_lens + pc.s) - pc.b * kBS : kBS;
      my_meta[2 * stage] = make_int4(pc.s, pc.h, pc.b, pc.n);
      my_meta[2 * stage + 1] = make_int4(p_uid, first | (last << 1), valid, 0);
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
    p_first = false;
    --p_left;
    ++produced;
    pa_advance(pc, s_cum, p.batch, hs);
  };

  for (int i = 0; i < kStages; ++i) produce();
  __syncwarp();

  // per-lane constant pieces of the ldmatrix addresses
  const int k_tok = (lane & 7) + ((lane >> 3) & 1) * 8;       // K: matrices 0/1 = tokens 0-7 / 8-15
  const int k_csel = lane >> 4;                               //    matrices 2/3 = next 8 dims
  const int v_tok = (lane & 7) + ((lane >> 4) & 1) * 8;       // V^T: matrices 2/3 = tokens 8-15
  const int v_csel = (lane >> 3) & 1;                         //      matrices 1/3 = next 8 dims
  const uint32_t k_row = k_tok * 128, v_row = v_tok * 128;
  const int k_x = k_tok & 7, v_x = v_tok & 7;

  const __nv_bfloat16* qbase = static_cast<const __nv_bfloat16*>(p.q);

  uint32_t qf[NT][KSTEPS][2];
  float acc[KSTEPS][NT][4];
  float m_run[NT][2], l_run[NT][2];

  for (int it = 0; it < produced; ++it) {
    const int stage = it % kStages;
    mbar_wait(my_bars + stage * 8, (it / kStages) & 1);
    if (p.dbg && it == dbg_static) asm volatile("mov.u64 %0, %globaltimer;" : "=l"(dbg_t1));
    const int4 md = my_meta[2 * stage];          // (s, h, b, n)
    const int4 mf = my_meta[2 * stage + 1];      // (range uid, first | last << 1, valid tokens, -)
    const int cs = md.x, ch = md.y, cn = md.w;

    if (mf.y & 1) {
      // new piece: Q^T B-fragments: b0 = Q[head = lane/4][dim = ks*16 + (lane%4)*2 + {0,1}], b1 = dim + 8
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) {
        const int hq = nt * 8 + (lane >> 2);
        const __nv_bfloat16* qp = qbase + (int64_t)cs * p.q_stride_b + (int64_t)(ch * g + hq) * p.q_stride_h +
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
    const int valid = mf.z;                          // >= 1; < 16 only in the last block
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
        mx = fmaxf(mx, __shfl_xor_sync(FULL, mx, 4));
        mx = fmaxf(mx, __shfl_xor_sync(FULL, mx, 8));
        mx = fmaxf(mx, __shfl_xor_sync(FULL, mx, 16));
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
    if (__any_sync(FULL, need_rescale)) {
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

    // stage (data and meta) consumed by every lane -> refill it
    __syncwarp();
    produce();
    __syncwarp();

    if (!(mf.y & 2)) continue;

    // ---- end of a piece = (range, segment) intersection --------------------------------------------
    float lsum[NT][2];
#pragma unroll
    for (int nt = 0; nt < NT; ++nt)
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        float l = l_run[nt][j];
        l += __shfl_xor_sync(FULL, l, 4);
        l += __shfl_xor_sync(FULL, l, 8);
        l += __shfl_xor_sync(FULL, l, 16);
        lsum[nt][j] = l;
      }
    const int seg = cs * hs + ch;
    const int lin_a = s_cum[cs] * hs + ch * cn;
    const SegInfo si = seg_info(pl, lin_a, lin_a + cn);
    const int uid = mf.x;
    const bool is_chunk = uid >= W;
    int gi = 0, cl = 0, eg = 1;                   // this chunk's group, first chunk of (group, segment), arrivals
    if (is_chunk) {
      gi = (uid - W) / kGroup;
      cl = max(gi * kGroup, si.c0);
      eg = min(gi * kGroup + kGroup - 1, si.c1) - cl + 1;
    }

    if (si.expected == 1 && eg == 1) {
      // the whole segment was streamed by this piece: normalise and write the result
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) {
        const int hq = nt * 8 + (lane & 3) * 2;      // this lane's two heads: hq, hq + 1
        const float inv0 = 1.f / lsum[nt][0], inv1 = 1.f / lsum[nt][1];
        __nv_bfloat16* o0 = obase + (int64_t)cs * p.o_stride_b + (int64_t)(ch * g + hq) * p.o_stride_h;
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
      }
      continue;
    }

    // partial: a lone chunk of its (group, segment) writes straight into the group's slot
    const bool grp_level = is_chunk && eg > 1;
    const int slot = (is_chunk && !grp_level) ? slot2_base + seg + gi : seg + uid;
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
      const int hq = nt * 8 + (lane & 3) * 2;
      float* po = p.part_o + ((int64_t)slot * g + hq) * HD;
#pragma unroll
      for (int mt = 0; mt < KSTEPS; ++mt) {
        const int d = mt * 16 + (lane >> 2);
        if (hq < g) { po[d] = acc[mt][nt][0]; po[d + 8] = acc[mt][nt][2]; }
        if (hq + 1 < g) { po[HD + d] = acc[mt][nt][1]; po[HD + d + 8] = acc[mt][nt][3]; }
      }
      if ((lane >> 2) == 0) {
        if (hq < g) p.part_ml[(int64_t)slot * g + hq] = make_float2(m_run[nt][0], lsum[nt][0]);
        if (hq + 1 < g) p.part_ml[(int64_t)slot * g + hq + 1] = make_float2(m_run[nt][1], lsum[nt][1]);
      }
    }
    __threadfence();
    __syncwarp();
    bool arrive_seg = !grp_level;
    if (grp_level) {
      int old = 0;
      if (lane == 0) old = atomicAdd(grp_cnt + seg + gi, 1);
      old = __shfl_sync(FULL, old, 0);
      if (old == eg - 1) {                           // last chunk of the group in: merge the group
        if (lane == 0) grp_cnt[seg + gi] = 0;        // self-cleaning: the header stays zero between launches
        __threadfence();
        const bool fin = si.expected == 1;
        merge_partials<HD>(p, seg + W + cl, eg, 0, 0, fin, slot2_base + seg + gi, cs, ch, lane);
        if (!fin) {
          __threadfence();
          __syncwarp();
          arrive_seg = true;
        }
      }
    }
    if (arrive_seg) {
      int old = 0;
      if (lane == 0) old = atomicAdd(seg_cnt + seg, 1);
      old = __shfl_sync(FULL, old, 0);
      if (old == si.expected - 1) {                  // every range of the segment is in: final merge
        if (lane == 0) seg_cnt[seg] = 0;
        __threadfence();
        merge_partials<HD>(p, seg + si.w0, si.ns, slot2_base + seg + si.g0, si.ng, true, 0, cs, ch, lane);
      }
    }
  }

  if (p.dbg && lane == 0) {
    unsigned long long t2;
    unsigned smid;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t2));
    asm volatile("mov.u32 %0, %smid;" : "=r"(smid));
    long long* d = p.dbg + 6 * (int64_t)worker;
    d[0] = smid; d[1] = (long long)dbg_t0; d[2] = (long long)dbg_t1; d[3] = (long long)t2;
    d[4] = dbg_static + 1; d[5] = produced;
  }
  // ---- exit: the last worker out re-arms the pool ticket for the next launch ----------------------
  if (lane == 0) {
    const int nworkers = gridDim.x * kWarps;
    if (atomicAdd(ticket + 1, 1) == nworkers - 1) {
      ticket[0] = 0;
      ticket[1] = 0;
    }
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
  const size_t smem = 1024 + (size_t)kWarps * kStages * STAGE_BYTES + kWarps * kStages * (8 + 32) + 40 * 4 +
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
int pa_decode_tma_nc_max(int num_sms) { return 8 * num_sms * kWarps; }
int pa_decode_tma_group() { return kGroup; }
int pa_decode_tma_header_ints() { return kHeaderInts; }

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
