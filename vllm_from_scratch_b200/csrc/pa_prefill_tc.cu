// Paged prefill / chunked-prefill attention on the 5th-generation tensor cores (tcgen05 + TMEM),
// bf16, block_size 16, head_dim 64 / 128, g = Hq / Hkv in {1, 2, 4, 8, 16}.
//
// Replaces the attention half of the reference prefill (01_paged_attention/paged_attention.py:
// 125-136: causal softmax(Q K^T / sqrt(D)) V with the g query heads of a group sharing one KV
// head) and the engine's per-request prefill attention (05_mini_vllm/mini_vllm.py:437-458),
// extended the way chunked prefill needs it (BLOG_PART2_V2.md:1104-1141): the query chunk of a
// sequence sits at positions [ctx - q_len, ctx) and attends to positions [0, pos] read THROUGH
// the block table, so history written by earlier chunks and the chunk's own K / V (written by
// reshape-and-cache just before) are one stream of paged tiles.  A batch may mix chunks of any
// length, including q_len = 1 (decode tokens).  This is the one dense contraction on the path.
//
// Work tile = (sequence, KV head, 128 query ROWS), a row being (token, query head of the group):
// R = 128 / g tokens x g heads, so the GQA heads that share a KV head share every K / V tile in
// shared memory and the MMA has the full M = 128.  One CTA per work tile, 6 warps:
//   warps 0-3  softmax: thread r owns row r = TMEM lane r.  Per KV tile of 128 tokens it reads the
//              S row from TMEM (tcgen05.ld 32x32b), masks (causal / ragged), runs the fp32 online
//              softmax in the exp2 domain, rescales the O accumulator IN TMEM (tcgen05.ld / st)
//              only when some row's maximum moved, and writes bf16 P into shared memory in the
//              K-major 128-byte-swizzled layout the PV MMA reads as its A operand.
//   warp 4     TMA producer: per KV tile 8 blocks x D/64 halves x {K, V} = up to 32
//              cp.async.bulk.tensor loads (one lane each) through the block table, laid out
//              slab-major [half][128 tokens][128 B] so a whole tile is ONE canonical UMMA operand;
//              2 stages each for K and V.
//   warp 5     MMA issuer (one elected lane): S[buf] = Q K^T (D/16 x tcgen05.mma M128 N128 K16,
//              K-major A and B) one tile AHEAD of the softmax (S is double-buffered in TMEM), then
//              O += P V (8 x tcgen05.mma M128 N=D K16, A = P K-major, B = V MN-major straight from
//              the [token][d] tile - no transpose anywhere).  tcgen05.commit -> mbarriers free the
//              K / V stages and publish S / O.
// TMEM: 512 columns = S0 [0,128) | S1 [128,256) | O [256, 256 + D), fp32.
// Shared memory: Q 32 KiB | P 32 KiB | K 2 x 32 KiB | V 2 x 32 KiB (D = 128) + barriers.
//
// Bound: tensor pipe (4 * D flops per (row, key) against 2 * D * 2 / (128 rows) bytes); with one
// softmax warpgroup per CTA the MUFU exp2 rate (16 / clk / SM) is co-critical, see DESIGN.md.
#include <cuda.h>
#include <math.h>
#include <stdio.h>
#include <string.h>

#include "pa_common.cuh"
#include "pa_tc_common.cuh"

int pa_get_kv_tensor_map(const void* base, const PaDecodeParams& p, int box_halves, CUtensorMap* out, int* pos_head);

namespace {

constexpr int kBS = 16;
constexpr int kTileN = 128;           // keys per KV tile (UMMA N of the S MMA)
constexpr int kBlocksPerTile = kTileN / kBS;
constexpr int kSoftmaxThreads = 128;
constexpr int kThreads = 192;
constexpr int kStages = 2;
constexpr uint32_t kSlab = kRows * 128;       // bytes of one [128 rows][64 bf16] swizzled slab
constexpr uint32_t kTmemCols = 512;
constexpr uint32_t kTmemO = 256;
constexpr uint32_t FULL = 0xffffffffu;

// Barrier slots in shared memory (8 bytes each).
enum : int {
  kBarKFull = 0, kBarKEmpty = 2, kBarVFull = 4, kBarVEmpty = 6, kBarSFull = 8, kBarSEmpty = 10,
  kBarPReady = 12, kBarPvDone = 13, kBarQReady = 14, kNumBars = 15
};

template <int HD>
__global__ void __launch_bounds__(kThreads, 1)
pa_prefill_tc_kernel(const __grid_constant__ CUtensorMap tm_k, const __grid_constant__ CUtensorMap tm_v,
                     const PaPrefillParams p) {
  constexpr int HALVES = HD / 64;
  constexpr uint32_t kOperand = HALVES * kSlab;        // Q tile, one K stage, one V stage
  constexpr uint32_t kPBytes = (kTileN / 64) * kSlab;

  // ---- which work tile ----
  const int s = blockIdx.z, h = blockIdx.y;
  const int g = p.group, R = kRows / g;
  int q_start, q_len, ctx;
  if (p.cu_q_lens) {
    q_start = __ldg(p.cu_q_lens + s);
    q_len = __ldg(p.cu_q_lens + s + 1) - q_start;
    ctx = __ldg(p.ctx_lens + s);
  } else {
    q_start = 0; q_len = p.num_tokens; ctx = p.ctx_len_single;
  }
  if (q_len <= 0 || ctx < q_len) return;               // nothing to do / malformed: leave `out` alone
  if (p.max_blocks > 0 && pa_ceil_div(ctx, kBS) > p.max_blocks) return;   // the table row is too short: never read past it
  const int q_tile = (int)gridDim.x - 1 - (int)blockIdx.x;      // long tiles first
  if (q_tile * R >= q_len) return;
  const int tok0 = q_tile * R;
  const int ntok = min(R, q_len - tok0);
  const int pos0 = ctx - q_len + tok0;                 // smallest query position of the tile
  const int maxpos = pos0 + ntok - 1;
  const int n_tiles = maxpos / kTileN + 1;
  const int nblk_need = maxpos / kBS + 1;

  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t sQ = base;
  const uint32_t sP = sQ + kOperand;
  const uint32_t sK = sP + kPBytes;
  const uint32_t sV = sK + kStages * kOperand;
  const uint32_t sBar = sV + kStages * kOperand;
  const uint32_t sTmem = sBar + kNumBars * 8;
  uint8_t* gen = smem_raw + (base - smem_u32(smem_raw));         // generic pointer to `base`
  auto bar = [&](int i) { return sBar + 8u * i; };

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

  if (warp == 4 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tm_k) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tm_v) : "memory");
    for (int i = 0; i < kNumBars; ++i) {
      const bool full_count = (i == kBarSEmpty || i == kBarSEmpty + 1 || i == kBarPReady || i == kBarQReady);
      mbar_init(bar(i), full_count ? kSoftmaxThreads : 1);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 5) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(sTmem), "n"(kTmemCols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *reinterpret_cast<volatile uint32_t*>(gen + (sTmem - base));

  if (warp < 4) {
    // =============================== softmax warps ===============================
    const int r = tid;                                   // row = TMEM lane
    // ---- Q tile -> shared memory, K-major, 128-byte swizzle (row = (token, head-in-group)) ----
    {
      const __nv_bfloat16* qb = static_cast<const __nv_bfloat16*>(p.q);
      constexpr int CH = HD / 8;                         // 16-byte chunks per row
      for (int idx = tid; idx < kRows * CH; idx += kSoftmaxThreads) {
        const int row = idx / CH, c = idx % CH;
        const int ti = row / g, j = row - ti * g;
        uint4 val = make_uint4(0, 0, 0, 0);
        if (ti < ntok)
          val = __ldg(reinterpret_cast<const uint4*>(qb + (int64_t)(q_start + tok0 + ti) * p.q_stride_t +
                                                     (int64_t)(h * g + j) * p.q_stride_h + c * 8));
        const uint32_t off = (uint32_t)(c >> 3) * kSlab + (uint32_t)row * 128u + (uint32_t)(((c & 7) ^ (row & 7)) << 4);
        *reinterpret_cast<uint4*>(gen + (sQ - base) + off) = val;
      }
      fence_proxy_async();
      mbar_arrive(bar(kBarQReady));
    }
    const int ti = r / g, j = r - ti * g;
    const bool valid = ti < ntok;
    const int qpos = pos0 + min(ti, ntok - 1);
    const float scale = p.scale_log2e;
    const uint32_t lane_base = tmem + ((uint32_t)(warp * 32) << 16);
    float m = -INFINITY, l = 0.f;

    for (int t = 0; t < n_tiles; ++t) {
      const int sb = t & 1;
      uint32_t v[kTileN];
      mbar_wait(bar(kBarSFull + sb), (t >> 1) & 1);
      tc_fence_after();
#pragma unroll
      for (int cc = 0; cc < kTileN / 32; ++cc) tmem_ld32(lane_base + sb * kTileN + cc * 32, v + cc * 32);
      tmem_wait_ld();
      tc_fence_before();
      mbar_arrive(bar(kBarSEmpty + sb));                 // S[sb] may be overwritten by tile t + 2

      const int kbase = t * kTileN;
      if (kbase + kTileN - 1 > pos0) {                   // causal / ragged mask (CTA-uniform test)
#pragma unroll
        for (int c = 0; c < kTileN; ++c)
          if (kbase + c > qpos) v[c] = 0xff800000u;      // -inf, also replaces NaN from stale keys
      }
      float mx = -INFINITY;
#pragma unroll
      for (int c = 0; c < kTileN; ++c) mx = fmaxf(mx, __uint_as_float(v[c]));
      const float m_new = fmaxf(m, mx * scale);          // finite from tile 0 on: key 0 is visible
      const float alpha = fast_exp2(m - m_new);
      uint32_t pk[kTileN / 2];
      float sum = 0.f;
#pragma unroll
      for (int c = 0; c < kTileN / 2; ++c) {
        const float p0 = fast_exp2(fmaf(__uint_as_float(v[2 * c]), scale, -m_new));
        const float p1 = fast_exp2(fmaf(__uint_as_float(v[2 * c + 1]), scale, -m_new));
        const uint32_t u = pack_bf16(p0, p1);
        sum += bf16lo(u) + bf16hi(u);                    // normaliser of the weights the MMA sees
        pk[c] = u;
      }
      l = l * alpha + sum;

      if (t > 0) {
        mbar_wait(bar(kBarPvDone), (t - 1) & 1);         // O holds tiles < t; P buffer is free
        tc_fence_after();
        if (__any_sync(FULL, m_new > m)) {
#pragma unroll
          for (int cc = 0; cc < HD / 32; ++cc) {
            uint32_t o[32];
            tmem_ld32(lane_base + kTmemO + cc * 32, o);
            tmem_wait_ld();
#pragma unroll
            for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
            tmem_st32(lane_base + kTmemO + cc * 32, o);
          }
          tmem_wait_st();
        }
      }
      m = m_new;
      // ---- P row -> shared memory (A operand of the PV MMA) ----
#pragma unroll
      for (int c = 0; c < kTileN / 8; ++c) {
        const uint32_t off = (uint32_t)(c >> 3) * kSlab + (uint32_t)r * 128u + (uint32_t)(((c & 7) ^ (r & 7)) << 4);
        *reinterpret_cast<uint4*>(gen + (sP - base) + off) = make_uint4(pk[4 * c], pk[4 * c + 1], pk[4 * c + 2], pk[4 * c + 3]);
      }
      // ---- last tile: V rows that were not loaded or lie beyond the context must be finite ----
      if (t == n_tiles - 1) {
        const int zero_from = min(nblk_need * kBS, ctx) - kbase;
        if (zero_from < kTileN) {
          mbar_wait(bar(kBarVFull + (t & 1)), (t >> 1) & 1);
          const int nz = (kTileN - zero_from) * HALVES * 8;
          uint8_t* vb = gen + (sV - base) + (uint32_t)(t & 1) * kOperand;
          for (int idx = tid; idx < nz; idx += kSoftmaxThreads) {
            const int row = zero_from + idx / (HALVES * 8), rem = idx % (HALVES * 8);
            *reinterpret_cast<uint4*>(vb + (uint32_t)(rem >> 3) * kSlab + (uint32_t)row * 128u + (uint32_t)((rem & 7) << 4)) =
                make_uint4(0, 0, 0, 0);
          }
        }
      }
      fence_proxy_async();
      tc_fence_before();
      mbar_arrive(bar(kBarPReady));
    }

    // ---- epilogue: O / l -> out ----
    mbar_wait(bar(kBarPvDone), (n_tiles - 1) & 1);
    tc_fence_after();
    const float inv = 1.f / l;
    __nv_bfloat16* ob = static_cast<__nv_bfloat16*>(p.out) + (int64_t)(q_start + tok0 + ti) * p.o_stride_t +
                        (int64_t)(h * g + j) * p.o_stride_h;
#pragma unroll
    for (int cc = 0; cc < HD / 32; ++cc) {
      uint32_t o[32];
      tmem_ld32(lane_base + kTmemO + cc * 32, o);
      tmem_wait_ld();
      if (valid) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          uint4 w;
          w.x = pack_bf16(__uint_as_float(o[8 * i]) * inv, __uint_as_float(o[8 * i + 1]) * inv);
          w.y = pack_bf16(__uint_as_float(o[8 * i + 2]) * inv, __uint_as_float(o[8 * i + 3]) * inv);
          w.z = pack_bf16(__uint_as_float(o[8 * i + 4]) * inv, __uint_as_float(o[8 * i + 5]) * inv);
          w.w = pack_bf16(__uint_as_float(o[8 * i + 6]) * inv, __uint_as_float(o[8 * i + 7]) * inv);
          *reinterpret_cast<uint4*>(ob + cc * 32 + i * 8) = w;
        }
      }
    }
    tc_fence_before();
  } else if (warp == 4) {
    // =============================== TMA producer ===============================
    const int32_t* bt = p.block_tables + (int64_t)s * p.bt_stride;
    const int b = lane / HALVES, hf = lane % HALVES;
    for (int t = 0; t < n_tiles; ++t) {
      const int st = t & 1;
      const uint32_t par = ((t >> 1) & 1) ^ 1;
      const int blk0 = t * kBlocksPerTile;
      const int nb = min(kBlocksPerTile, nblk_need - blk0);
      int blk = 0;
      const bool mine = lane < kBlocksPerTile * HALVES && b < nb;
      if (mine) blk = __ldg(bt + blk0 + b);
      const int c3 = p.pos_head == 3 ? h : blk, c4 = p.pos_head == 3 ? blk : h;
      const uint32_t dst_off = (uint32_t)st * kOperand + (uint32_t)hf * kSlab + (uint32_t)b * (kBS * 128);
      mbar_wait(bar(kBarKEmpty + st), par);
      if (lane == 0) mbar_expect_tx(bar(kBarKFull + st), (uint32_t)nb * HALVES * kBS * 128);
      __syncwarp();
      if (mine) tma_load_half(sK + dst_off, &tm_k, bar(kBarKFull + st), hf, c3, c4);
      mbar_wait(bar(kBarVEmpty + st), par);
      if (lane == 0) mbar_expect_tx(bar(kBarVFull + st), (uint32_t)nb * HALVES * kBS * 128);
      __syncwarp();
      if (mine) tma_load_half(sV + dst_off, &tm_v, bar(kBarVFull + st), hf, c3, c4);
    }
  } else {
    // =============================== MMA issuer ===============================
    constexpr uint32_t idesc_s = umma_idesc(kTileN, false);
    constexpr uint32_t idesc_o = umma_idesc(HD, true);
    auto issue_qk = [&](int t) {
      const int st = t & 1, sb = t & 1;
      mbar_wait(bar(kBarKFull + st), (t >> 1) & 1);
      mbar_wait(bar(kBarSEmpty + sb), ((t >> 1) & 1) ^ 1);
      tc_fence_after();
      if (lane == 0) {
#pragma unroll
        for (int kk = 0; kk < HD / 16; ++kk) {
          const uint32_t koff = (uint32_t)(kk >> 2) * kSlab + (uint32_t)(kk & 3) * 32u;
          umma_bf16(tmem + sb * kTileN, umma_desc(sQ + koff, 16, 1024),
                    umma_desc(sK + (uint32_t)st * kOperand + koff, 16, 1024), idesc_s, kk > 0);
        }
        umma_commit(bar(kBarKEmpty + st));
        umma_commit(bar(kBarSFull + sb));
      }
      __syncwarp();
    };
    auto issue_pv = [&](int t) {
      const int st = t & 1;
      mbar_wait(bar(kBarVFull + st), (t >> 1) & 1);
      mbar_wait(bar(kBarPReady), t & 1);
      tc_fence_after();
      if (lane == 0) {
#pragma unroll
        for (int ks = 0; ks < kTileN / 16; ++ks) {
          const uint64_t da = umma_desc(sP + (uint32_t)(ks >> 2) * kSlab + (uint32_t)(ks & 3) * 32u, 16, 1024);
          const uint64_t db = umma_desc(sV + (uint32_t)st * kOperand + (uint32_t)ks * (16 * 128), kSlab, 1024);
          umma_bf16(tmem + kTmemO, da, db, idesc_o, (t > 0 || ks > 0) ? 1u : 0u);
        }
        umma_commit(bar(kBarVEmpty + st));
        umma_commit(bar(kBarPvDone));
      }
      __syncwarp();
    };
    mbar_wait(bar(kBarQReady), 0);
    issue_qk(0);
    for (int t = 0; t < n_tiles; ++t) {
      if (t + 1 < n_tiles) issue_qk(t + 1);
      issue_pv(t);
    }
    tc_fence_before();
  }

  __syncthreads();
  if (warp == 5) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(kTmemCols) : "memory");
  }
}

template <int HD>
constexpr size_t smem_bytes() {
  return 1024 + (size_t)(HD / 64) * kSlab * (1 + 2 * kStages) + (kTileN / 64) * kSlab + kNumBars * 8 + 16;
}

template <int HD>
int launch(const PaPrefillParams& p, const CUtensorMap& mk, const CUtensorMap& mv, cudaStream_t stream) {
  auto kern = pa_prefill_tc_kernel<HD>;
  static PaPerDeviceFlag attr;
  const int dev = attr.need();
  if (dev != -2) {
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes<HD>()) != cudaSuccess)
      return -5;
    attr.done(dev);
  }
  const int R = kRows / p.group;
  dim3 grid((p.max_q_len + R - 1) / R, p.num_kv_heads, p.batch);
  kern<<<grid, kThreads, smem_bytes<HD>(), stream>>>(mk, mv, p);
  return cudaGetLastError() == cudaSuccess ? 0 : -5;
}

}  // namespace

bool pa_prefill_tc_supported(const PaPrefillParams& p, int dtype, int block_size) {
  if (dtype != 0 || block_size != kBS) return false;
  if (p.head_dim != 64 && p.head_dim != 128) return false;
  const int g = p.group;
  if (g < 1 || g > 16 || (g & (g - 1))) return false;
  if (p.batch < 1 || p.batch > 65535 || p.num_kv_heads > 65535 || p.max_q_len < 1) return false;
  if ((reinterpret_cast<uintptr_t>(p.k) | reinterpret_cast<uintptr_t>(p.v) | reinterpret_cast<uintptr_t>(p.q) |
       reinterpret_cast<uintptr_t>(p.out)) & 15)
    return false;
  if ((p.kv_stride_block | p.kv_stride_head | p.kv_stride_slot) & 7) return false;
  if (p.kv_stride_block <= 0 || p.kv_stride_head <= 0 || p.kv_stride_slot <= 0) return false;
  if ((p.q_stride_t | p.q_stride_h | p.o_stride_t | p.o_stride_h) & 7) return false;   // 16-byte row vectors
  return true;
}

int pa_launch_prefill_tc(const PaPrefillParams& p_in, cudaStream_t stream) {
  PaPrefillParams p = p_in;
  PaDecodeParams kv;
  memset(&kv, 0, sizeof(kv));
  kv.k = p.k; kv.v = p.v;
  kv.kv_stride_block = p.kv_stride_block; kv.kv_stride_head = p.kv_stride_head; kv.kv_stride_slot = p.kv_stride_slot;
  kv.num_kv_heads = p.num_kv_heads; kv.head_dim = p.head_dim; kv.num_blocks = p.num_blocks;
  CUtensorMap mk, mv;
  int ph = 3;
  int rc = pa_get_kv_tensor_map(p.k, kv, 1, &mk, &ph);
  if (rc) return rc;
  rc = pa_get_kv_tensor_map(p.v, kv, 1, &mv, &ph);
  if (rc) return rc;
  p.pos_head = ph;
  return p.head_dim == 128 ? launch<128>(p, mk, mv, stream) : launch<64>(p, mk, mv, stream);
}
