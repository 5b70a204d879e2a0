// Paged prefill / chunked-prefill attention, tcgen05 + TMEM, ping-pong version: TWO 128-row query
// tiles per CTA, each with its own softmax warpgroup, sharing every K / V tile; P stays in TMEM.
//
// Same contract as pa_prefill_tc.cu (reference: 01_paged_attention/paged_attention.py:125-136,
// 05_mini_vllm/mini_vllm.py:437-458; varlen batches of query chunks that attend causally to
// their whole paged history).  Why this shape (measurements: tools/umma_bench.cu and
// profiles/r01_prefill_*):
//   * a tcgen05.mma whose A operand is in SHARED memory costs >= 90.6 cycles on B200 whatever N is
//     (the 128 x 16 A slice is re-read every instruction); the math floor is 128 * N / 256 cycles,
//     so SS MMAs run at 70 % of peak for N = 128 and 35 % for N = 64.  Hence 128-key tiles for
//     S = Q K^T (SS, 90 cycles / instruction), and O += P V with P as the A operand IN TMEM (TS,
//     64 cycles / instruction at N = D = 128): 8 x 90 + 8 x 64 cycles per (tile, 128 keys)
//     against 8 x 64 + 8 x 64 of pure math - an 83 % ceiling.
//   * one softmax warpgroup cannot keep the pipe fed (pa_prefill_tc.cu: tensor pipe 29 % active):
//     a CTA owns rows [0,128) = tile A and [128,256) = tile B of one (sequence, KV head)
//     (2 * R tokens x g heads, R = 128 / g); warps 0-3 are A's softmax warpgroup, 4-7 B's
//     (warp % 4 = TMEM lane quarter), warp 8 the TMA producer, warp 9 the MMA issuer.  While A's
//     warpgroup exponentiates, the pipe runs B's P V and the next Q_B K^T, and vice versa.
//   * bf16 P overwrites the first 64 columns of its own S tile (tcgen05.st), so S is single-
//     buffered: the issuer queues PV_w(t) and then QK_w(t + 1), which the pipe executes in order.
//   * K / V tiles are 128 keys x D (32 KiB) in a 5-slot ring (K0 V0 K1 V1 ...), gathered by TMA
//     through the block table, slab-major so a whole tile is one canonical UMMA operand.
//   * softmax per thread = one row of 128 scores: one tcgen05.ld pass, FMNMX3 row max, FFMA2 /
//     ex2 / FADD2 (packed fp32x2), bf16 pack.  The O accumulator is rescaled only when a row's
//     maximum grew by more than 2^8 over the maximum in use (exact: l and O share that reference).
//   * persistent (round 2): one CTA per SM; warp 10 draws work tiles (2 * R tokens of one (sequence, KV head),
//     longest first) from a per-launch ticket, as late as it can, and publishes each tile's descriptor through
//     shared memory to the other roles.  Barriers, TMEM and the register split are set up once per CTA, the
//     K / V ring never drains across tiles, and a warpgroup's next Q tile is staged with cp.async under its
//     last softmax pass.  Per work tile the fixed cost went from 6.2 us + 3.5 us between CTAs to ~4.7 us + 0.6.
// TMEM (512 columns): S_A / P_A [0,128) | S_B / P_B [128,256) | O_A [256,384) | O_B [384,512).
// Shared memory (D = 128): Q_A, Q_B 32 KiB each | ring 5 x 32 KiB = 224 KiB.
#include <atomic>
#include <cuda.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "pa_common.cuh"
#include "pa_tc_common.cuh"

int pa_get_kv_tensor_map(const void* base, const PaDecodeParams& p, int box_halves, CUtensorMap* out, int* pos_head);

namespace {

constexpr int kBS = 16;
constexpr int kTileN = 128;                   // keys per KV tile
constexpr int kBlocksPerTile = kTileN / kBS;  // 8
constexpr int kWgThreads = 128;
constexpr int kThreads = 384;                 // 2 softmax warpgroups + a third warpgroup: producer warp, MMA warp, 2 idle
constexpr int kRing = 5;                      // K / V items in flight
constexpr uint32_t kSlab = kRows * 128;       // [128 rows][128 B]: one 64-column slab of a Q, K or V tile
constexpr uint32_t kTmemCols = 512;
constexpr uint32_t kTmemO = 256;
constexpr float kRescaleThreshold = 8.f;      // log2 units
constexpr uint32_t FULL = 0xffffffffu;

// Work tiles published ahead of the slowest reader: ONE, and drawn as late as possible (when the producer has
// two K / V tiles of the current work tile left to request): list scheduling is only as good as the moment a
// ticket is drawn.  Measured on full8k: ring of 4 filled at start-up 998 TFLOP/s (every CTA draws its tickets in
// the first microsecond - the list is dealt out before any work is done); ring of 1 refilled as soon as it is
// read 1083 (a CTA that starts a 45-tile unit has already committed to the one after it); a static
// boustrophedon deal 1121.
constexpr int kTileRing = 2;                  // entries: the tile being worked on and the next (an entry is released when its tile is DONE)
constexpr int kDrawAhead = 2;                 // K / V tiles of the current work tile still to be requested when the next ticket is drawn

// Barrier slots (8 bytes each).  Per-warpgroup groups are indexed [wg].
enum : int {
  kBarKvFull = 0, kBarKvEmpty = kRing,
  kBarSFull = 2 * kRing,            // [wg]  tcgen05.commit after Q_w K(t)^T (implies every earlier MMA is done)
  kBarPReady = kBarSFull + 2,       // [wg][2]  128 softmax threads: keys [0,64) / [64,128) of P(t) are in TMEM (and O is rescaled)
  kBarODone = kBarPReady + 4,       // [wg]  tcgen05.commit after the warpgroup's last P V
  kBarQReady = kBarODone + 2,       // [wg]  128 softmax threads: the Q tile is in shared memory
  kBarTFull = kBarQReady + 2,       // [kTileRing]  the scheduler warp published a work tile
  kBarTEmpty = kBarTFull + kTileRing,   // [kTileRing]  every reader (2 x 128 softmax threads, producer, issuer) has taken it
  kBarTReq = kBarTEmpty + kTileRing,    // the producer is about to run out of K / V tiles to request: draw the next work tile
  kNumBars = kBarTReq + 1
};
constexpr int kTileReaders = 2 * kWgThreads + 2;

// Work-tile tickets of the launches in flight: [slot][0] = next work tile of the launch, [1] = CTAs that have run
// out of tiles.  Zero when the module is loaded; the last CTA of a launch to run out puts its slot back to zero.
// A launch takes slot (launch number % kTicketSlots): two launches share a slot only if kTicketSlots launches
// apart AND running at the same time.
constexpr int kTicketSlots = 1024;
__device__ int g_tickets[kTicketSlots * 2];

// D[tmem] (+)= A[tmem] . B[smem]
__device__ __forceinline__ void umma_bf16_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t desc_b, uint32_t idesc,
                                             uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
      ::"r"(tmem_d), "r"(tmem_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}

__device__ __forceinline__ uint64_t pack2(float lo, float hi) {
  uint64_t r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void unpack2(uint64_t v, float& lo, float& hi) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ uint64_t ffma2(uint64_t a, uint64_t b, uint64_t c) {
  uint64_t r;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
  return r;
}
__device__ __forceinline__ uint64_t fadd2(uint64_t a, uint64_t b) {
  uint64_t r;
  asm("add.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
// exp2 of two values on the FMA / ALU pipes instead of the SFU: round-to-nearest split x = n + f
// (magic-number add), 2^f by a degree-3 minimax polynomial on [-0.5, 0.5] (max relative error 7.6e-5,
// far below the bf16 rounding P gets), 2^n by adding n to the exponent bits.  The softmax warpgroups
// are bound by the 4-lane-per-clock SFU (one ex2 per score); sending a share of the scores through
// this path runs both pipes side by side.
__device__ __forceinline__ void poly_exp2x2(uint64_t x, float& p0, float& p1) {
  float x0, x1;
  unpack2(x, x0, x1);
  x = pack2(fmaxf(x0, -126.f), fmaxf(x1, -126.f));           // masked scores are -inf; also keeps 2^n normal
  const float magic = 12582912.f;                            // 1.5 * 2^23: the low mantissa bits of x + magic hold n
  const uint64_t t = fadd2(x, pack2(magic, magic));
  const uint64_t n = fadd2(t, pack2(-magic, -magic));
  const uint64_t f = ffma2(n, pack2(-1.f, -1.f), x);
  uint64_t r = ffma2(f, pack2(0.05520550534f, 0.05520550534f), pack2(0.24261397123f, 0.24261397123f));
  r = ffma2(r, f, pack2(0.69325476885f, 0.69325476885f));
  r = ffma2(r, f, pack2(0.99992769957f, 0.99992769957f));
  float r0, r1, t0, t1;
  unpack2(r, r0, r1);
  unpack2(t, t0, t1);
  p0 = __uint_as_float(__float_as_uint(r0) + (__float_as_uint(t0) << 23));
  p1 = __uint_as_float(__float_as_uint(r1) + (__float_as_uint(t1) << 23));
}
constexpr uint32_t kPolyExpMask = 0x22u;      // score pairs with bit (c & 7) set take the polynomial path: 2 of 8 (measured: 0 -> 1025, 2/8 -> 1052, 3/8 -> 1051, 4/8 -> 1009 TFLOP/s on full8k)

__device__ __forceinline__ float fmax3(float a, float b, float c) {
  float r;
  asm("max.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c));
  return r;
}


constexpr int kTraceTiles = 48, kTraceSlots = 16;
#define PA_TRACE(slot, t)                                                                         \
  do {                                                                                            \
    if (trace_on && (t) < kTraceTiles) p.trace[(t) * kTraceSlots + (slot)] = clock64();           \
  } while (0)

// One unit of work: 2 * R tokens (tile A = the first R, tile B = the rest) of one (sequence, KV head).
struct __align__(16) WorkTile {                              // 16 ints: published through shared memory as four 16-byte words
  int s, h, q_start, tok0;
  int ntok_a, ntok_b, pos0, n_a;
  int n_b, n, nblk_need, ctx;
  int slot;                                                  // index in the debug stamp buffer (the old 3-D grid's linear block index)
  int item0;                                                 // K / V ring items the CTA streamed before this tile
  int L;                                                     // index in the launch's list; -1 = the CTA has no more tiles
  int pad;
};
static_assert(sizeof(WorkTile) == 64, "four 16-byte words");
// Work tile number L of the launch, longest tiles first: the query tile index is the slowest axis (a later
// query tile attends to more keys), (sequence, KV head) the fastest.  false = nothing to do there.
__device__ __forceinline__ bool work_tile(const PaPrefillParams& p, int L, int gx, int R, WorkTile& T) {
  const int per_x = p.num_kv_heads * p.batch;
  const int x = L / per_x, rem = L - x * per_x;
  T.s = rem / p.num_kv_heads;
  T.h = rem - T.s * p.num_kv_heads;
  T.slot = x + gx * (T.h + p.num_kv_heads * T.s);
  int q_len;
  if (p.cu_q_lens) {
    T.q_start = __ldg(p.cu_q_lens + T.s);
    q_len = __ldg(p.cu_q_lens + T.s + 1) - T.q_start;
    T.ctx = __ldg(p.ctx_lens + T.s);
  } else {
    T.q_start = 0; q_len = p.num_tokens; T.ctx = p.ctx_len_single;
  }
  if (q_len <= 0 || T.ctx < q_len) return false;             // nothing to do / malformed: leave `out` alone
  T.tok0 = (gx - 1 - x) * 2 * R;
  if (T.tok0 >= q_len) return false;
  const int ntok = min(2 * R, q_len - T.tok0);
  T.ntok_a = min(R, ntok);
  T.ntok_b = ntok - T.ntok_a;
  T.pos0 = T.ctx - q_len + T.tok0;                           // smallest query position of the tile
  T.n_a = (T.pos0 + T.ntok_a - 1) / kTileN + 1;
  T.n_b = T.ntok_b > 0 ? (T.pos0 + ntok - 1) / kTileN + 1 : 0;
  T.n = max(T.n_a, T.n_b);                                   // KV tiles streamed
  T.nblk_need = (T.pos0 + ntok - 1) / kBS + 1;
  if (p.max_blocks > 0 && pa_ceil_div(T.ctx, kBS) > p.max_blocks) return false;   // the table row is too short: never read past it
  return true;
}

// Persistent: one CTA per SM.  Warp 10 draws work tiles from the launch's ticket (the list is sorted longest first,
// so this is longest-processing-time-first list scheduling), drops the empty ones and publishes the rest through a
// small shared-memory ring that the four other roles read in the same order.  Barriers, TMEM and the register
// split are set up once; barrier phases run on across tiles (each role counts its own waits), the K / V ring never
// drains: the producer is already fetching the next tile's keys while the softmax warpgroups write the current
// tile's output, and a warpgroup stages its next Q tile before it waits for its last P V.
template <int HD>
__global__ void __launch_bounds__(kThreads, 1)
pa_prefill_tc2_kernel(const __grid_constant__ CUtensorMap tm_k, const __grid_constant__ CUtensorMap tm_v,
                      const PaPrefillParams p, const int gx, const int total_tiles, const int ticket_slot) {
  constexpr int HALVES = HD / 64;
  constexpr uint32_t kItem = HALVES * kSlab;                 // one Q, K or V tile: [half][128 rows][128 B]

  const int g = p.group, R = kRows / g;
  const int cta = (int)blockIdx.x;

  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t sQ = base;                                  // [2] tiles
  const uint32_t sRing = sQ + 2 * kItem;
  const uint32_t sBar = sRing + kRing * kItem;
  const uint32_t sTmem = sBar + kNumBars * 8;
  const uint32_t sTiles = (sTmem + 16 + 15u) & ~15u;         // [kTileRing] published work tiles, 64 bytes each
  uint8_t* gen = smem_raw + (base - smem_u32(smem_raw));     // generic pointer to `base`
  auto bar = [&](int i) { return sBar + 8u * i; };
  // The cnt-th work tile of this CTA, everything a role needs to know about it, computed ONCE by the scheduler
  // warp: every reader takes the entries in order (acquire), reads what it needs when it needs it (the copy costs a
  // few shared-memory loads - no global loads, no divisions) and lets go of the entry when it is DONE with the tile.
  auto acquire_tile = [&](int& cnt) {
    const int e = cnt % kTileRing;
    mbar_wait(bar(kBarTFull + e), (cnt / kTileRing) & 1);
    ++cnt;
    return e;
  };
  auto read_tile = [&](int e, WorkTile& T) {
    int4* d = reinterpret_cast<int4*>(&T);
#pragma unroll
    for (int i = 0; i < 4; ++i)
      asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(d[i].x), "=r"(d[i].y), "=r"(d[i].z), "=r"(d[i].w)
                   : "r"(sTiles + 64u * e + 16u * i) : "memory");
  };
  auto release_tile = [&](int e) { mbar_arrive(bar(kBarTEmpty + e)); };

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const bool trace_cta = p.trace != nullptr && cta == 0;     // stamps of this CTA's FIRST work tile only

  // debug: per-work-tile wall-clock stamps (thread 0): [0] smid [1] entry [2] prologue done [3] Q stored
  // [4] first S ready [5] O complete [6] outputs stored [7] KV tiles streamed
  auto stamp_at = [&](int slot_index, int slot) {
    if (p.cta_times != nullptr && (tid & 127) == 0) {
      unsigned long long t;
      asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
      p.cta_times[8 * (int64_t)slot_index + slot] = (long long)t;
    }
  };

  if (warp == 8 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tm_k) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tm_v) : "memory");
    for (int i = 0; i < kNumBars; ++i) {
      const bool by_warpgroup = (i >= kBarPReady && i < kBarODone) || (i >= kBarQReady && i < kBarTFull);
      mbar_init(bar(i), (i >= kBarTEmpty && i < kBarTReq) ? kTileReaders : by_warpgroup ? kWgThreads : 1);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 9) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(sTmem), "n"(kTmemCols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *reinterpret_cast<volatile uint32_t*>(gen + (sTmem - base));

  if (warp < 8) {
    // =============================== softmax warpgroups ===============================
    asm volatile("setmaxnreg.inc.sync.aligned.u32 232;");    // from 168: the 128 x 128 registers the third warpgroup gives up cover 64 x 256
    const int w = warp >> 2;                                 // 0 = tile A, 1 = tile B
    const int wt = tid & (kWgThreads - 1);                   // row within the tile = TMEM lane
    const float scale = p.scale_log2e;
    const uint64_t scale2 = pack2(scale, scale);
    const uint32_t lane_base = tmem + ((uint32_t)((warp & 3) * 32) << 16);
    const uint32_t tS = lane_base + (uint32_t)w * 128u, tO = lane_base + kTmemO + (uint32_t)w * 128u;
    const int ti = wt / g, j = wt - ti * g;

    // The warpgroup's own walk over the CTA's work tiles: `item0` = ring items (K / V tiles) the CTA streamed
    // before the current work tile - every valid tile counts, also those this warpgroup has no rows in.
    // Only the tile's ring entry lives across the KV loop; whatever a phase needs it reads from the entry again -
    // 232 registers hold a row of 128 scores and little else.
    int t_cnt = 0;
    int e_cur = -1;
    // The CTA's next work tile: 1 = it has rows for this warpgroup, 0 = it has none (tile B of a short tail), -1 = no
    // more tiles.  Reads exactly one entry of the ring: the entry after a tile is published while that tile is still
    // being worked on, the one after THAT only once the tile is nearly done - a warpgroup in the middle of a tile
    // must never wait for it (the issuer would wait for the warpgroup, the producer for the issuer).
    auto advance_one = [&](int& into_e) -> int {
      const int e = acquire_tile(t_cnt);
      WorkTile c;
      read_tile(e, c);
      if (c.L < 0 || (w == 0 ? c.ntok_a : c.ntok_b) == 0) {
        release_tile(e);
        return c.L < 0 ? -1 : 0;
      }
      into_e = e;
      return 1;
    };
    auto advance = [&](int& into_e) -> bool {                       // next work tile with rows for this warpgroup
      for (;;) {
        const int r = advance_one(into_e);
        if (r != 0) return r > 0;
      }
    };
    // ---- Q tile -> shared memory, K-major, 128-byte swizzle (row = (token, head-in-group)) ----
    // Asynchronous copies (cp.async, 16 bytes each, rows past the tile's tokens zero-filled): nothing passes through
    // registers, so the copies of the NEXT work tile's Q are issued at the top of the current tile's last softmax
    // pass - every Q K^T of the tile has completed by then - and their global round trip is over when it ends.
    auto stage_q_issue = [&](const int e) {
      WorkTile c;
      read_tile(e, c);
      const int ntok_w = w == 0 ? c.ntok_a : c.ntok_b;
      const int tokw0 = c.tok0 + w * R;
      const __nv_bfloat16* qb = static_cast<const __nv_bfloat16*>(p.q);
      constexpr int CH = HD / 8;                             // 16-byte chunks per row
      const uint32_t qdst = sQ + (uint32_t)w * kItem;
      constexpr int ITERS = kRows * CH / kWgThreads;         // 16 (D 128) / 8 (D 64)
#pragma unroll
      for (int it = 0; it < ITERS; ++it) {
        const int idx = wt + it * kWgThreads;
        const int row = idx / CH, c8 = idx % CH;
        const int rti = row / g, rj = row - rti * g;
        const bool live = rti < ntok_w;
        const __nv_bfloat16* src = qb + (int64_t)(c.q_start + tokw0 + (live ? rti : 0)) * p.q_stride_t +
                                   (int64_t)(c.h * g + rj) * p.q_stride_h + c8 * 8;
        const uint32_t off = (uint32_t)(c8 >> 3) * kSlab + (uint32_t)row * 128u + (uint32_t)(((c8 & 7) ^ (row & 7)) << 4);
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(qdst + off), "l"(src), "r"(live ? 16 : 0) : "memory");
      }
      asm volatile("cp.async.commit_group;" ::: "memory");
      return c.slot;
    };
    auto stage_q_commit = [&](const int slot_of) {
      asm volatile("cp.async.wait_group 0;" ::: "memory");
      fence_proxy_async();
      mbar_arrive(bar(kBarQReady + w));
      if (w == 0) stamp_at(slot_of, 3);
    };

    int s_cnt = 0, o_cnt = 0;                                // S tiles / work tiles this warpgroup has waited for so far
    bool have = advance(e_cur);
    if (have) stage_q_commit(stage_q_issue(e_cur));
    bool first_tile = true;
    while (have) {
      int n_w, n, posw0, qpos, slot, item0;
      bool valid;
      {
        WorkTile T;
        read_tile(e_cur, T);
        item0 = T.item0;
        const int ntok_w = w == 0 ? T.ntok_a : T.ntok_b;
        n_w = w == 0 ? T.n_a : T.n_b;
        n = T.n;
        posw0 = T.pos0 + w * R;
        valid = ti < ntok_w;
        qpos = posw0 + min(ti, ntok_w - 1);
        slot = T.slot;
        if (w == 0 && p.cta_times != nullptr && wt == 0) {
          unsigned smid;
          asm volatile("mov.u32 %0, %smid;" : "=r"(smid));
          p.cta_times[8 * (int64_t)slot] = smid;
          p.cta_times[8 * (int64_t)slot + 7] = n;
        }
        if (w == 0) { stamp_at(slot, 1); stamp_at(slot, 2); }
      }
      const bool trace_on = trace_cta && first_tile && (tid & 127) == 0;
      float m_used = -INFINITY, l = 0.f;
      int e_n = -1, slot_n = 0, next_state = 0;

      for (int t = 0; t < n_w; ++t) {
        uint32_t v[kTileN];
        // S(t) complete; the commit behind this barrier also covers P V(t - 1): O holds tiles < t
        mbar_wait(bar(kBarSFull + w), s_cnt & 1);
        ++s_cnt;
        if (t == 0 && w == 0) stamp_at(slot, 4);
        PA_TRACE(5 * w + 0, t);
        tc_fence_after();
        if (t == n_w - 1) {                                  // the tile's last S: its Q tile is not read any more
          next_state = advance_one(e_n);
          if (next_state > 0) slot_n = stage_q_issue(e_n);
        }
        // the scores arrive in two halves: the row maximum of keys [0, 64) is taken while keys [64, 128)
        // are still on their way from TMEM (tiles that need the causal / ragged mask wait for both first)
        tmem_ld32(tS, v);
        tmem_ld32(tS + 32, v + 32);
        tmem_wait_ld();
        tmem_ld32(tS + 64, v + 64);
        tmem_ld32(tS + 96, v + 96);
        PA_TRACE(5 * w + 1, t);

        const int kbase = t * kTileN;
        const bool masked = kbase + kTileN - 1 > posw0;      // warpgroup-uniform
        float mx4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};     // 4 independent FMNMX3 chains
        if (!masked) {
#pragma unroll
          for (int c = 0; c < kTileN / 2; c += 8) {
#pragma unroll
            for (int i = 0; i < 4; ++i) mx4[i] = fmax3(mx4[i], __uint_as_float(v[c + 2 * i]), __uint_as_float(v[c + 2 * i + 1]));
          }
        }
        tmem_wait_ld();
        if (masked) {
#pragma unroll
          for (int c = 0; c < kTileN; ++c)
            if (kbase + c > qpos) v[c] = 0xff800000u;        // -inf, also replaces NaN from stale keys
#pragma unroll
          for (int c = 0; c < kTileN / 2; c += 8) {
#pragma unroll
            for (int i = 0; i < 4; ++i) mx4[i] = fmax3(mx4[i], __uint_as_float(v[c + 2 * i]), __uint_as_float(v[c + 2 * i + 1]));
          }
        }
#pragma unroll
        for (int c = kTileN / 2; c < kTileN; c += 8) {
#pragma unroll
          for (int i = 0; i < 4; ++i) mx4[i] = fmax3(mx4[i], __uint_as_float(v[c + 2 * i]), __uint_as_float(v[c + 2 * i + 1]));
        }
        const float ms = fmaxf(fmax3(mx4[0], mx4[1], mx4[2]), mx4[3]) * scale;
        // Lazy rescale: keep exponentiating against the maximum in use until the true maximum has
        // grown by more than 2^8; p <= 256 then, and l / O use the same reference point.
        const bool grow = ms > m_used + kRescaleThreshold;   // always true at t == 0 (key 0 is visible)
        float alpha = 1.f;
        if (grow) {
          alpha = fast_exp2(m_used - ms);
          m_used = ms;
        }
        if (t > 0 && __any_sync(FULL, grow)) {               // rescale O before P V(t) accumulates into it
#pragma unroll
          for (int cc = 0; cc < HD / 32; ++cc) {
            uint32_t o[32];
            tmem_ld32(tO + cc * 32, o);
            tmem_wait_ld();
#pragma unroll
            for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
            tmem_st32(tO + cc * 32, o);
          }
        }
        // ---- the work tile's last KV tile: V rows that were not loaded or lie beyond the context must be finite ----
        if (t == n - 1) {
          WorkTile T;
          read_tile(e_cur, T);
          const int zero_from = min(T.nblk_need * kBS, T.ctx) - kbase;
          if (zero_from < kTileN) {
            const int item = item0 + 2 * t + 1;
            mbar_wait(bar(kBarKvFull + item % kRing), (item / kRing) & 1);
            const int nz = (kTileN - zero_from) * HALVES * 8;
            uint8_t* vb = gen + (sRing - base) + (uint32_t)(item % kRing) * kItem;
            for (int idx = wt; idx < nz; idx += kWgThreads) {
              const int row = zero_from + idx / (HALVES * 8), rem = idx % (HALVES * 8);
              *reinterpret_cast<uint4*>(vb + (uint32_t)(rem >> 3) * kSlab + (uint32_t)row * 128u + (uint32_t)((rem & 7) << 4)) =
                  make_uint4(0, 0, 0, 0);
            }
            fence_proxy_async();
          }
        }
        PA_TRACE(5 * w + 2, t);
        const uint64_t negm2 = pack2(-m_used, -m_used);
        uint64_t sum2[2] = {pack2(0.f, 0.f), pack2(0.f, 0.f)};
        // exponentiate in place: v[c] <- bf16x2(p[2c], p[2c + 1]); P = first 64 columns of S.  The two
        // 64-key halves are published separately so P V starts while the second half is exponentiated.
#pragma unroll
        for (int hf = 0; hf < 2; ++hf) {
#pragma unroll
          for (int c = hf * 32; c < hf * 32 + 32; ++c) {
            const uint64_t x2 = ffma2(pack2(__uint_as_float(v[2 * c]), __uint_as_float(v[2 * c + 1])), scale2, negm2);
            float p0, p1;
            if ((kPolyExpMask >> (c & 7)) & 1u) {
              poly_exp2x2(x2, p0, p1);
            } else {
              float x0, x1;
              unpack2(x2, x0, x1);
              p0 = fast_exp2(x0);
              p1 = fast_exp2(x1);
            }
            sum2[c & 1] = fadd2(sum2[c & 1], pack2(p0, p1));
            v[c] = pack_bf16(p0, p1);
          }
          tmem_st32(tS + hf * 32, v + hf * 32);
          if (hf == 1) PA_TRACE(5 * w + 3, t);
          tmem_wait_st();                                    // also covers the O rescale above
          tc_fence_before();
          mbar_arrive(bar(kBarPReady + 2 * w + hf));         // keys [0, 64) / [64, 128) of P(t) are in TMEM
          if (hf == 1) PA_TRACE(5 * w + 4, t);
        }
        float s0, s1;
        unpack2(fadd2(sum2[0], sum2[1]), s0, s1);
        l = l * alpha + (s0 + s1);
      }

      // ---- the next work tile's Q is in shared memory: the issuer may start its first Q K^T behind this tile's
      // last P V, while this warpgroup writes the tile's output ----
      if (next_state > 0) stage_q_commit(slot_n);

      // ---- epilogue: O / l -> out ----
      mbar_wait(bar(kBarODone + w), o_cnt & 1);
      ++o_cnt;
      if (w == 0) stamp_at(slot, 5);
      tc_fence_after();
      const float inv = 1.f / l;
      __nv_bfloat16* ob;
      {
        WorkTile T;
        read_tile(e_cur, T);
        ob = static_cast<__nv_bfloat16*>(p.out) + (int64_t)(T.q_start + T.tok0 + w * R + ti) * p.o_stride_t +
             (int64_t)(T.h * g + j) * p.o_stride_h;
      }
#pragma unroll
      for (int cc = 0; cc < HD / 32; ++cc) {
        uint32_t o[32];
        tmem_ld32(tO + cc * 32, o);
        tmem_wait_ld();
        if (valid) {
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            uint4 wv;
            wv.x = pack_bf16(__uint_as_float(o[8 * i]) * inv, __uint_as_float(o[8 * i + 1]) * inv);
            wv.y = pack_bf16(__uint_as_float(o[8 * i + 2]) * inv, __uint_as_float(o[8 * i + 3]) * inv);
            wv.z = pack_bf16(__uint_as_float(o[8 * i + 4]) * inv, __uint_as_float(o[8 * i + 5]) * inv);
            wv.w = pack_bf16(__uint_as_float(o[8 * i + 6]) * inv, __uint_as_float(o[8 * i + 7]) * inv);
            *reinterpret_cast<uint4*>(ob + cc * 32 + i * 8) = wv;
          }
        }
      }
      if (w == 0) stamp_at(slot, 6);
      // O is read: the next tile's first P V may overwrite it.  That MMA is issued behind this warpgroup's
      // next P-ready arrival, which every thread makes after this point.
      tc_fence_before();
      release_tile(e_cur);                                   // done with the tile: its entry may be overwritten
      bool have_n = next_state > 0;
      if (next_state == 0) {                                 // the next tile had no rows for this warpgroup: now it may wait
        have_n = advance(e_n);
        if (have_n) stage_q_commit(stage_q_issue(e_n));
      }
      e_cur = e_n;
      have = have_n;
      first_tile = false;
    }
  }
  else {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 40;");
  if (warp == 8) {
    // =============================== TMA producer ===============================
    const int b = lane / HALVES, hf = lane % HALVES;
    int t_cnt = 0;
    for (;;) {
      const int e = acquire_tile(t_cnt);                     // every lane waits: the entry is read by all of them
      WorkTile T;
      read_tile(e, T);
      __syncwarp();
      if (lane == 0) release_tile(e);                        // the copy is all this role needs
      if (T.L < 0) break;
      const int item0 = T.item0;
      const int32_t* bt = p.block_tables + (int64_t)T.s * p.bt_stride;
      for (int t = 0; t < T.n; ++t) {
        if (t == max(T.n - 1 - kDrawAhead, 0) && lane == 0) mbar_arrive(bar(kBarTReq));
        const int blk0 = t * kBlocksPerTile;
        const int nb = min(kBlocksPerTile, T.nblk_need - blk0);
        const bool mine = lane < kBlocksPerTile * HALVES && b < nb;
        int blk = 0;
        if (mine) blk = __ldg(bt + blk0 + b);
        const int c3 = p.pos_head == 3 ? T.h : blk, c4 = p.pos_head == 3 ? blk : T.h;
        const uint32_t in_item = (uint32_t)hf * kSlab + (uint32_t)b * (kBS * 128);
#pragma unroll
        for (int kv = 0; kv < 2; ++kv) {                     // item 2t = K(t), item 2t + 1 = V(t)
          const int item = item0 + 2 * t + kv, slot = item % kRing;
          mbar_wait(bar(kBarKvEmpty + slot), ((item / kRing) & 1) ^ 1);
          if (lane == 0) mbar_expect_tx(bar(kBarKvFull + slot), (uint32_t)nb * HALVES * kBS * 128);
          __syncwarp();
          if (mine) tma_load_half(sRing + (uint32_t)slot * kItem + in_item, kv ? &tm_v : &tm_k, bar(kBarKvFull + slot), hf, c3, c4);
        }
      }
    }
  } else if (warp == 9) {
    // =============================== MMA issuer ===============================
    // ONE elected thread runs the whole role (waits included): every tcgen05 operand then lives in
    // uniform registers and an MMA costs a handful of issue slots instead of an election loop.
    if (elect_one()) {
      constexpr uint32_t idesc_s = umma_idesc(kTileN, false);
      constexpr uint32_t idesc_o = umma_idesc(HD, true);
      constexpr uint64_t kDescK = ((uint64_t)(16 >> 4) << 16) | ((uint64_t)(1024 >> 4) << 32) | (1ull << 46) | (2ull << 61);
      constexpr uint64_t kDescV = ((uint64_t)(kSlab >> 4) << 16) | ((uint64_t)(1024 >> 4) << 32) | (1ull << 46) | (2ull << 61);
      int item0 = 0;
      uint32_t ph = 0;                                       // phase bits: 0, 1 = Q ready of A / B; 2, 3 = P ready of A / B
      bool tr = trace_cta;                                   // first work tile of CTA 0 only
      auto wait_item = [&](int item) { mbar_wait(bar(kBarKvFull + item % kRing), (item / kRing) & 1); };
      auto free_item = [&](int item) { umma_commit(bar(kBarKvEmpty + item % kRing)); };
      auto qk = [&](int w, int t) {                          // S_w = Q_w K(t)^T   (both operands in shared memory)
        if (tr && t < kTraceTiles) p.trace[t * kTraceSlots + 11 + 2 * w] = clock64();
        const uint64_t da = kDescK | (uint64_t)(((sQ + (uint32_t)w * kItem) & 0x3FFFFu) >> 4);
        const uint64_t db = kDescK | (uint64_t)(((sRing + (uint32_t)((item0 + 2 * t) % kRing) * kItem) & 0x3FFFFu) >> 4);
        const uint32_t d = tmem + (uint32_t)w * 128u;
#pragma unroll
        for (int kk = 0; kk < HD / 16; ++kk) {
          const uint64_t koff = (uint64_t)(((uint32_t)(kk >> 2) * kSlab + (uint32_t)(kk & 3) * 32u) >> 4);
          umma_bf16(d, da + koff, db + koff, idesc_s, kk > 0);
        }
        umma_commit(bar(kBarSFull + w));
      };
      auto pv = [&](int w, int t, bool last) {               // O_w += P_w V(t)      (P in TMEM, V in shared memory)
        // the first 64 keys' P V runs while the warpgroup still exponentiates the other 64
        // everything that can be computed before P is ready is: the issuer is ONE thread, and what it executes
        // between a barrier flipping and the MMA going out is added to every tile
        const uint32_t par = (ph >> (2 + w)) & 1u;
        ph ^= 4u << w;
        const uint64_t db = kDescV | (uint64_t)(((sRing + (uint32_t)((item0 + 2 * t + 1) % kRing) * kItem) & 0x3FFFFu) >> 4);
        const uint32_t d = tmem + kTmemO + (uint32_t)w * 128u, a = tmem + (uint32_t)w * 128u;
        mbar_wait(bar(kBarPReady + 2 * w), par);
        if (tr && t < kTraceTiles) p.trace[t * kTraceSlots + 10 + 2 * w] = clock64();
        tc_fence_after();
        umma_bf16_ts(d, a, db, idesc_o, t > 0 ? 1u : 0u);
#pragma unroll
        for (int ks = 1; ks < kTileN / 32; ++ks) umma_bf16_ts(d, a + (uint32_t)ks * 8u, db + (uint64_t)(ks * (16 * 128 >> 4)), idesc_o, 1u);
        mbar_wait(bar(kBarPReady + 2 * w + 1), par);
        tc_fence_after();
#pragma unroll
        for (int ks = kTileN / 32; ks < kTileN / 16; ++ks) umma_bf16_ts(d, a + (uint32_t)ks * 8u, db + (uint64_t)(ks * (16 * 128 >> 4)), idesc_o, 1u);
        if (last) umma_commit(bar(kBarODone + w));
      };
      int t_cnt = 0;
      for (;;) {
        int n, n_a, n_b, ntok_b;
        {
          const int e = acquire_tile(t_cnt);
          WorkTile T;
          read_tile(e, T);
          release_tile(e);
          if (T.L < 0) break;
          n = T.n; n_a = T.n_a; n_b = T.n_b; ntok_b = T.ntok_b;
          item0 = T.item0;
        }
        mbar_wait(bar(kBarQReady), ph & 1u);
        ph ^= 1u;
        if (ntok_b > 0) {
          mbar_wait(bar(kBarQReady + 1), (ph >> 1) & 1u);
          ph ^= 2u;
        }
        tc_fence_after();
        wait_item(item0);
        qk(0, 0);
        if (n_b > 0) qk(1, 0);
        free_item(item0);
        for (int t = 0; t < n; ++t) {
          wait_item(item0 + 2 * t + 1);                      // V(t)
          if (t + 1 < n) wait_item(item0 + 2 * t + 2);       // K(t + 1)
          if (t < n_a) {
            pv(0, t, t == n_a - 1);
            if (t + 1 < n_a) qk(0, t + 1);
          }
          if (t < n_b) {
            pv(1, t, t == n_b - 1);
            if (t + 1 < n_b) qk(1, t + 1);
          }
          free_item(item0 + 2 * t + 1);
          if (t + 1 < n) free_item(item0 + 2 * t + 2);
        }
        tr = false;
      }
      tc_fence_before();
    }
    __syncwarp();
  } else if (warp == 10) {
    // =============================== work-tile scheduler ===============================
    int* const tk = g_tickets + 2 * ticket_slot;
    int pub = 0;
    int items = 0;                                           // K / V ring items of the tiles published so far
    auto publish = [&](WorkTile& T, int L) {                 // lane 0
      const int e = pub % kTileRing;
      mbar_wait(bar(kBarTEmpty + e), ((pub / kTileRing) & 1) ^ 1);
      T.L = L;
      T.item0 = items;
      T.pad = 0;
      const int4* src = reinterpret_cast<const int4*>(&T);
#pragma unroll
      for (int i = 0; i < 4; ++i)
        asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(sTiles + 64u * e + 16u * i), "r"(src[i].x), "r"(src[i].y),
                     "r"(src[i].z), "r"(src[i].w) : "memory");
      mbar_arrive(bar(kBarTFull + e));
      ++pub;
    };
    const int hkv = p.num_kv_heads, pairs = total_tiles / hkv;   // a (query tile, sequence) pair = hkv consecutive tickets, all empty or none
    int drawn = 0;                                           // tickets with work published so far
    for (;;) {
      // the first draw is immediate, every later one waits for the producer's request
      if (drawn > 0 && lane == 0) mbar_wait(bar(kBarTReq), (drawn - 1) & 1);
      __syncwarp();
      int L = 0;
      if (lane == 0) L = atomicAdd(tk, 1);
      L = __shfl_sync(FULL, L, 0);
      if (L >= total_tiles) break;
      WorkTile T;
      if (work_tile(p, L, gx, R, T)) {
        if (lane == 0) publish(T, L);
        items += 2 * T.n;
        ++drawn;
        continue;
      }
      // An empty tile (a sequence shorter than the launch's longest, a masked row): the 32 lanes look for the
      // next pair that has work and move the launch's ticket there in one step - whoever gets there first.
      int pair = L / hkv + 1, found = pairs;
      while (pair < pairs) {
        bool v = false;
        if (pair + lane < pairs) {
          WorkTile U;
          v = work_tile(p, (pair + lane) * hkv, gx, R, U);
        }
        const unsigned m = __ballot_sync(FULL, v);
        if (m) {
          found = pair + __ffs(m) - 1;
          break;
        }
        pair += 32;
      }
      if (lane == 0) atomicMax(tk, found * hkv);
    }
    if (lane == 0) {
      WorkTile T;
      memset(&T, 0, sizeof(T));
      publish(T, -1);
      // every CTA draws exactly one ticket past the end, after its last touch of the slot: the last to do so clears it
      if (atomicAdd(tk + 1, 1) == (int)gridDim.x - 1) {
        tk[0] = 0;
        tk[1] = 0;
      }
    }
    __syncwarp();
  }
  }

  __syncthreads();
  if (warp == 9) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(kTmemCols) : "memory");
  }
}

template <int HD>
constexpr size_t smem_bytes() {
  return 1024 + (size_t)(HD / 64) * kSlab * (2 + kRing) + kNumBars * 8 + 16 + 16 + kTileRing * 64;
}

int sm_count_cached() {
  static int cached[64];
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) { cudaGetLastError(); return -1; }
  if (dev >= 0 && dev < 64 && cached[dev] > 0) return cached[dev];
  int n = 0;
  if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) { cudaGetLastError(); return -1; }
  if (dev >= 0 && dev < 64) cached[dev] = n;
  return n;
}

template <int HD>
int launch(const PaPrefillParams& p, const CUtensorMap& mk, const CUtensorMap& mv, cudaStream_t stream) {
  auto kern = pa_prefill_tc2_kernel<HD>;
  static PaPerDeviceFlag attr;
  const int dev = attr.need();
  if (dev != -2) {
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes<HD>()) != cudaSuccess)
      return -5;
    attr.done(dev);
  }
  const int R2 = 2 * (kRows / p.group);
  const int gx = (p.max_q_len + R2 - 1) / R2;
  const int64_t total = (int64_t)gx * p.num_kv_heads * p.batch;
  if (total <= 0) return 0;
  if (total > 0x7fffffff) return -2;
  const int sms = sm_count_cached();
  if (sms <= 0) return -5;
  const int ctas = (int)(total < sms ? total : sms);          // persistent: one CTA per SM walks the work tiles
  static std::atomic<unsigned> launches{0};
  const int slot = (int)(launches.fetch_add(1) % (unsigned)kTicketSlots);
  kern<<<ctas, kThreads, smem_bytes<HD>(), stream>>>(mk, mv, p, gx, (int)total, slot);
  return cudaGetLastError() == cudaSuccess ? 0 : -5;
}

}  // namespace

static long long* g_prefill_times = nullptr;
extern "C" int pa_prefill_debug_times(void* device_buffer) {      // 8 int64 per CTA of the next launches; null = off
  g_prefill_times = static_cast<long long*>(device_buffer);
  return 0;
}

int pa_launch_prefill_tc2(const PaPrefillParams& p_in, cudaStream_t stream) {
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
  p.trace = nullptr;
  p.cta_times = g_prefill_times;
  static const bool trace = getenv("PA_PREFILL_TRACE") != nullptr;
  if (!trace) return p.head_dim == 128 ? launch<128>(p, mk, mv, stream) : launch<64>(p, mk, mv, stream);
  // bring-up aid: clock64 stamps of the first tiles of CTA (0,0,0), printed as deltas to stderr
  static long long* dbuf = nullptr;
  const size_t nbytes = (size_t)kTraceTiles * kTraceSlots * sizeof(long long);
  if (!dbuf && cudaMalloc(&dbuf, nbytes) != cudaSuccess) return -5;
  cudaMemsetAsync(dbuf, 0, nbytes, stream);
  p.trace = dbuf;
  rc = p.head_dim == 128 ? launch<128>(p, mk, mv, stream) : launch<64>(p, mk, mv, stream);
  static int printed = 0;
  if (rc == 0 && printed++ == 3) {
    static long long h[kTraceTiles * kTraceSlots];
    cudaStreamSynchronize(stream);
    cudaMemcpy(h, dbuf, nbytes, cudaMemcpyDeviceToHost);
    const long long t0 = h[0];
    fprintf(stderr, "tile | A: sfull ld max exp arrive | B: sfull ld max exp arrive | MMA: pA qkA pB qkB  (cycles since A's first S)\n");
    for (int t = 0; t < kTraceTiles; ++t) {
      fprintf(stderr, "%3d |", t);
      for (int k = 0; k < 14; ++k)
        fprintf(stderr, " %7lld%s", h[t * kTraceSlots + k] ? h[t * kTraceSlots + k] - t0 : -1, (k == 4 || k == 9) ? " |" : "");
      fprintf(stderr, "\n");
    }
  }
  return rc;
}
