// Paged-attention decode, bf16, B200 fast path (block_size 16, head_dim 64 / 128, g <= 8).
//
// Replaces the per-sequence Python loop of the reference
// (01_paged_attention/paged_attention.py:80-101): for every (sequence, KV head) gather the
// K / V rows through the block table, softmax(q K^T / sqrt(D)) V for the g query heads of the
// group.  HBM-bound streaming gather at ~4 flop/byte; the design goal is to keep every SM's
// TMA queue full until the last byte and to read each K / V byte exactly once.
//
// Shape of the kernel
//  * split-K "flash-decoding" work units: a (sequence, KV head) segment is cut into units of C
//    consecutive KV blocks; C (a multiple of 8) is chosen on the device from the batch's total
//    block count so that there are ~units_per_sm units per SM.  Units are handed out through ONE
//    atomic ticket in linear (sequence, head, chunk) order.  B200 SMs do not all see the same HBM
//    bandwidth (the time to stream equal byte counts spreads 75 % .. 100 % across SMs, stable per
//    SM), so a static split leaves a fifth of the SM-time idle; with the ticket the fast SMs
//    simply take more units and every SM streams until the end.
//  * persistent grid, one CTA per SM, 20 warps in four roles.  Round 1 ran producer and consumer in the
//    same warp and merged every unit with two CTA-wide barriers; per-CTA clocks (tools/exp_balance.py,
//    profiles/r02_decode_timeline_c2_before.json) showed the consumer warps - not HBM - setting the pace
//    at C2: of 84 us a consumer warp waited for KV data only 6 us, but spent 14 us stalled in the TMA
//    issue path (back-pressure of a saturated memory system lands on whoever issues), 10 us in unit
//    merges (~1 us per unit) and 24 us in the math.  So the roles are now separate:
//      scheduler (warp 16): builds the plan, claims tickets, maps ticket -> (s, h, chunk)
//        by binary search over the per-sequence unit prefix sums in shared memory, reads the unit's
//        block-table entries with coalesced loads and publishes them, a unit descriptor and the unit's
//        Q rows (cp.async.bulk) into a ring of 8 descriptor slots (full / empty mbarriers).
//      producers (warps 8-15): producer c feeds consumer c.  Block j of a unit belongs to consumer
//        j % 8; each consumer owns a private ring of R stages (one stage = K tile + V tile of one
//        block for one KV head, [16 tokens x D] bf16 each).  The producer's one active lane walks the
//        descriptor ring, waits for a free stage (empty mbarrier) and issues the two TMA tensor loads
//        of the block (cp.async.bulk.tensor 5-D, 128-byte swizzle, one instruction per tile)
//        completing on the stage's full mbarrier - 8 independent issuers, R x 8 x 8 KiB in flight per
//        SM, and an issue stall costs nobody but the producer.
//      consumers (warps 0-7): math only.  Per block on legacy tensor cores, in the TRANSPOSED
//        formulation so the GQA query heads sit in the n = 8 dimension:
//            S^T[16 tok x 8 heads] = K[16 x D] . Q^T[D x 8]          (mma.sync m16n8k16 bf16,
//            O^T[D x 8 heads]     += V^T[D x 16 tok] . P^T[16 x 8]     fp32 accumulate)
//        A operands by ldmatrix (K) / ldmatrix.trans (V) straight from the swizzled tiles, P^T from
//        the S^T accumulator layout to the B layout by movmatrix, fp32 online softmax in the exp2
//        domain with warp-shuffle reductions.  tcgen05 needs M >= 64 and is not a fit for a <= 8-row
//        GQA group; the kernel is bound by HBM, not by the tensor pipe.  At the end of a unit every
//        consumer warp drops its (m, l, O) into a shared-memory hand-off area and goes straight on
//        to the next unit - no CTA-wide barrier.
//      merger (warp 17): merges the 8 warps' states of a unit and frees the hand-off area at once; a unit that
//        covers its whole segment is normalised and written to `out`, otherwise the merged partial goes to the
//        workspace as self-validating 16-byte words {value, epoch, value, epoch} (plain stores, no fence) and,
//        for the unit that owns the fold (a segment's first unit; for > 8 partials the first unit of every
//        group of 8, the last group's also for the final fold), a fold job is posted to the folders.
//      folders (warps 18-19): fold a split segment's partials in chunk order, half of the GQA heads each, 4
//        partials' loads in flight per lane, polling one sentinel word per partial until the batch has landed,
//        and write `out` - no second launch, no atomic and no fence on the hand-over, and the consumers never
//        wait for a global-memory round trip.
//      (scheduler: a ticket is claimed only when the CTA's backlog is short, and a unit is published in
//       16-block descriptors - see the scheduler branch; register budgets per role via setmaxnreg.)
//  * results are bit-reproducible: the block -> warp assignment and every merge order depend only
//    on the batch shape, never on which SM took which unit.
#include <atomic>
#include <cuda.h>
#include <mutex>
#include <string.h>

#include "pa_common.cuh"

namespace {

constexpr int kBS = 16;             // tokens per block on this path
constexpr int kConsumers = 8;       // consumer (math) warps per CTA: warps 0-7, two warpgroups
constexpr int kProducers = 8;       // TMA producer warps, one per consumer: warps 8-15
constexpr int kSchedWarp = kConsumers + kProducers;           // warp 16
constexpr int kFinWarps = 3;        // finisher warps per CTA: warp 17 merges units, warps 18-19 fold split segments together
constexpr int kMinUnitBlocks = 32;  // smallest split-K unit, in 8-KiB blocks (head_dim 128; twice as many blocks at head_dim 64)
constexpr int kFoldGroup = 8;       // segments of more partials are folded in two levels: groups of 8, then the groups
constexpr int kJobRing = 64;        // fold jobs in flight merger -> folders (a CTA has a handful pending; see the merger)
                                    //   (scheduler + finishers = one warpgroup)
constexpr int kThreads = (kConsumers + kProducers + 1 + kFinWarps) * 32;   // 640
constexpr int kDescRing = 8;        // unit descriptors in flight scheduler -> producers / consumers
constexpr int kNumBars = 2 * kConsumers * 8 /* kMaxStages */ + 3 * kDescRing + 2 + 2 * kJobRing;
#ifndef PA_DESC_BLOCKS
#define PA_DESC_BLOCKS 16
#endif
#ifndef PA_FOLD_DEPTH
#define PA_FOLD_DEPTH 4
#endif
constexpr int kDescBlocks = PA_DESC_BLOCKS;   // blocks per descriptor (a unit spans several); a multiple of 8, 16 / 32 / 64
constexpr int kMaxStages = 8;       // per-warp ring depth cap (kNumBars assumes 8)
constexpr int kMaxPartials = 32;    // split-K partials per (sequence, KV head) segment, at most
constexpr int kHeaderInts = 16;     // counter header: [0] unit ticket, [1] exited CTAs, [2] launch epoch
constexpr int kEpochSlot = 40;      // scratch word holding this launch's partial flag
constexpr int kDbgWords = 16;      // int64 words per CTA of the debug timeline (pa_decode_debug_times)
constexpr int kMaxBatch = 4096;     // seq_lens + unit prefix sums live in shared memory
constexpr int kSmemBudget = 227 * 1024;
constexpr uint32_t FULL = 0xffffffffu;

enum : int { kFirst = 1, kLast = 2 };   // descriptor flags: first / last descriptor of its unit

struct UnitDesc {                   // 32 bytes
  int s, h;                         // sequence, KV head
  int n, flags;                     // blocks in this descriptor (0: end of stream), kFirst | kLast
  int last_valid;                   // valid tokens in the unit's final block
  int nch;                          // units of this segment
  int slot, slot0;                  // this unit's ticket (= partial slot), ticket of the segment's first unit
};

// Shared-memory carve-up, identical on host and device.
struct SmemLayout {
  uint32_t ring, merge_o, merge_ml, qring, bt, desc, fin, bars, scratch, lens, cs, ucum, total;
};
__host__ __device__ inline SmemLayout smem_layout(int hd, int g, int r, int batch) {
  SmemLayout L;
  const uint32_t stage = 2u * (hd / 64) * kBS * 128;
  uint32_t o = 0;
  L.ring = o;      o += (uint32_t)kConsumers * r * stage;
  L.merge_o = o;   o += (uint32_t)kConsumers * g * (hd + 4) * 4;
  L.merge_ml = o;  o += kConsumers * 8 * 8;
  L.qring = o;     o += (uint32_t)kDescRing * g * hd * 2;
  L.bt = o;        o += kDescRing * kDescBlocks * 4;
  L.desc = o;      o += kDescRing * 32;
  L.fin = o;       o += 32 + kJobRing * 16;              // unit merge job (2 x int4) + fold job ring (int4 each)
  L.bars = o;      o += (uint32_t)kNumBars * 8;
  L.scratch = o;   o += 48 * 4;
  L.lens = o;      o += (uint32_t)batch * 4;
  L.cs = o;        o += (uint32_t)(batch + 1) * 4;
  L.ucum = o;      o += (uint32_t)(batch + 1) * 4;
  L.total = o + 1024;               // slack for the 1024-byte alignment of the ring
  return L;
}

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ bool mbar_test(uint32_t bar, uint32_t parity) {
  uint32_t done;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.b32 %0, 1, 0, p;\n\t}"
      : "=r"(done)
      : "r"(bar), "r"(parity)
      : "memory");
  return done != 0;
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
__device__ __forceinline__ void tma_load_5d(uint32_t dst, const CUtensorMap* map, uint32_t bar,
                                            int c3, int c4, uint64_t policy) {
  asm volatile(
      "cp.async.bulk.tensor.5d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint"
      " [%0], [%1, {%3, %3, %3, %4, %5}], [%2], %6;"
      ::"r"(dst), "l"(map), "r"(bar), "r"(0), "r"(c3), "r"(c4), "l"(policy)
      : "memory");
}
__device__ __forceinline__ void bulk_load(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
      ::"r"(dst), "l"(src), "r"(bytes), "r"(bar)
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
__device__ __forceinline__ void consumer_bar() {
  asm volatile("bar.sync 1, %0;" ::"n"(kConsumers * 32) : "memory");
}
__device__ __forceinline__ float4 ldcg4(const float* p) { return __ldcg(reinterpret_cast<const float4*>(p)); }
// A load that may be repeated until another SM's store shows up.  The __ldcv / __ldcg intrinsics are plain
// (non-volatile) asm to the compiler: two of them on one address are "the same value", and a polling loop
// built on them is deleted.  This one is volatile with a memory clobber, and relaxed.gpu reads at L2.
__device__ __forceinline__ float4 ld_poll(const float4* p) {
  float4 v;
  asm volatile("ld.relaxed.gpu.global.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
  return v;
}

// Exclusive scan of a[0 .. n) in place (shared memory), a[n] = total.  Called by ONE warp (the
// scheduler); lanes own contiguous slices.
__device__ __forceinline__ int warp_scan_inplace(int* a, int n, int lane) {
  const int per = pa_ceil_div(n, 32);
  const int lo = min(lane * per, n), hi = min(lo + per, n);
  int mine = 0;
  for (int i = lo; i < hi; ++i) mine += a[i];
  int incl = mine;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int t = __shfl_up_sync(FULL, incl, o);
    if (lane >= o) incl += t;
  }
  const int total = __shfl_sync(FULL, incl, 31);
  int run = incl - mine;
  for (int i = lo; i < hi; ++i) {
    const int v = a[i];
    a[i] = run;
    run += v;
  }
  if (lane == 0) a[n] = total;
  __syncwarp();
  return total;
}

// Launch plan, built by the scheduler warp of every CTA (nobody else needs it, so the consumers
// are already waiting for their first descriptor meanwhile).  Blocks per unit are chosen PER
// SEQUENCE by guided self-scheduling: C(s) ~ (blocks from sequence s to the end) / (guided_k *
// CTAs), so the units handed out first are large (few partials, few merges; whole segments while
// they are short enough) and the last ones small (every SM streams until the end), never below
// the uniform size that yields target_units units.  C(s) is a multiple of 8: every consumer warp
// gets the same share of a full unit.  Outputs: s_len[s] = max(seq_len, 0), s_c[s] = C(s),
// s_ucum = exclusive scan of ceil(nblk(s) / C(s)) (s_ucum[batch] = units per KV head).
__device__ __forceinline__ void build_plan(const int32_t* __restrict__ seq_lens, int batch, int hs, int target_units, int min_unit,
                                           int guided_div, int* s_len, int* s_c, int* s_ucum, int lane) {
  for (int i = lane; i < batch; i += 32) {
    const int len = max(__ldg(seq_lens + i), 0);
    s_len[i] = len;
    s_c[i] = pa_ceil_div(len, kBS);
  }
  __syncwarp();
  const int total = warp_scan_inplace(s_c, batch, lane) * hs;             // s_c[s] = blocks before sequence s
  // never below kMinUnitBlocks (256 KiB at head_dim 128): on small batches (tens of MiB of KV) finer units buy no
  // balance worth their folds - 4 x 4096 tokens took 38 us in 8-block units (32 partials per segment, a 15 us fold
  // backlog after 24 us of streaming)
  const int c_min = max(8 * max(1, pa_ceil_div(total, 8 * target_units)), min_unit);
  for (int i = lane; i < batch; i += 32) {
    const int nblk = pa_ceil_div(s_len[i], kBS);
    const int remaining = total - s_c[i] * hs;
    int c = max(c_min, 8 * (remaining / (8 * guided_div)));
    // at most kMaxPartials partials per segment: the merge of a segment is ONE CTA's job at the tail
    // of the kernel (256 partials of a 32k-token segment cost ~80 us of merging after 35 us of streaming)
    c = max(c, 8 * pa_ceil_div(nblk, 8 * kMaxPartials));
    c = min(c, 8 * pa_ceil_div(max(nblk, 1), 8));
    s_ucum[i] = pa_ceil_div(nblk, c);
    s_c[i] = c;                     // only this lane touches s_c[i] in this loop
  }
  __syncwarp();
  warp_scan_inplace(s_ucum, batch, lane);
}

// ---- fused output gather (pa_decode_gather) -------------------------------------------
// Replicate a final 4-byte output store into this rank's slot of every remote gather buffer
// (plain stores over NVLink: fire and forget).
__device__ __forceinline__ void peer_store32(const PaDecodeParams& p, int64_t off_elems, uint32_t v) {
  if (p.num_remote == 0) return;
  if (p.ll_epoch) {
    // this step's flag (never 0); loaded where it is used - output stores are rare - instead of
    // living in a register of every warp for the whole kernel
    const uint32_t ll_flag = __ldg(p.ll_epoch) + 1u;
    // LL: {data, flag} as one 8-byte store (single NVLink write, never torn); word index = off_elems / 2
    for (int r = 0; r < p.num_remote; ++r)
      *reinterpret_cast<uint2*>(static_cast<uint8_t*>(p.remote_out[r]) + off_elems * 4) = make_uint2(v, ll_flag);
    return;
  }
  for (int r = 0; r < p.num_remote; ++r)
    *reinterpret_cast<uint32_t*>(static_cast<__nv_bfloat16*>(p.remote_out[r]) + off_elems) = v;
}
// Count `n` completed segments into this rank's arrival counter on every rank.  Called by ONE
// thread after a CTA-scope barrier that follows the segment's stores: the system-scope fence is
// cumulative, so every store ordered before the barrier is visible to a peer that observes the count.
__device__ __forceinline__ void peer_signal(const PaDecodeParams& p, uint32_t n) {
  if (p.local_sig == nullptr) return;
  __threadfence_system();                 // ONE fence; the counts themselves are relaxed
  for (int r = 0; r < p.num_remote; ++r)
    asm volatile("red.relaxed.sys.global.add.u32 [%0], %1;" ::"l"(p.remote_sig[r]), "r"(n) : "memory");
  asm volatile("red.relaxed.sys.global.add.u32 [%0], %1;" ::"l"(p.local_sig), "r"(n) : "memory");
}

template <int HD>
__global__ void __launch_bounds__(kThreads, 1)
pa_decode_ws_kernel(const __grid_constant__ CUtensorMap tm_k, const __grid_constant__ CUtensorMap tm_v,
                    const PaDecodeParams p) {
  constexpr int HALVES = HD / 64;             // 128-byte column halves of a row
  constexpr int HALF_BYTES = kBS * 128;       // one [16 x 64] bf16 box
  constexpr int TILE_BYTES = HALVES * HALF_BYTES;
  constexpr int STAGE_BYTES = 2 * TILE_BYTES; // K tile + V tile
  constexpr int KSTEPS = HD / 16;
  constexpr int OPITCH = HD + 4;              // merge row pitch in floats (bank-conflict-free writes)

  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
  const int g = p.group, R = p.num_stages;
  const SmemLayout L = smem_layout(HD, g, R, p.batch);
  float* merge_o = reinterpret_cast<float*>(smem_gen + L.merge_o);
  float2* merge_ml = reinterpret_cast<float2*>(smem_gen + L.merge_ml);
  int* bt_ring = reinterpret_cast<int*>(smem_gen + L.bt);
  UnitDesc* desc = reinterpret_cast<UnitDesc*>(smem_gen + L.desc);
  int4* mjob = reinterpret_cast<int4*>(smem_gen + L.fin);               // [0] (s, h, nch, slot)  [1] (slot0, end, -, -)
  int* scratch = reinterpret_cast<int*>(smem_gen + L.scratch);
  int* s_len = reinterpret_cast<int*>(smem_gen + L.lens);
  int* s_c = reinterpret_cast<int*>(smem_gen + L.cs);
  int* s_ucum = reinterpret_cast<int*>(smem_gen + L.ucum);
  const uint32_t qring = smem_base + L.qring;
  const uint32_t bar_full = smem_base + L.bars;                          // [kConsumers][kMaxStages]  TMA -> consumer
  const uint32_t bar_empty = bar_full + kConsumers * kMaxStages * 8;     // [kConsumers][kMaxStages]  consumer -> producer
  const uint32_t bar_dfull = bar_empty + kConsumers * kMaxStages * 8;    // [kDescRing]
  const uint32_t bar_dempty = bar_dfull + kDescRing * 8;                 // [kDescRing]  8 consumers + 8 producers
  const uint32_t bar_qfull = bar_dempty + kDescRing * 8;                 // [kDescRing]
  const uint32_t bar_mfull = bar_qfull + kDescRing * 8;                  // a unit's state handed to the merger (8 warps arrive)
  const uint32_t bar_mempty = bar_mfull + 8;                             // hand-off area free again
  const uint32_t bar_jfull = bar_mempty + 8;                             // [kJobRing] fold job posted
  const uint32_t bar_jempty = bar_jfull + kJobRing * 8;                  // [kJobRing] read by both folders
  const uint32_t q_bytes = (uint32_t)g * HD * 2;

  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;

  // Programmatic dependent launch (the host sets the stream-serialisation attribute): the NEXT kernel of the
  // stream may become resident on an SM as soon as this CTA has left it, instead of after the whole grid has
  // drained, and a kernel launched that way runs its own prologue - barrier init, register re-allocation,
  // tensor-map prefetch - while its predecessor's last CTAs are still finishing.  pdl_wait() below is the
  // point where the predecessor's memory (ticket header, seq_lens, block tables, Q, KV, out) must be final.
  asm volatile("griddepcontrol.launch_dependents;");
  auto pdl_wait = [] { asm volatile("griddepcontrol.wait;" ::: "memory"); };

  unsigned long long dbg_t0 = 0;
  if (p.dbg && threadIdx.x == 0) asm volatile("mov.u64 %0, %globaltimer;" : "=l"(dbg_t0));

  // Register budget per role.  The pool setmaxnreg redistributes is what the CTA was LAUNCHED with - 640
  // threads x 96 registers = 61440 - not the SM's 64 Ki (asking for more blocks forever).  The producers'
  // single active lane needs next to nothing; what they give back goes to the consumers (fragments of Q,
  // the O accumulator, K / V fragment sets in flight) and to the finishers (8 partials' loads in flight per
  // lane): 256 x 144 + 256 x 24 + 128 x 144 = 61440.
  // Each role's code sits entirely inside the branch that executes its setmaxnreg (ptxas only grants the
  // larger budget to code dominated by the instruction), so the one CTA-wide barrier is issued per branch.
  auto cta_sync = [] { asm volatile("bar.sync 0, %0;" ::"n"(kThreads) : "memory"); };

  // mbarrier init spread over the CTA (154 barriers by one thread cost ~1 us of every launch)
  for (int i = threadIdx.x; i < kNumBars; i += kThreads) {
    const uint32_t b = bar_full + i * 8;
    uint32_t count = 1;
    if (b >= bar_dempty && b < bar_qfull) count = kConsumers + kProducers;
    if (b >= bar_mfull && b < bar_mempty) count = kConsumers;
    if (b >= bar_jempty) count = 2;                             // every other barrier has ONE arriving thread
    mbar_init(b, count);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (threadIdx.x == kThreads - 1) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tm_k) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tm_v) : "memory");
  }
  auto dbg_stamp = [&](int slot) {        // one thread of the calling role: globaltimer into the CTA's timeline
    if (p.dbg) {
      unsigned long long t;
      asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
      p.dbg[kDbgWords * (int64_t)blockIdx.x + slot] = (long long)t;
    }
  };
  const int hs = p.num_kv_heads;
  int* const ticket = p.counters;                            // [0] unit ticket, [1] exited CTAs, [2] launch epoch
  __nv_bfloat16* obase = static_cast<__nv_bfloat16*>(p.out);

  if (warp >= kSchedWarp) {
  asm volatile("setmaxnreg.inc.sync.aligned.u32 160;");
  if (warp == kSchedWarp) {
    // =========================== scheduler warp ==========================================
    // The plan and the first unit are prepared BEFORE the CTA-wide barrier below: nothing here needs the
    // mbarriers, and the other warps have nothing to do until the first descriptor is published anyway.
    // The first ticket is requested before the plan is built: its latency hides behind the plan.
    int tk = 0, epoch = 0;
    pdl_wait();
    if (lane == 0) {
      tk = atomicAdd(ticket, 1);
      // The partial flag of this launch: epoch + 1, never 0 (a fresh workspace is all zeros).  Stable while any
      // CTA may still read it: only the LAST scheduler to finish advances it, and every scheduler reads it here.
      epoch = *reinterpret_cast<volatile int*>(ticket + 2);
    }
    build_plan(p.seq_lens, p.batch, hs, p.target_units, kMinUnitBlocks * (128 / HD), p.guided_k * (int)gridDim.x, s_len, s_c,
               s_ucum, lane);
    if (lane == 0) scratch[kEpochSlot] = epoch + 1;                    // (the load's round trip hid behind the plan)
    const int NU = s_ucum[p.batch] * hs;
    if (lane == 0) dbg_stamp(6);                                       // plan built
    // `nx_*`: the unit claimed next - located, with its first 64 block-table entries in registers.  A ticket is
    // claimed only when the CTA's backlog (blocks published but not yet consumed + blocks of the current unit
    // not yet published) has become short: every ticket a CTA holds beyond what keeps its TMA rings busy is
    // work the other SMs cannot take at the tail of the kernel (round 1 / early round 2 claimed two units
    // ahead: the consumers' end times spread over 7 us = 1.3 tail units).
    int nx_s = 0, nx_h = 0, nx_b0 = 0, nx_n = 0, nx_nch = 0, nx_t = 0, nx_c = 0, nx_bt0 = 0, nx_bt1 = 0;
    auto setup_next = [&](bool first) {
      if (!first && lane == 0) tk = atomicAdd(ticket, 1);
      nx_t = __shfl_sync(FULL, tk, 0);
      if (nx_t >= NU) { nx_n = 0; if (lane == 0) dbg_stamp(9); return; }     // tickets exhausted
      const PaCursor cu = pa_locate(s_ucum, p.batch, hs, nx_t);     // (s, h, chunk, chunks)
      nx_s = cu.s; nx_h = cu.h; nx_c = cu.b; nx_nch = cu.n;
      const int C = s_c[cu.s];
      nx_b0 = cu.b * C;
      nx_n = min(C, pa_ceil_div(s_len[cu.s], kBS) - nx_b0);
      const int32_t* row = p.block_tables + (int64_t)cu.s * p.bt_stride + nx_b0;
      nx_bt0 = lane < nx_n ? __ldg(row + lane) : 0;
      nx_bt1 = lane + 32 < nx_n ? __ldg(row + lane + 32) : 0;
    };
    setup_next(true);
    cta_sync();                                                // mbarriers are initialised

    // sequences with no tokens: defined result is zeros (the reference raises, paged_attention.py:93)
    for (int s = blockIdx.x; s < p.batch; s += gridDim.x)
      if (s_len[s] == 0) {
        for (int e = lane * 2; e < p.num_q_heads * HD; e += 64) {
          const int64_t off = (int64_t)s * p.o_stride_b + (int64_t)(e / HD) * p.o_stride_h + e % HD;
          *reinterpret_cast<uint32_t*>(obase + off) = 0u;
          peer_store32(p, off, 0u);
        }
        __syncwarp();
        if (lane == 0) peer_signal(p, (uint32_t)hs);
      }
    const __nv_bfloat16* qbase = static_cast<const __nv_bfloat16*>(p.q);
    int u = 0;                                                       // descriptors published so far
    int rel = 0, outstanding = 0;                                    // descriptors [rel, u) not known released
    int* const hist = scratch;                                       // blocks of descriptor slot i
    auto wait_release = [&]() {
      if (lane == 0) mbar_wait(bar_dempty + (rel % kDescRing) * 8, (rel / kDescRing) & 1);
      __syncwarp();
      outstanding -= hist[rel % kDescRing];
      ++rel;
    };
    // A unit is published in descriptors of <= kDescBlocks blocks, one at a time, while fewer than `claim_limit`
    // published blocks are unconsumed: deep enough to keep the 8 x R stage rings full, shallow enough that the
    // release of one descriptor is a fine-grained progress signal.  The next ticket is claimed when the backlog
    // (outstanding + unpublished rest of the current unit) drops below claim_limit + kDescBlocks.
    // claim_limit is quoted in 8-KiB blocks (head_dim 128); a head_dim-64 block is half the bytes
    const int pub_cap = p.claim_limit * (128 / HD), claim_at = pub_cap + kDescBlocks;
    int cs = 0, ch = 0, b0 = 0, n = 0, nch = 0, t = 0, c = 0, bt0 = 0, bt1 = 0, j0 = 0, last_valid = kBS;
    bool have_cur = false, next_ready = true;                        // the first unit is already in nx_*
    for (;;) {
      const int unpub = have_cur ? n - j0 : 0;
      if (!next_ready && outstanding + unpub < claim_at) {
        setup_next(false);
        next_ready = true;
        continue;
      }
      if (have_cur) {
        if (outstanding >= pub_cap || u - rel >= kDescRing) { wait_release(); continue; }
        const int nd = min(kDescBlocks, n - j0);
        const int o = j0 & 63;
        if (o == 0 && j0 > 0) {                                      // the unit's next 64 block-table entries
          const int32_t* row = p.block_tables + (int64_t)cs * p.bt_stride + b0 + j0;
          bt0 = lane < n - j0 ? __ldg(row + lane) : 0;
          bt1 = lane + 32 < n - j0 ? __ldg(row + lane + 32) : 0;
        }
        const int ds = u % kDescRing;
        hist[ds] = nd;
        outstanding += nd;
#pragma unroll
        for (int i0 = 0; i0 < kDescBlocks; i0 += 32) {
          const int e = o + i0 + lane;                               // entry of the 64-entry chunk
          const int v0 = __shfl_sync(FULL, bt0, e & 31), v1 = __shfl_sync(FULL, bt1, e & 31);
          if (i0 + lane < kDescBlocks) bt_ring[ds * kDescBlocks + i0 + lane] = (e & 32) ? v1 : v0;
        }
        __syncwarp();
        if (lane == 0) {
          UnitDesc d;
          d.s = cs; d.h = ch; d.n = nd;
          d.flags = (j0 == 0 ? kFirst : 0) | (j0 + nd == n ? kLast : 0);
          d.last_valid = last_valid; d.nch = nch; d.slot = t; d.slot0 = t - c;
          desc[ds] = d;
          mbar_arrive(bar_dfull + ds * 8);                           // the producers may start loading K / V
          if (u == 0) dbg_stamp(7);                                  // first descriptor published
          if (j0 == 0) {                                             // Q rides its own barrier: needed at math time only
            mbar_expect_tx(bar_qfull + ds * 8, q_bytes);
            bulk_load(qring + ds * q_bytes, qbase + (int64_t)cs * p.q_stride_b + (int64_t)(ch * g) * p.q_stride_h,
                      q_bytes, bar_qfull + ds * 8);
          } else {
            mbar_arrive(bar_qfull + ds * 8);
          }
        }
        ++u;
        j0 += nd;
        if (j0 == n) have_cur = false;
        continue;
      }
      if (!next_ready) { wait_release(); continue; }                 // backlog still long: wait for consumption
      if (nx_n == 0) break;                                          // tickets exhausted, everything published
      cs = nx_s; ch = nx_h; b0 = nx_b0; n = nx_n; nch = nx_nch; t = nx_t; c = nx_c; bt0 = nx_bt0; bt1 = nx_bt1;
      const int len = s_len[cs];
      last_valid = (b0 + n == pa_ceil_div(len, kBS)) ? len - (b0 + n - 1) * kBS : kBS;
      j0 = 0;
      have_cur = true;
      next_ready = false;
    }
    while (u - rel >= kDescRing) wait_release();
    if (lane == 0) {
      const int ds = u % kDescRing;
      desc[ds].n = 0;                                   // end of stream
      mbar_arrive(bar_dfull + ds * 8);
      // the last CTA out re-arms the ticket for the next launch (the header stays zero between launches)
      if (atomicAdd(ticket + 1, 1) == (int)gridDim.x - 1) {
        ticket[0] = 0;
        ticket[1] = 0;
        ticket[2] = (epoch + 1) & 0x3fffffff;
      }
    }
    return;
  }

    // =========================== finisher warps ==========================================
    pdl_wait();
    cta_sync();                                                // mbarriers are initialised
    // Warp 17, the MERGER, takes the units of the CTA in the order the consumers finish them: it merges the 8
    // consumer warps' (m, l, O) from the hand-off area and frees the area at once - the consumers only ever wait
    // for THIS part, ~0.3 us per unit (per-head weights are computed once, not per element).  A unit that is its
    // segment's only one is normalised and written to `out`.  Any other unit's merged state goes to the unit's
    // partial slot as self-validating 16-byte words {value, flag, value, flag} (flag = this launch's epoch):
    // plain stores, no fence.  The unit also checks in at its segment's counter (the round trip hides behind the
    // merge), and when it is the last to check in the merger posts a FOLD JOB and goes back to the hand-off.
    // Warps 18 / 19, the FOLDERS, work on every fold job together, each on half of the GQA heads: they fold the
    // segment's partials in chunk order with D partials' loads in flight per lane (one L2 round trip per D
    // partials instead of one per partial), polling each word until its flag shows this launch, and write `out`.
    // Earlier in round 2 each of three finisher warps did merge AND fold for "its" units: a fold of n partials
    // was n dependent round trips (~0.7 us each while HBM is saturated) during which every third unit's hand-off
    // stalled - the consumers of a CTA with 16 units waited 16 us for the hand-off area (32 units: 25 us).
    // The result does not depend on which SM merges or folds: every element is folded in chunk order.
    const int fw = warp - kSchedWarp - 1;
    const int fli = scratch[kEpochSlot];
    const float flf = __int_as_float(fli);
    float4* const W = reinterpret_cast<float4*>(p.part_o);
    float4* const W2 = reinterpret_cast<float4*>(p.part_o2);   // level-2 partials (folds of kFoldGroup partials)
    int4* const jobs = mjob + 2;                                // [kJobRing] (s, h, nch, slot0); nch == 0: end of stream
    constexpr int WPR = HD / 2 + 1;                             // words per partial row: (m, l) + HD / 2 value pairs
    constexpr int IPH = HD / 64;                                // 32-element rows of value words per head
    if (fw == 0) {
      int njobs = 0;
      auto post = [&](int4 job) {
        const int js = njobs % kJobRing;
        if (lane == 0) {
          mbar_wait(bar_jempty + js * 8, ((njobs / kJobRing) & 1) ^ 1);
          jobs[js] = job;
          mbar_arrive(bar_jfull + js * 8);
        }
        ++njobs;
      };
      for (int mi = 0;; ++mi) {
        mbar_wait(bar_mfull, mi & 1);
        const int4 j0 = mjob[0], j1 = mjob[1];
        const int cs = j0.x, ch = j0.y, nch = j0.z, slot = j0.w, slot0 = j1.x, end = j1.y;
        if (!end) {
          for (int head = 0; head < g; ++head) {
            float mw[kConsumers], wg[kConsumers];
            float M = -INFINITY, Lsum = 0.f;
#pragma unroll
            for (int w = 0; w < kConsumers; ++w) {
              const float2 ml = merge_ml[w * 8 + head];
              mw[w] = ml.x;
              wg[w] = ml.y;
              M = fmaxf(M, ml.x);
            }
#pragma unroll
            for (int w = 0; w < kConsumers; ++w) {
              const float l = wg[w];
              wg[w] = 0.f;
              if (mw[w] != -INFINITY) {
                wg[w] = exp2f(mw[w] - M);
                Lsum = fmaf(l, wg[w], Lsum);
              }
            }
            const float inv = 1.f / Lsum;
            float4* row = W + ((int64_t)slot * g + head) * WPR;
#pragma unroll
            for (int r = 0; r < IPH; ++r) {
              const int dp = (r * 32 + lane) * 2;
              float o0 = 0.f, o1 = 0.f;
#pragma unroll
              for (int w = 0; w < kConsumers; ++w) {
                if (mw[w] != -INFINITY) {                       // a warp without blocks left its rows unwritten
                  const float2 v = *reinterpret_cast<const float2*>(merge_o + (w * g + head) * OPITCH + dp);
                  o0 = fmaf(v.x, wg[w], o0);
                  o1 = fmaf(v.y, wg[w], o1);
                }
              }
              if (nch == 1) {
                const int64_t off = (int64_t)cs * p.o_stride_b + (int64_t)(ch * g + head) * p.o_stride_h + dp;
                const uint32_t w0 = pack_bf16(o0 * inv, o1 * inv);
                *reinterpret_cast<uint32_t*>(obase + off) = w0;
                peer_store32(p, off, w0);
              } else {
                __stcg(row + 1 + (dp >> 1), make_float4(o0, flf, o1, flf));
                if (dp == 0) __stcg(row, make_float4(M, flf, Lsum, flf));
              }
            }
          }
        }
        __syncwarp();                                             // hand-off area read
        if (lane == 0) mbar_arrive(bar_mempty);                   // the consumers may drop their next unit (or end marker)
        if (!end && lane == 0) dbg_stamp(10);                     // (last) unit merged
        if (end) {
          post(make_int4(0, 0, 0, 0));
          break;
        }
        if (nch == 1) {
          if (lane == 0) peer_signal(p, 1u);                      // whole segment in one unit: its rows are stored
          continue;
        }
        // The CTA that ran a split segment's FIRST unit folds the segment: its folders start polling the partial
        // slots right away and finish one L2 round trip after the last partial lands.  (Electing the LAST unit to
        // check in at an atomic counter - rounds 1 and early 2 - piles the folds of the final segments onto the
        // few CTAs that finish last: with 16-block tail units one CTA folded for 13 us after its consumers' end.)
        // Units of a segment hold consecutive tickets, so they are claimed by different CTAs at about the same
        // time and the folds spread evenly over all SMs' folder warps.
        // A segment of more than kFoldGroup partials is folded in two levels (a fold of n partials costs n / DEPTH
        // dependent L2 round trips on ONE warp pair): the CTA that ran chunk 8 j folds partials [8 j, 8 j + 8) into a
        // level-2 partial (second half of the partial area, same slot index), and the last group's CTA folds those.
        // job = (s, h, count | stride << 8 | source level << 16 | writes a level-2 partial << 17, first slot).
        const int c = slot - slot0;
        if (nch <= kFoldGroup) {
          if (c == 0) post(make_int4(cs, ch, nch | (1 << 8), slot0));
        } else if (c % kFoldGroup == 0) {
          post(make_int4(cs, ch, min(kFoldGroup, nch - c) | (1 << 8) | (1 << 17), slot));
          // The final fold belongs to the CTA of the LAST group: every job then depends only on units and on jobs
          // posted for smaller tickets, and a CTA posts in ticket order, so its FIFO can never hold a job behind
          // one that waits for it (with the chunk-0 CTA folding, a CTA that also ran chunk 8 of the same segment
          // queued the group fold BEHIND the final fold that polls for its result: deadlock).
          if (c + kFoldGroup >= nch)
            post(make_int4(cs, ch, pa_ceil_div(nch, kFoldGroup) | (kFoldGroup << 8) | (1 << 16), slot0));
        }
      }
    } else {
      constexpr int FU = 2 * IPH;                                 // rows (value words per lane) of one pass: two heads
      constexpr int HPG = FU / IPH;                               // heads of one pass
      constexpr int DEPTH = PA_FOLD_DEPTH;                    // partials in flight per lane: DEPTH * (FU + HPG) words
      const int h0 = fw == 1 ? 0 : (g + 1) / 2, h1 = fw == 1 ? (g + 1) / 2 : g;   // this folder's GQA heads
      auto ready = [&](const float4& w) { return __float_as_int(w.y) == fli && __float_as_int(w.w) == fli; };
      for (int j = 0;; ++j) {
        const int js = j % kJobRing;
        mbar_wait(bar_jfull + js * 8, (j / kJobRing) & 1);
        const int4 job = jobs[js];
        __syncwarp();
        if (lane == 0) mbar_arrive(bar_jempty + js * 8);
        const int cs = job.x, ch = job.y, nch = job.z & 0xff, kstride = (job.z >> 8) & 0xff, slot0 = job.w;
        const bool to_partial = (job.z >> 17) & 1;
        if (nch == 0) break;
        if (fw == 1 && lane == 0) dbg_stamp(11);                  // (last) fold job taken
        const float4* const Wsrc = ((job.z >> 16) & 1) ? W2 : W;
        for (int hg = h0; hg < h1; hg += HPG) {
          float M[HPG], Ls[HPG], o[FU][2];
#pragma unroll
          for (int h = 0; h < HPG; ++h) { M[h] = -INFINITY; Ls[h] = 0.f; }
#pragma unroll
          for (int u = 0; u < FU; ++u) o[u][0] = o[u][1] = 0.f;
          for (int k0 = 0; k0 < nch; k0 += DEPTH) {
            float4 wv[DEPTH][FU], wm[DEPTH][HPG];
            // Wait for this batch of partials on ONE sentinel word per partial (lane d polls the LAST word its writer
            // stores of the heads folded in this pass, partial k0 + d): the
            // folders start polling long before the partials land, and a re-poll of every word of every partial
            // would put terabytes per second of L2 traffic next to the KV stream.  (Re-polling word by word, as
            // rounds 1 and early 2 did, is worse still: 24 dependent round trips per batch - a 13 us fold tail.)
            for (;;) {
              bool ok = true;
              if (lane < DEPTH && k0 + lane < nch)
                ok = ready(ld_poll(Wsrc + ((int64_t)(slot0 + (k0 + lane) * kstride) * g + min(hg + HPG, h1)) * WPR - 1));
              if (__all_sync(FULL, ok)) break;
              __nanosleep(100);
            }
            if (fw == 1 && lane == 0) dbg_stamp(12);              // (last) batch of partials seen
            // all of the batch's words in flight at once; the stores of a partial are not ordered among
            // themselves, so every word is validated and the whole batch is reloaded in the (rare) miss
            for (;;) {
              bool ok = true;
#pragma unroll
              for (int d = 0; d < DEPTH; ++d) {
                const float4* const base = Wsrc + (int64_t)(slot0 + (k0 + d) * kstride) * g * WPR;
#pragma unroll
                for (int u = 0; u < FU; ++u) {
                  const int head = hg + u / IPH;
                  wv[d][u] = make_float4(0.f, flf, 0.f, flf);
                  if (u % IPH == 0) wm[d][u / IPH] = make_float4(-INFINITY, flf, 0.f, flf);
                  if (head < h1 && k0 + d < nch) {
                    const float4* src = base + (int64_t)head * WPR;
                    wv[d][u] = ld_poll(src + 1 + (u % IPH) * 32 + lane);
                    if (u % IPH == 0) wm[d][u / IPH] = ld_poll(src);
                  }
                }
              }
#pragma unroll
              for (int d = 0; d < DEPTH; ++d) {
#pragma unroll
                for (int u = 0; u < FU; ++u) {
                  ok = ok && ready(wv[d][u]);
                  if (u % IPH == 0) ok = ok && ready(wm[d][u / IPH]);
                }
              }
              if (__all_sync(FULL, ok)) break;
              __nanosleep(60);
            }
#pragma unroll
            for (int d = 0; d < DEPTH; ++d) {
              if (k0 + d < nch) {
#pragma unroll
                for (int h = 0; h < HPG; ++h) {
                  // every partial holds >= 1 valid token (finite maximum); a head beyond h1 keeps weight 0
                  const float Mn = fmaxf(M[h], wm[d][h].x);
                  const float a = Mn == -INFINITY ? 0.f : exp2f(M[h] - Mn);
                  const float bw = Mn == -INFINITY ? 0.f : exp2f(wm[d][h].x - Mn);
                  Ls[h] = Ls[h] * a + wm[d][h].z * bw;
                  M[h] = Mn;
#pragma unroll
                  for (int i = 0; i < IPH; ++i) {
                    o[h * IPH + i][0] = o[h * IPH + i][0] * a + wv[d][h * IPH + i].x * bw;
                    o[h * IPH + i][1] = o[h * IPH + i][1] * a + wv[d][h * IPH + i].z * bw;
                  }
                }
              }
            }
          }
#pragma unroll
          for (int u = 0; u < FU; ++u) {
            const int head = hg + u / IPH;
            if (head < h1 && to_partial) {                      // level-2 partial: same row format as a unit's
              float4* row = W2 + ((int64_t)slot0 * g + head) * WPR;
              __stcg(row + 1 + (u % IPH) * 32 + lane, make_float4(o[u][0], flf, o[u][1], flf));
              if (u % IPH == 0 && lane == 0) __stcg(row, make_float4(M[u / IPH], flf, Ls[u / IPH], flf));
            } else if (head < h1) {
              const float inv = 1.f / Ls[u / IPH];
              const int dp = ((u % IPH) * 32 + lane) * 2;
              const int64_t off = (int64_t)cs * p.o_stride_b + (int64_t)(ch * g + head) * p.o_stride_h + dp;
              const uint32_t w0 = pack_bf16(o[u][0] * inv, o[u][1] * inv);
              *reinterpret_cast<uint32_t*>(obase + off) = w0;
              peer_store32(p, off, w0);
            }
          }
        }
        if (!to_partial && p.local_sig != nullptr) {
          asm volatile("bar.sync 2, 64;" ::: "memory");           // both folders have stored their heads of the segment
          if (fw == 1 && lane == 0) peer_signal(p, 1u);
        }
      }
    }
    if (p.dbg && lane == 0) {                                   // the last of the three warps to leave
      unsigned long long t;
      asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
      atomicMax(reinterpret_cast<unsigned long long*>(p.dbg) + kDbgWords * (int64_t)blockIdx.x + 4, t);
    }
    return;
  }

  if (warp >= kConsumers) {
    // =========================== producer warps ==========================================
    // Producer c feeds consumer c: block j of a descriptor belongs to consumer j % 8.  One lane walks the
    // descriptor ring, takes the next free stage of the consumer's private ring and issues the block's two
    // tensor loads.  Waiting here - for a descriptor, for a free stage, or in the issue path itself when
    // the memory system pushes back - delays nobody's math.
    asm volatile("setmaxnreg.dec.sync.aligned.u32 24;");
    pdl_wait();
    cta_sync();                                                // mbarriers are initialised
    if (lane != 0) return;
    const int c = warp - kConsumers;
    uint64_t policy;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(policy));
    const bool head3 = p.pos_head == 3;                         // which of tensor-map dims 3 / 4 is the head axis
    const uint32_t my_ring = smem_base + L.ring + c * (R * STAGE_BYTES);
    const uint32_t my_full = bar_full + c * (kMaxStages * 8), my_empty = bar_empty + c * (kMaxStages * 8);
    int stage = 0, phase = 0;
    bool first = true;
    for (int pu = 0;; ++pu) {
      const int ds = pu % kDescRing;
      mbar_wait(bar_dfull + ds * 8, (pu / kDescRing) & 1);
      const int n = desc[ds].n;
      if (n == 0) break;
      const int ph = desc[ds].h;
      for (int j = c; j < n; j += kConsumers) {
        const int phys = bt_ring[ds * kDescBlocks + j];
        mbar_wait(my_empty + stage * 8, phase ^ 1);             // first pass over the ring: free by construction
        const uint32_t bar = my_full + stage * 8;
        const uint32_t dst = my_ring + stage * STAGE_BYTES;
        const int c3 = head3 ? ph : phys, c4 = head3 ? phys : ph;
        mbar_expect_tx(bar, STAGE_BYTES);
        if (first && c == 0) dbg_stamp(8);                      // this CTA's first TMA load goes out
        first = false;
        tma_load_5d(dst, &tm_k, bar, c3, c4, policy);
        tma_load_5d(dst + TILE_BYTES, &tm_v, bar, c3, c4, policy);
        if (++stage == R) { stage = 0; phase ^= 1; }
      }
      mbar_arrive(bar_dempty + ds * 8);                         // this descriptor's block ids are no longer needed here
    }
    return;
  }

  // ============================= consumer warps ============================================
  asm volatile("setmaxnreg.inc.sync.aligned.u32 136;");
  pdl_wait();
  cta_sync();                                                  // mbarriers are initialised
  const uint32_t my_ring = smem_base + L.ring + warp * (R * STAGE_BYTES);
  const uint32_t my_full = bar_full + warp * (kMaxStages * 8), my_empty = bar_empty + warp * (kMaxStages * 8);
  // per-lane constant pieces of the ldmatrix addresses
  const int k_tok = (lane & 7) + ((lane >> 3) & 1) * 8;       // K: matrices 0/1 = tokens 0-7 / 8-15
  const int k_csel = lane >> 4;                               //    matrices 2/3 = next 8 dims
  const int v_tok = (lane & 7) + ((lane >> 4) & 1) * 8;       // V^T: matrices 2/3 = tokens 8-15
  const int v_csel = (lane >> 3) & 1;                         //      matrices 1/3 = next 8 dims
  const uint32_t k_row = k_tok * 128, v_row = v_tok * 128;
  const int k_x = k_tok & 7, v_x = v_tok & 7;
  const int hq_b = lane >> 2;                                 // query head of this lane's B-fragment column
  const int hq_c = (lane & 3) * 2;                            // first of this lane's two accumulator heads
  float* my_mo = merge_o + warp * g * OPITCH;

  int c_stage = 0, c_phase = 0;
  uint32_t qf[KSTEPS][2];
  float acc[KSTEPS][4];
  float m_run[2], l_run[2];
  int units_done = 0;
  bool dbg_seen = false;
  long long dbg_desc_wait = 0, dbg_data_wait = 0, dbg_hand_wait = 0;   // cycles warp 0 waited for units / KV data / the hand-off area

  for (int cu = 0;; ++cu) {
    const int ds = cu % kDescRing;
    {
      const long long w0 = p.dbg ? clock64() : 0;
      mbar_wait(bar_dfull + ds * 8, (cu / kDescRing) & 1);
      if (p.dbg && cu > 0) dbg_desc_wait += clock64() - w0;
    }
    const int4 d0 = reinterpret_cast<const int4*>(desc + ds)[0];      // s, h, n, flags
    const int4 d1 = reinterpret_cast<const int4*>(desc + ds)[1];      // last_valid, nch, slot, slot0
    if (d0.z == 0) break;
    const int cs = d0.x, ch = d0.y, n = d0.z, flags = d0.w;
    const int last_valid = d1.x, nch = d1.y, slot = d1.z, slot0 = d1.w;

    if (flags & kFirst) {
      mbar_wait(bar_qfull + ds * 8, (cu / kDescRing) & 1);
      // Q^T B-fragments: b0 = Q[head = lane/4][dim = ks*16 + (lane%4)*2 + {0,1}], b1 = dim + 8
      const uint32_t qs = qring + ds * q_bytes + hq_b * (HD * 2) + (lane & 3) * 4;
#pragma unroll
      for (int ks = 0; ks < KSTEPS; ++ks) {
        uint32_t b0 = 0, b1 = 0;
        if (hq_b < g) {
          asm volatile("ld.shared.b32 %0, [%1];" : "=r"(b0) : "r"(qs + ks * 32));
          asm volatile("ld.shared.b32 %0, [%1];" : "=r"(b1) : "r"(qs + ks * 32 + 16));
        }
        qf[ks][0] = b0;
        qf[ks][1] = b1;
      }
      m_run[0] = m_run[1] = -INFINITY;
      l_run[0] = l_run[1] = 0.f;
#pragma unroll
      for (int mt = 0; mt < KSTEPS; ++mt) acc[mt][0] = acc[mt][1] = acc[mt][2] = acc[mt][3] = 0.f;
    }

    for (int cj = warp; cj < n; cj += kConsumers) {
      {
        const long long w0 = p.dbg ? clock64() : 0;
        mbar_wait(my_full + c_stage * 8, c_phase);
        if (p.dbg && dbg_seen) dbg_data_wait += clock64() - w0;
      }
      if (p.dbg && threadIdx.x == 0 && !dbg_seen) dbg_stamp(5);      // this CTA's first KV block has landed
      dbg_seen = true;
      const uint32_t kt = my_ring + c_stage * STAGE_BYTES;
      const uint32_t vt = kt + TILE_BYTES;

      // ---- S^T = K . Q^T ----------------------------------------------------------
      float sc[2][4];
#pragma unroll
      for (int c = 0; c < 2; ++c) sc[c][0] = sc[c][1] = sc[c][2] = sc[c][3] = 0.f;
#pragma unroll
      for (int ks = 0; ks < KSTEPS; ++ks) {
        uint32_t a[4];
        const int chunk = (ks & 3) * 2 + k_csel;
        ldsm_x4(a, kt + (ks >> 2) * HALF_BYTES + k_row + ((chunk ^ k_x) << 4));
        mma_bf16(sc[ks & 1], a, qf[ks][0], qf[ks][1]);
      }

      // ---- online softmax over this block's 16 tokens ----------------------------------
      const int valid = ((flags & kLast) && cj == n - 1) ? last_valid : kBS;   // < 16 only in a sequence's last block
      const int t0 = lane >> 2;
      uint32_t pb[2];
      bool need_rescale = false;
      float alpha[2];
      {
        float s0[2], s1[2];
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          s0[j] = (sc[0][j] + sc[1][j]) * p.scale_log2e;           // token t0,     head hq_c + j
          s1[j] = (sc[0][j + 2] + sc[1][j + 2]) * p.scale_log2e;   // token t0 + 8
          if (t0 >= valid) s0[j] = -INFINITY;
          if (t0 + 8 >= valid) s1[j] = -INFINITY;
          float mx = fmaxf(s0[j], s1[j]);
          mx = fmaxf(mx, __shfl_xor_sync(FULL, mx, 4));
          mx = fmaxf(mx, __shfl_xor_sync(FULL, mx, 8));
          mx = fmaxf(mx, __shfl_xor_sync(FULL, mx, 16));
          const float m_new = fmaxf(m_run[j], mx);       // finite: token 0 of the block is always valid
          alpha[j] = exp2f(m_run[j] - m_new);            // exp2(-inf) = 0 on the first block
          need_rescale |= (m_new != m_run[j]);
          m_run[j] = m_new;
          s0[j] = exp2f(s0[j] - m_new);
          s1[j] = exp2f(s1[j] - m_new);
        }
        const uint32_t p0 = pack_bf16(s0[0], s0[1]);       // token t0:     heads (hq_c, hq_c + 1)
        const uint32_t p1 = pack_bf16(s1[0], s1[1]);       // token t0 + 8
        // the normaliser sums the SAME bf16-rounded weights the PV product uses
        l_run[0] = l_run[0] * alpha[0] + bf16lo(p0) + bf16lo(p1);
        l_run[1] = l_run[1] * alpha[1] + bf16hi(p0) + bf16hi(p1);
        pb[0] = movm_t(p0);                                // -> B fragment: k = tokens 0-7,  n = head
        pb[1] = movm_t(p1);                                //                k = tokens 8-15
      }
      if (__any_sync(FULL, need_rescale)) {
#pragma unroll
        for (int mt = 0; mt < KSTEPS; ++mt) {
          acc[mt][0] *= alpha[0];
          acc[mt][1] *= alpha[1];
          acc[mt][2] *= alpha[0];
          acc[mt][3] *= alpha[1];
        }
      }

      // A V row of a token past seq_len may hold anything (Inf/NaN bit patterns in a recycled
      // block); 0 * NaN would poison the product, so clear those rows in shared memory.
      if (valid < kBS) {
        for (int e = lane; e < (kBS - valid) * (HD / 8); e += 32) {
          const int row = valid + e / (HD / 8), c = e % (HD / 8);
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
        mma_bf16(acc[mt], a, pb[0], pb[1]);
      }

      // stage consumed by every lane -> the producer may refill it
      __syncwarp();
      if (lane == 0) mbar_arrive(my_empty + c_stage * 8);
      if (++c_stage == R) { c_stage = 0; c_phase ^= 1; }
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(bar_dempty + ds * 8);        // descriptor slot handed back
    if (!(flags & kLast)) continue;

    // ---- end of the unit: drop this warp's (m, l, O) into the hand-off area; the finishers merge ---------
    float lsum[2];
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      float l = l_run[j];
      l += __shfl_xor_sync(FULL, l, 4);
      l += __shfl_xor_sync(FULL, l, 8);
      l += __shfl_xor_sync(FULL, l, 16);
      lsum[j] = l;
    }
    {
      const long long w0 = p.dbg ? clock64() : 0;
      mbar_wait(bar_mempty, (units_done & 1) ^ 1);            // the previous unit's state has been merged
      if (p.dbg) dbg_hand_wait += clock64() - w0;
    }
    if (m_run[0] != -INFINITY) {          // warp-uniform: this warp had at least one block of the unit
#pragma unroll
      for (int mt = 0; mt < KSTEPS; ++mt) {
        const int d = mt * 16 + (lane >> 2);
        if (hq_c < g) { my_mo[hq_c * OPITCH + d] = acc[mt][0]; my_mo[hq_c * OPITCH + d + 8] = acc[mt][2]; }
        if (hq_c + 1 < g) { my_mo[(hq_c + 1) * OPITCH + d] = acc[mt][1]; my_mo[(hq_c + 1) * OPITCH + d + 8] = acc[mt][3]; }
      }
    }
    if ((lane >> 2) == 0) {
      if (hq_c < g) merge_ml[warp * 8 + hq_c] = make_float2(m_run[0], lsum[0]);
      if (hq_c + 1 < g) merge_ml[warp * 8 + hq_c + 1] = make_float2(m_run[1], lsum[1]);
    }
    if (threadIdx.x == 0) {
      mjob[0] = make_int4(cs, ch, nch, slot);
      mjob[1] = make_int4(slot0, 0, 0, 0);
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(bar_mfull);
    ++units_done;
  }
  // end of stream: one marker for the merger (which passes it on to the folders), through the same hand-off protocol
  const int units_total = units_done;
  {
    mbar_wait(bar_mempty, (units_done & 1) ^ 1);
    if (threadIdx.x == 0) mjob[1] = make_int4(0, 1, 0, 0);
    __syncwarp();
    if (lane == 0) mbar_arrive(bar_mfull);
    ++units_done;
  }

  if (p.dbg && threadIdx.x == 0) {
    unsigned long long t1;
    unsigned smid;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t1));
    asm volatile("mov.u32 %0, %smid;" : "=r"(smid));
    long long* d = p.dbg + kDbgWords * (int64_t)blockIdx.x;
    d[0] = smid; d[1] = (long long)dbg_t0; d[2] = (long long)t1; d[3] = units_total;
    d[13] = dbg_desc_wait; d[14] = dbg_data_wait; d[15] = dbg_hand_wait;
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
  int32_t nh, hd;
  int32_t box_halves, nb;
  int32_t dev, pad;
  bool operator==(const MapKey& o) const { return memcmp(this, &o, sizeof(MapKey)) == 0; }
};
struct MapEntry {
  MapKey key;
  CUtensorMap map;
  int pos_head;
  bool used;
  uint64_t stamp;                   // last use (LRU)
};
// Hashed, set-associative LRU cache: a 32-layer model has 64 decode maps + 64 prefill maps (K and V per
// layer, two box shapes) and an engine loop cycles through all of them every step - a round-robin
// cache smaller than that working set would re-encode on every launch.  1024 entries in 8-way sets.
constexpr int kMapSets = 128, kMapWays = 8;
MapEntry g_maps[kMapSets * kMapWays];
uint64_t g_map_clock = 0;
std::mutex g_map_mu;

uint32_t map_hash(const MapKey& k) {
  const uint64_t* w = reinterpret_cast<const uint64_t*>(&k);
  uint64_t h = 0x9e3779b97f4a7c15ull;
  for (size_t i = 0; i < sizeof(MapKey) / 8; ++i) h = (h ^ w[i]) * 0xff51afd7ed558ccdull, h ^= h >> 32;
  return (uint32_t)h;
}

// Builds (or fetches) the 5-D tensor map of one K or V plane:
//   dim0 = 64 head_dim columns (128 bytes, SWIZZLE_128B), dim1 = the 16 slots of a block,
//   dim2 = the D / 64 column halves of a row (stride 128 bytes), dims 3, 4 = {head, block} by
//   ascending stride.  One box = {64, 16, D / 64, 1, 1} = the whole [16 x D] tile of one
//   (block, head), landing in shared memory as [half][slot][128 B] - one TMA instruction per tile.
//   box_halves < D / 64 (the prefill kernel passes 1) makes one box a single 64-column half, so the
//   halves of consecutive blocks can be laid out slab-major for UMMA.
int get_map(const void* base, const PaDecodeParams& p, CUtensorMap* out, int* pos_head, int box_halves = 0) {
  MapKey key;
  memset(&key, 0, sizeof(key));
  key.base = base;
  key.sb = p.kv_stride_block; key.sh = p.kv_stride_head; key.ss = p.kv_stride_slot;
  key.nh = p.num_kv_heads; key.hd = p.head_dim;
  if (box_halves <= 0) box_halves = p.head_dim / 64;
  key.box_halves = box_halves;
  key.nb = p.num_blocks;
  if (cudaGetDevice(&key.dev) != cudaSuccess) { cudaGetLastError(); key.dev = -1; }
  static_assert(sizeof(MapKey) % 8 == 0, "hashed as 64-bit words");
  static_assert(kThreads * 96 == 256 * 136 + 256 * 24 + 128 * 160, "setmaxnreg budgets must add up to the launch allocation");
  std::lock_guard<std::mutex> lock(g_map_mu);
  MapEntry* set = g_maps + (map_hash(key) % kMapSets) * kMapWays;
  MapEntry* victim = set;
  for (int i = 0; i < kMapWays; ++i) {
    if (set[i].used && set[i].key == key) {
      set[i].stamp = ++g_map_clock;
      *out = set[i].map; *pos_head = set[i].pos_head;
      return 0;
    }
    if (!set[i].used) { if (victim->used) victim = set + i; }
    else if (victim->used && set[i].stamp < victim->stamp) victim = set + i;
  }
  EncodeTiledFn enc = get_encode_fn();
  if (!enc) return -5;
  // The block axis is bounded by the cache's own block count: a block id outside [0, NB) (a
  // corrupted block table) is zero-filled by TMA instead of reading memory outside the cache.
  if (p.num_blocks <= 0) return -1;
  const bool head_first = p.kv_stride_head <= p.kv_stride_block;
  const uint64_t halves = (uint64_t)p.head_dim / 64;
  cuuint64_t gdim[5] = {64, (cuuint64_t)kBS, halves, 0, 0};
  cuuint64_t gstr[4] = {(cuuint64_t)p.kv_stride_slot * 2, 128, 0, 0};
  const cuuint64_t head_dim_sz = (cuuint64_t)p.num_kv_heads, blk_dim_sz = (cuuint64_t)p.num_blocks;
  if (head_first) {
    gdim[3] = head_dim_sz; gstr[2] = (cuuint64_t)p.kv_stride_head * 2;
    gdim[4] = blk_dim_sz;  gstr[3] = (cuuint64_t)p.kv_stride_block * 2;
  } else {
    gdim[3] = blk_dim_sz;  gstr[2] = (cuuint64_t)p.kv_stride_block * 2;
    gdim[4] = head_dim_sz; gstr[3] = (cuuint64_t)p.kv_stride_head * 2;
  }
  cuuint32_t box[5] = {64, (cuuint32_t)kBS, (cuuint32_t)box_halves, 1, 1};
  cuuint32_t estr[5] = {1, 1, 1, 1, 1};
  CUtensorMap m;
  CUresult r = enc(&m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, const_cast<void*>(base), gdim, gstr, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return -5;
  MapEntry& e = *victim;
  e.key = key; e.map = m; e.pos_head = head_first ? 3 : 4; e.used = true; e.stamp = ++g_map_clock;
  *out = m; *pos_head = e.pos_head;
  return 0;
}

// per-warp ring depth that fits the shared-memory budget
int stages_for(int hd, int g, int batch) {
  const int stage = 2 * (hd / 64) * kBS * 128;
  const int fixed = (int)smem_layout(hd, g, 0, batch).total;
  const int r = (kSmemBudget - fixed) / (kConsumers * stage);
  return r > kMaxStages ? kMaxStages : r;
}

std::atomic<int> g_pdl{1};          // launch with programmatic stream serialisation (pa_decode_set_pdl)

template <int HD>
int launch(const PaDecodeParams& p, const CUtensorMap& mk, const CUtensorMap& mv, int num_sms, cudaStream_t stream) {
  const size_t smem = smem_layout(HD, p.group, p.num_stages, p.batch).total;
  auto kern = pa_decode_ws_kernel<HD>;
  static PaPerDeviceFlag attr;
  const int dev = attr.need();
  if (dev != -2) {
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBudget) != cudaSuccess) return -5;
    attr.done(dev);
  }
  if (smem > (size_t)kSmemBudget) return -2;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(num_sms);
  cfg.blockDim = dim3(kThreads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = g_pdl.load() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, mk, mv, p) == cudaSuccess ? 0 : -5;
}

}  // namespace

int pa_decode_tma_header_ints() { return kHeaderInts; }

int pa_decode_tma_set_pdl(int on) { return g_pdl.exchange(on != 0); }

int pa_get_kv_tensor_map(const void* base, const PaDecodeParams& p, int box_halves, CUtensorMap* out, int* pos_head) {
  return get_map(base, p, out, pos_head, box_halves);
}

// Upper bound on work units (= partial slots) for any seq_lens: sum_s Hkv * ceil(nblk_s / C) <=
// total / C + segments <= target_units + segments.
int64_t pa_decode_tma_max_units(int batch, int num_kv_heads, int target_units) {
  return (int64_t)target_units + (int64_t)batch * num_kv_heads;
}

bool pa_decode_tma_supported(const PaDecodeParams& p, int dtype) {
  if (dtype != 0) return false;                                   // bf16 only
  if (p.block_size != kBS) return false;
  if (p.head_dim != 64 && p.head_dim != 128) return false;
  if (p.group > 8) return false;
  if (p.batch > kMaxBatch) return false;
  if ((reinterpret_cast<uintptr_t>(p.k) | reinterpret_cast<uintptr_t>(p.v)) & 15) return false;
  if ((p.kv_stride_block | p.kv_stride_head | p.kv_stride_slot) & 7) return false;   // 16-byte strides
  if (p.kv_stride_block <= 0 || p.kv_stride_head <= 0 || p.kv_stride_slot <= 0) return false;
  // Q rows of a KV head are fetched with one bulk copy: contiguous heads, 16-byte aligned groups
  if ((reinterpret_cast<uintptr_t>(p.q) & 15) || p.q_stride_h != p.head_dim || (p.q_stride_b & 7)) return false;
  if ((reinterpret_cast<uintptr_t>(p.out) & 3) || ((p.o_stride_b | p.o_stride_h) & 1)) return false;
  if (stages_for(p.head_dim, p.group, p.batch) < 2) return false;
  return true;
}

int pa_launch_decode_tma(const PaDecodeParams& p_in, int num_sms, cudaStream_t stream) {
  PaDecodeParams p = p_in;
  CUtensorMap mk, mv;
  int ph;
  int rc = get_map(p.k, p, &mk, &ph);
  if (rc) return rc;
  rc = get_map(p.v, p, &mv, &ph);
  if (rc) return rc;
  p.pos_head = ph;
  const int fit = stages_for(p.head_dim, p.group, p.batch);
  p.num_stages = (p.num_stages > 0 && p.num_stages < fit) ? p.num_stages : fit;
  if (p.head_dim == 128) return launch<128>(p, mk, mv, num_sms, stream);
  return launch<64>(p, mk, mv, num_sms, stream);
}
