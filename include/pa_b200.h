/*
 * pa_b200.h - C ABI of libpa_b200.so: B200 (sm_100a) paged-attention hot path.
 *
 * This is the drop-in boundary for ONE path of HeduAiDev/vllm-from-scratch:
 * paged-attention decode over a block table, the KV-cache slot write
 * (reshape-and-cache) and block copy / swap.  The reference has no FFI of its
 * own (it is pure Python); the entry points below are what a binding for the
 * module-level functions of 01_paged_attention/paged_attention.py would bind.
 * Each entry cites the reference interface it replaces (paths relative to the
 * reference root).  INTEGRATION.md shows the reference-side stub.
 *
 * Conventions
 *  - plain pointers and sizes only; no torch / C++ types cross this boundary;
 *  - every pointer is a DEVICE pointer unless the name says "host";
 *  - all launches are asynchronous on the caller's `stream`; nothing here
 *    synchronises, allocates device memory or keeps caller pointers;
 *  - return value: 0 (PA_OK) or a negative PaStatus; never throws;
 *  - strides are in ELEMENTS of the cache dtype; the head_dim axis is
 *    contiguous everywhere;
 *  - thread-safety: stateless launchers (one internal mutex guards a small
 *    tensor-map cache); one process per GPU in multi-GPU use.  The process-global
 *    test / bring-up / measurement hooks (kernel-path forcing, split-K tuning,
 *    per-CTA timestamps, per-kernel event timing) are NOT part of this ABI: they
 *    live in pa_b200_debug.h.
 */
#ifndef PA_B200_H_
#define PA_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PA_B200_ABI_VERSION 3

typedef enum PaStatus {
  PA_OK = 0,
  PA_ERR_BAD_ARG = -1,      /* null pointer, negative size, g not integral ...   */
  PA_ERR_UNSUPPORTED = -2,  /* shape / dtype outside the supported set           */
  PA_ERR_MISALIGNED = -3,   /* pointer or stride alignment requirement not met   */
  PA_ERR_WORKSPACE = -4,    /* workspace too small                                */
  PA_ERR_CUDA = -5,         /* a CUDA runtime / driver call failed                */
  PA_ERR_NO_DEVICE = -6,    /* no sm_100 device is current                        */
  PA_ERR_STATIC_LIMIT = -7  /* pa_decode_static: the shape needs more scratch than
                             * PA_DECODE_STATIC_BYTES; use pa_decode + a workspace */
} PaStatus;

typedef enum PaDtype { PA_BF16 = 0, PA_F32 = 1 } PaDtype;

/* A paged KV cache for ONE layer, described by strides so that both reference
 * layouts are accepted without a transposing copy:
 *   'hbsd'  kv_cache[2, NB, Hkv, BS, D]   (01_paged_attention/paged_attention.py:26-30, :69)
 *           k = kv_cache[0], v = kv_cache[1]; stride_block = Hkv*BS*D, stride_head = BS*D, stride_slot = D
 *   'bshd'  k_cache / v_cache [NB, BS, H, D]   (05_mini_vllm/mini_vllm.py:544-552)
 *           stride_block = BS*H*D, stride_slot = H*D, stride_head = D                          */
typedef struct PaKvCache {
  void* k;                /* K plane base (block 0, head 0, slot 0)               */
  void* v;                /* V plane base                                          */
  int64_t stride_block;
  int64_t stride_head;
  int64_t stride_slot;
  int32_t num_blocks;     /* NB                                                    */
  int32_t num_kv_heads;   /* Hkv                                                   */
  int32_t block_size;     /* BS, tokens per block                                  */
  int32_t head_dim;       /* D                                                     */
  int32_t dtype;          /* PaDtype of the cache AND of q / out / key / value     */
  int32_t reserved;
} PaKvCache;

/* ---- introspection ------------------------------------------------------ */
int pa_abi_version(void);
const char* pa_status_string(int status);
/* Number of persistent decode CTAs the fast path runs on the current device
 * (one per SM); the workspace size depends on it.  <0 on error.              */
int pa_decode_num_workers(void);

/* ---- decode (replaces paged_attention_decode, paged_attention.py:67-101,
 *      batched: semantics = that function looped over the batch) -----------
 * q   [batch, num_q_heads, D]   (q_stride_b / q_stride_h in elements)
 * out [batch, num_q_heads, D]   same dtype as q
 * block_tables int32 [batch, bt_stride]; entries beyond ceil(seq_len/BS) are
 *   never dereferenced.  seq_lens int32 [batch]; seq_len == 0 -> zeros.
 * softmax scale is 1/sqrt(D) (paged_attention.py:80).
 * Split-K ("flash-decoding") over (sequence, KV head, context partition) work
 * units, followed by a cross-partition merge.
 * Two launches on the generic path (fp32 caches, odd shapes); ONE on the fast
 * path (bf16, block_size 16, head_dim 64 / 128, g <= 8), whose kernel merges
 * its own split-K partials (the CTA that ran a split segment's first unit folds
 * them, polling self-validating words: no second launch, no atomic, no fence).
 * workspace: pa_decode_workspace_bytes() bytes, 16-byte aligned, owned by the
 *   caller.  Its first 1 MiB is a counter header that must be ZERO on entry:
 *   call pa_decode_workspace_init once after allocating it (or memset it);
 *   every pa_decode / pa_prefill call leaves the header zero again, so the same
 *   workspace serves any sequence of stream-ordered calls of any shape.  Two
 *   calls that may run concurrently (different streams) need two workspaces.  */
size_t pa_decode_workspace_bytes(int32_t batch, int32_t num_q_heads,
                                 int32_t num_kv_heads, int32_t head_dim);
int pa_decode_workspace_init(void* workspace, size_t workspace_bytes, void* stream);

int pa_decode(const void* q, int64_t q_stride_b, int64_t q_stride_h,
              void* out, int64_t o_stride_b, int64_t o_stride_h,
              const PaKvCache* cache,
              const int32_t* block_tables, int32_t bt_stride,
              const int32_t* seq_lens,
              int32_t batch, int32_t num_q_heads,
              void* workspace, size_t workspace_bytes,
              void* stream /* cudaStream_t */);

/* ---- decode without a workspace, for callers that may not allocate -------
 * No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA-graph
 * capture, tight serving loops, harnesses that forbid it).  Same arguments,
 * kernels and results as pa_decode (bit-identical outputs); the workspace is a
 * file-scope device array of PA_DECODE_STATIC_BYTES inside the library, held by
 * every GPU that loads the library until the process exits.  The call, and
 * everything it reaches, calls no cudaMalloc / cudaFree / allocating routine.
 * Limit: a shape whose pa_decode_workspace_bytes() exceeds
 *   PA_DECODE_STATIC_BYTES returns PA_ERR_STATIC_LIMIT and launches nothing
 *   (96 MiB holds e.g. batch <= 256 at 32 query / 8 KV heads of dim 128; the
 *   need grows with batch * num_q_heads * head_dim) - use pa_decode then.
 * Sharing: ALL calls on a device use the same array, which keeps the previous
 *   call's contents (the two counters that must start at zero are cleared,
 *   stream-ordered, before every launch; partials carry a per-launch epoch, so
 *   stale ones never validate).  Calls
 *   that can overlap on the GPU - different streams without an ordering event -
 *   RACE on it: keep them on one stream, or use pa_decode with one workspace
 *   per stream.                                                               */
#define PA_DECODE_STATIC_BYTES (96u << 20)
int pa_decode_static(const void* q, int64_t q_stride_b, int64_t q_stride_h,
                     void* out, int64_t o_stride_b, int64_t o_stride_h,
                     const PaKvCache* cache,
                     const int32_t* block_tables, int32_t bt_stride,
                     const int32_t* seq_lens,
                     int32_t batch, int32_t num_q_heads,
                     void* stream /* cudaStream_t */);

/* ---- decode fused with the output gather of KV-head-sharded decode ---------
 * (rank r owns KV heads [r Hkv/P, (r+1) Hkv/P) of every
 * block; the only exchange on the path is the gather of the [B, Hq/P, D]
 * outputs.)  Same as pa_decode, and in the SAME kernel every final output row
 * is also stored - plain stores over NVLink peer memory - into this rank's slot
 * of each remote rank's gather buffer (remote_out[r]: same strides as `out`),
 * and every completed (sequence, KV head) segment is counted, release at system
 * scope, into this rank's arrival counter on every rank (remote_sig[r], and
 * local_sig on this rank).  No collective call, no second kernel on the sender.
 * The consumer orders itself with pa_peer_wait: counters are monotonic, after
 * step n (1-based) each holds n * batch * num_kv_heads.  Buffers must alternate
 * between two sets from step to step (a fast rank may run one step ahead).
 * Fast path only (bf16, block_size 16, head_dim 64 / 128): PA_ERR_UNSUPPORTED
 * otherwise.                                                                  */
#define PA_MAX_REMOTE_PEERS 7
typedef struct {
  int32_t num_remote;                          /* ranks other than this one, <= 7 */
  void* remote_out[PA_MAX_REMOTE_PEERS];
  uint32_t* remote_sig[PA_MAX_REMOTE_PEERS];
  uint32_t* local_sig;
  /* LL ("flag in data") form, selected by a non-null ll_epoch: remote_out[r] then
   * point at 8-byte-word areas (2x the slot size) and every 4-byte output word
   * travels as {data, step number} in ONE 8-byte store - nothing is fenced or
   * counted on the sender (remote_sig / local_sig unused).  *ll_epoch (device
   * uint32) holds the number of completed steps; pa_peer_unpack advances it.
   * `out` must be a contiguous [B, Hq, D] slot.                               */
  const uint32_t* ll_epoch;
} PaPeerGather;

int pa_decode_gather(const void* q, int64_t q_stride_b, int64_t q_stride_h,
                     void* out, int64_t o_stride_b, int64_t o_stride_h,
                     const PaKvCache* cache,
                     const int32_t* block_tables, int32_t bt_stride,
                     const int32_t* seq_lens,
                     int32_t batch, int32_t num_q_heads,
                     void* workspace, size_t workspace_bytes,
                     const PaPeerGather* peers, void* stream);

/* One small kernel on `stream`: returns once local_sig[r] >= expected for every
 * r < world (ld.acquire.sys), so later work on the stream sees all ranks' rows.
 * Bounded: after timeout_us it writes 1 + r (the late rank) to *status (device
 * int32, may be null) and returns instead of hanging the GPU.                 */
int pa_peer_wait(const uint32_t* local_sig, int32_t world, uint32_t expected,
                 int64_t timeout_us, int32_t* status, void* stream);

/* Same, with the step number kept on the device: waits for (*epoch + 1) *
 * segments, then advances *epoch (uint32 in device memory, zero at the start).
 * No per-step host argument, so decode + wait can be captured in a CUDA graph. */
int pa_peer_wait_next(const uint32_t* local_sig, int32_t world, uint32_t* epoch,
                      uint32_t segments, int64_t timeout_us, int32_t* status,
                      void* stream);

/* Receiver side of the LL form, on the consumer's stream after pa_decode_gather:
 * polls every 8-byte word the remote ranks stored into this rank's LL area
 * ([world][words_per_slot] x {data, flag}) until its flag is *epoch + 1, writes the
 * data words into `gathered` ([world][words_per_slot] uint32; the local slot is
 * left alone - the decode kernel wrote it in place), then advances *epoch.
 * `done` is a zero-initialised device word (block counter).  Bounded like
 * pa_peer_wait.  Graph-capturable (no per-step host argument).  Launched with
 * programmatic stream serialisation: behind pa_decode_gather its CTAs poll the
 * incoming words while the local decode kernel is still running, and only its
 * tail (step number, completion) waits for that kernel.                        */
int pa_peer_unpack(const void* ll_local, void* gathered, int32_t world, int32_t rank,
                   int32_t words_per_slot, uint32_t* epoch, uint32_t* done,
                   int64_t timeout_us, int32_t* status, void* stream);

/* Peer-visible device memory for the gather buffers and counters (the one place
 * the library allocates: cudaMalloc, zero-filled, plus its 64-byte CUDA IPC
 * handle).  Another process of the same node maps it with pa_peer_open (peer
 * access is enabled on first use) and unmaps it with pa_peer_close.           */
int pa_peer_alloc(size_t bytes, void** ptr, void* ipc_handle_64);
int pa_peer_open(const void* ipc_handle_64, void** ptr);
int pa_peer_close(void* ptr);
int pa_peer_free(void* ptr);

/* ---- slot mapping on the device (replaces compute_slot_mapping,
 *      paged_attention.py:54-64, and Scheduler._compute_slot_mapping,
 *      mini_vllm.py:209-219): for i in [0, num_tokens)
 *        pos = start_pos + i;  slots[i] = block_table[pos / BS] * BS + pos % BS
 * block_table int32 (device), slots int64 (device).  Integer-exact.          */
int pa_slot_mapping(const int32_t* block_table, int32_t start_pos, int32_t num_tokens,
                    int32_t block_size, int64_t* slots, void* stream);

/* ---- persistent block tables (replaces the per-step
 *      rebuild of every request's table from Python lists, mini_vllm.py:729-739,
 *      fed by the scheduler's appends, :195-207) ---------------------------------
 * table int32 [num_rows, table_stride] lives on the device across steps, one row
 * per live request.  One call per engine step, two small launches:
 *  1. apply the scheduler's delta: updates int32 [*num_updates][3] = (row, column,
 *     block id); *num_updates is read from DEVICE memory (at most max_updates
 *     are applied, out-of-range entries are skipped), so the call has no per-step
 *     host argument and can sit in a CUDA graph;
 *  2. gather the scheduled requests' rows in scheduling order:
 *     step_table[j, 0:n) = table[row_of[j], 0:n), n = ceil(ctx_lens[j] / BS), for
 *     j < num_gather; row_of[j] < 0 (padding) leaves row j alone.
 * step_table int32 [num_gather, step_stride] is what pa_decode / pa_prefill_varlen
 * take as block_tables.  Integer-exact.  max_updates == 0 / num_gather == 0 skip
 * the respective launch.                                                        */
int pa_block_table_step(int32_t* table, int32_t table_stride, int32_t num_rows,
                        const int32_t* updates, const int32_t* num_updates, int32_t max_updates,
                        const int32_t* row_of, const int32_t* ctx_lens, int32_t num_gather,
                        int32_t block_size, int32_t* step_table, int32_t step_stride,
                        void* stream);

/* ---- KV slot write (replaces the write loops paged_attention.py:48-51 and
 *      :119-123, and the scatter mini_vllm.py:424-427) ----------------------
 * key/value [num_tokens, Hkv, D] with token / head strides in elements.
 * slot_mapping int64 [num_tokens]: slot = block*BS + offset; slot < 0 skips
 * the token (padding).  Pure byte copy: bit-exact.
 * Out-of-range input (no host sync is possible on this path): a slot >=
 * NB*BS is skipped like a negative one - the reference's index_put would
 * raise IndexError (mini_vllm.py:424-427); the host mirror's scheduler-side
 * slot computation (compute_slot_mapping) is where that error is raised.     */
int pa_reshape_and_cache(const void* key, const void* value,
                         int64_t kv_stride_token, int64_t kv_stride_head,
                         const PaKvCache* cache,
                         const int64_t* slot_mapping, int32_t num_tokens,
                         void* stream);

/* ---- fused RoPE + QKV split + KV slot write (replaces mini_vllm.py:410-427:
 *      chunk(3), rope(positions) (:372-375), q*cos + rotate_half(q)*sin
 *      (:355-357, :418-419) and the two cache scatters (:424-427)) --------------
 * qkv [num_tokens, (Hq + 2 Hkv) * D]: per token the Q heads, then the K heads,
 *   then the V heads (qkv_stride_t in elements).  cos_table / sin_table
 *   [max_positions, D] are the module's own cached tables (cos_cached /
 *   sin_cached, :366-370) in the same dtype; positions int64 [num_tokens].
 * Writes rotated Q to q_out [num_tokens, Hq, D], rotated K and plain V to the
 *   cache slots (slot < 0 or >= NB*BS skips the cache write; a position outside
 *   [0, max_positions) is clamped into the table where the reference's table
 *   gather would raise - positions come from the scheduler, which bounds them).
 *   Rounds exactly where torch's
 *   op-by-op evaluation rounds (no FMA contraction): bit-exact in fp32 and bf16. */
int pa_rope_and_cache(const void* qkv, int64_t qkv_stride_t, int32_t num_q_heads,
                      const void* cos_table, const void* sin_table,
                      int64_t table_stride, int32_t max_positions,
                      const int64_t* positions, const int64_t* slot_mapping,
                      int32_t num_tokens,
                      void* q_out, int64_t q_stride_t, int64_t q_stride_h,
                      const PaKvCache* cache, void* stream);

/* ---- block copy (no reference code; definitional:
 *      plane[p][dst] = plane[p][src] for every plane and pair;
 *      prose: BLOG_PART1_V2.md:748-777) --------------------------------------
 * A pool is `num_planes` planes (2 x layers for [2, L, NB, ...]) of blocks;
 * `pairs` int64 [num_pairs, 2] = (src, dst) on the DEVICE.  The dst set must
 * be disjoint from the src set.  block_bytes must be a multiple of 16.       */
int pa_copy_blocks(void* base, int64_t plane_stride_bytes, int32_t num_planes,
                   int64_t block_stride_bytes, int64_t block_bytes,
                   const int64_t* pairs, int32_t num_pairs, void* stream);

/* ---- swap (no reference code; same definition between a device pool and a
 *      pinned-host pool of identical per-block layout) -----------------------
 * `pairs_host` int64 [num_pairs, 2] = (src_block, dst_block) in HOST memory
 * (it is walked on the host to issue the copies).  direction: 0 = device ->
 * host (swap-out), 1 = host -> device (swap-in).  Copies are issued as
 * cudaMemcpy2DAsync (all planes of a run of consecutive blocks per call) on
 * `stream`, which should be a side stream; the caller records an event.      */
int pa_swap_blocks(void* dst_base, int64_t dst_plane_stride_bytes, int64_t dst_block_stride_bytes,
                   const void* src_base, int64_t src_plane_stride_bytes, int64_t src_block_stride_bytes,
                   int32_t num_planes, int64_t block_bytes,
                   const int64_t* pairs_host, int32_t num_pairs, int32_t direction,
                   void* stream);

/* Staged swap for scattered device blocks: a gather kernel packs the device
 * blocks of the pairs into `staging` (device memory, laid out
 * [plane][pair][block_bytes], >= num_planes * num_pairs * block_bytes) at HBM
 * speed, then ONE pinned cudaMemcpyAsync per (plane, run of consecutive HOST
 * blocks) crosses the link; swap-in is the mirror image (copies, then a scatter
 * kernel).  Everything is ordered on `stream`; nothing synchronises.  Host
 * blocks must be dense (host_block_stride_bytes == block_bytes).  `pairs_host`
 * and `pairs_dev` hold the same int64 [num_pairs, 2] (src_block, dst_block) in
 * host and device memory.  direction as for pa_swap_blocks.                   */
int pa_swap_blocks_staged(void* dev_base, int64_t dev_plane_stride_bytes, int64_t dev_block_stride_bytes,
                          void* host_base, int64_t host_plane_stride_bytes, int64_t host_block_stride_bytes,
                          int32_t num_planes, int64_t block_bytes,
                          const int64_t* pairs_host, const int64_t* pairs_dev, int32_t num_pairs,
                          int32_t direction, void* staging, int64_t staging_bytes, void* stream);

/* ---- prefill attention (replaces the attention half of
 *      paged_attention_prefill, paged_attention.py:125-136, extended with
 *      history: query chunk at positions [start_pos, start_pos + num_tokens)
 *      attends causally to positions [0, start_pos + i] read THROUGH the block
 *      table - the KV of the chunk must already be in the cache, i.e. call
 *      pa_reshape_and_cache first) ------------------------------------------------
 * q / out [num_tokens, num_q_heads, D].  One sequence per call.
 * Two implementations behind this entry:
 *  - tensor-core path (bf16, block_size 16, head_dim 64 / 128, g a power of
 *    two <= 16, 16-byte aligned strides): pa_prefill_tc2_kernel, tcgen05.mma
 *    with TMEM accumulators, K / V tiles gathered by TMA through the block
 *    table (csrc/pa_prefill_tc2.cu).  Persistent, one CTA per SM; the CTAs
 *    draw their work tiles from a per-launch ticket kept in a 1024-slot
 *    device array inside the library (slot = launch number mod 1024, put back
 *    to zero by the launch's last CTA): no workspace, any number of launches
 *    in flight on any streams - two launches share a slot only if 1024
 *    launches apart AND running at the same time;
 *  - everything else (fp32 caches, the engine's toy head sizes): every query
 *    token is a decode "sequence" of length start_pos + i + 1 over the shared
 *    block table (same kernels and numerics as pa_decode).                   */
size_t pa_prefill_workspace_bytes(int32_t num_tokens, int32_t num_q_heads,
                                  int32_t num_kv_heads, int32_t head_dim);

int pa_prefill(const void* q, int64_t q_stride_t, int64_t q_stride_h,
               void* out, int64_t o_stride_t, int64_t o_stride_h,
               const PaKvCache* cache, const int32_t* block_table,
               int32_t start_pos, int32_t num_tokens, int32_t num_q_heads,
               void* workspace, size_t workspace_bytes,
               void* stream);

/* ---- varlen prefill: a BATCH of query chunks in one launch (the engine's mixed
 *      chunked-prefill + decode step, mini_vllm.py:437-481; varlen layout as in
 *      BLOG_PART2_V2.md:1104-1141) ---------------------------------------------
 * q / out [total_tokens, num_q_heads, D], the chunks of the sequences packed
 * back to back: sequence s owns tokens [cu_q_lens[s], cu_q_lens[s + 1]) which
 * sit at positions [ctx_lens[s] - q_len, ctx_lens[s]) of its context and
 * attend causally to [0, pos] through block_tables[s].  q_len = 1 is a decode
 * token.  cu_q_lens int32 [batch + 1], ctx_lens int32 [batch] on the DEVICE;
 * max_q_len (host) >= every q_len sizes the grid.  Tensor-core path only:
 * PA_ERR_UNSUPPORTED for shapes outside it (see pa_prefill); the caller then
 * loops pa_prefill per sequence.                                              
 * A request with ctx_lens[s] < its q_len (e.g. ctx_lens[s] = 0) is skipped and
 * its rows of `out` are left alone: callers mask requests out that way.  So is a
 * request whose context needs more than `max_blocks` block-table entries (the
 * row's readable length; the lengths live on the device, so the host cannot
 * raise the reference's IndexError without a sync): the table is never read
 * past max_blocks.                                                            */
int pa_prefill_varlen(const void* q, int64_t q_stride_t, int64_t q_stride_h,
                      void* out, int64_t o_stride_t, int64_t o_stride_h,
                      const PaKvCache* cache,
                      const int32_t* block_tables, int32_t bt_stride, int32_t max_blocks,
                      const int32_t* cu_q_lens, const int32_t* ctx_lens,
                      int32_t batch, int32_t max_q_len, int32_t num_q_heads,
                      void* stream);
/* 1 when pa_prefill / pa_prefill_varlen take the tensor-core path for this
 * cache and head count (pointer alignment is checked per call), else 0.       */
int pa_prefill_tc_available(const PaKvCache* cache, int32_t num_q_heads);
#ifdef __cplusplus
}
#endif
#endif /* PA_B200_H_ */
