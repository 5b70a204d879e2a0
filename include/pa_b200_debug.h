/*
 * pa_b200_debug.h - test, bring-up and measurement hooks of libpa_b200.so.
 *
 * NOT part of the drop-in ABI (include/pa_b200.h): these entry points set
 * process-global state (atomics) that the launchers read, so they are not safe
 * to flip while another host thread is launching, and a binding of the reference
 * API never needs them.  Used by tests/ (kernel-path forcing), tools/ (per-CTA
 * timelines, split-K tuning sweeps) and bench.py (per-kernel CUDA-event timing
 * for the roofline line).
 */
#ifndef PA_B200_DEBUG_H_
#define PA_B200_DEBUG_H_

#include "pa_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

/* 0 = automatic (default: the two-tile ping-pong tcgen05 kernel); 1 = force the
 * decode-kernel formulation in pa_prefill even where a tensor-core kernel
 * applies; 2 = the one-tile-per-CTA tcgen05 kernel.  Returns the old value.     */
int pa_prefill_force_path(int path);

/* ---- test / tuning hook ---------------------------------------------------
 * 0 = automatic kernel choice (default); 1 = force the generic CUDA-core
 * decode path even for shapes the TMA kernel takes (the cp.async streaming kernel where rows
 * are 16-byte aligned); 2 = force the plain generic kernel (no shared-memory staging) everywhere.
 * Returns the old value. */
int pa_decode_force_path(int path);
/* Split-K granularity of the fast path (guided self-scheduling): a unit handed out when
 * `remaining` blocks are left is about remaining / (guided_k * SMs) blocks (default 3), never
 * smaller than total / (units_per_sm * SMs) (1 .. 64, default 16).  stages: per-warp TMA ring
 * depth, 0 = as deep as shared memory allows.  claim_limit: a CTA claims its next unit
 * only while its backlog (published-unconsumed + unpublished blocks) is below claim_limit + 16, and
 * publishes a unit's 16-block descriptors only while fewer than claim_limit are unconsumed (default 32). */
int pa_decode_set_tuning(int units_per_sm, int guided_k, int stages, int claim_limit);
/* Launch the TMA decode kernel with programmatic stream serialisation (default on): back-to-back launches on
 * one stream overlap the next kernel's prologue with the previous kernel's tail.  Returns the old setting. */
int pa_decode_set_pdl(int on);
/* Debug: when non-null, every fast-path CTA writes its timeline, 16 int64 at device_buffer[16 * cta]
 * (globaltimer ns): [0] SM id, [1] CTA start, [2] consumers' end, [3] units processed, [4] merger / folders'
 * end, [5] first KV block landed, [6] plan built, [7] first unit published, [8] first TMA load issued,
 * [9] ticket counter found exhausted, [10] last unit merged, [11] last fold job taken by the folders, [12] its last
 * batch of partials seen; (SM cycles, consumer warp 0:) [13] waited for unit descriptors,
 * [14] for KV data, [15] for the hand-off area to the merger.  pa_decode_num_workers() entries.   */
int pa_decode_debug_times(void* device_buffer);
/* Same for the tcgen05 prefill kernel: 8 int64 per CTA (grid order x, y, z): SM id,
 * globaltimer at entry / prologue done / Q stored / first S / O complete / outputs
 * stored, KV tiles streamed.  Null switches it off.                              */
int pa_prefill_debug_times(void* device_buffer);

/* ---- per-kernel device timing (bench.py's roofline line) ---------------------
 * pa_profile_begin(n): the next n pa_decode calls record CUDA events on their
 * stream before the split-K kernel, between it and the reduce kernel, and after.
 * pa_profile_end: synchronises on those events, writes the per-call durations
 * (milliseconds) and returns how many calls were recorded (<0 on error).        */
int pa_profile_begin(int32_t max_calls);
int pa_profile_end(float* decode_ms_host, float* reduce_ms_host, int32_t capacity);

#ifdef __cplusplus
}
#endif
#endif /* PA_B200_DEBUG_H_ */
