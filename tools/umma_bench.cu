// Micro-benchmark: sustained rate of back-to-back tcgen05.mma (cta_group::1, kind::f16, M = 128)
// with both operands in shared memory, for the operand shapes the prefill kernels use.
// Build + run on a B200:
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I vllm_from_scratch_b200/csrc \
//        tools/umma_bench.cu -o /tmp/umma_bench && /tmp/umma_bench
// Prints cycles per MMA instruction per configuration, on 1 CTA and on 148 CTAs (one per SM).
#include <cstdio>
#include <cuda_runtime.h>

#include "pa_tc_common.cuh"

namespace {

struct Cfg {
  int n;            // UMMA N
  int b_mn_major;   // 0: B K-major (QK^T style), 1: B MN-major (PV style)
  int a_step;       // bytes added to the A descriptor start per instruction (cycled over 4)
  int distinct;     // 1: cycle through 8 different operand offsets, 0: same operands every time
  int a_tmem;       // 1: A operand in tensor memory (TS), 0: in shared memory (SS)
  int alt;          // 1: alternate SS (N = 128, K-major B) and TS (MN-major B) instructions, as the prefill kernel does
};

__device__ __forceinline__ void umma_bf16_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t desc_b, uint32_t idesc,
                                             uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
      ::"r"(tmem_d), "r"(tmem_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}

__global__ void __launch_bounds__(128, 1) umma_bench_kernel(Cfg c, int iters, long long* out) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  __shared__ uint32_t tmem_slot;
  __shared__ uint64_t barv;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // zero the operands (values do not matter for timing; avoid NaN slow paths just in case)
  for (uint32_t i = threadIdx.x * 16; i < 160 * 1024; i += blockDim.x * 16)
    *reinterpret_cast<uint4*>(smem_raw + (base - smem_u32(smem_raw)) + i) = make_uint4(0, 0, 0, 0);
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_slot)), "n"(512) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  if (threadIdx.x == 0) {
    mbar_init(smem_u32(&barv), 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;
  if (warp == 1 && elect_one()) {
    // descriptors are precomputed and the loop is unrolled by 8: the issuing thread must not be the limit
    const uint32_t idesc = umma_idesc(c.n, c.b_mn_major != 0);
    const uint32_t idesc_ss = umma_idesc(128, false), idesc_ts = umma_idesc(128, true);
    const uint32_t sA = base, sB = base + 64 * 1024;
    uint64_t da[8], db[8], dbk[8], dbm[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      da[j] = umma_desc(sA + (uint32_t)(j >> 2) * 16384u + (uint32_t)(j & 3) * 32u, 16, 1024);
      dbk[j] = umma_desc(sB + (uint32_t)(j >> 2) * 16384u + (uint32_t)(j & 3) * 32u, 16, 1024);
      dbm[j] = umma_desc(sB + (uint32_t)j * 2048u, 16384, 1024);
      db[j] = c.b_mn_major ? dbm[j] : dbk[j];
    }
    long long t0 = clock64();
    if (c.alt) {
      for (int i = 0; i < iters; i += 16) {
#pragma unroll
        for (int j = 0; j < 8; ++j) umma_bf16(tmem, da[j], dbk[j], idesc_ss, 1);
#pragma unroll
        for (int j = 0; j < 8; ++j) umma_bf16_ts(tmem + 256, tmem + 128 + (uint32_t)j * 8u, dbm[j], idesc_ts, 1);
      }
    } else if (c.a_tmem) {
      for (int i = 0; i < iters; i += 8) {
#pragma unroll
        for (int j = 0; j < 8; ++j) umma_bf16_ts(tmem, tmem + 256 + (uint32_t)j * 8u, db[j], idesc, 1);
      }
    } else {
      for (int i = 0; i < iters; i += 8) {
#pragma unroll
        for (int j = 0; j < 8; ++j) umma_bf16(tmem, da[j], db[j], idesc, 1);
      }
    }
    umma_commit(smem_u32(&barv));
    mbar_wait(smem_u32(&barv), 0);
    long long t1 = clock64();
    if (blockIdx.x == 0) out[0] = t1 - t0;
  }
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(512) : "memory");
}

}  // namespace

int main() {
  long long* d_out;
  cudaMalloc(&d_out, 8);
  cudaFuncSetAttribute(umma_bench_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  const Cfg cfgs[] = {
      {128, 0, 32, 1, 0, 0}, {64, 0, 32, 1, 0, 0}, {32, 0, 32, 1, 0, 0}, {256, 0, 32, 1, 0, 0}, {128, 1, 32, 1, 0, 0}, {256, 1, 32, 1, 0, 0},
      {128, 1, 32, 1, 1, 0}, {64, 1, 32, 1, 1, 0}, {32, 1, 32, 1, 1, 0}, {256, 1, 32, 1, 1, 0}, {128, 0, 32, 1, 1, 0}, {64, 0, 32, 1, 1, 0},
      {128, 0, 32, 1, 0, 1},
  };
  const int iters = 4096;
  for (const Cfg& c : cfgs) {
    for (int grid : {148}) {
      long long h = 0;
      for (int rep = 0; rep < 2; ++rep) {
        umma_bench_kernel<<<grid, 128, 200 * 1024>>>(c, iters, d_out);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("CUDA error: %s\n", cudaGetErrorString(e)); return 1; }
      }
      cudaMemcpy(&h, d_out, 8, cudaMemcpyDeviceToHost);
      const double cyc = (double)h / iters;
      printf("%s N=%3d B=%s distinct=%d grid=%3d : %.1f cycles/MMA  (floor %d)  -> %.0f flop/clk/SM\n", c.alt ? "SS/TS alternating" : (c.a_tmem ? "TS" : "SS"), c.n,
             c.b_mn_major ? "MN" : "K ", c.distinct, grid, cyc, 128 * c.n / 256, 2.0 * 128 * c.n * 16 / cyc);
    }
  }
  return 0;
}
