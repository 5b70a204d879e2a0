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
};

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
  if (warp == 1 && lane == 0) {
    const uint32_t idesc = umma_idesc(c.n, c.b_mn_major != 0);
    const uint32_t sA = base, sB = base + 64 * 1024;
    long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
      const int j = c.distinct ? (i & 7) : 0;
      const uint64_t da = umma_desc(sA + (uint32_t)(j >> 2) * 16384u + (uint32_t)(j & 3) * c.a_step, 16, 1024);
      const uint64_t db = c.b_mn_major ? umma_desc(sB + (uint32_t)j * 2048u, 16384, 1024)
                                       : umma_desc(sB + (uint32_t)(j >> 2) * 16384u + (uint32_t)(j & 3) * 32u, 16, 1024);
      umma_bf16(tmem, da, db, idesc, 1);
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
      {128, 0, 32, 1}, {128, 0, 32, 0}, {64, 0, 32, 1}, {256, 0, 32, 1}, {16, 0, 32, 1}, {32, 0, 32, 1},
      {128, 1, 32, 1}, {128, 1, 32, 0}, {64, 1, 32, 1}, {256, 1, 32, 1},
  };
  const int iters = 4096;
  for (const Cfg& c : cfgs) {
    for (int grid : {1, 148}) {
      long long h = 0;
      for (int rep = 0; rep < 2; ++rep) {
        umma_bench_kernel<<<grid, 128, 200 * 1024>>>(c, iters, d_out);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("CUDA error: %s\n", cudaGetErrorString(e)); return 1; }
      }
      cudaMemcpy(&h, d_out, 8, cudaMemcpyDeviceToHost);
      const double cyc = (double)h / iters;
      printf("N=%3d B=%s distinct=%d grid=%3d : %.1f cycles/MMA  (floor %d)  -> %.0f flop/clk/SM\n", c.n,
             c.b_mn_major ? "MN" : "K ", c.distinct, grid, cyc, 128 * c.n / 256, 2.0 * 128 * c.n * 16 / cyc);
    }
  }
  return 0;
}
