// Hardware experiment (bring-up, not product code): tensor-pipe cost of one tcgen05.mma
// (M=128, K=16, bf16 -> fp32, cta_group::1, both operands from shared memory) for different operand
// layouts and N, measured as cycles per MMA over a back-to-back chain issued by one thread.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -std=c++17 tools/exp_mma_rate.cu -o tools/exp_mma_rate.bin
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// layout: 0 = SW128 aligned, 1 = SW128 A start +1 row (unaligned), 2 = SW64, 3 = no-swizzle Toeplitz
// (LBO 16, SBO 176) with no-swizzle B, 4 = SW128 A start +8 rows (aligned shift)
__global__ void __launch_bounds__(128, 1)
rate_kernel(long long* out_cycles, int layout, int n, int chain, int k_adv) {
  extern __shared__ uint8_t raw[];
  const uint32_t base = (smem_u32(raw) + 1023u) & ~1023u;
  uint8_t* gen = raw + (base - smem_u32(raw));
  uint32_t* w = reinterpret_cast<uint32_t*>(gen);
  for (int i = threadIdx.x; i < (64 * 1024 + 32 * 1024) / 4; i += 128) w[i] = 0x3c003c00u + (i & 7);   // small finite bf16
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  const uint32_t a_addr = base, b_addr = base + 64 * 1024, bar = base + 96 * 1024, slot = bar + 8;
  volatile uint32_t* slot_ptr = (volatile uint32_t*)(gen + 96 * 1024 + 8);
  const int warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 256;" ::"r"(slot));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = *slot_ptr;
  if (threadIdx.x == 0) {
    uint64_t da, db;
    const uint64_t ver = 1ull << 46;
    if (layout == 0 || layout == 1 || layout == 4) {
      const uint32_t a0 = a_addr + (layout == 1 ? 128u : layout == 4 ? 1024u : 0u);
      da = ((uint64_t)((a0 & 0x3ffff) >> 4)) | (1ull << 16) | ((uint64_t)(1024 >> 4) << 32) | ver | (2ull << 61);
      db = ((uint64_t)((b_addr & 0x3ffff) >> 4)) | (1ull << 16) | ((uint64_t)(1024 >> 4) << 32) | ver | (2ull << 61);
    } else if (layout == 2) {
      da = ((uint64_t)((a_addr & 0x3ffff) >> 4)) | (1ull << 16) | ((uint64_t)(512 >> 4) << 32) | ver | (4ull << 61);
      db = ((uint64_t)((b_addr & 0x3ffff) >> 4)) | (1ull << 16) | ((uint64_t)(512 >> 4) << 32) | ver | (4ull << 61);
    } else {
      da = ((uint64_t)((a_addr & 0x3ffff) >> 4)) | ((uint64_t)(16 >> 4) << 16) | ((uint64_t)(176 >> 4) << 32) | ver;
      db = ((uint64_t)((b_addr & 0x3ffff) >> 4)) | ((uint64_t)(1024 >> 4) << 16) | ((uint64_t)(128 >> 4) << 32) | ver;
    }
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | (((uint32_t)n >> 3) << 17) | ((128u >> 4) << 24);
    const long long t0 = clock64();
    // fully unrolled groups of 8 MMAs with constant descriptor increments (what a tuned issue loop does)
    for (int i = 0; i < chain; i += 8) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const uint64_t adv = (uint64_t)(2 * (j % 4 < k_adv ? j % 4 : 0));
        asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n}"
                     ::"r"(tmem), "l"(da + adv), "l"(db + adv), "r"(idesc), "r"(1) : "memory");
      }
    }
    const long long t1 = clock64();
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
    uint32_t ok = 0;
    while (!ok) asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\nselp.u32 %0,1,0,p;\n}" : "=r"(ok) : "r"(bar) : "memory");
    const long long t2 = clock64();
    out_cycles[0] = t1 - t0;      // issue time
    out_cycles[1] = t2 - t0;      // until the last MMA retired
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 256;" ::"r"(tmem));
}

int main() {
  long long* d; CK(cudaMalloc(&d, 16));
  const int smem = 96 * 1024 + 64 + 1024;
  CK(cudaFuncSetAttribute(rate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  const char* names[5] = {"SW128 aligned", "SW128 start+1 row", "SW64", "no-swizzle Toeplitz", "SW128 start+8 rows"};
  const int kadv[5] = {4, 4, 2, 2, 4};
  for (int layout = 0; layout < 5; ++layout)
    for (int n = 64; n <= 256; n *= 2) {
      if (layout == 3 && n > 64) continue;
      for (int rep = 0; rep < 2; ++rep) {
        rate_kernel<<<1, 128, smem>>>(d, layout, n, 512, kadv[layout]);
        CK(cudaDeviceSynchronize());
      }
      long long h[2]; CK(cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost));
      printf("%-22s N=%3d : issue %6.1f cyc/MMA, retire %6.1f cyc/MMA  (math floor %d)\n", names[layout], n,
             h[0] / 512.0, h[1] / 512.0, n / 2);
    }
  return 0;
}
