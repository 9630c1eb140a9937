// Hardware experiment (bring-up, not product code): how fast can ONE SM take data into shared memory?
//  1. TMA: a persistent CTA per SM streams 16 KiB boxes (128 rows x 128 bytes, SWIZZLE_128B) through a ring of `stages`
//     buffers; the consumer releases a buffer as soon as it is full (no MMA, no reads).  Swept: SMs active, ring depth
//     (= bytes in flight), source (L2-resident 32 MiB / HBM 2 GiB, every box read once) and layout (box rows 128 B
//     apart = one contiguous chunk / 8 KiB apart = a K-major weight matrix).
//  2. the same with other box shapes / swizzles and with the 1-D cp.async.bulk copy;
//  3. TMA multicast inside clusters of 2 and 4 (every CTA receives every box; this GPU's memory only);
//  4. cp.async (LDGSTS, 16 B per thread) from 16 warps with commit groups, alone and beside a TMA stream;
//  5. cp.async rings shaped like a GEMM operand pipeline (completion booked per stage) -- these stop at one stage per
//     L2 round trip however many stages are in flight, which is why the product's FC layers stayed on TMA.
// Result on B200 (profiles/r02_exp_tma_ingest.log): TMA tops out at 42.3 B/clk/SM whatever the depth, SM count, layout
// or source (HBM: 26 B/clk/SM with all 148 SMs = 7 TB/s); 1-D bulk copies the same; cp.async reaches 73 B/clk/SM.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -std=c++17 tools/exp_tma_ingest.cu -o tools/exp_tma_ingest.bin -lcuda
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <cstdlib>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ bool try_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
               : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
  return ok != 0;
}

constexpr int kBoxBytes = 16384;
constexpr int kMaxStages = 12;

__global__ void __launch_bounds__(64, 1)
ingest_kernel(const __grid_constant__ CUtensorMap tm, int stages, int iters, int strided, int row_blocks,
              long long* out_cycles, const uint8_t* bulk_src, int box_w) {
  extern __shared__ uint8_t raw[];
  const uint32_t base = (smem_u32(raw) + 1023u) & ~1023u;
  const uint32_t bars = base + kMaxStages * kBoxBytes;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int s = 0; s < 2 * kMaxStages; ++s) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bars + 8u * s));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const long long t0 = clock64();
  if (warp == 0 && lane == 0) {
    int s = 0; uint32_t ph = 0;
    for (int i = 0; i < iters; ++i) {
      while (!try_wait(bars + 8u * (kMaxStages + s), ph ^ 1u)) {}
      const uint32_t full = bars + 8u * s;
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(full), "r"(kBoxBytes) : "memory");
      int x, y;
      if (strided) {            // [R][4096] matrix: column block i % 64, row block walks down
        x = (i & 63) * 64;
        y = (int)(((long long)blockIdx.x + (long long)(i >> 6) * gridDim.x) % row_blocks) * 128;
      } else {                  // [rows][64]: box b = 128 consecutive 128-byte rows
        x = 0;
        y = (int)(((long long)blockIdx.x * iters + i) % row_blocks) * 128;
      }
      if (bulk_src != nullptr)      // plain 1-D bulk copy of the same 16 KiB (contiguous layout only)
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(base + s * kBoxBytes), "l"(bulk_src + (size_t)(y / 128) * kBoxBytes), "r"(kBoxBytes), "r"(full) : "memory");
      else
        asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                     ::"r"(base + s * kBoxBytes), "l"(&tm), "r"(box_w == 64 ? x : 0), "r"(box_w == 64 ? y : y * 64 / box_w), "r"(full) : "memory");
      if (++s == stages) { s = 0; ph ^= 1u; }
    }
  } else if (warp == 1 && lane == 0) {
    int s = 0; uint32_t ph = 0;
    for (int i = 0; i < iters; ++i) {
      while (!try_wait(bars + 8u * s, ph)) {}
      asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bars + 8u * (kMaxStages + s)) : "memory");
      if (++s == stages) { s = 0; ph ^= 1u; }
    }
    out_cycles[blockIdx.x] = clock64() - t0;
  }
}


// Cluster of CS CTAs: every CTA receives ALL boxes; CTA r issues box i when i % CS == r, multicast to the whole cluster.
// Each full barrier expects the box once (whoever issued it); a buffer is free again when every CTA of the cluster has
// released it (remote arrives on the issuer's... simpler: on every CTA's own empty barrier, count CS).
template <int CS>
__global__ void __launch_bounds__(64, 1)
ingest_mc_kernel(const __grid_constant__ CUtensorMap tm, int stages, int iters, int row_blocks, long long* out_cycles) {
  extern __shared__ uint8_t raw[];
  const uint32_t base = (smem_u32(raw) + 1023u) & ~1023u;
  const uint32_t bars = base + kMaxStages * kBoxBytes;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint32_t rank;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(rank));
  if (threadIdx.x == 0) {
    for (int s = 0; s < kMaxStages; ++s) {
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bars + 8u * s));
      asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bars + 8u * (kMaxStages + s)), "r"(CS));
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  asm volatile("barrier.cluster.arrive.release.aligned;\nbarrier.cluster.wait.acquire.aligned;" ::: "memory");
  const long long t0 = clock64();
  const int cluster_id = blockIdx.x / CS;
  if (warp == 0 && lane == 0) {
    int s = 0; uint32_t ph = 0;
    for (int i = 0; i < iters; ++i) {
      // every CTA arms its own full barrier for the box; the issuer waits until ALL CTAs have freed the buffer
      { unsigned n = 0; while (!try_wait(bars + 8u * (kMaxStages + s), ph ^ 1u)) { if (++n > 20000000u) return; } }
      const uint32_t full = bars + 8u * s;
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(full), "r"(kBoxBytes) : "memory");
      if ((uint32_t)(i % CS) == rank) {
        const int y = (int)(((long long)cluster_id * iters + i) % row_blocks) * 128;
        const uint16_t mask = (uint16_t)((1u << CS) - 1);
        asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster [%0], [%1, {%2, %3}], [%4], %5;"
                     ::"r"(base + s * kBoxBytes), "l"(&tm), "r"(0), "r"(y), "r"(full), "h"(mask) : "memory");
      }
      if (++s == stages) { s = 0; ph ^= 1u; }
    }
  } else if (warp == 1 && lane == 0) {
    int s = 0; uint32_t ph = 0;
    for (int i = 0; i < iters; ++i) {
      { unsigned n = 0; while (!try_wait(bars + 8u * s, ph)) { if (++n > 20000000u) return; } }
      // release the buffer in EVERY CTA of the cluster (the issuer of the next box into it may be any of them,
      // and a multicast writes into all of them)
      for (int r = 0; r < CS; ++r) {
        uint32_t remote;
        asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(bars + 8u * (kMaxStages + s)), "r"(r));
        asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(remote) : "memory");
      }
      if (++s == stages) { s = 0; ph ^= 1u; }
    }
    out_cycles[blockIdx.x] = clock64() - t0;
  }
  asm volatile("barrier.cluster.arrive.release.aligned;\nbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}

template <int CS>
static int run_mc(const CUtensorMap& tm, int sms, int smem, int row_blocks, long long* d_cyc, cudaEvent_t e0, cudaEvent_t e1) {
  CK(cudaFuncSetAttribute(ingest_mc_kernel<CS>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  const int iters = 800;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(sms / CS * CS); cfg.blockDim = dim3(64); cfg.dynamicSmemBytes = smem;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = CS; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  cfg.attrs = at; cfg.numAttrs = 1;
  for (int rep = 0; rep < 2; ++rep) {
    CK(cudaEventRecord(e0));
    CK(cudaLaunchKernelEx(&cfg, ingest_mc_kernel<CS>, tm, 8, iters, row_blocks, d_cyc));
    CK(cudaEventRecord(e1));
    CK(cudaDeviceSynchronize());
  }
  float ms = 0;
  CK(cudaEventElapsedTime(&ms, e0, e1));
  long long cyc[148];
  const int n = sms / CS * CS;
  CK(cudaMemcpy(cyc, d_cyc, n * sizeof(long long), cudaMemcpyDeviceToHost));
  double avg = 0;
  for (int i = 0; i < n; ++i) avg += (double)cyc[i];
  avg /= n;
  printf("multicast, cluster of %d: every CTA receives all boxes  %4d  %8.1f received B/clk/SM  (%.0f GB/s received in total)\n", CS, n,
         (double)iters * kBoxBytes / avg, (double)n * iters * kBoxBytes / (ms * 1e-3) / 1e9);
  return 0;
}

// Second path into shared memory: cp.async (LDGSTS, 16 bytes per thread) issued by 16 warps, with or without the
// TMA stream of ingest_kernel running beside it in the same CTA (warps 16 / 17).  Do the two paths add up?
__global__ void __launch_bounds__(576, 1)
ingest_lsu_kernel(const __grid_constant__ CUtensorMap tm, const uint8_t* src, size_t src_bytes, int iters, int use_tma,
                  int use_lsu, int row_blocks, long long* out_cycles) {
  extern __shared__ uint8_t raw[];
  const uint32_t base = (smem_u32(raw) + 1023u) & ~1023u;
  constexpr int kTmaStages = 4;
  const uint32_t lsu_base = base + kTmaStages * kBoxBytes;            // 16 warps x 4 groups x 2 KiB = 128 KiB
  const uint32_t bars = lsu_base + 16 * 4 * 2048;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int s = 0; s < 2 * kTmaStages; ++s) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bars + 8u * s));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const long long t0 = clock64();
  if (warp < 16) {
    if (use_lsu) {
      // iteration i of the CTA = 32 KiB: warp w takes 2 KiB of it as 4 x (32 lanes x 16 B)
      const size_t cta_off = ((size_t)blockIdx.x * iters * 32768) % src_bytes;
      for (int i = 0; i < iters; ++i) {
        const uint8_t* g = src + (cta_off + (size_t)i * 32768) % src_bytes + warp * 2048 + lane * 16;
        const uint32_t d = lsu_base + (uint32_t)(warp * 4 + (i & 3)) * 2048u + lane * 16;
#pragma unroll
        for (int j = 0; j < 4; ++j)
          asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d + j * 512), "l"(g + j * 512) : "memory");
        asm volatile("cp.async.commit_group;" ::: "memory");
        asm volatile("cp.async.wait_group 3;" ::: "memory");
      }
      asm volatile("cp.async.wait_group 0;" ::: "memory");
    }
    if (threadIdx.x == 0) out_cycles[blockIdx.x] = clock64() - t0;
  } else if (warp == 16 && lane == 0 && use_tma) {
    int s = 0; uint32_t ph = 0;
    for (int i = 0; i < iters; ++i) {
      while (!try_wait(bars + 8u * (kTmaStages + s), ph ^ 1u)) {}
      const uint32_t full = bars + 8u * s;
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(full), "r"(kBoxBytes) : "memory");
      const int y = (int)(((long long)blockIdx.x * iters + i) % row_blocks) * 128;
      asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                   ::"r"(base + s * kBoxBytes), "l"(&tm), "r"(0), "r"(y), "r"(full) : "memory");
      if (++s == kTmaStages) { s = 0; ph ^= 1u; }
    }
  } else if (warp == 17 && lane == 0 && use_tma) {
    int s = 0; uint32_t ph = 0;
    for (int i = 0; i < iters; ++i) {
      while (!try_wait(bars + 8u * s, ph)) {}
      asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bars + 8u * (kTmaStages + s)) : "memory");
      if (++s == kTmaStages) { s = 0; ph ^= 1u; }
    }
    out_cycles[148 + blockIdx.x] = clock64() - t0;
  }
}

// Shapes of a cp.async ring for 16 KiB stages (the weight box of one k-block).  `loaders` threads copy
// 1024 / loaders pieces of 16 bytes each per stage.  mode 0: every thread books completion with
// cp.async.mbarrier.arrive.noinc (barrier count = loaders);  mode 1: commit groups; a warp issues while a buffer is free
// and otherwise waits for its OLDEST group and arrives once per warp (barrier count = loaders / 32).
template <int N> __device__ __forceinline__ void cp_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ void cp_wait_dyn(int n) {
  switch (n) { case 0: cp_wait<0>(); break; case 1: cp_wait<1>(); break; case 2: cp_wait<2>(); break; case 3: cp_wait<3>(); break;
               case 4: cp_wait<4>(); break; default: cp_wait<5>(); break; }
}
__global__ void __launch_bounds__(576, 1)
ingest_ring_kernel(const uint8_t* src, size_t src_bytes, int stages, int iters, int loaders, int mode, long long* out_cycles,
                   int linear, int nowait) {
  extern __shared__ uint8_t raw[];
  const uint32_t base = (smem_u32(raw) + 1023u) & ~1023u;
  const int stage_bytes = 16384, per_thread = 1024 / loaders;
  const uint32_t bars = base + 6 * 32768;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int s = 0; s < kMaxStages; ++s) {
      asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bars + 8u * s), "r"(mode ? loaders / 32 : loaders));
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bars + 8u * (kMaxStages + s)));
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const long long t0 = clock64();
  if ((int)threadIdx.x < loaders) {
    const size_t cta_off = ((size_t)blockIdx.x * iters * stage_bytes) % src_bytes;
    int s = 0, sig = 0, inflight = 0; uint32_t ph = 0;
    auto signal_oldest = [&]() {
      cp_wait_dyn(inflight - 1);
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      __syncwarp();
      if (lane == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bars + 8u * sig) : "memory");
      if (++sig == stages) sig = 0;
      --inflight;
    };
    for (int i = 0; i < iters; ++i) {
      if (nowait) {
      } else if (mode == 0) {
        while (!try_wait(bars + 8u * (kMaxStages + s), ph ^ 1u)) {}
      } else {
        while (!try_wait(bars + 8u * (kMaxStages + s), ph ^ 1u)) { if (inflight > 0) signal_oldest(); }
      }
      const uint8_t* g = src + (cta_off + (size_t)i * stage_bytes) % src_bytes + threadIdx.x * 16;
      const uint32_t d = base + (uint32_t)s * stage_bytes;
      for (int j = 0; j < per_thread; ++j) {
        const int e = threadIdx.x + loaders * j, row = e >> 3, c = e & 7;
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(linear ? d + e * 16 : d + row * 128 + ((c ^ (row & 7)) << 4)), "l"(g + j * loaders * 16) : "memory");
      }
      if (mode == 0) asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(bars + 8u * s) : "memory");
      else { asm volatile("cp.async.commit_group;" ::: "memory"); if (++inflight > 5) signal_oldest(); }
      if (++s == stages) { s = 0; ph ^= 1u; }
    }
    while (mode && inflight > 0) signal_oldest();
  } else if (warp == 16 && lane == 0) {
    int s = 0; uint32_t ph = 0;
    for (int i = 0; i < iters; ++i) {
      while (!try_wait(bars + 8u * s, ph)) {}
      asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bars + 8u * (kMaxStages + s)) : "memory");
      if (++s == stages) { s = 0; ph ^= 1u; }
    }
    out_cycles[blockIdx.x] = clock64() - t0;
  }
}

typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                             const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                             CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main() {
  void* fp = nullptr;
  cudaDriverEntryPointQueryResult q;
  CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fp, cudaEnableDefault, &q));
  EncodeFn enc = (EncodeFn)fp;
  const size_t big = 2ull << 30;
  uint8_t* buf = nullptr;
  CK(cudaMalloc(&buf, big));
  CK(cudaMemset(buf, 1, big));
  long long* d_cyc = nullptr;
  CK(cudaMalloc(&d_cyc, 2 * 148 * sizeof(long long)));
  const int smem = kMaxStages * kBoxBytes + 2 * kMaxStages * 8 + 1024;
  CK(cudaFuncSetAttribute(ingest_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  printf("source layout   SMs stages  B/clk/SM  aggregate GB/s\n");
  for (int src = 0; src < 2; ++src)            // 0: 64 MiB (L2), 1: 2 GiB (HBM)
    for (int strided = 0; strided < 2; ++strided) {
      const size_t bytes = src ? big : (32ull << 20);
      CUtensorMap tm;
      cuuint64_t gdim[2], gstr[1];
      int row_blocks;
      if (strided) { gdim[0] = 4096; gdim[1] = bytes / 8192; gstr[0] = 8192; row_blocks = (int)(gdim[1] / 128); }
      else { gdim[0] = 64; gdim[1] = bytes / 128; gstr[0] = 128; row_blocks = (int)(gdim[1] / 128); }
      cuuint32_t box[2] = {64, 128}, estr[2] = {1, 1};
      CUresult r = enc(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, buf, gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                       CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
      if (r != CUDA_SUCCESS) { printf("encode failed %d\n", (int)r); return 1; }
      const int sms_list[4] = {8, 37, 74, 148};
      const int st_list[5] = {2, 4, 6, 8, 12};
      for (int a = 0; a < 4; ++a)
        for (int b = 0; b < 5; ++b) {
          const int sms = sms_list[a], stages = st_list[b];
          // HBM: every box once (148 * 800 boxes * 16 KiB = 1.9 GiB); L2: 32 MiB = 2048 boxes, re-read
          const int iters = 800;
          for (int rep = 0; rep < (src ? 1 : 2); ++rep) {      // L2 case: first pass warms
            CK(cudaEventRecord(e0));
            ingest_kernel<<<sms, 64, smem>>>(tm, stages, iters, strided, row_blocks, d_cyc, nullptr, 64);
            CK(cudaEventRecord(e1));
            CK(cudaDeviceSynchronize());
          }
          float ms = 0;
          CK(cudaEventElapsedTime(&ms, e0, e1));
          long long cyc[148];
          CK(cudaMemcpy(cyc, d_cyc, sms * sizeof(long long), cudaMemcpyDeviceToHost));
          double avg = 0;
          for (int i = 0; i < sms; ++i) avg += (double)cyc[i];
          avg /= sms;
          printf("%-6s %-8s %4d %6d  %8.1f  %10.0f\n", src ? "HBM" : "L2", strided ? "8K-pitch" : "contig", sms, stages,
                 (double)iters * kBoxBytes / avg, (double)sms * iters * kBoxBytes / (ms * 1e-3) / 1e9);
        }
    }
  // ---- what sets the 42 B/clk?  L2-resident source, contiguous layout, 148 SMs, 8 stages: box shape / swizzle / 1-D bulk
  printf("\nvariant (L2, contiguous, 8 stages)            SMs  B/clk/SM  aggregate GB/s\n");
  struct Var { const char* name; int box_w, box_h, swz, bulk; };
  const Var vars[] = {{"tensor 2d  64 x 128  SWIZZLE_128B", 64, 128, 3, 0}, {"tensor 2d  64 x 128  SWIZZLE_NONE", 64, 128, 0, 0},
                      {"tensor 2d 128 x  64  SWIZZLE_NONE", 128, 64, 0, 0}, {"tensor 2d 256 x  32  SWIZZLE_NONE", 256, 32, 0, 0},
                      {"tensor 2d  32 x 256  SWIZZLE_64B ", 32, 256, 2, 0}, {"cp.async.bulk 1-D 16 KiB         ", 64, 128, 3, 1}};
  for (const Var& v : vars)
    for (int sms : {8, 148}) {
      const size_t bytes = 32ull << 20;
      CUtensorMap tm;
      cuuint64_t gdim[2] = {(cuuint64_t)v.box_w, bytes / (2 * v.box_w)}, gstr[1] = {(cuuint64_t)(2 * v.box_w)};
      cuuint32_t box[2] = {(cuuint32_t)v.box_w, (cuuint32_t)v.box_h}, estr[2] = {1, 1};
      CUresult r = enc(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, buf, gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                       (CUtensorMapSwizzle)v.swz, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
      if (r != CUDA_SUCCESS) { printf("%s: encode failed %d\n", v.name, (int)r); continue; }
      const int iters = 800, row_blocks = (int)(bytes / kBoxBytes);
      for (int rep = 0; rep < 2; ++rep) {
        CK(cudaEventRecord(e0));
        ingest_kernel<<<sms, 64, smem>>>(tm, 8, iters, 0, row_blocks, d_cyc, v.bulk ? buf : nullptr, v.box_w);
        CK(cudaEventRecord(e1));
        CK(cudaDeviceSynchronize());
      }
      float ms = 0;
      CK(cudaEventElapsedTime(&ms, e0, e1));
      long long cyc[148];
      CK(cudaMemcpy(cyc, d_cyc, sms * sizeof(long long), cudaMemcpyDeviceToHost));
      double avg = 0;
      for (int i = 0; i < sms; ++i) avg += (double)cyc[i];
      avg /= sms;
      printf("%s  %4d  %8.1f  %10.0f\n", v.name, sms, (double)iters * kBoxBytes / avg, (double)sms * iters * kBoxBytes / (ms * 1e-3) / 1e9);
    }
  {
    const size_t bytes = 32ull << 20;
    CUtensorMap tm;
    cuuint64_t gdim[2] = {64, bytes / 128}, gstr[1] = {128};
    cuuint32_t box[2] = {64, 128}, estr[2] = {1, 1};
    CUresult r = enc(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, buf, gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { printf("encode failed\n"); return 1; }
    printf("\n");
    for (int sms : {8, 148}) {
      if (run_mc<2>(tm, sms, smem, (int)(bytes / kBoxBytes), d_cyc, e0, e1)) return 1;
      if (run_mc<4>(tm, sms, smem, (int)(bytes / kBoxBytes), d_cyc, e0, e1)) return 1;
    }
    printf("\ncp.async (16 warps, 4 x 2 KiB groups in flight per warp) and TMA (4 stages) in one CTA, L2-resident source, 148 SMs\n");
    const int lsmem = 4 * kBoxBytes + 16 * 4 * 2048 + 64 + 1024;
    CK(cudaFuncSetAttribute(ingest_lsu_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, lsmem));
    for (int mode = 1; mode <= 3; ++mode) {
      const int use_tma = mode & 1, use_lsu = (mode >> 1) & 1, iters = 800;
      for (int rep = 0; rep < 2; ++rep) {
        ingest_lsu_kernel<<<148, 576, lsmem>>>(tm, buf, bytes, iters, use_tma, use_lsu, (int)(bytes / kBoxBytes), d_cyc);
        CK(cudaDeviceSynchronize());
      }
      long long cyc[296];
      CK(cudaMemcpy(cyc, d_cyc, sizeof(cyc), cudaMemcpyDeviceToHost));
      double a_lsu = 0, a_tma = 0;
      for (int i = 0; i < 148; ++i) { a_lsu += (double)cyc[i]; a_tma += (double)cyc[148 + i]; }
      printf("  tma %d  cp.async %d :  TMA %6.1f B/clk/SM   cp.async %6.1f B/clk/SM\n", use_tma, use_lsu,
             use_tma ? (double)iters * kBoxBytes / (a_tma / 148) : 0.0, use_lsu ? (double)iters * 32768 / (a_lsu / 148) : 0.0);
    }
    printf("\ncp.async ring, 16 KiB stages, 6 stages, 148 SMs: who books completion\n");
    const int rsmem = 6 * 32768 + 2 * kMaxStages * 8 + 1024;
    CK(cudaFuncSetAttribute(ingest_ring_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, rsmem));
    for (int srcsel = 0; srcsel < 2; ++srcsel)
      for (int mode = 0; mode < 2; ++mode)
        for (int loaders : {128, 256, 512}) {
          const int iters = 800;
          const size_t sb = srcsel ? big : bytes;
          for (int rep = 0; rep < (srcsel ? 1 : 2); ++rep) {
            ingest_ring_kernel<<<148, 576, rsmem>>>(buf, sb, 6, iters, loaders, mode, d_cyc, 0, 0);
            CK(cudaDeviceSynchronize());
          }
          long long cyc[148];
          CK(cudaMemcpy(cyc, d_cyc, sizeof(cyc), cudaMemcpyDeviceToHost));
          double avg = 0;
          for (int i = 0; i < 148; ++i) avg += (double)cyc[i];
          printf("  %s  %s  %3d loader threads :  %6.1f B/clk/SM\n", srcsel ? "HBM" : "L2 ",
                 mode ? "commit groups, one arrive per warp   " : "mbarrier.arrive.noinc by every thread", loaders,
                 (double)iters * 16384 / (avg / 148));
        }
    for (int v = 0; v < 3; ++v) {
      const int linear = v == 0, nowait = v >= 1, loaders = v == 2 ? 128 : 512;
      for (int rep = 0; rep < 2; ++rep) {
        ingest_ring_kernel<<<148, 576, rsmem>>>(buf, bytes, 6, 800, loaders, 0, d_cyc, linear, nowait);
        CK(cudaDeviceSynchronize());
      }
      long long cyc[148];
      CK(cudaMemcpy(cyc, d_cyc, sizeof(cyc), cudaMemcpyDeviceToHost));
      double avg = 0;
      for (int i = 0; i < 148; ++i) avg += (double)cyc[i];
      printf("  L2   noinc, %d loaders, %s, %s :  %6.1f B/clk/SM\n", loaders, linear ? "linear destination" : "swizzled destination",
             nowait ? "loaders never wait for a free buffer" : "with back-pressure", 800.0 * 16384 / (avg / 148));
    }
  }
  return 0;
}
