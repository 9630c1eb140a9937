// Hardware experiment (bring-up, not product code): can a tcgen05.mma A operand start at a row that
// is NOT a multiple of 8 inside a 128B-swizzled tile written by one TMA box (halo reuse across
// filter taps)?  For every row offset 0..11 and three encodings of the descriptor's base_offset
// field, D = A[off : off+128] * I is compared with the expected rows.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -std=c++17 tools/exp_base_offset.cu -o /tmp/exp_bo
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <cstring>
#include <vector>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

constexpr int ROWS = 256;     // rows in the staged A tile (two 128-row boxes)

__global__ void __launch_bounds__(128, 1)
exp_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
           float* out, int off, int bo_mode, int k_steps) {
  extern __shared__ uint8_t raw[];
  const uint32_t base = (smem_u32(raw) + 1023u) & ~1023u;
  const uint32_t a_addr = base, b_addr = base + ROWS * 128, bar = b_addr + 64 * 128, bar2 = bar + 8, slot = bar + 16;
  volatile uint32_t* slot_ptr = (volatile uint32_t*)(raw + (slot - smem_u32(raw)));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar));
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar2));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 64;" ::"r"(slot));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = *slot_ptr;
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(ROWS * 128 + 64 * 128) : "memory");
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                 ::"r"(a_addr), "l"(&tmA), "r"(bar), "r"(0), "r"(0) : "memory");
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                 ::"r"(a_addr + 128 * 128), "l"(&tmA), "r"(bar), "r"(0), "r"(128) : "memory");
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                 ::"r"(b_addr), "l"(&tmB), "r"(bar), "r"(0), "r"(0) : "memory");
  }
  // everyone waits for the data
  {
    uint32_t ok = 0;
    while (!ok) asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\nselp.u32 %0,1,0,p;\n}" : "=r"(ok) : "r"(bar) : "memory");
  }
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  if (threadIdx.x == 0) {
    const uint32_t a_start = a_addr + off * 128;
    uint32_t bo = 0;
    if (bo_mode == 1) bo = (a_start >> 7) & 7;
    if (bo_mode == 2) bo = (8 - ((a_start >> 7) & 7)) & 7;
    uint64_t da = ((uint64_t)((a_start & 0x3ffff) >> 4)) | (1ull << 16) | ((uint64_t)(1024 >> 4) << 32) | (1ull << 46) |
                  ((uint64_t)bo << 49) | (2ull << 61);
    uint64_t db = ((uint64_t)((b_addr & 0x3ffff) >> 4)) | (1ull << 16) | ((uint64_t)(1024 >> 4) << 32) | (1ull << 46) | (2ull << 61);
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((64u >> 3) << 17) | ((128u >> 4) << 24);
    for (int k = 0; k < k_steps; ++k) {
      asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n}"
                   ::"r"(tmem), "l"(da + 2 * k), "l"(db + 2 * k), "r"(idesc), "r"(k) : "memory");
    }
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar2) : "memory");
  }
  {
    uint32_t ok = 0;
    while (!ok) asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\nselp.u32 %0,1,0,p;\n}" : "=r"(ok) : "r"(bar2) : "memory");
  }
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  uint32_t v[32];
  for (int c0 = 0; c0 < 64; c0 += 32) {
    const uint32_t taddr = tmem + ((uint32_t)(warp * 32) << 16) + c0;
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
                 : "=r"(v[0]),"=r"(v[1]),"=r"(v[2]),"=r"(v[3]),"=r"(v[4]),"=r"(v[5]),"=r"(v[6]),"=r"(v[7]),"=r"(v[8]),"=r"(v[9]),"=r"(v[10]),"=r"(v[11]),"=r"(v[12]),"=r"(v[13]),"=r"(v[14]),"=r"(v[15]),
                   "=r"(v[16]),"=r"(v[17]),"=r"(v[18]),"=r"(v[19]),"=r"(v[20]),"=r"(v[21]),"=r"(v[22]),"=r"(v[23]),"=r"(v[24]),"=r"(v[25]),"=r"(v[26]),"=r"(v[27]),"=r"(v[28]),"=r"(v[29]),"=r"(v[30]),"=r"(v[31])
                 : "r"(taddr) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    for (int j = 0; j < 32; ++j) out[(warp * 32 + lane) * 64 + c0 + j] = __uint_as_float(v[j]);
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 64;" ::"r"(tmem));
}

typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                             const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                             CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main() {
  void* fp = nullptr; cudaDriverEntryPointQueryResult q;
  CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fp, cudaEnableDefault, &q));
  EncodeFn enc = (EncodeFn)fp;
  // A[ROWS][64]: value = row + col/64 (exact in bf16 for small rows? use row*0.5 + col*0.001 rounded)
  std::vector<__nv_bfloat16> ha(ROWS * 64), hb(64 * 64);
  std::vector<float> fa(ROWS * 64);
  for (int r = 0; r < ROWS; ++r) for (int c = 0; c < 64; ++c) {
    float v = (float)((r * 7 + c * 3) % 251) - 125.0f;       // integers: exact in bf16
    ha[r * 64 + c] = __float2bfloat16(v); fa[r * 64 + c] = v;
  }
  for (int n = 0; n < 64; ++n) for (int k = 0; k < 64; ++k) hb[n * 64 + k] = __float2bfloat16(n == k ? 1.0f : 0.0f);
  __nv_bfloat16 *da, *db; float* dout;
  CK(cudaMalloc(&da, ha.size() * 2)); CK(cudaMalloc(&db, hb.size() * 2)); CK(cudaMalloc(&dout, 128 * 64 * 4));
  CK(cudaMemcpy(da, ha.data(), ha.size() * 2, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(db, hb.data(), hb.size() * 2, cudaMemcpyHostToDevice));
  CUtensorMap tmA, tmB;
  cuuint64_t gd[2] = {64, ROWS}, gs[1] = {128}; cuuint32_t box[2] = {64, 128}, es[2] = {1, 1};
  if (enc(&tmA, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, da, gd, gs, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE)) { printf("encode A failed\n"); return 1; }
  cuuint64_t gdb[2] = {64, 64}; cuuint32_t boxb[2] = {64, 64};
  if (enc(&tmB, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, db, gdb, gs, boxb, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE)) { printf("encode B failed\n"); return 1; }
  const int smem = ROWS * 128 + 64 * 128 + 64 + 1024;
  CK(cudaFuncSetAttribute(exp_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  std::vector<float> ho(128 * 64);
  for (int off = 0; off <= 11; ++off) {
    for (int mode = 0; mode < 3; ++mode) {
      CK(cudaMemset(dout, 0, 128 * 64 * 4));
      exp_kernel<<<1, 128, smem>>>(tmA, tmB, dout, off, mode, 4);
      CK(cudaDeviceSynchronize());
      CK(cudaMemcpy(ho.data(), dout, ho.size() * 4, cudaMemcpyDeviceToHost));
      int bad = 0, first = -1;
      for (int m = 0; m < 128; ++m) for (int c = 0; c < 64; ++c)
        if (ho[m * 64 + c] != fa[(m + off) * 64 + c]) { if (first < 0) first = m * 64 + c; ++bad; }
      printf("off %2d  base_offset mode %d : %s (%d mismatches%s)\n", off, mode, bad ? "WRONG" : "ok", bad,
             bad ? "" : "");
      if (bad && first >= 0) {
        int m = first / 64, c = first % 64;
        printf("        first mismatch row %d col %d: got %.1f want %.1f; row got[0..7] =", m, c, ho[first], fa[(m + off) * 64 + c]);
        for (int j = 0; j < 8; ++j) printf(" %.0f", ho[m * 64 + j]);
        printf(" | want"); for (int j = 0; j < 8; ++j) printf(" %.0f", fa[(m + off) * 64 + j]); printf("\n");
      }
    }
  }
  return 0;
}
