// Hardware experiment (bring-up, not product code): a NO-SWIZZLE K-major UMMA descriptor whose
// 8x16-byte core matrices OVERLAP (leading-dimension byte offset = 16): A[m][k] = raw[m*8 + k],
// i.e. im2col of a stride-2, 4-channel image row done by the descriptor itself.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -std=c++17 tools/exp_toeplitz.cu -o tools/exp_toeplitz.bin
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <vector>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// raw: strips of `strip_elems` bf16 each, 16 strips; A row m = 8g + i  ->  strip g, element i*8 + k
// (sbo = strip pitch in bytes; sbo = 128 with one long strip gives the pure Toeplitz matrix)
__global__ void __launch_bounds__(128, 1)
exp_kernel(const __nv_bfloat16* raw, int raw_elems, const __nv_bfloat16* bmat /* [64][32] */, float* out,
           int sbo_bytes, int lbo_bytes) {
  extern __shared__ uint8_t rawsm[];
  const uint32_t base = (smem_u32(rawsm) + 1023u) & ~1023u;
  uint8_t* gen = rawsm + (base - smem_u32(rawsm));
  __nv_bfloat16* a_s = reinterpret_cast<__nv_bfloat16*>(gen);                 // raw strip(s), 16 KiB
  __nv_bfloat16* b_s = reinterpret_cast<__nv_bfloat16*>(gen + 16384);          // B, no-swizzle canonical
  const uint32_t bar = base + 16384 + 8192, slot = bar + 8;
  volatile uint32_t* slot_ptr = (volatile uint32_t*)(gen + 16384 + 8192 + 8);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < raw_elems; i += 128) a_s[i] = raw[i];
  // B canonical no-swizzle K-major: element (n, k) at ((k/8) * 64 rows + n) * 16 bytes + (k%8)*2
  for (int i = threadIdx.x; i < 64 * 32; i += 128) {
    const int n = i / 32, k = i % 32;
    b_s[((k / 8) * 64 + n) * 8 + (k % 8)] = bmat[i];
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar));
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
    const uint32_t a_addr = base, b_addr = base + 16384;
    // no swizzle (layout type 0): LBO = bytes between core matrices along K, SBO = between 8-row groups
    uint64_t da = ((uint64_t)((a_addr & 0x3ffff) >> 4)) | ((uint64_t)(lbo_bytes >> 4) << 16) |
                  ((uint64_t)(sbo_bytes >> 4) << 32) | (1ull << 46);
    uint64_t db = ((uint64_t)((b_addr & 0x3ffff) >> 4)) | ((uint64_t)(1024 >> 4) << 16) |
                  ((uint64_t)(128 >> 4) << 32) | (1ull << 46);
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((64u >> 3) << 17) | ((128u >> 4) << 24);
    for (int k = 0; k < 2; ++k) {      // K = 32 = 2 x 16; second step: A +32 bytes, B +2 core matrices
      asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n}"
                   ::"r"(tmem), "l"(da + (uint64_t)(2 * k)), "l"(db + (uint64_t)((2048 >> 4) * k)), "r"(idesc), "r"(k) : "memory");
    }
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
  }
  { uint32_t ok = 0; long spins = 0;
    while (!ok && spins < 100000000) { ++spins; asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\nselp.u32 %0,1,0,p;\n}" : "=r"(ok) : "r"(bar) : "memory"); } }
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

int main() {
  const int RAW = 8192;   // elements
  std::vector<__nv_bfloat16> hraw(RAW), hb(64 * 32);
  std::vector<float> fraw(RAW);
  for (int i = 0; i < RAW; ++i) { float v = (float)((i * 5 + 3) % 127) - 63.0f; hraw[i] = __float2bfloat16(v); fraw[i] = v; }
  for (int n = 0; n < 64; ++n) for (int k = 0; k < 32; ++k) hb[n * 32 + k] = __float2bfloat16((k == n % 32) ? (n < 32 ? 1.0f : 2.0f) : 0.0f);
  __nv_bfloat16 *draw, *db; float* dout;
  CK(cudaMalloc(&draw, RAW * 2)); CK(cudaMalloc(&db, 64 * 32 * 2)); CK(cudaMalloc(&dout, 128 * 64 * 4));
  CK(cudaMemcpy(draw, hraw.data(), RAW * 2, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(db, hb.data(), 64 * 32 * 2, cudaMemcpyHostToDevice));
  const int smem = 16384 + 8192 + 64 + 1024;
  CK(cudaFuncSetAttribute(exp_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  std::vector<float> ho(128 * 64);
  // (sbo, lbo): (128,16) pure Toeplitz stride 8 elements; (176,16) and (192,16): 16 strips of 8 output columns
  const int cfg[3][2] = {{128, 16}, {176, 16}, {192, 16}};
  for (int c = 0; c < 3; ++c) {
    const int sbo = cfg[c][0], lbo = cfg[c][1];
    CK(cudaMemset(dout, 0, 128 * 64 * 4));
    exp_kernel<<<1, 128, smem>>>(draw, RAW, db, dout, sbo, lbo);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("sbo %d lbo %d: kernel failed: %s\n", sbo, lbo, cudaGetErrorString(e)); return 1; }
    CK(cudaMemcpy(ho.data(), dout, ho.size() * 4, cudaMemcpyDeviceToHost));
    int bad = 0, first = -1;
    for (int m = 0; m < 128; ++m) for (int n = 0; n < 64; ++n) {
      const int g = m / 8, i = m % 8;
      const float want = fraw[g * (sbo / 2) + i * 8 + (n % 32)] * (n < 32 ? 1.0f : 2.0f);
      if (ho[m * 64 + n] != want) { if (first < 0) first = m * 64 + n; ++bad; }
    }
    printf("toeplitz sbo %3d lbo %2d : %s (%d mismatches)\n", sbo, lbo, bad ? "WRONG" : "ok", bad);
    if (bad) { int m = first / 64, n = first % 64; printf("   first mismatch m %d n %d got %.1f; row got[0..7] =", m, n, ho[first]);
      for (int j = 0; j < 8; ++j) printf(" %.0f", ho[m * 64 + j]); printf(" | raw[m*8..] ="); for (int j = 0; j < 8; ++j) printf(" %.0f", fraw[(m/8)*(sbo/2) + (m%8)*8 + j]); printf("\n"); }
  }
  return 0;
}
