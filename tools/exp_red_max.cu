#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ void red_max_v4(void* p, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("red.global.v4.bf16x2.max.noftz [%0], {%1, %2, %3, %4};" :: "l"(p), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__global__ void k_red(uint4* out, size_t n16, int reps, uint32_t v) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (int r = 0; r < reps; ++r)
    for (size_t j = i; j < n16; j += stride) red_max_v4(out + j, v + r, v, v, v);
}
__global__ void k_st(uint4* out, size_t n16, int reps, uint32_t v) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (int r = 0; r < reps; ++r)
    for (size_t j = i; j < n16; j += stride) out[j] = make_uint4(v + r, v, v, v);
}
int main() {
  const size_t bytes = 160ull << 20;           // the pooled ResNet stem output of 400 images
  uint4* d; cudaMalloc(&d, bytes); cudaMemset(d, 0, bytes);
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  for (int mode = 0; mode < 2; ++mode) {
    for (int warm = 0; warm < 2; ++warm) {
      cudaEventRecord(a);
      if (mode == 0) k_red<<<148 * 8, 256>>>(d, bytes / 16, 2, 0x3f803f80u);
      else k_st<<<148 * 8, 256>>>(d, bytes / 16, 2, 0x3f803f80u);
      cudaEventRecord(b); cudaEventSynchronize(b);
      float ms; cudaEventElapsedTime(&ms, a, b);
      if (warm) printf("%s: %.3f ms for 2 x %zu MB -> %.1f GB/s of 16-byte ops (%.1f Gops/s)  err=%s\n", mode ? "st.v4" : "red.v4.bf16x2.max",
                       ms, bytes >> 20, 2.0 * bytes / ms / 1e6, 2.0 * bytes / 16 / ms / 1e6, cudaGetErrorString(cudaGetLastError()));
    }
  }
  // semantics: max per bf16 half
  uint32_t h[4]; cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost); printf("word0 %08x word1 %08x\n", h[0], h[1]);
  return 0;
}
