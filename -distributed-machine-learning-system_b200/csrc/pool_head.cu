// Tail kernels: 3x3/stride-2 max-pool between layouts, ResNet18 head (global average pool + fp32 FC +
// softmax/top-1 in ONE kernel on clusters of 4 CTAs; the three-kernel form is kept as its cross-check) and
// softmax/top-1 over fp32 logits.
#include "common.cuh"

namespace idn {

// ------------------------------------------------------------------------------------- max-pool
// One thread = one output pixel x 8 channels (one 128-bit load per window tap, one 128-bit store).
// Replaces nn.MaxPool2d(3, 2[, 1]) of tv:models/alexnet.py:24,27,34 and tv:models/resnet.py:201.
// Window taps outside the valid H x W extent are skipped (PyTorch pads with -inf).
__device__ __forceinline__ uint32_t bf16x2_max(uint32_t a, uint32_t b) {
  __nv_bfloat162 x = *reinterpret_cast<__nv_bfloat162*>(&a), y = *reinterpret_cast<__nv_bfloat162*>(&b);
  __nv_bfloat162 m = __hmax2(x, y);
  return *reinterpret_cast<uint32_t*>(&m);
}

// grid = (ceil(OH*OW*C/8 / 256), n); 32-bit index math, packed bf16x2 max (inputs are finite)
__global__ void __launch_bounds__(256) maxpool3x3s2_kernel(const PoolParams p) {
  pdl_wait();
  pdl_launch_dependents();
  const int cvec = p.C >> 3;
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= p.OH * p.OW * cvec) return;
  const int n = blockIdx.y;
  const int cv = t % cvec;
  const int pix = t / cvec;
  const int oh = pix / p.OW, ow = pix - oh * p.OW;
  const int h0 = oh * 2 - p.pad, w0 = ow * 2 - p.pad;
  const uint4* base = reinterpret_cast<const uint4*>(p.in + (long long)n * p.s_img_rows * p.C) + cv;
  const uint32_t ninf = 0xff80ff80u;                       // (-inf, -inf) in bf16
  uint4 m = make_uint4(ninf, ninf, ninf, ninf);
#pragma unroll
  for (int dy = 0; dy < 3; ++dy) {
    const int h = h0 + dy;
    if (h < 0 || h >= p.H) continue;
#pragma unroll
    for (int dx = 0; dx < 3; ++dx) {
      const int w = w0 + dx;
      if (w < 0 || w >= p.W) continue;
      const uint4 v = __ldg(base + (h * p.s_wp + w) * cvec);
      m.x = bf16x2_max(m.x, v.x); m.y = bf16x2_max(m.y, v.y);
      m.z = bf16x2_max(m.z, v.z); m.w = bf16x2_max(m.w, v.w);
    }
  }
  reinterpret_cast<uint4*>(p.out + ((long long)n * p.d_img_rows + oh * p.d_wp + ow) * p.C)[cv] = m;
}

int launch_maxpool(const PoolParams& p, cudaStream_t st) {
  if (p.C % 8 != 0) { set_error("maxpool: C % 8 != 0"); return IDN_EINVAL; }
  const int per_img = p.OH * p.OW * (p.C / 8);
  if (per_img <= 0 || p.nimg <= 0) return IDN_OK;
  dim3 grid((per_img + 255) / 256, p.nimg);
  IDN_CUDA_TRY(launch_pdl(maxpool3x3s2_kernel, grid, dim3(256), 0, st, p));
  return IDN_OK;
}

// ------------------------------------------------------------------------- softmax / top-1 tail
// Warp-level: max + first argmax, then sum of exp(l - max); top-1 probability = 1 / sum.
// Replaces softmax + topk(.,1) of alexnet_resnet.py:80-84; exact ties -> lowest class index.
// The 1000 logits of an image sit in registers (8 float4 per lane, all loads issued before the first use),
// so the row is read once and the two passes cost one memory latency instead of ~60.
__device__ __forceinline__ void warp_softmax_top1_regs(float4 (&v)[8], int lane, int32_t* idx_out,
                                                       float* prob_out, float* copy_out);
__device__ __forceinline__ void warp_softmax_top1(const float* logits, int lane, int32_t* idx_out,
                                                  float* prob_out, float* copy_out) {
  constexpr int kVec = IDN_NUM_CLASSES / 4;          // 250 float4
  float4 v[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int q = lane + 32 * i;
    v[i] = q < kVec ? *(reinterpret_cast<const float4*>(logits) + q)       // global or shared memory
                    : make_float4(-INFINITY, -INFINITY, -INFINITY, -INFINITY);
  }
  warp_softmax_top1_regs(v, lane, idx_out, prob_out, copy_out);
}
// the row's 1000 logits in registers: lane l holds float4 groups l, l + 32, ... (-inf beyond 250)
__device__ __forceinline__ void warp_softmax_top1_regs(float4 (&v)[8], int lane, int32_t* idx_out,
                                                       float* prob_out, float* copy_out) {
  constexpr int kVec = IDN_NUM_CLASSES / 4;
  float best = -INFINITY;
  int best_i = 0x7fffffff;
#pragma unroll
  for (int i = 0; i < 8; ++i) {                        // ascending index within the lane: '>' keeps the first
    const int j = 4 * (lane + 32 * i);
    if (v[i].x > best) { best = v[i].x; best_i = j; }
    if (v[i].y > best) { best = v[i].y; best_i = j + 1; }
    if (v[i].z > best) { best = v[i].z; best_i = j + 2; }
    if (v[i].w > best) { best = v[i].w; best_i = j + 3; }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float ob = __shfl_xor_sync(0xffffffffu, best, o);
    const int oi = __shfl_xor_sync(0xffffffffu, best_i, o);
    if (ob > best || (ob == best && oi < best_i)) { best = ob; best_i = oi; }
  }
  float sum = 0.0f;
#pragma unroll
  for (int i = 0; i < 8; ++i) {                        // expf(-inf) = 0 for the padding lanes
    sum += expf(v[i].x - best); sum += expf(v[i].y - best);
    sum += expf(v[i].z - best); sum += expf(v[i].w - best);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
  if (copy_out != nullptr) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int q = lane + 32 * i;
      if (q < kVec) reinterpret_cast<float4*>(copy_out)[q] = v[i];
    }
  }
  if (lane == 0) {
    *idx_out = best_i;
    *prob_out = 1.0f / sum;
  }
}

__global__ void __launch_bounds__(128) softmax_top1_kernel(const float* __restrict__ logits, int ld,
                                                           int nimg, float* __restrict__ logits_out,
                                                           int32_t* __restrict__ idx,
                                                           float* __restrict__ prob) {
  pdl_wait();
  pdl_launch_dependents();
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (warp >= nimg) return;
  const float* l = logits + (long long)warp * ld;       // ld % 4 == 0: rows are float4-aligned
  warp_softmax_top1(l, lane, idx + warp, prob + warp,
                    logits_out != nullptr ? logits_out + (long long)warp * IDN_NUM_CLASSES : nullptr);
}

// AlexNet's FC8 with split-K leaves `splits` fp32 partial tiles per 128 x 256 tile half
// ([tile half][split][128 rows][256 columns], tile half = (m_pair * num_n_tiles + n_tile) * 2 + CTA rank, see
// conv_tc_pair_kernel); this kernel adds them in split order, applies scale/shift (the bias) -- the arithmetic and
// order of the in-kernel reduction it replaces, so the logits are the same bits -- and goes on to softmax/top-1.
// No CTA of the FC kernel has to wait for another one.
__global__ void __launch_bounds__(128) softmax_top1_splitk_kernel(const float* __restrict__ ws, int splits, int num_n_tiles,
                                                                  const float* __restrict__ scale, const float* __restrict__ shift,
                                                                  int nimg, float* __restrict__ logits_out,
                                                                  int32_t* __restrict__ idx, float* __restrict__ prob) {
  pdl_wait();
  pdl_launch_dependents();
  // one image per block: 128 threads add the partials of two float4 groups each (16 loads in flight) into the image's
  // logits in shared memory, then warp 0 runs the same warp-level softmax/top-1 as softmax_top1_kernel on them
  __shared__ __align__(16) float s_logits[IDN_NUM_CLASSES];
  const int img = blockIdx.x;
  constexpr int kVec = IDN_NUM_CLASSES / 4, kTile = 128 * 256;
  const int m_pair = img >> 8, rank = (img >> 7) & 1, row = img & 127;
  for (int q = threadIdx.x; q < kVec; q += 128) {
    const int n_tile = q >> 6, col = (q & 63) * 4;
    const float* src = ws + ((size_t)((m_pair * num_n_tiles + n_tile) * 2 + rank) * splits) * kTile + row * 256 + col;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int sp0 = 0; sp0 < splits; sp0 += 8) {          // eight partials in flight, added in split order
      float4 t[8];
#pragma unroll
      for (int j = 0; j < 8; ++j)
        t[j] = sp0 + j < splits ? __ldcg(reinterpret_cast<const float4*>(src + (size_t)(sp0 + j) * kTile))
                                : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int j = 0; j < 8; ++j)
        if (sp0 + j < splits) { acc.x += t[j].x; acc.y += t[j].y; acc.z += t[j].z; acc.w += t[j].w; }
    }
    const float4 sc = __ldg(reinterpret_cast<const float4*>(scale) + q), sh = __ldg(reinterpret_cast<const float4*>(shift) + q);
    reinterpret_cast<float4*>(s_logits)[q] =
        make_float4(fmaf(acc.x, sc.x, sh.x), fmaf(acc.y, sc.y, sh.y), fmaf(acc.z, sc.z, sh.z), fmaf(acc.w, sc.w, sh.w));
  }
  __syncthreads();
  if (threadIdx.x < 32)
    warp_softmax_top1(s_logits, (int)threadIdx.x, idx + img, prob + img,
                      logits_out != nullptr ? logits_out + (long long)img * IDN_NUM_CLASSES : nullptr);
}

int launch_softmax_top1_splitk(const float* ws, int splits, int num_n_tiles, const float* scale, const float* shift, int nimg,
                               float* logits_out, int32_t* idx, float* prob, cudaStream_t st) {
  if (nimg <= 0) return IDN_OK;
  IDN_CUDA_TRY(launch_pdl(softmax_top1_splitk_kernel, dim3(nimg), dim3(128), 0, st, ws, splits, num_n_tiles, scale, shift,
                          nimg, logits_out, idx, prob));
  return IDN_OK;
}

int launch_softmax_top1(const float* logits, int ld, int nimg, float* logits_out, int32_t* idx,
                        float* prob, cudaStream_t st) {
  if (nimg <= 0) return IDN_OK;
  IDN_CUDA_TRY(launch_pdl(softmax_top1_kernel, dim3((nimg + 3) / 4), dim3(128), 0, st, logits, ld, nimg, logits_out, idx, prob));
  return IDN_OK;
}

__global__ void __launch_bounds__(256)
copy_words_kernel(uint32_t* __restrict__ dst, const uint32_t* __restrict__ src, size_t n) {
  pdl_wait();
  pdl_launch_dependents();
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    dst[i] = src[i];
}

int launch_copy_words(void* dst, const void* src, size_t n_words, cudaStream_t st) {
  if (n_words == 0) return IDN_OK;
  const unsigned blocks = (unsigned)std::min<size_t>((n_words + 255) / 256, 64);
  IDN_CUDA_TRY(launch_pdl(copy_words_kernel, dim3(blocks), dim3(256), 0, st, static_cast<uint32_t*>(dst),
                          static_cast<const uint32_t*>(src), n_words));
  return IDN_OK;
}

// ---------------------------------------------------------------------------------- ResNet head
// avgpool (tv:models/resnet.py:278) -> fp32 features, fc (:280) in fp32 on fp32 weights, then the
// warp-shuffle softmax/top-1 kernel above (alexnet_resnet.py:80-84).  The FC stays fp32
// (1 MFLOP/image): the last-layer rounding is what decides near-tie top-1s.
__global__ void __launch_bounds__(256)
avgpool_kernel(const bf16* __restrict__ act, int nimg, int img_rows, int wp, int H, int W, int C,
               float* __restrict__ feats) {
  pdl_wait();
  pdl_launch_dependents();
  const int cpairs = C >> 1;
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= nimg * cpairs) return;
  const int n = t / cpairs, cp = t - n * cpairs;
  const uint32_t* base = reinterpret_cast<const uint32_t*>(act + ((long long)n * img_rows) * C) + cp;
  float s0 = 0.0f, s1 = 0.0f;
  for (int h = 0; h < H; ++h) {
#pragma unroll 7
    for (int w = 0; w < W; ++w) {
      const uint32_t u = __ldg(base + (long long)(h * wp + w) * cpairs);
      s0 += __uint_as_float(u << 16);
      s1 += __uint_as_float(u & 0xffff0000u);
    }
  }
  const float inv = 1.0f / (float)(H * W);
  reinterpret_cast<float2*>(feats)[t] = make_float2(s0 * inv, s1 * inv);
}

// logits[n][j] = bias[j] + sum_c feats[n][c] * fc_w[j][c], fp32, c ascending (deterministic).  A
// register-tiled SIMT GEMM: 32 images x 64 classes per CTA (13 x 16 = 208 CTAs at batch 400, two per SM:
// the kernel is latency-bound, so more and smaller CTAs beat fewer big ones), 4 x 4 outputs per thread, K in
// slabs of 32 staged k-major in shared memory: one k step costs a thread two 128-bit shared loads for
// 16 FMAs.  The next slab is fetched into registers while the current one is multiplied.
constexpr int kFcTileM = 32, kFcTileN = 64, kFcK = 32, kFcThreads = (kFcTileM / 4) * (kFcTileN / 4);

__global__ void __launch_bounds__(kFcThreads)
fc_fp32_kernel(const float* __restrict__ feats, int nimg, int C, const float* __restrict__ fc_w,
               const float* __restrict__ fc_b, float* __restrict__ logits) {
  pdl_wait();
  pdl_launch_dependents();
  __shared__ __align__(16) float As[kFcK][kFcTileM + 4];     // [k][image]
  __shared__ __align__(16) float Bs[kFcK][kFcTileN + 4];     // [k][class]
  const int m0 = blockIdx.x * kFcTileM, n0 = blockIdx.y * kFcTileN;
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  constexpr int kLoadsA = kFcTileM * (kFcK / 4) / kFcThreads, kLoadsB = kFcTileN * (kFcK / 4) / kFcThreads;
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.0f;
  float4 ra[kLoadsA], rb[kLoadsB];
  auto fetch = [&](int k0) {
#pragma unroll
    for (int j = 0; j < kLoadsA; ++j) {
      const int idx = threadIdx.x + kFcThreads * j;
      const int row = idx % kFcTileM, kq = (idx / kFcTileM) * 4;
      const int m = m0 + row;
      ra[j] = m < nimg ? __ldg(reinterpret_cast<const float4*>(feats + (long long)m * C + k0 + kq))
                       : make_float4(0.f, 0.f, 0.f, 0.f);
    }
#pragma unroll
    for (int j = 0; j < kLoadsB; ++j) {
      const int idx = threadIdx.x + kFcThreads * j;
      const int row = idx % kFcTileN, kq = (idx / kFcTileN) * 4;
      const int n = n0 + row;
      rb[j] = n < IDN_NUM_CLASSES ? __ldg(reinterpret_cast<const float4*>(fc_w + (long long)n * C + k0 + kq))
                                  : make_float4(0.f, 0.f, 0.f, 0.f);
    }
  };
  fetch(0);
  for (int k0 = 0; k0 < C; k0 += kFcK) {
#pragma unroll
    for (int j = 0; j < kLoadsA; ++j) {
      const int idx = threadIdx.x + kFcThreads * j;
      const int row = idx % kFcTileM, kq = (idx / kFcTileM) * 4;
      As[kq + 0][row] = ra[j].x; As[kq + 1][row] = ra[j].y; As[kq + 2][row] = ra[j].z; As[kq + 3][row] = ra[j].w;
    }
#pragma unroll
    for (int j = 0; j < kLoadsB; ++j) {
      const int idx = threadIdx.x + kFcThreads * j;
      const int row = idx % kFcTileN, kq = (idx / kFcTileN) * 4;
      Bs[kq + 0][row] = rb[j].x; Bs[kq + 1][row] = rb[j].y; Bs[kq + 2][row] = rb[j].z; Bs[kq + 3][row] = rb[j].w;
    }
    __syncthreads();
    if (k0 + kFcK < C) fetch(k0 + kFcK);
#pragma unroll
    for (int k = 0; k < kFcK; ++k) {
      const float4 a = *reinterpret_cast<const float4*>(&As[k][ty * 4]);
      const float4 b = *reinterpret_cast<const float4*>(&Bs[k][tx * 4]);
      const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
    }
    __syncthreads();
  }
  const int n = n0 + tx * 4;
  if (n < IDN_NUM_CLASSES) {                         // 1000 = 15 full tiles + 40: whole float4s either way
    const float4 bias = __ldg(reinterpret_cast<const float4*>(fc_b + n));
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int m = m0 + ty * 4 + i;
      if (m < nimg)
        *reinterpret_cast<float4*>(logits + (long long)m * IDN_NUM_CLASSES + n) =
            make_float4(acc[i][0] + bias.x, acc[i][1] + bias.y, acc[i][2] + bias.z, acc[i][3] + bias.w);
    }
  }
}

// ------------------------------------------------------------------------ ResNet head, ONE kernel
// avgpool -> fp32 FC -> softmax / top-1 for 23 images per cluster of 8 CTAs (18 clusters = 144 CTAs at batch
// 400, one wave), nothing but the 8 result bytes per image (and, on request, the logits) leaves the chip:
//   1. every CTA averages an eighth of the channels of the cluster's 23 images (49 pixels, h then w ascending,
//      then * 1/49: the order of avgpool_kernel; the 49 loads of an item are all in flight before the first add);
//      the eight slices meet in every CTA's shared memory by way of a small global scratch (L2) and a cluster barrier;
//   2. CTA r owns classes 125 r .. 125 r + 124; a thread accumulates 4 classes x 6 images in registers, k
//      ascending with fmaf, bias added last -- the arithmetic and order of fc_fp32_kernel, so logits are
//      identical bit for bit.  The CTA's slice of the transposed fp32 weights ([512][128], one contiguous block
//      per cluster rank, built at load time) streams through a three-stage shared-memory ring, ONE 16 KB bulk
//      copy (cp.async.bulk + mbarrier) per 32 rows: the first version read weights with plain loads and was bound
//      by L2 latency -- every cluster asks for the same lines at the same moment (ncu: 70 % of the samples on
//      FFMAs waiting for their weight).  Logits go to the shared memory of the CTA that owns the image (i % 8);
//   3. one warp per owned image runs the same warp-shuffle softmax / top-1 as softmax_top1_kernel on the 1000
//      logits in shared memory.
constexpr int kHeadImgs = 23, kHeadCluster = 8, kHeadClasses = IDN_NUM_CLASSES / kHeadCluster, kHeadThreads = 384;
constexpr int kHeadHW = 7, kHeadC = 512, kHeadPad = 24, kHeadGroup = 6, kHeadIGroups = 4;
constexpr int kHeadOwn = (kHeadImgs + kHeadCluster - 1) / kHeadCluster;           // 3
constexpr int kHeadKStage = 32, kHeadStages = 3, kHeadWPitch = 128;
constexpr int kHeadStageBytes = kHeadKStage * kHeadWPitch * 4;                      // 16 KB: one bulk copy
constexpr int kHeadSmemBytes = (kHeadC * kHeadPad + kHeadOwn * IDN_NUM_CLASSES + kHeadStages * kHeadKStage * kHeadWPitch) * 4 + 64;
static_assert(kHeadClasses * kHeadCluster == IDN_NUM_CLASSES && kHeadClasses <= kHeadWPitch, "class split");
static_assert(kHeadIGroups * kHeadGroup >= kHeadImgs && kHeadIGroups * kHeadGroup <= kHeadPad && kHeadIGroups * 32 <= kHeadThreads, "image groups");

__device__ __forceinline__ uint32_t head_mapa(const void* p, uint32_t rank) {
  uint32_t a = (uint32_t)__cvta_generic_to_shared(p), r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(rank));
  return r;
}
__device__ __forceinline__ void head_st_remote(uint32_t addr, float v) {
  asm volatile("st.shared::cluster.f32 [%0], %1;" :: "r"(addr), "f"(v) : "memory");
}
__device__ __forceinline__ void head_cluster_sync() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void head_bulk_load(float* smem_dst, const float* gsrc, uint32_t bytes, uint64_t* bar) {
  const uint32_t b = (uint32_t)__cvta_generic_to_shared(bar);
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(b), "r"(bytes) : "memory");
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               :: "r"((uint32_t)__cvta_generic_to_shared(smem_dst)), "l"(gsrc), "r"(bytes), "r"(b) : "memory");
}
__device__ __forceinline__ void head_bar_wait(uint64_t* bar, uint32_t parity) {
  const uint32_t b = (uint32_t)__cvta_generic_to_shared(bar);
  uint32_t ok = 0;
  while (!ok)
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(ok) : "r"(b), "r"(parity) : "memory");
}

__global__ void __cluster_dims__(kHeadCluster, 1, 1) __launch_bounds__(kHeadThreads)
resnet_head_kernel(const bf16* __restrict__ act, int nimg, int img_rows, int wp,
                   const float* __restrict__ wt8 /* [8 ranks][512][128]: transposed, one slice per cluster rank */,
                   const float* __restrict__ fc_b, float* feats /* [n][512] scratch */, float* __restrict__ logits_out,
                   int32_t* __restrict__ idx, float* __restrict__ prob) {
  extern __shared__ __align__(16) float head_smem[];
  float (*s_feat)[kHeadPad] = reinterpret_cast<float (*)[kHeadPad]>(head_smem);                       // [k][image]
  float (*s_logit)[IDN_NUM_CLASSES] = reinterpret_cast<float (*)[IDN_NUM_CLASSES]>(head_smem + kHeadC * kHeadPad);
  float* s_w = head_smem + kHeadC * kHeadPad + kHeadOwn * IDN_NUM_CLASSES;                            // [stage][k][class]
  uint64_t* s_bar = reinterpret_cast<uint64_t*>(s_w + kHeadStages * kHeadKStage * kHeadWPitch);        // full[stage]
  uint32_t rank;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(rank));
  const int img0 = (int)(blockIdx.x / kHeadCluster) * kHeadImgs;
  const int t = threadIdx.x;
  const float* wslice = wt8 + (size_t)rank * kHeadC * kHeadWPitch;
  constexpr int kNumStages = kHeadC / kHeadKStage;                 // 16 stages of 32 rows = 16 KB each
  if (t == 0) {
    for (int i = 0; i < kHeadStages; ++i)
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"((uint32_t)__cvta_generic_to_shared(&s_bar[i])) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    for (int i = 0; i < kHeadStages; ++i)                  // weights do not depend on the previous kernel: start now
      head_bulk_load(s_w + i * kHeadKStage * kHeadWPitch, wslice + (size_t)i * kHeadKStage * kHeadWPitch, kHeadStageBytes, &s_bar[i]);
  }
  head_cluster_sync();                                     // every CTA of the cluster runs: its shared memory may be written
  pdl_wait();
  pdl_launch_dependents();
  // ---- 1. global average pool (7 x 7) of channels [64 rank, 64 rank + 64): item = (channel pair, image)
  for (int item = t; item < 32 * kHeadImgs; item += kHeadThreads) {
    const int cp = item & 31, im = item >> 5;
    const int k0 = 64 * (int)rank + 2 * cp;
    const int n = img0 + im;
    float s0 = 0.0f, s1 = 0.0f;
    if (n < nimg) {
      const uint32_t* base = reinterpret_cast<const uint32_t*>(act + ((long long)n * img_rows) * kHeadC) + (k0 >> 1);
      uint32_t u[kHeadHW * kHeadHW];
#pragma unroll
      for (int h = 0; h < kHeadHW; ++h)
#pragma unroll
        for (int w = 0; w < kHeadHW; ++w) u[h * kHeadHW + w] = __ldg(base + (long long)(h * wp + w) * (kHeadC / 2));
#pragma unroll
      for (int j = 0; j < kHeadHW * kHeadHW; ++j) {                 // h, then w, ascending: avgpool_kernel's order
        s0 += __uint_as_float(u[j] << 16);
        s1 += __uint_as_float(u[j] & 0xffff0000u);
      }
    }
    const float inv = 1.0f / (float)(kHeadHW * kHeadHW);
    if (n < nimg) *reinterpret_cast<float2*>(feats + (long long)n * kHeadC + k0) = make_float2(s0 * inv, s1 * inv);
  }
  // The eight CTAs exchange their channel slices through global memory (L2): a warp writes 256 contiguous
  // bytes and reads them back coalesced.  (Storing each value into eight remote shared memories cost one
  // 4-byte DSMEM transaction per lane and store -- ~12 k of them arriving per CTA -- and was most of the kernel.)
  __threadfence();
  head_cluster_sync();
  for (int i = t; i < kHeadImgs * kHeadC; i += kHeadThreads) {
    const int im = i / kHeadC, k = i - im * kHeadC;
    const int n = img0 + im;
    s_feat[k][im] = n < nimg ? __ldcg(feats + (long long)n * kHeadC + k) : 0.0f;
  }
  __syncthreads();
  // ---- 2. fp32 FC: thread = 4 classes x 6 images (128 threads: 32 class groups x 4 image groups).  The loop is
  //         bound by shared-memory wavefronts (a warp-wide load costs one per 128 requested bytes, broadcast or
  //         not), so each thread reuses what it loads: 10 words for 24 FMAs per k.  k ascending, fmaf, bias last.
  {
    const int cg = t & 31, ig = t >> 5;                              // ig < 4 computes; the other warps only wait
    const bool worker = ig < kHeadIGroups;
    const int i0 = ig * kHeadGroup;
    float acc[kHeadGroup][4];
#pragma unroll
    for (int i = 0; i < kHeadGroup; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[i][j] = 0.0f;
    for (int st = 0; st < kNumStages; ++st) {
      const int slot = st % kHeadStages;
      if (worker) {
        head_bar_wait(&s_bar[slot], (uint32_t)((st / kHeadStages) & 1));
        const float* w = s_w + slot * kHeadKStage * kHeadWPitch + 4 * cg;
#pragma unroll 4
        for (int kk = 0; kk < kHeadKStage; ++kk) {
          const float4 wv = *reinterpret_cast<const float4*>(w + kk * kHeadWPitch);
          const float* f = &s_feat[st * kHeadKStage + kk][i0];       // i0 * 4 bytes is a multiple of 8
          const float2 f0 = *reinterpret_cast<const float2*>(f), f1 = *reinterpret_cast<const float2*>(f + 2),
                       f2 = *reinterpret_cast<const float2*>(f + 4);
          const float fv[kHeadGroup] = {f0.x, f0.y, f1.x, f1.y, f2.x, f2.y};
#pragma unroll
          for (int i = 0; i < kHeadGroup; ++i) {
            acc[i][0] = fmaf(fv[i], wv.x, acc[i][0]); acc[i][1] = fmaf(fv[i], wv.y, acc[i][1]);
            acc[i][2] = fmaf(fv[i], wv.z, acc[i][2]); acc[i][3] = fmaf(fv[i], wv.w, acc[i][3]);
          }
        }
      }
      __syncthreads();                                               // everybody is done with this ring slot
      if (t == 0 && st + kHeadStages < kNumStages)
        head_bulk_load(s_w + slot * kHeadKStage * kHeadWPitch, wslice + (size_t)(st + kHeadStages) * kHeadKStage * kHeadWPitch,
                       kHeadStageBytes, &s_bar[slot]);
    }
    if (worker) {
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int cl = 4 * cg + j;
        if (cl < kHeadClasses) {
          const int c = kHeadClasses * (int)rank + cl;
          const float b = __ldg(fc_b + c);
#pragma unroll
          for (int i = 0; i < kHeadGroup; ++i) {
            const int im = i0 + i;
            if (im < kHeadImgs)
              head_st_remote(head_mapa(&s_logit[im / kHeadCluster][c], (uint32_t)(im % kHeadCluster)), acc[i][j] + b);
          }
        }
      }
    }
  }
  head_cluster_sync();
  // ---- 3. softmax / top-1 of the images this CTA owns: image rank + 8 w by warp w
  const int warp = t >> 5, lane = t & 31;
  if (warp < kHeadOwn) {
    const int im = (int)rank + kHeadCluster * warp;
    const int n = img0 + im;
    if (im < kHeadImgs && n < nimg)
      warp_softmax_top1(&s_logit[warp][0], lane, idx + n, prob + n,
                        logits_out != nullptr ? logits_out + (long long)n * IDN_NUM_CLASSES : nullptr);
  }
}

int launch_resnet_head_fused(const bf16* act, int nimg, int img_rows, int wp, int H, int W, int C,
                             const float* fc_wt, const float* fc_b, float* feats, float* logits_out, int32_t* idx,
                             float* prob, cudaStream_t st) {
  if (nimg <= 0) return IDN_OK;
  if (C != kHeadC || H != kHeadHW || W != kHeadHW || fc_wt == nullptr) {
    set_error("resnet_head: the fused head is built for 512 features on a 7 x 7 map");
    return IDN_EINVAL;
  }
  static bool configured[64] = {false};
  int dev = 0;
  cudaGetDevice(&dev);
  if (!configured[dev & 63]) {
    IDN_CUDA_TRY(cudaFuncSetAttribute(resnet_head_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kHeadSmemBytes));
    configured[dev & 63] = true;
  }
  const int clusters = (nimg + kHeadImgs - 1) / kHeadImgs;
  IDN_CUDA_TRY(launch_pdl(resnet_head_kernel, dim3(clusters * kHeadCluster), dim3(kHeadThreads), (size_t)kHeadSmemBytes, st,
                          act, nimg, img_rows, wp, fc_wt, fc_b, feats, logits_out, idx, prob));
  return IDN_OK;
}

int launch_resnet_head(const bf16* act, int nimg, int img_rows, int wp, int H, int W, int C,
                       const float* fc_w, const float* fc_b, float* feats, float* logits,
                       int32_t* idx, float* prob, cudaStream_t st) {
  if (nimg <= 0) return IDN_OK;
  if (C % kFcK != 0 || logits == nullptr || feats == nullptr) { set_error("resnet_head: bad arguments"); return IDN_EINVAL; }
  const int items = nimg * (C / 2);
  IDN_CUDA_TRY(launch_pdl(avgpool_kernel, dim3((items + 255) / 256), dim3(256), 0, st, act, nimg, img_rows, wp, H, W, C, feats));
  dim3 grid((nimg + kFcTileM - 1) / kFcTileM, (IDN_NUM_CLASSES + kFcTileN - 1) / kFcTileN);
  IDN_CUDA_TRY(launch_pdl(fc_fp32_kernel, grid, dim3(kFcThreads), 0, st, feats, nimg, C, fc_w, fc_b, logits));
  return launch_softmax_top1(logits, IDN_NUM_CLASSES, nimg, nullptr, idx, prob, st);
}

}  // namespace idn
