// uint8 HWC image batch -> Resize(256) / CenterCrop(224) / ToTensor / Normalize -> bf16.
//
// Replaces alexnet_resnet.py:57-66.  The resize is Pillow's BILINEAR resample restated in its own
// 22-bit fixed point (horizontal pass -> uint8 -> vertical pass -> uint8), so the uint8 stage is
// bit-exact; only the 224 x 224 cropped outputs are evaluated.  Per-axis coefficient tables
// depend on (source size, resized size, crop offset) only and are built on the host in double
// precision exactly as Pillow does; the kernel performs integer MACs.  u8 -> (x/255 - mean)/std
// is a 768-entry table of bf16 values computed in IEEE fp32 in torchvision's operation order.
#include <cmath>
#include <vector>

#include "common.cuh"
#include <type_traits>

namespace idn {

static const float kMean[3] = {0.485f, 0.456f, 0.406f};   // alexnet_resnet.py:61
static const float kStd[3] = {0.229f, 0.224f, 0.225f};

void build_norm_lut(uint16_t* lut768) {
  for (int c = 0; c < 3; ++c)
    for (int v = 0; v < 256; ++v) {
      volatile float x = (float)v / 255.0f;     // ToTensor: tv:transforms/functional.py:175-176
      volatile float y = x - kMean[c];          // Normalize: sub then div,
      volatile float z = y / kStd[c];           //   tv:transforms/_functional_tensor.py:928
      lut768[c * 256 + v] = f32_to_bf16_bits(z);
    }
}

// tv:transforms/functional.py:368-370: shorter side -> 256, longer -> int(256 * long / short)
void resized_size(int h, int w, int* nh, int* nw) {
  const int s = w <= h ? w : h, l = w <= h ? h : w;
  const int new_s = 256;
  const int new_l = (int)(256.0 * (double)l / (double)s);
  if (w <= h) { *nw = new_s; *nh = new_l; } else { *nh = new_s; *nw = new_l; }
}

// tv:transforms/functional.py:592-593: int(round((size - 224) / 2.0)), banker's rounding
int crop_offset(int resized) {
  const double v = (resized - IDN_CROP) / 2.0;
  return (int)std::nearbyint(v);   // default rounding mode: to nearest even
}

// Table layout (int32): [ksize][xmin[224]][xcnt[224]][k[224][ksize]]
void build_axis_table(int in_size, int out_size, int crop_off, std::string* out) {
  const int prec = 22;
  std::vector<int> xmin(IDN_CROP), xcnt(IDN_CROP);
  std::vector<int> kk;
  int ksize;
  if (in_size == out_size) {                       // Pillow skips the pass: identity taps
    ksize = 1;
    kk.assign(IDN_CROP, 1 << prec);
    for (int i = 0; i < IDN_CROP; ++i) { xmin[i] = i + crop_off; xcnt[i] = 1; }
  } else {
    const double scale = (double)in_size / (double)out_size;
    const double filterscale = scale < 1.0 ? 1.0 : scale;
    const double support = 1.0 * filterscale;      // bilinear filter support
    ksize = (int)std::ceil(support) * 2 + 1;
    kk.assign((size_t)IDN_CROP * ksize, 0);
    const double ss = 1.0 / filterscale;
    std::vector<double> w(ksize);
    for (int i = 0; i < IDN_CROP; ++i) {
      const int xx = i + crop_off;
      const double center = (xx + 0.5) * scale;
      int x0 = (int)(center - support + 0.5);
      if (x0 < 0) x0 = 0;
      int x1 = (int)(center + support + 0.5);
      if (x1 > in_size) x1 = in_size;
      const int n = x1 - x0;
      double ww = 0.0;
      for (int x = 0; x < n; ++x) {
        double a = (x + x0 - center + 0.5) * ss;
        if (a < 0.0) a = -a;
        const double v = a < 1.0 ? 1.0 - a : 0.0;
        w[x] = v;
        ww += v;
      }
      for (int x = 0; x < n; ++x) {
        if (ww != 0.0) w[x] /= ww;
        const double v = w[x] * (double)(1 << prec);
        kk[(size_t)i * ksize + x] = w[x] < 0 ? (int)(v - 0.5) : (int)(v + 0.5);
      }
      xmin[i] = x0;
      xcnt[i] = n;
    }
  }
  std::vector<int> flat;
  flat.push_back(ksize);
  flat.insert(flat.end(), xmin.begin(), xmin.end());
  flat.insert(flat.end(), xcnt.begin(), xcnt.end());
  flat.insert(flat.end(), kk.begin(), kk.end());
  out->assign(reinterpret_cast<const char*>(flat.data()), flat.size() * sizeof(int));
}

__device__ __forceinline__ int clip8(int v) { return v < 0 ? 0 : (v > 255 ? 255 : v); }

// ---------------------------------------------------------------------------------------------
// One CTA = kPreRows consecutive output rows of one image.
//   1. stage the source rows those outputs touch into shared memory (coalesced 32-bit loads);
//   2. horizontal pass: thread x owns output column x, keeps its <= 4 coefficients in registers,
//      reads its 3*xn source bytes as aligned words + funnel shifts -> uint8 RGBx in smem;
//   3. vertical pass over the uint8 intermediate (Pillow's order) -> uint8 RGBx tile in smem;
//   4. writer: normalise through the 768-entry bf16 table and store 128 bits at a time in one of
//      three layouts: planar NCHW (public entry point), 4-channel NHWC with the stem's left pad
//      (ResNet18 stem image), or the 4x4 space-to-depth image AlexNet's stem reads (57x57x64).
// Source tiles that do not fit (huge downscales) take the direct per-pixel path of step 2+3; the
// arithmetic is the same integer arithmetic either way.
// ---------------------------------------------------------------------------------------------
constexpr int kPreRows = 8;
constexpr int kPreMaxSrcRows = 26;                      // upper limits of the staged tile; the launch
constexpr int kPreMaxRawWords = 9 * 1024;               // sizes shared memory for the batch at hand

__device__ __forceinline__ void pre_direct_pixel(const uint8_t* __restrict__ src, int W,
                                                 const int* __restrict__ ht, const int* __restrict__ vt,
                                                 int x, int y, int out3[3]) {
  const int hks = ht[0], vks = vt[0];
  const int x0 = ht[1 + x], xn = ht[1 + IDN_CROP + x];
  const int* hk = ht + 1 + 2 * IDN_CROP + x * hks;
  const int y0 = vt[1 + y], yn = vt[1 + IDN_CROP + y];
  const int* vk = vt + 1 + 2 * IDN_CROP + y * vks;
  int acc[3] = {1 << 21, 1 << 21, 1 << 21};
  for (int j = 0; j < yn; ++j) {
    const uint8_t* rowp = src + ((long long)(y0 + j) * W + x0) * 3;
    int h0 = 1 << 21, h1 = 1 << 21, h2 = 1 << 21;
    for (int i = 0; i < xn; ++i) {
      const int k = __ldg(hk + i);
      h0 += (int)__ldg(rowp + 3 * i + 0) * k;
      h1 += (int)__ldg(rowp + 3 * i + 1) * k;
      h2 += (int)__ldg(rowp + 3 * i + 2) * k;
    }
    const int kv = __ldg(vk + j);
    acc[0] += clip8(h0 >> 22) * kv;
    acc[1] += clip8(h1 >> 22) * kv;
    acc[2] += clip8(h2 >> 22) * kv;
  }
  out3[0] = clip8(acc[0] >> 22); out3[1] = clip8(acc[1] >> 22); out3[2] = clip8(acc[2] >> 22);
}

__device__ __forceinline__ int bsel(uint32_t v, int i) {          // byte i of v (one PRMT)
  return (int)__byte_perm(v, 0u, 0x4440u + (uint32_t)i);
}

template <int OUT_MODE>
__global__ void __launch_bounds__(256)
preprocess_kernel(const uint8_t* __restrict__ imgs, const PreMeta* __restrict__ meta,
                  const int* __restrict__ tables, const uint16_t* __restrict__ lut, void* out,
                  int left_pad, int ws, int raw_cap_words, int src_rows_cap,
                  uint4* __restrict__ zero_ptr, int zero_vec_per_img) {
  pdl_wait();
  pdl_launch_dependents();
  // Optional side job: clear this image's slice of a buffer a later kernel accumulates into with
  // red.max (the pooled ResNet stem output).  This kernel is bound by instruction issue, its store
  // path is idle, so the 415 KB per image ride along for free; stream order keeps them behind the
  // previous pass's readers.
  if (zero_ptr != nullptr) {
    const int per_cta = (zero_vec_per_img + (int)gridDim.x - 1) / (int)gridDim.x;
    const int lo = per_cta * (int)blockIdx.x;
    const int hi = min(lo + per_cta, zero_vec_per_img);
    uint4* zp = zero_ptr + (long long)blockIdx.y * zero_vec_per_img;
    for (int i = lo + (int)threadIdx.x; i < hi; i += (int)blockDim.x) zp[i] = make_uint4(0u, 0u, 0u, 0u);
  }
  extern __shared__ uint32_t sm[];
  uint32_t* s_raw = sm;                                        // staged source bytes
  uint32_t* s_h = sm + raw_cap_words;                          // [src rows][224] RGBx after pass 1
  uint16_t* s_lut = reinterpret_cast<uint16_t*>(s_h + src_rows_cap * IDN_CROP);
  const int tid = threadIdx.x, n = blockIdx.y;
  const int warp = tid >> 5, lane = tid & 31;
  for (int i = tid; i < 768; i += blockDim.x) s_lut[i] = lut[i];
  const PreMeta m = meta[n];
  const int* __restrict__ ht = tables + m.htab;
  const int* __restrict__ vt = tables + m.vtab;
  const uint8_t* __restrict__ src = imgs + m.offset;
  // crop rows of this CTA: 8 rows; the S2D4 tiles are shifted by the stem's padding (2) so that a CTA
  // owns two complete rows of 4x4 blocks (29 CTAs per image)
  const int ybase = OUT_MODE == PRE_OUT_S2D4 ? kPreRows * (int)blockIdx.x - 2 : kPreRows * (int)blockIdx.x;
  const int ylo = ybase < 0 ? 0 : ybase;
  const int yhi = ybase + kPreRows > IDN_CROP ? IDN_CROP : ybase + kPreRows;
  const int hks = ht[0], vks = vt[0];
  const int ys0 = vt[1 + ylo];
  const int nrows = vt[1 + yhi - 1] + vt[1 + IDN_CROP + yhi - 1] - ys0;
  const int c0 = ht[1], c1 = ht[1 + IDN_CROP - 1] + ht[1 + 2 * IDN_CROP - 1];
  const int rowbytes = (c1 - c0) * 3;
  const int pitchw = (rowbytes + 3) / 4 + 4;                              // words per staged row
  const bool tiled = hks <= 4 && vks <= 4 && nrows <= src_rows_cap && nrows * pitchw <= raw_cap_words;
  const int rowstep = m.W * 3;
  const uint8_t* g_first = src + ((long long)ys0 * m.W + c0) * 3;         // first needed byte

  // ---- per-pixel output (bf16 bits of the three channels), shared by both paths
  auto emit = [&](int y, int x, uint32_t rr, uint32_t gg, uint32_t bb) {
    const uint32_t r = s_lut[rr], g = s_lut[256 + gg], b = s_lut[512 + bb];
    if (OUT_MODE == PRE_OUT_NCHW) {
      uint16_t* o = reinterpret_cast<uint16_t*>(out) + (long long)n * 3 * IDN_CROP * IDN_CROP + y * IDN_CROP + x;
      o[0] = (uint16_t)r; o[IDN_CROP * IDN_CROP] = (uint16_t)g; o[2 * IDN_CROP * IDN_CROP] = (uint16_t)b;
    } else if (OUT_MODE == PRE_OUT_STEM) {
      reinterpret_cast<uint2*>(out)[((long long)n * IDN_CROP + y) * ws + left_pad + x] = make_uint2(r | (g << 16), b);
    }
  };
  // S2D4 writer: s2d pixel (by, bx) = the 4x4 block of crop pixels starting at (4by-2, 4bx-2), channel
  // (dy*4+dx)*3 + c, 57 x 57 pixels of 64 channels per image.  One item = one block row (by, bx, dy):
  // 4 pixels x 3 channels = 24 contiguous bytes, so the four dy of a block fill 96 contiguous bytes.
  // Channels 48..63 are never written (the buffer is zeroed when it is allocated); positions outside
  // the crop are written as zeros.
  auto write_s2d4 = [&](auto row_ctx, auto pixel) {
    const int by0 = 2 * (int)blockIdx.x;
    for (int it = tid; it < 2 * 57 * 4; it += blockDim.x) {
      const int dy = it & 3, bx = (it >> 2) % 57, byl = it / (4 * 57);
      const int by = by0 + byl, y = ybase + 4 * byl + dy;
      if (by >= 57) continue;
      uint32_t w[6] = {0u, 0u, 0u, 0u, 0u, 0u};
      if (y >= 0 && y < IDN_CROP) {
        const auto ctx = row_ctx(y);                  // everything that depends on y only, once per item
#pragma unroll
        for (int dx = 0; dx < 4; ++dx) {
          const int x = 4 * bx - 2 + dx;
          if (x < 0 || x >= IDN_CROP) continue;
          const uint32_t px = pixel(ctx, x);
          const uint32_t c[3] = {s_lut[px & 255u], s_lut[256 + ((px >> 8) & 255u)], s_lut[512 + (px >> 16)]};
#pragma unroll
          for (int ch = 0; ch < 3; ++ch) {
            const int e = dx * 3 + ch;
            w[e >> 1] |= c[ch] << (16 * (e & 1));
          }
        }
      }
      uint2* o = reinterpret_cast<uint2*>(reinterpret_cast<uint16_t*>(out) +
                                         (((long long)n * 57 + by) * 57 + bx) * 64 + dy * 12);
      o[0] = make_uint2(w[0], w[1]); o[1] = make_uint2(w[2], w[3]); o[2] = make_uint2(w[4], w[5]);
    }
  };

  if (tiled) {
    // ---- 1. stage the source rows: warp w takes rows w, w+8, ...; aligned 32-bit loads, the two
    //         edge words of a row byte by byte so that nothing outside the row is touched
    for (int r = warp; r < nrows; r += 8) {
      const uint8_t* g = g_first + (long long)r * rowstep;
      const int shift = (int)(reinterpret_cast<uintptr_t>(g) & 3);
      const int nw = (shift + rowbytes + 3) >> 2;
      const uint32_t* gw = reinterpret_cast<const uint32_t*>(g - shift);
      uint32_t* dst = s_raw + r * pitchw;
      for (int w = lane + 1; w < nw - 1; w += 32) dst[w] = __ldg(gw + w);
      if (lane < 2) {
        const int w = lane == 0 ? 0 : nw - 1;
        uint32_t v = 0;
#pragma unroll
        for (int b = 0; b < 4; ++b) {
          const int off = 4 * w + b - shift;
          if (off >= 0 && off < rowbytes) v |= (uint32_t)__ldg(g + off) << (8 * b);
        }
        dst[w] = v;
      }
    }
    __syncthreads();
    // Upscales (every source whose short side is < 256, e.g. the 224 x 224 benchmark set) have 3-tap
    // kernels in Pillow's tables, mild downscales 3-4; the table format allows up to 4.  The branch is uniform per image, the arithmetic is the same sum with the zero
    // terms left out.
    auto passes = [&](auto taps_c) {
      constexpr int NT = decltype(taps_c)::value;
      if (tid < IDN_CROP) {
        const int x = tid;
        // ---- 2. horizontal pass: <= NT taps, coefficients in registers, 3*NT source bytes per row
        const int x0 = ht[1 + x], xn = ht[1 + IDN_CROP + x];
        const int* hk = ht + 1 + 2 * IDN_CROP + x * hks;
        const int k0 = __ldg(hk), k1 = xn > 1 ? __ldg(hk + 1) : 0;
        const int k2 = (NT > 2 && xn > 2) ? __ldg(hk + 2) : 0, k3 = (NT > 3 && xn > 3) ? __ldg(hk + 3) : 0;
        const int boff = (x0 - c0) * 3;
        int shift = (int)(reinterpret_cast<uintptr_t>(g_first) & 3);
        const uint32_t* rowp = s_raw;
        uint32_t* hp = s_h + x;
#pragma unroll 2
        for (int r = 0; r < nrows; ++r) {
          const int bstart = shift + boff;
          const uint32_t* wp = rowp + (bstart >> 2);
          const int sh = (bstart & 3) * 8;
          const uint32_t a = wp[0], b = wp[1], c = wp[2];
          const uint32_t p0 = __funnelshift_r(a, b, sh), p1 = __funnelshift_r(b, c, sh);
          int h0 = 1 << 21, h1 = 1 << 21, h2 = 1 << 21;
          h0 += bsel(p0, 0) * k0; h1 += bsel(p0, 1) * k0; h2 += bsel(p0, 2) * k0;
          h0 += bsel(p0, 3) * k1; h1 += bsel(p1, 0) * k1; h2 += bsel(p1, 1) * k1;
          if (NT > 2) {
            // 9 bytes (3 pixels) end inside word c whatever the shift; the 4th pixel needs word d
            const uint32_t p2 = __funnelshift_r(c, NT > 3 ? wp[3] : 0u, sh);
            h0 += bsel(p1, 2) * k2; h1 += bsel(p1, 3) * k2; h2 += bsel(p2, 0) * k2;
            if (NT > 3) { h0 += bsel(p2, 1) * k3; h1 += bsel(p2, 2) * k3; h2 += bsel(p2, 3) * k3; }
          }
          // coefficients are >= 0 and sum to 2^22 (+- rounding), so the result is already in [0, 255]
          *hp = (uint32_t)(h0 >> 22) | ((uint32_t)(h1 >> 22) << 8) | ((uint32_t)(h2 >> 22) << 16);
          hp += IDN_CROP;
          rowp += pitchw;
          shift = (shift + rowstep) & 3;
        }
      }
      __syncthreads();
      // ---- 3. vertical pass over the uint8 intermediate (Pillow's order) -> packed uint8 RGB
      struct VRow { int r0, r1, r2, r3, k0, k1, k2, k3; };
      auto vrow = [&](int y) -> VRow {
        const int y0 = vt[1 + y] - ys0, yn = vt[1 + IDN_CROP + y];
        const int* vk = vt + 1 + 2 * IDN_CROP + y * vks;
        const int last = nrows - 1;
        VRow v;
        v.r0 = y0 * IDN_CROP; v.r1 = min(y0 + 1, last) * IDN_CROP;
        v.r2 = min(y0 + 2, last) * IDN_CROP; v.r3 = min(y0 + 3, last) * IDN_CROP;
        v.k0 = __ldg(vk); v.k1 = yn > 1 ? __ldg(vk + 1) : 0;
        v.k2 = (NT > 2 && yn > 2) ? __ldg(vk + 2) : 0; v.k3 = (NT > 3 && yn > 3) ? __ldg(vk + 3) : 0;
        return v;
      };
      auto vpix = [&](const VRow& v, int x) -> uint32_t {
        const uint32_t q0 = s_h[v.r0 + x], q1 = s_h[v.r1 + x];
        int a0 = 1 << 21, a1 = 1 << 21, a2 = 1 << 21;
        a0 += bsel(q0, 0) * v.k0; a1 += bsel(q0, 1) * v.k0; a2 += bsel(q0, 2) * v.k0;
        a0 += bsel(q1, 0) * v.k1; a1 += bsel(q1, 1) * v.k1; a2 += bsel(q1, 2) * v.k1;
        if (NT > 2) {
          const uint32_t q2 = s_h[v.r2 + x];
          a0 += bsel(q2, 0) * v.k2; a1 += bsel(q2, 1) * v.k2; a2 += bsel(q2, 2) * v.k2;
        }
        if (NT > 3) {
          const uint32_t q3 = s_h[v.r3 + x];
          a0 += bsel(q3, 0) * v.k3; a1 += bsel(q3, 1) * v.k3; a2 += bsel(q3, 2) * v.k3;
        }
        return (uint32_t)(a0 >> 22) | ((uint32_t)(a1 >> 22) << 8) | ((uint32_t)(a2 >> 22) << 16);
      };
      if (OUT_MODE == PRE_OUT_S2D4) {
        write_s2d4(vrow, vpix);
      } else if (tid < IDN_CROP) {
#pragma unroll 2
        for (int y = ylo; y < yhi; ++y) {
          const uint32_t px = vpix(vrow(y), tid);
          emit(y, tid, px & 255u, (px >> 8) & 255u, px >> 16);
        }
      }
    };
    if (hks <= 2 && vks <= 2) passes(std::integral_constant<int, 2>{});
    else if (hks <= 3 && vks <= 3) passes(std::integral_constant<int, 3>{});   // every upscale: Pillow's ksize is 3
    else passes(std::integral_constant<int, 4>{});
  } else {
    __syncthreads();                              // s_lut
    auto direct = [&](int y, int x) -> uint32_t {
      int o[3];
      pre_direct_pixel(src, m.W, ht, vt, x, y, o);
      return (uint32_t)o[0] | ((uint32_t)o[1] << 8) | ((uint32_t)o[2] << 16);
    };
    if (OUT_MODE == PRE_OUT_S2D4) {
      write_s2d4([](int y) { return y; }, direct);
    } else if (tid < IDN_CROP) {
      for (int y = ylo; y < yhi; ++y) {
        const uint32_t px = direct(y, tid);
        emit(y, tid, px & 255u, (px >> 8) & 255u, px >> 16);
      }
    }
  }
  if (OUT_MODE == PRE_OUT_STEM && tid >= IDN_CROP) {      // the stem layout's horizontal zero padding
    const int padpx = tid - IDN_CROP;
    if (padpx < ws - IDN_CROP) {
      const int px = padpx < left_pad ? padpx : IDN_CROP + padpx;
      for (int y = ylo; y < yhi; ++y)
        reinterpret_cast<uint2*>(out)[((long long)n * IDN_CROP + y) * ws + px] = make_uint2(0u, 0u);
    }
  }
}

int launch_preprocess(const uint8_t* d_imgs, const PreMeta* d_meta, const int* d_tables,
                      const uint16_t* d_lut, int n, int out_mode, void* out, int stem_left_pad,
                      int stem_ws, int raw_words, int src_rows, cudaStream_t st,
                      void* zero_ptr, size_t zero_bytes_per_img) {
  uint4* zp = static_cast<uint4*>(zero_ptr);
  const int zv = (int)(zero_bytes_per_img / 16);
  if (zp != nullptr && (zero_bytes_per_img % 16 != 0 || (reinterpret_cast<uintptr_t>(zp) & 15) != 0)) {
    set_error("preprocess: the buffer to clear must be 16-byte aligned");
    return IDN_EINVAL;
  }
  if (n <= 0) return IDN_OK;
  if (out_mode == PRE_OUT_STEM && stem_ws - IDN_CROP > 32) {
    set_error("preprocess: stem row pitch too wide");
    return IDN_EINVAL;
  }
  // shared memory for the largest staged tile of this batch (images beyond the caps take the
  // direct path inside the kernel)
  const int raw_cap = raw_words < 64 ? 64 : (raw_words > kPreMaxRawWords ? kPreMaxRawWords : raw_words);
  const int rows_cap = src_rows < 1 ? 1 : (src_rows > kPreMaxSrcRows ? kPreMaxSrcRows : src_rows);
  const int smem = (raw_cap + rows_cap * IDN_CROP) * 4 + 768 * 2;
  if (out_mode == PRE_OUT_NCHW) {
    if (smem > 48 * 1024) IDN_CUDA_TRY(cudaFuncSetAttribute(preprocess_kernel<PRE_OUT_NCHW>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    IDN_CUDA_TRY(launch_pdl(preprocess_kernel<PRE_OUT_NCHW>, dim3(IDN_CROP / kPreRows, n), dim3(256), (size_t)smem, st,
                            d_imgs, d_meta, d_tables, d_lut, out, 0, 0, raw_cap, rows_cap, zp, zv));
  } else if (out_mode == PRE_OUT_STEM) {
    if (smem > 48 * 1024) IDN_CUDA_TRY(cudaFuncSetAttribute(preprocess_kernel<PRE_OUT_STEM>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    IDN_CUDA_TRY(launch_pdl(preprocess_kernel<PRE_OUT_STEM>, dim3(IDN_CROP / kPreRows, n), dim3(256), (size_t)smem, st,
                            d_imgs, d_meta, d_tables, d_lut, out, stem_left_pad, stem_ws, raw_cap, rows_cap, zp, zv));
  } else {
    if (smem > 48 * 1024) IDN_CUDA_TRY(cudaFuncSetAttribute(preprocess_kernel<PRE_OUT_S2D4>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    IDN_CUDA_TRY(launch_pdl(preprocess_kernel<PRE_OUT_S2D4>, dim3(29, n), dim3(256), (size_t)smem, st,
                            d_imgs, d_meta, d_tables, d_lut, out, 0, 0, raw_cap, rows_cap, zp, zv));
  }
  IDN_CUDA_TRY(cudaGetLastError());
  return IDN_OK;
}

// Shared-memory need of the staged tile for one image: (staged words, source rows) for the worst
// window of kPreRows consecutive crop rows (any alignment: the S2D4 tiles are shifted by 2).
void preprocess_tile_need(const int* htab, const int* vtab, int* raw_words, int* src_rows) {
  const int c0 = htab[1], c1 = htab[1 + IDN_CROP - 1] + htab[1 + 2 * IDN_CROP - 1];
  const int pitchw = ((c1 - c0) * 3 + 3) / 4 + 4;
  int span = 1;
  for (int y = 0; y < IDN_CROP; ++y) {
    const int last = y + kPreRows - 1 < IDN_CROP ? y + kPreRows - 1 : IDN_CROP - 1;
    const int s = vtab[1 + last] + vtab[1 + IDN_CROP + last] - vtab[1 + y];
    if (s > span) span = s;
  }
  *src_rows = span;
  *raw_words = span * pitchw;
}

}  // namespace idn
