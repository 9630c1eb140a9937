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
// Source rows (or columns) [*first, *last) that crop positions 0..IDN_CROP-1 of an axis read: the bounds
// build_axis_table computes for the first and the last crop position (same arithmetic).
void needed_source_range(int in_size, int out_size, int crop_off, int* first, int* last) {
  if (in_size == out_size) { *first = crop_off; *last = crop_off + IDN_CROP; return; }
  const double scale = (double)in_size / (double)out_size;
  const double support = scale < 1.0 ? 1.0 : scale;
  const double c0 = (crop_off + 0.5) * scale, c1 = (IDN_CROP - 1 + crop_off + 0.5) * scale;
  int x0 = (int)(c0 - support + 0.5);
  if (x0 < 0) x0 = 0;
  int x1 = (int)(c1 + support + 0.5);
  if (x1 > in_size) x1 = in_size;
  *first = x0; *last = x1;
}

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
  // Pillow sizes its coefficient rows for the filter's support (3 for any upscale) but a row only uses xcnt of them
  // (2 when upscaling): keep the widest row actually used, so that the kernel's register window is no wider.
  int kuse = 1;
  for (int i = 0; i < IDN_CROP; ++i) kuse = std::max(kuse, xcnt[i]);
  if (kuse < ksize) {
    std::vector<int> k2((size_t)IDN_CROP * kuse, 0);
    for (int i = 0; i < IDN_CROP; ++i)
      for (int x = 0; x < xcnt[i]; ++x) k2[(size_t)i * kuse + x] = kk[(size_t)i * ksize + x];
    kk.swap(k2);
    ksize = kuse;
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
// One CTA = kPreBand consecutive crop rows of one image; thread x owns crop column x for the whole band.
//
// The kernel is bound by instruction issue, not by bytes (the fixed-point arithmetic is ~36 multiply-adds and
// byte extractions per pixel), so everything is arranged to execute as few instructions per pixel as
// possible and to need no staging pass and no barrier inside the band:
//   * a thread walks down its column and keeps the last NT horizontally filtered source rows (packed RGB
//     uint8, Pillow's intermediate) in registers: every source row is filtered exactly once per column,
//     the vertical pass is NT x 3 multiply-adds on registers;
//   * its <= NT horizontal coefficients live in registers for the whole band; the window is shifted left at
//     the right image edge (coefficients re-aligned) so that it never leaves the row;
//   * source bytes come straight from global memory as the aligned words under the thread's 3 NT bytes,
//     fetched one source row ahead; neighbouring threads read overlapping words, so a warp's request is one
//     or two 128-byte lines and every line leaves DRAM once;
//   * per crop row the band's vertical coefficients are re-aligned to the register window once per CTA
//     (shared memory, read as two broadcast vector loads per row);
//   * u8 -> (x / 255 - mean) / std -> bf16 is the 768-entry table in shared memory;
//   * writers: 4-channel NHWC with the stem's left pad (one 8-byte store per pixel, a warp writes 256
//     contiguous bytes), the 4x4 space-to-depth image AlexNet's stem reads, planar NCHW (public entry).
//     Padding pixels / channels of the two internal layouts are zero from allocation and never written.
// NT = 3 covers every upscale (Pillow's bilinear tables have 3 taps when the short side is < 256) and the
// identity; NT = 4 mild downscales; anything else (or images narrower than NT) takes the direct per-pixel
// path below -- the same integer arithmetic either way.
// ---------------------------------------------------------------------------------------------
constexpr int kPreBand = 32;                            // crop rows per CTA: 7 bands per image
constexpr int kPreThreads = IDN_CROP;

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

// Writes pixel (y, x) of image n: r, g, b are the resized uint8 values.
template <int OUT_MODE>
__device__ __forceinline__ void pre_emit(void* out, const uint16_t* s_lut, int n, int y, int x, uint32_t rr,
                                         uint32_t gg, uint32_t bb, int left_pad, int ws) {
  const uint32_t r = s_lut[rr], g = s_lut[256 + gg], b = s_lut[512 + bb];
  if (OUT_MODE == PRE_OUT_NCHW) {
    uint16_t* o = reinterpret_cast<uint16_t*>(out) + (long long)n * 3 * IDN_CROP * IDN_CROP + y * IDN_CROP + x;
    o[0] = (uint16_t)r; o[IDN_CROP * IDN_CROP] = (uint16_t)g; o[2 * IDN_CROP * IDN_CROP] = (uint16_t)b;
  } else if (OUT_MODE == PRE_OUT_STEM) {
    reinterpret_cast<uint2*>(out)[((long long)n * IDN_CROP + y) * ws + left_pad + x] = make_uint2(r | (g << 16), b);
  } else {
    // s2d pixel (by, bx) = the 4x4 block of crop pixels starting at (4 by - 2, 4 bx - 2), channel (dy*4+dx)*3 + c
    const int by = (y + 2) >> 2, dy = (y + 2) & 3, bx = (x + 2) >> 2, dx = (x + 2) & 3;
    uint16_t* o = reinterpret_cast<uint16_t*>(out) + (((long long)n * 57 + by) * 57 + bx) * 64 + (dy * 4 + dx) * 3;
    o[0] = (uint16_t)r; o[1] = (uint16_t)g; o[2] = (uint16_t)b;
  }
}

struct PreVRow {            // per crop row of the band, aligned to the register window (shared memory)
  int nnew;                 // source rows to filter before this crop row can be computed
  int c[4];                 // vertical coefficient of window slot j (0 where the slot's row is not a tap)
  int pad[3];
};

template <int OUT_MODE, int NT>
__device__ __forceinline__ void pre_fast_band(const uint8_t* __restrict__ src, const PreMeta& m,
                                              const int* __restrict__ ht, const PreVRow* s_v, const uint16_t* s_lut,
                                              void* out, int n, int ylo, int yhi, int first_row, int left_pad, int ws) {
  // Column of this thread.  For the space-to-depth writer the columns are rotated by two so that lanes
  // 4 j .. 4 j + 3 hold the four pixels (dx = 0..3) of ONE s2d pixel row: (x + 2) % 4 == lane % 4.
  const int x = OUT_MODE == PRE_OUT_S2D4 ? (int)((threadIdx.x + IDN_CROP - 2) % IDN_CROP) : (int)threadIdx.x;
  const int hks = ht[0];
  // horizontal taps, window shifted left at the right edge so that [x0, x0 + NT) stays inside the row
  int x0 = ht[1 + x];
  const int xn = ht[1 + IDN_CROP + x];
  const int* hk = ht + 1 + 2 * IDN_CROP + x * hks;
  int kk[NT];
  {
    const int x0s = min(x0, m.W - NT), d = x0 - x0s;
#pragma unroll
    for (int j = 0; j < NT; ++j) {
      const int t = j - d;
      kk[j] = (t >= 0 && t < xn) ? __ldg(hk + t) : 0;
    }
    x0 = x0s;
  }
  const long long rowstep = (long long)m.W * 3;
  const uint8_t* colp = src + (long long)x0 * 3;
  constexpr int NW = NT <= 3 ? 3 : 4;                     // aligned words under 3 NT bytes at any alignment
  uint32_t pw[NW];
  int psh;
  auto fetch = [&](int r) {                               // words of source row r (clamped: the last rows are
    r = min(r, m.H - 1);                                  // fetched ahead of need and may never be used)
    const uint8_t* a = colp + (long long)r * rowstep;
    const uint32_t* wp = reinterpret_cast<const uint32_t*>(reinterpret_cast<uintptr_t>(a) & ~(uintptr_t)3);
    psh = (int)(reinterpret_cast<uintptr_t>(a) & 3) * 8;
    pw[0] = __ldg(wp); pw[1] = __ldg(wp + 1); pw[2] = __ldg(wp + 2);
    if (NT == 4) pw[3] = psh != 0 ? __ldg(wp + 3) : 0u;   // 12 bytes from an aligned start end inside word 2
  };
  uint32_t hrow[NT];
#pragma unroll
  for (int j = 0; j < NT; ++j) hrow[j] = 0u;
  int nr = first_row;                                     // next source row to filter
  fetch(nr);
  for (int y = ylo; y < yhi; ++y) {
    const PreVRow v = s_v[y - ylo];
    for (int f = 0; f < v.nnew; ++f) {
      // ---- horizontal pass of source row nr -> packed uint8 RGB, into the youngest window slot
      const uint32_t p0 = __funnelshift_r(pw[0], pw[1], psh), p1 = __funnelshift_r(pw[1], pw[2], psh);
      const uint32_t p2 = NT >= 3 ? __funnelshift_r(pw[2], NT == 4 ? pw[3] : 0u, psh) : 0u;
      ++nr;
      fetch(nr);
      int h0 = 1 << 21, h1 = 1 << 21, h2 = 1 << 21;
      h0 += bsel(p0, 0) * kk[0]; h1 += bsel(p0, 1) * kk[0]; h2 += bsel(p0, 2) * kk[0];
      h0 += bsel(p0, 3) * kk[1]; h1 += bsel(p1, 0) * kk[1]; h2 += bsel(p1, 1) * kk[1];
      if (NT >= 3) { h0 += bsel(p1, 2) * kk[NT >= 3 ? 2 : 0]; h1 += bsel(p1, 3) * kk[NT >= 3 ? 2 : 0]; h2 += bsel(p2, 0) * kk[NT >= 3 ? 2 : 0]; }
      if (NT == 4) { h0 += bsel(p2, 1) * kk[3]; h1 += bsel(p2, 2) * kk[3]; h2 += bsel(p2, 3) * kk[3]; }
      // coefficients are >= 0 and sum to 2^22 (+- rounding), so the result is already in [0, 255]
#pragma unroll
      for (int j = 0; j + 1 < NT; ++j) hrow[j] = hrow[j + 1];
      hrow[NT - 1] = (uint32_t)(h0 >> 22) | ((uint32_t)(h1 >> 22) << 8) | ((uint32_t)(h2 >> 22) << 16);
    }
    // ---- vertical pass over the uint8 intermediate (Pillow's order)
    int a0 = 1 << 21, a1 = 1 << 21, a2 = 1 << 21;
#pragma unroll
    for (int j = 0; j < NT; ++j) {
      a0 += bsel(hrow[j], 0) * v.c[j]; a1 += bsel(hrow[j], 1) * v.c[j]; a2 += bsel(hrow[j], 2) * v.c[j];
    }
    if (OUT_MODE != PRE_OUT_S2D4) {
      pre_emit<OUT_MODE>(out, s_lut, n, y, x, (uint32_t)(a0 >> 22), (uint32_t)(a1 >> 22), (uint32_t)(a2 >> 22), left_pad, ws);
    } else {
      // The 4 pixels x 3 channels of a lane quad are 24 contiguous bytes of the s2d image: three lanes store 8
      // bytes each after two shuffles (three 2-byte stores per lane made the kernel LSU-bound).  The quad of
      // threads 0..3 straddles the row's last and first s2d pixel (two real pixels each): plain stores there.
      const uint32_t r = s_lut[(uint32_t)(a0 >> 22)], g = s_lut[256 + (uint32_t)(a1 >> 22)], b = s_lut[512 + (uint32_t)(a2 >> 22)];
      const uint32_t rg = r | (g << 16);
      const uint32_t rg_n = __shfl_down_sync(0xffffffffu, rg, 1), b_n = __shfl_down_sync(0xffffffffu, b, 1);
      const int by = (y + 2) >> 2, dy = (y + 2) & 3, q = (int)threadIdx.x & 3;
      if (threadIdx.x >= 4) {
        const int bx = (int)threadIdx.x >> 2;
        uint32_t w0, w1;
        if (q == 0) { w0 = rg; w1 = b | (rg_n << 16); }                          // r0 g0 | b0 r1
        else if (q == 1) { w0 = (rg >> 16) | (b << 16); w1 = rg_n; }             // g1 b1 | r2 g2
        else { w0 = b | (rg_n << 16); w1 = (rg_n >> 16) | (b_n << 16); }         // b2 r3 | g3 b3   (lane q == 2)
        if (q != 3) {
          uint2* o = reinterpret_cast<uint2*>(reinterpret_cast<uint16_t*>(out) +
                                              (((long long)n * 57 + by) * 57 + bx) * 64 + dy * 12) + q;
          *o = make_uint2(w0, w1);
        }
      } else {
        const int bx = (x + 2) >> 2;
        uint16_t* o = reinterpret_cast<uint16_t*>(out) + (((long long)n * 57 + by) * 57 + bx) * 64 + (dy * 4 + q) * 3;
        o[0] = (uint16_t)r; o[1] = (uint16_t)g; o[2] = (uint16_t)b;
      }
    }
  }
}

template <int OUT_MODE>
__global__ void __launch_bounds__(kPreThreads)
preprocess_kernel(const uint8_t* __restrict__ imgs, const PreMeta* __restrict__ meta,
                  const int* __restrict__ tables, const uint16_t* __restrict__ lut, void* out,
                  int left_pad, int ws) {
  pdl_wait();
  pdl_launch_dependents();
  __shared__ uint16_t s_lut[768];
  __shared__ __align__(16) PreVRow s_v[kPreBand];
  const int tid = threadIdx.x, n = blockIdx.y;
  for (int i = tid; i < 768; i += kPreThreads) s_lut[i] = lut[i];
  const PreMeta m = meta[n];
  const int* __restrict__ ht = tables + m.htab;
  const int* __restrict__ vt = tables + m.vtab;
  const uint8_t* __restrict__ src = imgs + m.offset;
  const int ylo = kPreBand * (int)blockIdx.x;
  const int yhi = min(ylo + kPreBand, IDN_CROP);
  const int hks = ht[0], vks = vt[0];
  const int nt = (hks <= 2 && vks <= 2) ? 2 : (hks <= 3 && vks <= 3) ? 3 : (hks <= 4 && vks <= 4) ? 4 : 0;
  const bool fast = nt != 0 && m.W >= nt;
  const int first_row = vt[1 + ylo];
  if (fast && tid < yhi - ylo) {
    // vertical taps of crop row y, re-aligned to the register window after the row's new source rows
    // have been filtered: slot j holds source row y0 + yn - nt + j
    const int y = ylo + tid;
    const int y0 = vt[1 + y], yn = vt[1 + IDN_CROP + y];
    const int prev = tid == 0 ? y0 : vt[1 + y - 1] + vt[1 + IDN_CROP + y - 1];
    const int* vk = vt + 1 + 2 * IDN_CROP + y * vks;
    PreVRow v;
    v.nnew = y0 + yn - prev;
    v.pad[0] = v.pad[1] = v.pad[2] = 0;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int t = yn - nt + j;                          // tap index of slot j
      v.c[j] = (j < nt && t >= 0 && t < yn) ? __ldg(vk + t) : 0;
    }
    s_v[tid] = v;
  }
  __syncthreads();
  if (fast) {
    if (nt == 2) pre_fast_band<OUT_MODE, 2>(src, m, ht, s_v, s_lut, out, n, ylo, yhi, first_row, left_pad, ws);
    else if (nt == 3) pre_fast_band<OUT_MODE, 3>(src, m, ht, s_v, s_lut, out, n, ylo, yhi, first_row, left_pad, ws);
    else pre_fast_band<OUT_MODE, 4>(src, m, ht, s_v, s_lut, out, n, ylo, yhi, first_row, left_pad, ws);
  } else {                                              // huge downscales, images narrower than the window
    for (int y = ylo; y < yhi; ++y) {
      int o[3];
      pre_direct_pixel(src, m.W, ht, vt, tid, y, o);
      pre_emit<OUT_MODE>(out, s_lut, n, y, tid, (uint32_t)o[0], (uint32_t)o[1], (uint32_t)o[2], left_pad, ws);
    }
  }
}

int launch_preprocess(const uint8_t* d_imgs, const PreMeta* d_meta, const int* d_tables,
                      const uint16_t* d_lut, int n, int out_mode, void* out, int stem_left_pad,
                      int stem_ws, cudaStream_t st) {
  if (n <= 0) return IDN_OK;
  const dim3 grid((IDN_CROP + kPreBand - 1) / kPreBand, n), block(kPreThreads);
  if (out_mode == PRE_OUT_NCHW) {
    IDN_CUDA_TRY(launch_pdl(preprocess_kernel<PRE_OUT_NCHW>, grid, block, 0, st, d_imgs, d_meta, d_tables, d_lut, out, 0, 0));
  } else if (out_mode == PRE_OUT_STEM) {
    IDN_CUDA_TRY(launch_pdl(preprocess_kernel<PRE_OUT_STEM>, grid, block, 0, st, d_imgs, d_meta, d_tables, d_lut, out,
                            stem_left_pad, stem_ws));
  } else {
    IDN_CUDA_TRY(launch_pdl(preprocess_kernel<PRE_OUT_S2D4>, grid, block, 0, st, d_imgs, d_meta, d_tables, d_lut, out, 0, 0));
  }
  IDN_CUDA_TRY(cudaGetLastError());
  return IDN_OK;
}

}  // namespace idn
