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

// One CTA = one output row y of one image; thread x < 224 = one output pixel (3 channels).
// grid = (224, n).  Threads 224..231 write the stem layout's horizontal zero padding.
template <int OUT_MODE>
__global__ void __launch_bounds__(256)
preprocess_kernel(const uint8_t* __restrict__ imgs, const PreMeta* __restrict__ meta,
                  const int* __restrict__ tables, const uint16_t* __restrict__ lut, void* out,
                  int left_pad, int ws) {
  __shared__ uint16_t s_lut[768];
  for (int i = threadIdx.x; i < 768; i += blockDim.x) s_lut[i] = lut[i];
  __syncthreads();
  const int y = blockIdx.x, n = blockIdx.y, x = threadIdx.x;
  const PreMeta m = meta[n];
  if (x < IDN_CROP) {
    const int* ht = tables + m.htab;
    const int* vt = tables + m.vtab;
    const int hks = ht[0], vks = vt[0];
    const int x0 = ht[1 + x], xn = ht[1 + IDN_CROP + x];
    const int* hk = ht + 1 + 2 * IDN_CROP + x * hks;
    const int y0 = vt[1 + y], yn = vt[1 + IDN_CROP + y];
    const int* vk = vt + 1 + 2 * IDN_CROP + y * vks;
    const uint8_t* src = imgs + m.offset;
    int acc[3] = {1 << 21, 1 << 21, 1 << 21};
    for (int j = 0; j < yn; ++j) {
      const uint8_t* rowp = src + ((long long)(y0 + j) * m.W + x0) * 3;
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
    const uint16_t r = s_lut[clip8(acc[0] >> 22)];
    const uint16_t g = s_lut[256 + clip8(acc[1] >> 22)];
    const uint16_t b = s_lut[512 + clip8(acc[2] >> 22)];
    if (OUT_MODE == PRE_OUT_NCHW) {
      uint16_t* o = reinterpret_cast<uint16_t*>(out) + (long long)n * 3 * IDN_CROP * IDN_CROP +
                    (long long)y * IDN_CROP + x;
      o[0] = r;
      o[IDN_CROP * IDN_CROP] = g;
      o[2 * IDN_CROP * IDN_CROP] = b;
    } else {
      uint2* o = reinterpret_cast<uint2*>(out) + ((long long)n * IDN_CROP + y) * ws + left_pad + x;
      *o = make_uint2((uint32_t)r | ((uint32_t)g << 16), (uint32_t)b);
    }
  } else if (OUT_MODE == PRE_OUT_STEM) {
    const int padpx = x - IDN_CROP;               // 0 .. ws-224-1
    if (padpx < ws - IDN_CROP) {
      const int px = padpx < left_pad ? padpx : IDN_CROP + padpx;
      uint2* o = reinterpret_cast<uint2*>(out) + ((long long)n * IDN_CROP + y) * ws + px;
      *o = make_uint2(0u, 0u);
    }
  }
}

int launch_preprocess(const uint8_t* d_imgs, const PreMeta* d_meta, const int* d_tables,
                      const uint16_t* d_lut, int n, int out_mode, void* out, int stem_left_pad,
                      int stem_ws, cudaStream_t st) {
  if (n <= 0) return IDN_OK;
  if (out_mode == PRE_OUT_STEM && stem_ws - IDN_CROP > 32) {
    set_error("preprocess: stem row pitch too wide");
    return IDN_EINVAL;
  }
  dim3 grid(IDN_CROP, n);
  if (out_mode == PRE_OUT_NCHW)
    preprocess_kernel<PRE_OUT_NCHW><<<grid, 256, 0, st>>>(d_imgs, d_meta, d_tables, d_lut, out, 0, 0);
  else
    preprocess_kernel<PRE_OUT_STEM><<<grid, 256, 0, st>>>(d_imgs, d_meta, d_tables, d_lut, out,
                                                         stem_left_pad, stem_ws);
  IDN_CUDA_TRY(cudaGetLastError());
  return IDN_OK;
}

}  // namespace idn
