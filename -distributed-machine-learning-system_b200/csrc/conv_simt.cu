// CUDA-core cross-check of the shifted-GEMM convolution (common.cuh).  Same operand layouts, tap
// tables, tile decode and epilogue semantics as conv_tc.cu, but operands are fetched with plain
// loads that imitate the TMA boxes (out-of-range rows read as zero).  Used by tests and bring-up
// (IDN_IMPL_SIMT); never selected implicitly.
#include "common.cuh"

namespace idn {

__device__ __forceinline__ float simt_load_a(const ConvParams& p, int m_tile, int i, int k) {
  if (p.mode == MODE_STEM) {           // k = r * 32 + (window pixel * 4 + channel)
    const int per_img = p.tiles_w * p.tiles_h;
    const int n = m_tile / per_img;
    const int t = m_tile - n * per_img;
    const int th = t / p.tiles_w, tw = t - th * p.tiles_w;
    const int oh = th * kStemTH + i / kStemTW;
    const int ow = tw * kStemTW + i % kStemTW;
    const int r = k / kStemWin, e = k - r * kStemWin;
    const int grp = oh + p.tap_off[r];
    if (grp < 0 || grp >= p.st_nrow) return 0.0f;
    const long long row = (long long)n * p.st_h + (long long)grp * 2 + p.tap_par[r];
    return __bfloat162float(p.a_ptr[(row * p.st_ws + 2LL * ow) * 4 + e]);
  }
  const int tap = k / (p.c_blocks * kBlockK);
  const int c = k - tap * (p.c_blocks * kBlockK);
  {
    const long long row = (long long)m_tile * kBlockM + i + p.tap_off[tap];
    if (row < 0 || row >= p.a_rows) return 0.0f;
    return __bfloat162float(p.a_ptr[row * p.a_ld + c]);
  }
  return 0.0f;
}

struct SimtRow {
  bool valid;
  long long dst_row, res_row;
};
__device__ __forceinline__ SimtRow simt_decode(const ConvParams& p, int m_tile, int i) {
  SimtRow r;
  int n, h, w;
  if (p.mode == MODE_FLAT) {
    const long long q = (long long)m_tile * kBlockM + i;
    n = (int)(q / p.img_rows);
    const int rem = (int)(q - (long long)n * p.img_rows);
    h = rem / p.wp;
    w = rem - h * p.wp;
    r.res_row = q;
  } else {
    const int per_img = p.tiles_w * p.tiles_h;
    n = m_tile / per_img;
    const int t = m_tile - n * per_img;
    const int th = t / p.tiles_w, tw = t - th * p.tiles_w;
    h = th * kStemTH + i / kStemTW;
    w = tw * kStemTW + i % kStemTW;
    r.res_row = 0;
  }
  r.valid = (n < p.nimg) && (h < p.H) && (w < p.W);
  if (p.dest_mode == DEST_GEOM)
    r.dst_row = (long long)n * p.d_img_rows + (long long)h * p.d_wp + w;
  else
    r.dst_row = (long long)((h & 1) * 2 + (w & 1)) * p.d_plane_rows + (long long)n * p.d_img_rows +
                (long long)(h >> 1) * p.d_wp + (w >> 1);
  return r;
}

// block: 256 threads -> 128 rows x 64 cols; thread (ty, tx): rows ty*8..+7, cols tx*4..+3
__global__ void __launch_bounds__(256) conv_simt_kernel(const ConvParams p) {
  pdl_wait();
  pdl_launch_dependents();
  __shared__ float As[16][kBlockM + 1];
  __shared__ float Bs[16][64 + 1];
  const int m_tile = blockIdx.x;
  const int n0 = blockIdx.y * 64;
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  float acc[8][4];
#pragma unroll
  for (int a = 0; a < 8; ++a)
#pragma unroll
    for (int b = 0; b < 4; ++b) acc[a][b] = 0.0f;

  for (int k0 = 0; k0 < p.ktot; k0 += 16) {
    for (int e = threadIdx.x; e < kBlockM * 16; e += 256) {
      const int i = e >> 4, kk = e & 15;
      As[kk][i] = simt_load_a(p, m_tile, i, k0 + kk);
    }
    for (int e = threadIdx.x; e < 64 * 16; e += 256) {
      const int j = e >> 4, kk = e & 15;
      Bs[kk][j] = __bfloat162float(p.b_ptr[(long long)(n0 + j) * p.ktot + k0 + kk]);
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < 16; ++kk) {
      float a[8], b[4];
#pragma unroll
      for (int x = 0; x < 8; ++x) a[x] = As[kk][ty * 8 + x];
#pragma unroll
      for (int x = 0; x < 4; ++x) b[x] = Bs[kk][tx * 4 + x];
#pragma unroll
      for (int x = 0; x < 8; ++x)
#pragma unroll
        for (int z = 0; z < 4; ++z) acc[x][z] = fmaf(a[x], b[z], acc[x][z]);
    }
    __syncthreads();
  }

#pragma unroll
  for (int x = 0; x < 8; ++x) {
    const SimtRow ri = simt_decode(p, m_tile, ty * 8 + x);
    if (!ri.valid) continue;
#pragma unroll
    for (int z = 0; z < 4; ++z) {
      const int col = n0 + tx * 4 + z;
      float y = fmaf(acc[x][z], p.scale[col], p.shift[col]);
      if (p.residual != nullptr) y += __bfloat162float(p.residual[ri.res_row * p.ldc + col]);
      if (p.relu) y = fmaxf(y, 0.0f);
      if (p.out_f32)
        reinterpret_cast<float*>(p.out)[ri.dst_row * p.ldc + col] = y;
      else
        reinterpret_cast<bf16*>(p.out)[ri.dst_row * p.ldc + col] = __float2bfloat16_rn(y);
    }
  }
}

int launch_conv_simt(const ConvParams& p, cudaStream_t st) {
  if (p.n_total % 64 != 0) { set_error("launch_conv_simt: n_total % 64 != 0"); return IDN_EINVAL; }
  if (p.num_m_tiles <= 0) return IDN_OK;
  dim3 grid(p.num_m_tiles, p.n_total / 64);
  IDN_CUDA_TRY(launch_pdl(conv_simt_kernel, dim3(grid), dim3(256), 0, st, p));
  IDN_CUDA_TRY(cudaGetLastError());
  return IDN_OK;
}

}  // namespace idn
