// Shared declarations of the idunno_b200 CUDA library (sm_100a only).
#pragma once

#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>

#include "../../include/idunno_b200.h"

namespace idn {

typedef __nv_bfloat16 bf16;

void set_error(const std::string& msg);
const char* last_error_cstr();

#define IDN_CUDA_TRY(expr)                                                                   \
  do {                                                                                       \
    cudaError_t _e = (expr);                                                                 \
    if (_e != cudaSuccess) {                                                                 \
      ::idn::set_error(std::string(#expr) + " failed: " + cudaGetErrorString(_e) + " (" +    \
                       __FILE__ + ":" + std::to_string(__LINE__) + ")");                     \
      return IDN_ECUDA;                                                                      \
    }                                                                                        \
  } while (0)

#define IDN_TRY(expr)                 \
  do {                                \
    int _r = (expr);                  \
    if (_r != IDN_OK) return _r;      \
  } while (0)

// ---- programmatic dependent launch -------------------------------------------------------------
// Every kernel of a plan is launched with cudaLaunchAttributeProgrammaticStreamSerialization: its CTAs may
// become resident while the previous kernel of the stream is still draining (one CTA per SM at a time for
// the large-shared-memory kernels), run their prologue -- barrier init, TMEM allocation, tensor-map
// prefetch, resident weights -- and then block in pdl_wait() until the previous kernel has completed and
// its writes are visible.  Rules kept by every kernel: (1) pdl_wait() before the first access to anything
// another kernel wrote or reads, (2) pdl_launch_dependents() only AFTER pdl_wait(), so that at most one
// kernel runs ahead and completion stays transitive along the stream.
#if defined(__CUDACC__)
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
#endif
bool pdl_enabled();         // IDN_NO_PDL=1 switches the launch attribute off (engine.cu)

template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                              Args&&... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
  cfg.attrs = attr; cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}

inline uint16_t f32_to_bf16_bits(float f) {  // round to nearest even, as torch .to(bfloat16)
  uint32_t u;
  std::memcpy(&u, &f, 4);
  if ((u & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((u >> 16) | 0x40);  // NaN
  u += 0x7fffu + ((u >> 16) & 1u);
  return (uint16_t)(u >> 16);
}
inline float bf16_bits_to_f32(uint16_t b) {
  uint32_t u = (uint32_t)b << 16;
  float f;
  std::memcpy(&f, &u, 4);
  return f;
}

// ---------------------------------------------------------------------------------------------
// Convolution / FC as "shifted GEMM":  out[q][n] = sum_taps sum_c A[q + off(tap)][c] * Wt[n][tap, c]
//
// Activations live in HBM as bf16 "flat padded NHWC": image n occupies img_rows = Hp*Wp rows of C
// channels; pixel (h, w) is row n*img_rows + h*Wp + w; rows with h >= H or w >= W are zero and stay
// zero (they are the convolution's zero padding: the row before pixel (h, 0) is pixel (h-1, Wp-1),
// a pad column).  A filter tap (r, s) of a stride-1 convolution is then the constant row offset
// (r-pad)*Wp + (s-pad), i.e. a plain 2-D TMA box over the [rows][C] matrix.  Stride-2 convolutions
// read "parity planes" (the producing layer scatters pixel (h,w) to plane (h&1, w&1)), which makes
// their taps constant row offsets as well.  ResNet18's 7x7/s2 stem reads a 4-channel NHWC image
// through a no-swizzle descriptor whose core matrices overlap (stem_toeplitz_kernel, conv_tc.cu);
// AlexNet's 11x11/s4 stem is a 3x3 convolution over a 4x4 space-to-depth image.
// ---------------------------------------------------------------------------------------------
constexpr int kMaxTaps = 32;
constexpr int kBlockM = 128;   // GEMM rows (pixels) per tile == TMEM lanes
constexpr int kBlockK = 64;    // bf16 channels per k-block == one 128-byte swizzle row
constexpr int kStemTW = 8;     // ResNet stem tile: 16 rows x 8 columns of output pixels
constexpr int kStemTH = 16;
constexpr int kStemWin = 32;   // elements per filter row: 8-pixel window x 4 channels (7 taps used)

enum { MODE_FLAT = 0, MODE_STEM = 1 };
enum { DEST_GEOM = 0, DEST_PARITY = 1 };

// Division by a run-time constant for the row decode of the epilogues: q / d == umulhi(q, mul) >> shift
// for 0 <= q < 2^31 (d = 1: mul = 0 means "q itself").
struct FastDiv {
  unsigned int mul, shift, d;
};
inline FastDiv make_fastdiv(unsigned int d) {
  FastDiv f; f.d = d; f.mul = 0; f.shift = 0;
  if (d > 1) {
    unsigned int lg = 0;
    while ((1ull << lg) < d) ++lg;
    const unsigned int p = 31 + lg;
    f.mul = (unsigned int)(((1ull << p) + d - 1) / d);
    f.shift = p - 32;
  }
  return f;
}
__device__ __forceinline__ unsigned int fastdiv(unsigned int q, const FastDiv& f) {
  return f.mul == 0 ? q : (__umulhi(q, f.mul) >> f.shift);
}

struct ConvParams {
  // K loop
  int num_taps;
  int c_blocks;               // 64-channel blocks per tap
  int tap_off[kMaxTaps];      // flat: row offset of the tap; stem: row-group offset `a` of the tap
  int tap_par[kMaxTaps];      // stem: row phase `b` of the tap (input row = stride*(oh+a)+b)
  // tiles
  int num_m_tiles, num_n_tiles;
  // tile (source) geometry
  int mode;
  int nimg;                   // images valid in this pass
  int img_rows, wp, H, W;     // flat: rows per image, row pitch, valid extents (stem: H=OH, W=OW)
  int tiles_w, tiles_h;       // stem: tiles per image
  // destination
  int dest_mode;
  int d_img_rows, d_wp, d_plane_rows;
  int ldc;                    // elements per destination row
  int n_total;                // padded Cout (multiple of the N tile)
  void* out;
  const bf16* residual;       // same geometry as the tile rows (flat only), ld = ldc
  const float* scale;
  const float* shift;
  int relu;
  int out_f32;
  // raw operand views for the SIMT cross-check kernel
  const bf16* a_ptr;
  const bf16* b_ptr;
  long long a_rows;           // flat: number of rows of A (reads outside are zero, like TMA)
  int a_ld;                   // flat: elements per A row
  int ktot;                   // elements per B row (= num_taps * c_blocks * 64)
  int st_stride, st_ws, st_h, st_ow, st_nrow;   // stem image: conv stride, row pitch (px), rows,
                                                // window starts per row, row groups per image
  const bf16* stem_w;         // stem weights in the shared-memory order of stem_toeplitz_kernel
  // halo kernel (conv_halo_kernel): one staged A tile per 64-channel block serves every tap
  int halo_lo;                // smallest tap offset (rows)
  int halo_r1, halo_r2;       // rows of the two TMA boxes that make up the staged tile (r2 may be 0)
  int a_stage_bytes;          // (r1 + r2) * 128 rounded up to 1024
  int a_stages, b_stages;     // ring depths chosen by the host from the shared-memory budget
  int msub;                   // 128-row sub-tiles per work item (1 or 2)
  FastDiv fd_img_rows, fd_wp, fd_per_img, fd_tiles_w;   // row decode without integer division
  int profile;                // 1: book pipeline-wait cycles per role (bench.py timing mode)
  int tc_pair;                // 1: conv_tc_pair_kernel (per-k-block ring on CTA pairs); weight boxes are N/2 rows
  int pair;                   // 1: conv_pair_kernel (CTA pairs, cta_group::2); weight boxes are N/2 rows
  int b_group;                // halo kernel, streamed weights: filter taps per weight stage (1 or 3).  One
                              // barrier round-trip of the MMA warp costs ~500 cycles whatever it covers
  // stem only: 3x3 stride-2 pad-1 max-pool fused into the epilogue (ResNet18 conv1 -> maxpool): the conv
  // output is not stored, stem_pool_kernel writes the pooled pixels to pool_out with plain stores.
  bf16* pool_out;             // nullptr: plain stem
  int pool_img_rows, pool_wp, pool_oh, pool_ow;
  bf16* pool_ring;            // conv_pair_kernel<.., POOL>: [pairs][1024 rows][n_total] scratch for conv rows (L2-resident)
  // conv_tc_pair_kernel, FC layers: split-K.  A function of the LAYER's shape only, never of the batch, so
  // that an image's result does not depend on the engine's capacity or on the tensor batch it falls in.
  int k_steps;                // conv_halo_kernel, resident weights: MMA steps (16 channels) issued per channel block;
                              // the steps left out have all-zero weights (0 is read as 4)
  int split_defer;            // split-K only: 1 = write the partial tiles and stop; the kernel after this one adds them
                              // (AlexNet's FC8: softmax_top1_splitk_kernel) -- no CTA waits for another
  int split_k;                // 0 / 1: off
  float* splitk_ws;           // fp32 partial tiles [tile half][split][128][N tile]
  unsigned int* splitk_cnt;   // per tile half: {arrivals, slices reduced}; zero between launches
};

// launchers (conv_tc.cu / conv_simt.cu)
// A second convolution over the same output tiling that rides in the first one's launch (conv_tc.cu).
struct ConvSecond { const CUtensorMap* tmA; const CUtensorMap* tmB; const ConvParams* p; };
int launch_conv_tc(const CUtensorMap& tmA, const CUtensorMap& tmB, const ConvParams& p, int block_n,
                   int num_sms, cudaStream_t st, const ConvSecond* second = nullptr);
int launch_conv_simt(const ConvParams& p, cudaStream_t st);
// Chooses the halo-kernel configuration for a stride-1 multi-tap convolution; returns false when
// the layer should stay on conv_tc_kernel.  Fills p.halo_* / stages / msub, *block_n, *b_resident.
bool plan_tc_pair(int num_taps, int n_total, long long m_tiles_cap, int num_sms, int* block_n);   // conv_tc_pair_kernel
bool plan_conv_halo(ConvParams* p, int n_total, long long m_rows_cap, int num_sms, int* block_n,
                    bool* b_resident);
int launch_conv_halo(const CUtensorMap& tmA1, const CUtensorMap& tmA2, const CUtensorMap& tmB,
                     const ConvParams& p, int block_n, bool b_resident, int num_sms, cudaStream_t st);
int read_tc_watchdog();
int read_tc_wait_cycles(unsigned long long* out8);
int conv_tc_configure();   // cudaFuncSetAttribute for every instantiation; once per process/device

// tensor maps (conv_tc.cu)
int make_tmap_2d(CUtensorMap* tm, const void* base, uint64_t inner_elems, uint64_t rows,
                 uint64_t row_stride_bytes, uint32_t box_inner, uint32_t box_rows);
int make_tmap_stem(CUtensorMap* tm, const void* base, int ws_px, int img_h, int nimg_cap);
int make_tmap_stem_row(CUtensorMap* tm, const void* base, int ws_px, int img_h, int nimg_cap);
int make_tmap_stem_band(CUtensorMap* tm, const void* base, int ws_px, int img_h, int nimg_cap);
int launch_stem_tc(const CUtensorMap& tmA, const CUtensorMap& tmRow, const CUtensorMap& tmBand, const ConvParams& p,
                   int num_sms, cudaStream_t st);

// other kernels (pool_head.cu)
struct PoolParams {
  const bf16* in; bf16* out;
  int nimg, C;
  int s_img_rows, s_wp, H, W;      // source geometry (flat padded or dense)
  int d_img_rows, d_wp, OH, OW;    // destination geometry
  int pad;                          // 3x3 window, stride 2
};
int launch_maxpool(const PoolParams& p, cudaStream_t st);
int launch_resnet_head(const bf16* act, int nimg, int img_rows, int wp, int H, int W, int C,
                       const float* fc_w, const float* fc_b, float* feats, float* logits,
                       int32_t* idx, float* prob, cudaStream_t st);
int launch_resnet_head_fused(const bf16* act, int nimg, int img_rows, int wp, int H, int W, int C,
                             const float* fc_wt, const float* fc_b, float* feats, float* logits_out_or_null,
                             int32_t* idx, float* prob, cudaStream_t st);
int launch_softmax_top1_splitk(const float* ws, int splits, int num_n_tiles, const float* scale, const float* shift, int nimg,
                               float* logits_out_or_null, int32_t* idx, float* prob, cudaStream_t st);
int launch_softmax_top1(const float* logits, int ld, int nimg, float* logits_out_or_null,
                        int32_t* idx, float* prob, cudaStream_t st);
// Copies n_words 32-bit words with SM loads/stores.  Used for the few KB of per-batch records that come
// from page-locked host memory: a cudaMemcpyAsync would queue behind the 60 MB image copy of the NEXT
// batch on the host-to-device copy engine and stall this batch's kernels for a millisecond.
int launch_copy_words(void* dst, const void* src, size_t n_words, cudaStream_t st);

// preprocess (preprocess.cu)
struct PreMeta {            // one per image, device resident
  long long offset;         // byte offset of the image in the packed buffer
  int H, W;                 // source size
  int htab, vtab;           // offsets (in int32 units) of the resample tables in the table pool
};
enum { PRE_OUT_NCHW = 0, PRE_OUT_STEM = 1, PRE_OUT_S2D4 = 2 };
int launch_preprocess(const uint8_t* d_imgs, const PreMeta* d_meta, const int* d_tables,
                      const uint16_t* d_lut, int n, int out_mode, void* out, int stem_left_pad,
                      int stem_ws, cudaStream_t st);
// Builds the resample table (Pillow fixed point) of the 224 cropped outputs of one axis.
void build_axis_table(int in_size, int out_size, int crop_off, std::string* ints /* int32 bytes */);
void resized_size(int h, int w, int* nh, int* nw);
int crop_offset(int resized);
void needed_source_range(int in_size, int out_size, int crop_off, int* first, int* last);
void build_norm_lut(uint16_t* lut768);

}  // namespace idn
