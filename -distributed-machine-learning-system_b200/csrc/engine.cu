// Model plans (AlexNet, ResNet18), weight folding / re-layout, buffers and the C ABI.
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <map>
#include <memory>
#include <tuple>
#include <vector>

#include "common.cuh"

namespace idn {

// ------------------------------------------------------------------------------ error plumbing
static thread_local std::string g_last_error;
void set_error(const std::string& msg) { g_last_error = msg; }
bool pdl_enabled() {
  static const bool on = [] { const char* v = std::getenv("IDN_NO_PDL"); return !(v && v[0] == '1'); }();
  return on;
}
const char* last_error_cstr() { return g_last_error.c_str(); }

// ------------------------------------------------------------------------------------ geometry
struct Geom {            // flat padded NHWC geometry of one activation buffer
  int H = 0, W = 0, Hp = 0, Wp = 0;
  int img_rows() const { return Hp * Wp; }
};
static Geom make_geom(int h, int w, int pad) { Geom g; g.H = h; g.W = w; g.Hp = h + pad; g.Wp = w + pad; return g; }

struct Act {             // a device activation buffer
  bf16* ptr = nullptr;
  Geom g;
  int C = 0;
  int planes = 1;        // 4 = parity planes of the geometry g
  long long plane_rows = 0;   // rows per plane (= cap * img_rows)
  long long rows() const { return plane_rows * planes; }
};

struct ConvWeights {     // device, kernel layout
  bf16* wt = nullptr;    // [n_total][ktot] K-major
  float* scale = nullptr;
  float* shift = nullptr;
  int n_total = 0, ktot = 0, taps = 0, c_blocks = 0;
  int k_steps = 4;       // 16-channel MMA steps of a channel block that carry non-zero weights (single-block layers)
  bf16* stem_flat = nullptr;   // ResNet stem only: weights in stem_toeplitz_kernel's smem order
};

enum StepKind { STEP_PRE = 0, STEP_CONV, STEP_POOL, STEP_RHEAD, STEP_SOFTMAX };

struct Step {
  StepKind kind;
  std::string name;
  // conv
  CUtensorMap tmA, tmA2, tmB;
  ConvParams cp;
  int block_n = 0;
  bool use_halo = false, b_resident = false;   // conv_halo_kernel instead of conv_tc_kernel
  int second = -1;         // plan step of a convolution that rides in this step's launch (1x1 downsample)
  int merged_into = -1;    // ... and, on that step, the step whose launch carries it
  // pool
  PoolParams pp;
  // what the step wrote (for idn_read_activation)
  const void* out_ptr = nullptr;
  int out_kind = 0;      // 0 none, 1 bf16 geom, 2 bf16 parity planes, 3 f32 rows (FC logits)
  Geom out_g;            // for parity: geometry of one plane
  int out_C = 0, out_ld = 0;
  long long out_plane_rows = 0;
  int out_H = 0, out_W = 0;   // logical image extent (parity: 2x plane extent)
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  unsigned long long wait_cycles[8] = {0, 0, 0, 0, 0, 0, 0, 0};   // filled in timing mode
};

}  // namespace idn

using namespace idn;

constexpr int kSlots = 3;     // submissions that may be in flight (IDN_MAX_IN_FLIGHT)

struct idn_engine {
  int device = 0, model = 0, cap = 0, num_sms = 148, impl = IDN_IMPL_TC;
  cudaStream_t own_stream = nullptr, stream = nullptr;
  std::vector<void*> dev_allocs;
  std::vector<Step> steps;
  // preprocess
  PreMeta* d_meta = nullptr;
  PreMeta* h_meta = nullptr;          // pinned, kMetaRing x cap records (device-resident entry point)
  cudaEvent_t meta_ev[4] = {nullptr, nullptr, nullptr, nullptr};   // "ring entry has been read"
  int meta_ring = 1, meta_next = 0;
  int* d_tables = nullptr;
  int* h_tables = nullptr;                  // page-locked mirror of the pool (upload source, tile sizing)
  size_t tables_cap = 0, tables_used = 0;   // int32 units
  std::map<std::tuple<int, int, int>, int> table_index;
  uint16_t* d_lut = nullptr;
  void* xs = nullptr;                 // stem input image
  int stem_left = 0, stem_ws = 232;
  // Submission slots of the host-buffer path: each has its own staging buffer, page-locked and device
  // records, page-locked result buffers and events, so that the H2D copies of the next tensor batches
  // run (on the copy stream) while the kernels of the current one run (on the handle's stream).
  struct Slot {
    uint8_t* d_stage = nullptr; size_t d_stage_cap = 0;
    PreMeta* h_meta = nullptr; PreMeta* d_meta = nullptr;
    float* d_logits = nullptr;
    int32_t* h_idx = nullptr; float* h_prob = nullptr;       // pinned; the last kernel stores into them
    cudaEvent_t ev_copied = nullptr, ev_done = nullptr;
    int n = 0; bool busy = false, logits = false;
  } slot[kSlots];
  cudaStream_t copy_stream = nullptr;
  int next_slot = 0;
  float* d_logits = nullptr;                           // ResNet logits when the caller passes none
  // head
  const bf16* head_act = nullptr; Geom head_g; int head_C = 0;
  float* fc_w = nullptr; float* fc_b = nullptr;        // ResNet fp32 FC
  float* fc_wt = nullptr;                              // the same weights transposed, [8][512][128] (fused head)
  bool no_head_fusion = false;                         // IDN_NO_HEAD_FUSION=1: avgpool, FC, softmax as three kernels
  bool no_ds_fusion = false;                           // IDN_NO_DS_FUSION=1: 1x1 downsamples as launches of their own
  float* feats = nullptr;                              // ResNet pooled features [cap][512] fp32
  int pre_mode = PRE_OUT_STEM;
  float* logits_pad = nullptr;                         // AlexNet [cap][1024] fp32
  int64_t launches = 0;
  bool stem_simt = false;   // bring-up switch IDN_STEM_SIMT=1: run the two C=3 stem layers on the
                            // CUDA-core kernel (explicit, never automatic)
  bool timing = false;
  bool wait_profile = false;  // timing mode 2: also book per-role pipeline waits (slower kernels)
  int stop_after = -1;      // diagnostic: run the plan up to and including this step
  // ResNet18: conv1 and the max-pool behind it run as ONE kernel on the product path (stem_pool_kernel:
  // pooling windows are resolved in shared memory, pooled pixels leave with plain stores).
  int fuse_stem_step = -1, fuse_pool_step = -1, pool_last_reader = -1;
  bf16* pool_buf = nullptr; size_t pool_img_bytes = 0;
  int pool_img_rows = 0, pool_wp = 0, pool_oh = 0, pool_ow = 0;
  // further conv -> max-pool pairs fused on the CTA-pair kernel (AlexNet features.3/5, features.10/12)
  struct FusePair { int conv_step, pool_step; bf16* buf; int img_rows, wp, oh, ow; };
  std::vector<FusePair> fuse_pairs;
  bf16* pool_ring = nullptr;           // [num_sms / 2][1024 rows][256] bf16: the pairs' conv-row rings
  bool no_pool_fusion = false;
  bool full_h2d = false;               // IDN_FULL_H2D=1: always copy whole images
  bool no_split_defer = false;         // IDN_NO_SPLIT_DEFER=1: FC8's splits are added inside its own kernel (CTAs wait)
  int64_t h2d_bytes = 0;               // image bytes queued host-to-device so far (idn_h2d_bytes)
  int last_n = 0;
  // split-K scratch of the FC layers, shared by all of them (they run one after the other)
  float* splitk_ws = nullptr; size_t splitk_ws_floats = 0;
  unsigned int* splitk_cnt = nullptr; size_t splitk_cnt_n = 0;
};

namespace idn {

template <class T>
static int dev_alloc(idn_engine* e, T** p, size_t count, bool zero = true) {
  void* q = nullptr;
  IDN_CUDA_TRY(cudaMalloc(&q, count * sizeof(T) + 256));
  if (zero) IDN_CUDA_TRY(cudaMemset(q, 0, count * sizeof(T) + 256));
  if (e) e->dev_allocs.push_back(q);
  *p = reinterpret_cast<T*>(q);
  return IDN_OK;
}

static int alloc_act(idn_engine* e, Act* a, Geom g, int C, int cap, int planes = 1) {
  a->g = g; a->C = C; a->planes = planes;
  a->plane_rows = (long long)cap * g.img_rows();
  return dev_alloc(e, &a->ptr, (size_t)a->rows() * C);
}

static int pick_block_n(int n_total, long long m_tiles, int num_sms) {
  const int cands[4] = {256, 192, 128, 64};
  if (const char* v = std::getenv("IDN_BLOCK_N")) {     // test hook: force an N tile that divides Cout
    const int f = std::atoi(v);
    for (int c : cands) if (c == f && n_total % c == 0) return c;
  }
  for (int c : cands)
    if (n_total % c == 0 && m_tiles * (n_total / c) * 10 >= (long long)num_sms * 8) return c;
  if (n_total % 128 == 0) return 128;
  return 64;
}

// Uploads weights [cout][cin][R][S] (fp32, host) as bf16 [n_total][taps*cpad] with the tap-major /
// channel-minor K order of the kernels, plus fp32 scale/shift padded to n_total.
static int upload_conv_weights(idn_engine* e, ConvWeights* cw, const float* w, int cout, int cin,
                               int R, int S, const float* scale, const float* shift) {
  const int cpad = (cin + 63) / 64 * 64;
  cw->taps = R * S; cw->c_blocks = cpad / 64; cw->ktot = R * S * cpad;
  cw->n_total = (cout + 63) / 64 * 64;
  std::vector<uint16_t> h((size_t)cw->n_total * cw->ktot, 0);
  for (int n = 0; n < cout; ++n)
    for (int c = 0; c < cin; ++c)
      for (int r = 0; r < R; ++r)
        for (int s = 0; s < S; ++s)
          h[(size_t)n * cw->ktot + (size_t)(r * S + s) * cpad + c] =
              f32_to_bf16_bits(w[(((size_t)n * cin + c) * R + r) * S + s]);
  if (cw->c_blocks == 1) {
    // trailing all-zero channels (AlexNet's space-to-depth stem fills 48 of its 64): their MMA steps add nothing
    int top = 0;
    for (size_t i = 0; i < h.size(); ++i)
      if ((h[i] & 0x7fffu) != 0) top = std::max(top, (int)(i % 64));
    cw->k_steps = top / 16 + 1;
  }
  IDN_TRY(dev_alloc(e, &cw->wt, h.size(), false));
  IDN_CUDA_TRY(cudaMemcpy(cw->wt, h.data(), h.size() * 2, cudaMemcpyHostToDevice));
  std::vector<float> sc(cw->n_total, 0.0f), sh(cw->n_total, 0.0f);
  for (int n = 0; n < cout; ++n) { sc[n] = scale ? scale[n] : 1.0f; sh[n] = shift ? shift[n] : 0.0f; }
  IDN_TRY(dev_alloc(e, &cw->scale, sc.size(), false));
  IDN_TRY(dev_alloc(e, &cw->shift, sh.size(), false));
  IDN_CUDA_TRY(cudaMemcpy(cw->scale, sc.data(), sc.size() * 4, cudaMemcpyHostToDevice));
  IDN_CUDA_TRY(cudaMemcpy(cw->shift, sh.data(), sh.size() * 4, cudaMemcpyHostToDevice));
  return IDN_OK;
}

// ResNet stem weights [64][3][7][7] -> K-major [64][7*32] with k = r*32 + s*4 + c (window pixel s,
// channel c of the 4-channel pixel; s = 7 and c = 3 are zero), plus the same values in the
// shared-memory order of stem_toeplitz_kernel: [r][k chunk of 8][n][8].
static int upload_stem_weights(idn_engine* e, ConvWeights* cw, const float* w, int cout, int R,
                               int S, const float* scale, const float* shift) {
  if (cout != 64 || R != 7 || S != 7) { set_error("stem: only the 7x7 stride-2 ResNet stem"); return IDN_EINVAL; }
  cw->taps = R; cw->c_blocks = 1; cw->ktot = R * kStemWin; cw->n_total = 64;
  std::vector<uint16_t> h((size_t)64 * cw->ktot, 0), flat((size_t)64 * cw->ktot, 0);
  for (int n = 0; n < cout; ++n)
    for (int c = 0; c < 3; ++c)
      for (int r = 0; r < R; ++r)
        for (int s = 0; s < S; ++s) {
          const uint16_t v = f32_to_bf16_bits(w[(((size_t)n * 3 + c) * R + r) * S + s]);
          const int kk = s * 4 + c;                       // 0..31 inside the filter row
          h[(size_t)n * cw->ktot + r * kStemWin + kk] = v;
          flat[((size_t)(r * 4 + kk / 8) * 64 + n) * 8 + kk % 8] = v;
        }
  IDN_TRY(dev_alloc(e, &cw->wt, h.size(), false));
  IDN_CUDA_TRY(cudaMemcpy(cw->wt, h.data(), h.size() * 2, cudaMemcpyHostToDevice));
  IDN_TRY(dev_alloc(e, &cw->stem_flat, flat.size(), false));
  IDN_CUDA_TRY(cudaMemcpy(cw->stem_flat, flat.data(), flat.size() * 2, cudaMemcpyHostToDevice));
  std::vector<float> sc(64), sh(64);
  for (int n = 0; n < 64; ++n) { sc[n] = scale ? scale[n] : 1.0f; sh[n] = shift ? shift[n] : 0.0f; }
  IDN_TRY(dev_alloc(e, &cw->scale, 64, false));
  IDN_TRY(dev_alloc(e, &cw->shift, 64, false));
  IDN_CUDA_TRY(cudaMemcpy(cw->scale, sc.data(), 256, cudaMemcpyHostToDevice));
  IDN_CUDA_TRY(cudaMemcpy(cw->shift, sh.data(), 256, cudaMemcpyHostToDevice));
  return IDN_OK;
}

struct ConvSpec {
  int R = 1, S = 1, stride = 1, pad = 0;
  bool relu = false, out_f32 = false;
};

static void fill_dest(Step* st, const Act& dst, bool out_f32, int ldc) {
  ConvParams& cp = st->cp;
  cp.ldc = ldc;
  cp.out = dst.ptr;
  cp.out_f32 = out_f32 ? 1 : 0;
  cp.d_img_rows = dst.g.img_rows();
  cp.d_wp = dst.g.Wp;
  cp.d_plane_rows = (int)dst.plane_rows;
  cp.dest_mode = dst.planes == 4 ? DEST_PARITY : DEST_GEOM;
  st->out_ptr = dst.ptr;
  st->out_kind = out_f32 ? 3 : (dst.planes == 4 ? 2 : 1);
  st->out_g = dst.g;
  st->out_C = dst.C;
  st->out_ld = ldc;
  st->out_plane_rows = dst.plane_rows;
  st->out_H = dst.planes == 4 ? dst.g.H * 2 : dst.g.H;
  st->out_W = dst.planes == 4 ? dst.g.W * 2 : dst.g.W;
}

// Flat convolution step: src (geometry or parity planes) -> dst.  The M tiling runs over the
// OUTPUT pixel grid, which for stride 1 is src.g and for stride 2 the plane geometry of src.
static int add_conv(idn_engine* e, const std::string& name, const Act& src, const ConvWeights& cw,
                    const ConvSpec& spec, const Act& dst, const Act* residual, int cap) {
  Step st;
  st.kind = STEP_CONV;
  st.name = name;
  ConvParams& cp = st.cp;
  std::memset(&cp, 0, sizeof(cp));
  const Geom& tg = src.g;                    // tile geometry (== plane geometry for stride 2)
  cp.mode = MODE_FLAT;
  cp.num_taps = cw.taps;
  cp.c_blocks = cw.c_blocks;
  if (cw.taps != spec.R * spec.S || cw.taps > kMaxTaps) { set_error(name + ": tap count"); return IDN_EINVAL; }
  if (spec.stride == 1) {
    if (src.planes != 1 || tg.Wp - tg.W < spec.pad || tg.Hp - tg.H < spec.pad) {
      set_error(name + ": source layout lacks the zero padding this convolution needs");
      return IDN_EINVAL;
    }
    for (int r = 0; r < spec.R; ++r)
      for (int s = 0; s < spec.S; ++s)
        cp.tap_off[r * spec.S + s] = (r - spec.pad) * tg.Wp + (s - spec.pad);
  } else {
    if (src.planes != 4 || !((spec.R == 3 && spec.S == 3 && spec.pad == 1) ||
                             (spec.R == 1 && spec.S == 1 && spec.pad == 0))) {
      set_error(name + ": stride-2 convolutions read parity planes (3x3 p1 or 1x1 p0)");
      return IDN_EINVAL;
    }
    for (int r = 0; r < spec.R; ++r)
      for (int s = 0; s < spec.S; ++s) {
        int ph = 0, pw = 0, di = 0, dj = 0;
        if (spec.R == 3) {                   // input row 2*oh + r - 1
          ph = (r == 1) ? 0 : 1; di = (r == 0) ? -1 : 0;
          pw = (s == 1) ? 0 : 1; dj = (s == 0) ? -1 : 0;
        }
        cp.tap_off[r * spec.S + s] = (int)((ph * 2 + pw) * src.plane_rows) + di * tg.Wp + dj;
      }
  }
  cp.img_rows = tg.img_rows(); cp.wp = tg.Wp; cp.H = tg.H; cp.W = tg.W;
  cp.n_total = cw.n_total;
  cp.scale = cw.scale; cp.shift = cw.shift;
  cp.relu = spec.relu ? 1 : 0;
  cp.residual = residual ? residual->ptr : nullptr;
  if (residual && (residual->g.img_rows() != tg.img_rows() || residual->g.Wp != tg.Wp ||
                   residual->C != dst.C || residual->planes != 1)) {
    set_error(name + ": residual geometry differs from the tile geometry");
    return IDN_EINVAL;
  }
  fill_dest(&st, dst, spec.out_f32, dst.C);
  if (dst.C != cw.n_total) { set_error(name + ": dst channels != padded Cout"); return IDN_EINVAL; }
  cp.k_steps = cw.k_steps;
  cp.a_ptr = src.ptr; cp.b_ptr = cw.wt;
  cp.a_rows = src.rows(); cp.a_ld = src.C; cp.ktot = cw.ktot;
  if (src.C != cw.c_blocks * 64) { set_error(name + ": src channels != weight channels"); return IDN_EINVAL; }
  const long long m_tiles_cap = ((long long)cap * tg.img_rows() + kBlockM - 1) / kBlockM;
  st.use_halo = plan_conv_halo(&cp, cw.n_total, (long long)cap * tg.img_rows(), e->num_sms,
                               &st.block_n, &st.b_resident);
  if (st.use_halo) {
    cp.num_n_tiles = cw.n_total / st.block_n;
    IDN_TRY(make_tmap_2d(&st.tmA, src.ptr, (uint64_t)src.C, (uint64_t)src.rows(),
                         (uint64_t)src.C * 2, kBlockK, (uint32_t)cp.halo_r1));
    IDN_TRY(make_tmap_2d(&st.tmA2, src.ptr, (uint64_t)src.C, (uint64_t)src.rows(),
                         (uint64_t)src.C * 2, kBlockK, (uint32_t)(cp.halo_r2 > 0 ? cp.halo_r2 : 8)));
    IDN_TRY(make_tmap_2d(&st.tmB, cw.wt, (uint64_t)cw.ktot, (uint64_t)cw.n_total,
                         (uint64_t)cw.ktot * 2, kBlockK, (uint32_t)(cp.pair ? st.block_n / 2 : st.block_n)));
    e->steps.push_back(st);
    return IDN_OK;
  }
  st.block_n = pick_block_n(cw.n_total, m_tiles_cap, e->num_sms);
  cp.tc_pair = plan_tc_pair(cp.num_taps, cw.n_total, m_tiles_cap, e->num_sms, &st.block_n) ? 1 : 0;
  // Narrow FC layers (one row per image, long K, N <= 1024: AlexNet's FC8 is 4 x 2 tile pairs of 64 k-blocks for
  // 74 CTA pairs): 256-wide pair tiles with split-K.  The split is a function of the layer's shape alone (never
  // of the capacity or the batch), so every engine adds the same partial sums in the same order.  Wider FC
  // layers fill the machine without it and lost a little to the extra pass (FC7 27 -> 31 us).
  static const bool no_split = [] { const char* v = std::getenv("IDN_NO_FC_SPLIT"); return v && v[0] == '1'; }();
  const int k_blocks = cp.num_taps * cp.c_blocks;
  if (!no_split && cp.num_taps == 1 && tg.img_rows() == 1 && residual == nullptr && k_blocks >= 32 &&
      cw.n_total % 256 == 0 && cw.n_total <= 1024 && e->num_sms % 2 == 0) {
    int split = 1;
    while (split < 8 && split * 2 * (cw.n_total / 256) <= 32 && k_blocks / (split * 2) >= 8 &&
           k_blocks % (split * 2) == 0) split *= 2;
    if (split > 1) {
      cp.tc_pair = 1;
      st.block_n = 256;
      cp.split_k = split;
      const size_t halves = (size_t)((m_tiles_cap + 1) / 2) * 2 * (cw.n_total / 256);
      const size_t ws_floats = halves * split * kBlockM * 256;
      if (ws_floats > e->splitk_ws_floats) {
        IDN_TRY(dev_alloc(e, &e->splitk_ws, ws_floats, false));      // earlier, smaller one stays in dev_allocs
        e->splitk_ws_floats = ws_floats;
      }
      if (halves > e->splitk_cnt_n) {
        IDN_TRY(dev_alloc(e, &e->splitk_cnt, 2 * halves));
        e->splitk_cnt_n = halves;
      }
    }
  }
  cp.num_n_tiles = cw.n_total / st.block_n;
  IDN_TRY(make_tmap_2d(&st.tmA, src.ptr, (uint64_t)src.C, (uint64_t)src.rows(),
                       (uint64_t)src.C * 2, kBlockK, kBlockM));
  IDN_TRY(make_tmap_2d(&st.tmB, cw.wt, (uint64_t)cw.ktot, (uint64_t)cw.n_total,
                       (uint64_t)cw.ktot * 2, kBlockK, (uint32_t)(cp.tc_pair ? st.block_n / 2 : st.block_n)));
  e->steps.push_back(st);
  return IDN_OK;
}

static int add_stem(idn_engine* e, const std::string& name, const void* xs, int ws, int left_pad,
                    const ConvWeights& cw, int R, int stride, int pad, int OH, int OW,
                    const Act& dst, int cap) {
  (void)left_pad;
  Step st;
  st.kind = STEP_CONV;
  st.name = name;
  ConvParams& cp = st.cp;
  std::memset(&cp, 0, sizeof(cp));
  cp.mode = MODE_STEM;
  if (R != 7 || stride != 2 || pad != 3 || OH != 112 || OW != 112) { set_error(name + ": unsupported stem"); return IDN_EINVAL; }
  cp.num_taps = R; cp.c_blocks = 1;
  cp.stem_w = cw.stem_flat;
  for (int r = 0; r < R; ++r) {             // input row = stride*oh + r - pad = stride*(oh+a) + b
    const int d = r - pad;
    const int a = (int)std::floor((double)d / stride);
    cp.tap_off[r] = a;
    cp.tap_par[r] = d - a * stride;
  }
  cp.H = OH; cp.W = OW;
  cp.tiles_w = (OW + kStemTW - 1) / kStemTW;
  cp.tiles_h = (OH + kStemTH - 1) / kStemTH;
  cp.n_total = 64; cp.num_n_tiles = 1;
  cp.scale = cw.scale; cp.shift = cw.shift; cp.relu = 1;
  fill_dest(&st, dst, false, 64);
  cp.a_ptr = reinterpret_cast<const bf16*>(xs); cp.b_ptr = cw.wt; cp.ktot = cw.ktot;
  cp.st_stride = stride; cp.st_ws = ws; cp.st_h = IDN_CROP; cp.st_ow = OW;
  cp.st_nrow = IDN_CROP / stride;
  st.block_n = 64;
  if (!e->stem_simt) {
    IDN_TRY(make_tmap_stem(&st.tmA, xs, ws, IDN_CROP, cap));
    IDN_TRY(make_tmap_stem_row(&st.tmA2, xs, ws, IDN_CROP, cap));
    IDN_TRY(make_tmap_stem_band(&st.tmB, xs, ws, IDN_CROP, cap));
  }
  e->steps.push_back(st);
  return IDN_OK;
}

static void add_pool(idn_engine* e, const std::string& name, const Act& src, const Act& dst, int pad) {
  Step st;
  st.kind = STEP_POOL;
  st.name = name;
  std::memset(&st.cp, 0, sizeof(st.cp));
  PoolParams& pp = st.pp;
  pp.in = src.ptr; pp.out = dst.ptr; pp.C = src.C; pp.nimg = 0;
  pp.s_img_rows = src.g.img_rows(); pp.s_wp = src.g.Wp; pp.H = src.g.H; pp.W = src.g.W;
  pp.d_img_rows = dst.g.img_rows(); pp.d_wp = dst.g.Wp; pp.OH = dst.g.H; pp.OW = dst.g.W;
  pp.pad = pad;
  st.out_ptr = dst.ptr; st.out_kind = 1; st.out_g = dst.g; st.out_C = dst.C; st.out_ld = dst.C;
  st.out_plane_rows = dst.plane_rows; st.out_H = dst.g.H; st.out_W = dst.g.W;
  e->steps.push_back(st);
}

// BatchNorm (eval, eps 1e-5) folded to scale/shift: tv:models/resnet.py:197,208-213
static void fold_bn(const float* g, const float* b, const float* rm, const float* rv, int n,
                    std::vector<float>* scale, std::vector<float>* shift) {
  scale->resize(n); shift->resize(n);
  for (int i = 0; i < n; ++i) {
    const float s = g[i] / std::sqrt(rv[i] + 1e-5f);
    (*scale)[i] = s;
    (*shift)[i] = b[i] - rm[i] * s;
  }
}

struct BlobReader {
  const float* p; size_t left;
  const float* take(size_t n) {
    if (n > left) return nullptr;
    const float* r = p; p += n; left -= n; return r;
  }
};

static int try_merge_conv_steps(idn_engine* e, int a, int b, const ConvWeights& wb);

static int build_resnet18(idn_engine* e, const float* blob, size_t count) {
  const int cap = e->cap;
  BlobReader br{blob, count};
  auto conv_bn = [&](ConvWeights* cw, int cout, int cin, int R, bool stem) -> int {
    const float* w = br.take((size_t)cout * cin * R * R);
    const float* g = br.take(cout); const float* b = br.take(cout);
    const float* rm = br.take(cout); const float* rv = br.take(cout);
    if (!w || !g || !b || !rm || !rv) { set_error("resnet18: weight blob too short"); return IDN_EWEIGHTS; }
    std::vector<float> sc, sh;
    fold_bn(g, b, rm, rv, cout, &sc, &sh);
    if (stem) return upload_stem_weights(e, cw, w, cout, R, R, sc.data(), sh.data());
    return upload_conv_weights(e, cw, w, cout, cin, R, R, sc.data(), sh.data());
  };
  // stem
  e->stem_left = 3; e->stem_ws = 232;
  bf16* xs = nullptr;
  IDN_TRY(dev_alloc(e, &xs, (size_t)cap * IDN_CROP * e->stem_ws * 4 + 4096));
  e->xs = xs;
  { Step st; st.kind = STEP_PRE; st.name = "preprocess"; std::memset(&st.cp, 0, sizeof(st.cp)); e->steps.push_back(st); }
  ConvWeights w_stem;
  IDN_TRY(conv_bn(&w_stem, 64, 3, 7, true));
  Act s1; IDN_TRY(alloc_act(e, &s1, make_geom(112, 112, 0), 64, cap));
  IDN_TRY(add_stem(e, "conv1+bn1+relu", xs, e->stem_ws, e->stem_left, w_stem, 7, 2, 3, 112, 112, s1, cap));
  const int chans[4] = {64, 128, 256, 512};
  const int sizes[4] = {56, 28, 14, 7};
  Act a, t, o;       // block input, mid, block output (same geometry)
  IDN_TRY(alloc_act(e, &a, make_geom(56, 56, 1), 64, cap));
  add_pool(e, "maxpool", s1, a, 1);
  e->fuse_stem_step = (int)e->steps.size() - 2; e->fuse_pool_step = (int)e->steps.size() - 1;
  e->pool_buf = a.ptr; e->pool_img_rows = a.g.img_rows(); e->pool_wp = a.g.Wp;
  e->pool_img_bytes = (size_t)a.g.img_rows() * 64 * sizeof(bf16);
  Act planes;        // parity planes feeding the next stage
  for (int li = 0; li < 4; ++li) {
    const int C = chans[li];
    const Geom g = make_geom(sizes[li], sizes[li], 1);
    IDN_TRY(alloc_act(e, &t, g, C, cap));
    IDN_TRY(alloc_act(e, &o, g, C, cap));
    Act nextp;
    const bool last = li == 3;
    if (!last) IDN_TRY(alloc_act(e, &nextp, make_geom(sizes[li + 1], sizes[li + 1], 1), C, cap, 4));
    const std::string L = "layer" + std::to_string(li + 1);
    ConvSpec c3r; c3r.R = 3; c3r.S = 3; c3r.pad = 1; c3r.relu = true;
    // ---- block 0
    ConvWeights w1, w2, wd;
    if (li == 0) {
      IDN_TRY(conv_bn(&w1, C, C, 3, false));
      IDN_TRY(add_conv(e, L + ".0.conv1", a, w1, c3r, t, nullptr, cap));
      IDN_TRY(conv_bn(&w2, C, C, 3, false));
      IDN_TRY(add_conv(e, L + ".0.conv2+add", t, w2, c3r, o, &a, cap));
      e->pool_last_reader = (int)e->steps.size() - 1;
    } else {
      Act d; IDN_TRY(alloc_act(e, &d, g, C, cap));
      if (last) IDN_TRY(alloc_act(e, &a, g, C, cap));
      ConvSpec s2 = c3r; s2.stride = 2;
      IDN_TRY(conv_bn(&w1, C, chans[li - 1], 3, false));
      IDN_TRY(add_conv(e, L + ".0.conv1", planes, w1, s2, t, nullptr, cap));
      IDN_TRY(conv_bn(&w2, C, C, 3, false));
      ConvSpec ds; ds.R = 1; ds.S = 1; ds.stride = 2; ds.pad = 0; ds.relu = false;
      IDN_TRY(conv_bn(&wd, C, chans[li - 1], 1, false));
      IDN_TRY(add_conv(e, L + ".0.downsample", planes, wd, ds, d, nullptr, cap));
      IDN_TRY(try_merge_conv_steps(e, (int)e->steps.size() - 2, (int)e->steps.size() - 1, wd));
      IDN_TRY(add_conv(e, L + ".0.conv2+add", t, w2, c3r, o, &d, cap));
    }
    // ---- block 1: input o, output -> parity planes of the next stage (or `a` for the last)
    ConvWeights w3, w4;
    IDN_TRY(conv_bn(&w3, C, C, 3, false));
    IDN_TRY(add_conv(e, L + ".1.conv1", o, w3, c3r, t, nullptr, cap));
    IDN_TRY(conv_bn(&w4, C, C, 3, false));
    if (!last) {
      IDN_TRY(add_conv(e, L + ".1.conv2+add", t, w4, c3r, nextp, &o, cap));
      planes = nextp;
    } else {
      IDN_TRY(add_conv(e, L + ".1.conv2+add", t, w4, c3r, a, &o, cap));
      e->head_act = a.ptr; e->head_g = g; e->head_C = C;
    }
  }
  const float* fw = br.take((size_t)1000 * 512);
  const float* fb = br.take(1000);
  if (!fw || !fb || br.left != 0) { set_error("resnet18: weight blob has the wrong size"); return IDN_EWEIGHTS; }
  IDN_TRY(dev_alloc(e, &e->fc_w, (size_t)1000 * 512, false));
  IDN_TRY(dev_alloc(e, &e->fc_b, 1000, false));
  IDN_TRY(dev_alloc(e, &e->feats, (size_t)cap * 512));
  IDN_CUDA_TRY(cudaMemcpy(e->fc_w, fw, (size_t)1000 * 512 * 4, cudaMemcpyHostToDevice));
  IDN_CUDA_TRY(cudaMemcpy(e->fc_b, fb, 4000, cudaMemcpyHostToDevice));
  {
    // transposed, one contiguous [512][128] slice per cluster rank of the fused head (classes 125 r .. 125 r + 124)
    std::vector<float> wt((size_t)8 * 512 * 128, 0.0f);
    for (int c = 0; c < 1000; ++c)
      for (int k = 0; k < 512; ++k) wt[((size_t)(c / 125) * 512 + k) * 128 + (c % 125)] = fw[(size_t)c * 512 + k];
    IDN_TRY(dev_alloc(e, &e->fc_wt, wt.size(), false));
    IDN_CUDA_TRY(cudaMemcpy(e->fc_wt, wt.data(), wt.size() * 4, cudaMemcpyHostToDevice));
  }
  { Step st; st.kind = STEP_RHEAD; st.name = "avgpool+fc+softmax+top1"; std::memset(&st.cp, 0, sizeof(st.cp)); e->steps.push_back(st); }
  return IDN_OK;
}

static int build_alexnet(idn_engine* e, const float* blob, size_t count) {
  const int cap = e->cap;
  BlobReader br{blob, count};
  // The 11x11 / stride-4 / pad-2 stem (tv:models/alexnet.py:22) runs as a 3x3 stride-1 convolution
  // over the 4x4 space-to-depth image the preprocess kernel writes (57 x 57 pixels of 48+16
  // channels, block origin at crop pixel (-2, -2) so the stem's padding is part of the layout):
  // input row 4*oh - 2 + r with r = 4a + dy is s2d row oh + a, channel (dy*4+dx)*3 + c.
  e->pre_mode = PRE_OUT_S2D4;
  Act xs;
  { Geom g; g.H = 55; g.W = 55; g.Hp = 57; g.Wp = 57; IDN_TRY(alloc_act(e, &xs, g, 64, cap)); }
  e->xs = xs.ptr;
  { Step st; st.kind = STEP_PRE; st.name = "preprocess"; std::memset(&st.cp, 0, sizeof(st.cp)); e->steps.push_back(st); }
  auto take2 = [&](size_t nw, size_t nb, const float** w, const float** b) -> int {
    *w = br.take(nw); *b = br.take(nb);
    if (!*w || !*b) { set_error("alexnet: weight blob too short"); return IDN_EWEIGHTS; }
    return IDN_OK;
  };
  const float *w, *b;
  ConvWeights w0;
  IDN_TRY(take2((size_t)64 * 3 * 121, 64, &w, &b));
  {
    std::vector<float> w2((size_t)64 * 64 * 9, 0.0f);
    for (int n = 0; n < 64; ++n)
      for (int c = 0; c < 3; ++c)
        for (int r = 0; r < 11; ++r)
          for (int s = 0; s < 11; ++s) {
            const int a = r / 4, dy = r % 4, bb = s / 4, dx = s % 4, ch = (dy * 4 + dx) * 3 + c;
            w2[(((size_t)n * 64 + ch) * 3 + a) * 3 + bb] = w[(((size_t)n * 3 + c) * 11 + r) * 11 + s];
          }
    IDN_TRY(upload_conv_weights(e, &w0, w2.data(), 64, 64, 3, 3, nullptr, b));
  }
  Act c1; IDN_TRY(alloc_act(e, &c1, make_geom(55, 55, 0), 64, cap));
  { ConvSpec s0; s0.R = 3; s0.S = 3; s0.pad = 0; s0.relu = true;
    IDN_TRY(add_conv(e, "features.0+relu", xs, w0, s0, c1, nullptr, cap)); }
  Act p1; IDN_TRY(alloc_act(e, &p1, make_geom(27, 27, 2), 64, cap));
  add_pool(e, "features.2 maxpool", c1, p1, 0);
  // features.2 runs in features.0's epilogue (conv_halo_kernel<64, 2, true, 9, true>); the stand-alone pool is the
  // cross-check (IDN_NO_POOL_FUSION=1) and serves activation taps of features.0
  e->fuse_stem_step = (int)e->steps.size() - 2; e->fuse_pool_step = (int)e->steps.size() - 1;
  e->pool_buf = p1.ptr; e->pool_img_rows = p1.g.img_rows(); e->pool_wp = p1.g.Wp;
  e->pool_oh = 27; e->pool_ow = 27;
  // features.3: 5x5 p2, 64 -> 192
  ConvWeights w3;
  IDN_TRY(take2((size_t)192 * 64 * 25, 192, &w, &b));
  IDN_TRY(upload_conv_weights(e, &w3, w, 192, 64, 5, 5, nullptr, b));
  Act c2; IDN_TRY(alloc_act(e, &c2, make_geom(27, 27, 2), 192, cap));
  ConvSpec s5; s5.R = 5; s5.S = 5; s5.pad = 2; s5.relu = true;
  IDN_TRY(add_conv(e, "features.3+relu", p1, w3, s5, c2, nullptr, cap));
  Act p2; IDN_TRY(alloc_act(e, &p2, make_geom(13, 13, 1), 192, cap));
  add_pool(e, "features.5 maxpool", c2, p2, 0);
  // (features.5 inside features.3 is correct but does not pay: the 192-wide layer's epilogue has no slack for the
  //  hand-over and the pooling, 0.1538 ms fused against 0.1235 + 0.0313 -- it stays behind IDN_FUSE_FEATURES5=1)
  { const char* v = std::getenv("IDN_FUSE_FEATURES5");
    if (v && v[0] == '1')
      e->fuse_pairs.push_back({(int)e->steps.size() - 2, (int)e->steps.size() - 1, p2.ptr, p2.g.img_rows(), p2.g.Wp, p2.g.H, p2.g.W}); }
  ConvSpec s3; s3.R = 3; s3.S = 3; s3.pad = 1; s3.relu = true;
  ConvWeights w6, w8, w10;
  IDN_TRY(take2((size_t)384 * 192 * 9, 384, &w, &b));
  IDN_TRY(upload_conv_weights(e, &w6, w, 384, 192, 3, 3, nullptr, b));
  Act c3; IDN_TRY(alloc_act(e, &c3, make_geom(13, 13, 1), 384, cap));
  IDN_TRY(add_conv(e, "features.6+relu", p2, w6, s3, c3, nullptr, cap));
  IDN_TRY(take2((size_t)256 * 384 * 9, 256, &w, &b));
  IDN_TRY(upload_conv_weights(e, &w8, w, 256, 384, 3, 3, nullptr, b));
  Act c4; IDN_TRY(alloc_act(e, &c4, make_geom(13, 13, 1), 256, cap));
  IDN_TRY(add_conv(e, "features.8+relu", c3, w8, s3, c4, nullptr, cap));
  IDN_TRY(take2((size_t)256 * 256 * 9, 256, &w, &b));
  IDN_TRY(upload_conv_weights(e, &w10, w, 256, 256, 3, 3, nullptr, b));
  Act c5; IDN_TRY(alloc_act(e, &c5, make_geom(13, 13, 1), 256, cap));
  IDN_TRY(add_conv(e, "features.10+relu", c4, w10, s3, c5, nullptr, cap));
  Act p3; IDN_TRY(alloc_act(e, &p3, make_geom(6, 6, 0), 256, cap));
  add_pool(e, "features.12 maxpool", c5, p3, 0);
  e->fuse_pairs.push_back({(int)e->steps.size() - 2, (int)e->steps.size() - 1, p3.ptr, p3.g.img_rows(), p3.g.Wp, p3.g.H, p3.g.W});
  IDN_TRY(dev_alloc(e, &e->pool_ring, (size_t)(e->num_sms / 2 + 1) * 1024 * 256, false));
  // classifier: the flatten order of torch.flatten(x, 1) on [B,256,6,6] is (c, h, w)
  // (tv:models/alexnet.py:50); activations here are (h, w, c), so FC6's columns are permuted once.
  IDN_TRY(take2((size_t)4096 * 9216, 4096, &w, &b));
  ConvWeights f6, f7, f8;
  {
    std::vector<float> wp((size_t)4096 * 9216);
    for (int n = 0; n < 4096; ++n)
      for (int c = 0; c < 256; ++c)
        for (int hw = 0; hw < 36; ++hw)
          wp[(size_t)n * 9216 + (size_t)hw * 256 + c] = w[(size_t)n * 9216 + (size_t)c * 36 + hw];
    IDN_TRY(upload_conv_weights(e, &f6, wp.data(), 4096, 9216, 1, 1, nullptr, b));
  }
  Act x6; x6.ptr = p3.ptr; x6.g = make_geom(1, 1, 0); x6.C = 9216; x6.planes = 1; x6.plane_rows = cap;
  Act y6; IDN_TRY(alloc_act(e, &y6, make_geom(1, 1, 0), 4096, cap));
  ConvSpec fc; fc.relu = true;
  IDN_TRY(add_conv(e, "classifier.1+relu", x6, f6, fc, y6, nullptr, cap));
  IDN_TRY(take2((size_t)4096 * 4096, 4096, &w, &b));
  IDN_TRY(upload_conv_weights(e, &f7, w, 4096, 4096, 1, 1, nullptr, b));
  Act y7; IDN_TRY(alloc_act(e, &y7, make_geom(1, 1, 0), 4096, cap));
  IDN_TRY(add_conv(e, "classifier.4+relu", y6, f7, fc, y7, nullptr, cap));
  IDN_TRY(take2((size_t)1000 * 4096, 1000, &w, &b));
  if (br.left != 0) { set_error("alexnet: weight blob has the wrong size"); return IDN_EWEIGHTS; }
  IDN_TRY(upload_conv_weights(e, &f8, w, 1000, 4096, 1, 1, nullptr, b));   // n_total = 1024
  IDN_TRY(dev_alloc(e, &e->logits_pad, (size_t)cap * 1024));
  Act y8; y8.ptr = reinterpret_cast<bf16*>(e->logits_pad); y8.g = make_geom(1, 1, 0); y8.C = 1024;
  y8.planes = 1; y8.plane_rows = cap;
  ConvSpec fc8; fc8.relu = false; fc8.out_f32 = true;
  IDN_TRY(add_conv(e, "classifier.6", y7, f8, fc8, y8, nullptr, cap));
  { Step st; st.kind = STEP_SOFTMAX; st.name = "softmax+top1"; std::memset(&st.cp, 0, sizeof(st.cp)); e->steps.push_back(st); }
  return IDN_OK;
}

// ------------------------------------------------------------------------------- preprocess io
// Resample tables live in one device pool with a page-locked mirror on the host.  A table that a batch
// needs for the first time is written into the mirror and copied to the pool by copy_words_kernel on the
// stream that carries the batch's records (`copy_on`), so a previously unseen image size costs one small
// kernel and NO synchronisation: earlier batches keep running (a cudaMemcpyAsync from pageable memory
// would first drain the stream, once per new size -- on a folder of mixed-size JPEGs that used to collapse
// the submit/collect overlap until the cache was warm).  Every reader of the pool runs on the handle's
// stream, which is either `copy_on` itself or waits for an event recorded on it after the upload.
// The pool holds ~180 distinct (source size, resized size, crop) axes before it has to grow; growing is the
// one slow path (device-wide synchronise, reallocate, copy).
constexpr size_t kTablePoolInts = 1u << 20;

static int ensure_table(idn_engine* e, int in_size, int out_size, int crop, int* off, cudaStream_t copy_on) {
  auto key = std::make_tuple(in_size, out_size, crop);
  auto it = e->table_index.find(key);
  if (it != e->table_index.end()) { *off = it->second; return IDN_OK; }
  std::string bytes;
  build_axis_table(in_size, out_size, crop, &bytes);
  const size_t n = bytes.size() / 4;
  if (e->tables_used + n > e->tables_cap) {
    const size_t ncap = std::max({e->tables_cap * 2, e->tables_used + n + (1u << 16), kTablePoolInts});
    int* nt = nullptr;
    int* nh = nullptr;
    IDN_CUDA_TRY(cudaMalloc(&nt, ncap * 4));
    if (cudaMallocHost(&nh, ncap * 4) != cudaSuccess) { cudaFree(nt); set_error("table pool: cudaMallocHost failed"); return IDN_ENOMEM; }
    if (e->d_tables) {
      IDN_CUDA_TRY(cudaDeviceSynchronize());            // kernels in flight may be reading the old pool
      IDN_CUDA_TRY(cudaMemcpy(nt, e->d_tables, e->tables_used * 4, cudaMemcpyDeviceToDevice));
      std::memcpy(nh, e->h_tables, e->tables_used * 4);
      cudaFree(e->d_tables);
      cudaFreeHost(e->h_tables);
    }
    e->d_tables = nt; e->h_tables = nh; e->tables_cap = ncap;
  }
  std::memcpy(e->h_tables + e->tables_used, bytes.data(), bytes.size());
  IDN_TRY(launch_copy_words(e->d_tables + e->tables_used, e->h_tables + e->tables_used, n, copy_on));
  e->launches++;
  *off = (int)e->tables_used;
  e->table_index[key] = *off;
  e->tables_used += n;
  return IDN_OK;
}

static int fill_meta(idn_engine* e, const int64_t* offsets, const int32_t* hw, int n, int64_t base,
                     PreMeta* hm = nullptr, PreMeta* d_dst = nullptr, cudaStream_t copy_on = nullptr) {
  if (d_dst == nullptr) { d_dst = e->d_meta; copy_on = e->stream; }
  int ring = -1;
  if (hm == nullptr) {
    // The records are read by a kernel when the stream gets there, not now: rotate through a few
    // buffers and wait (normally not at all) until the previous reader of this one has run.
    ring = e->meta_next;
    e->meta_next = (e->meta_next + 1) % e->meta_ring;
    if (e->meta_ev[ring]) IDN_CUDA_TRY(cudaEventSynchronize(e->meta_ev[ring]));
    hm = e->h_meta + (size_t)ring * (e->meta_ring > 1 ? e->cap : 0);
  }
  for (int i = 0; i < n; ++i) {
    const int H = hw[2 * i], W = hw[2 * i + 1];
    if (H <= 0 || W <= 0) { set_error("image with non-positive size"); return IDN_EINVAL; }
    int nh, nw;
    resized_size(H, W, &nh, &nw);
    PreMeta& m = hm[i];
    m.offset = offsets[i] - base; m.H = H; m.W = W;
    IDN_TRY(ensure_table(e, W, nw, crop_offset(nw), &m.htab, copy_on));
    IDN_TRY(ensure_table(e, H, nh, crop_offset(nh), &m.vtab, copy_on));
  }
  static_assert(sizeof(PreMeta) % 4 == 0, "PreMeta is copied in 32-bit words");
  IDN_TRY(launch_copy_words(d_dst, hm, sizeof(PreMeta) / 4 * (size_t)n, copy_on));
  e->launches++;
  if (ring >= 0) {
    if (!e->meta_ev[ring]) IDN_CUDA_TRY(cudaEventCreateWithFlags(&e->meta_ev[ring], cudaEventDisableTiming));
    IDN_CUDA_TRY(cudaEventRecord(e->meta_ev[ring], copy_on));
  }
  return IDN_OK;
}

// --------------------------------------------------------------------------------- forward pass
static void prepare_conv_params(idn_engine* e, ConvParams& cp, int n) {
  cp.nimg = n;
  cp.profile = (e->timing && e->wait_profile) ? 1 : 0;
  cp.fd_img_rows = make_fastdiv((unsigned)std::max(cp.img_rows, 1));
  cp.fd_wp = make_fastdiv((unsigned)std::max(cp.wp, 1));
  cp.fd_per_img = make_fastdiv((unsigned)std::max(cp.tiles_w * cp.tiles_h, 1));
  cp.fd_tiles_w = make_fastdiv((unsigned)std::max(cp.tiles_w, 1));
  const long long rows = (long long)n * cp.img_rows;
  if (cp.mode == MODE_STEM) cp.num_m_tiles = n * cp.tiles_w * cp.tiles_h;
  else cp.num_m_tiles = (int)((rows + kBlockM - 1) / kBlockM);
}

// Whether the 1x1 downsample steps ride in the launches of their 3x3 stride-2 partners in this pass.
static bool ds_fusion_active(const idn_engine* e, int impl) { return impl == IDN_IMPL_TC && !e->no_ds_fusion; }

static int launch_conv_step(idn_engine* e, Step& s, int n, int impl, cudaStream_t st) {
  ConvParams& cp = s.cp;
  prepare_conv_params(e, cp, n);
  const long long rows = (long long)n * cp.img_rows;
  if (impl != IDN_IMPL_TC) return launch_conv_simt(cp, st);
  if (cp.mode == MODE_STEM) return e->stem_simt ? launch_conv_simt(cp, st) : launch_stem_tc(s.tmA, s.tmA2, s.tmB, cp, e->num_sms, st);
  if (s.use_halo) {
    const int item_rows = kBlockM * cp.msub;
    cp.num_m_tiles = (int)((rows + item_rows - 1) / item_rows);
    return launch_conv_halo(s.tmA, s.tmA2, s.tmB, cp, s.block_n, s.b_resident, e->num_sms, st);
  }
  if (cp.split_k > 1) { cp.splitk_ws = e->splitk_ws; cp.splitk_cnt = e->splitk_cnt; }
  if (s.second >= 0 && ds_fusion_active(e, impl)) {
    Step& t = e->steps[s.second];
    prepare_conv_params(e, t.cp, n);
    const ConvSecond sec{&t.tmA, &t.tmB, &t.cp};
    return launch_conv_tc(s.tmA, s.tmB, cp, s.block_n, e->num_sms, st, &sec);
  }
  return launch_conv_tc(s.tmA, s.tmB, cp, s.block_n, e->num_sms, st);
}

// Lets convolution step `b` (the block's 1x1 / stride-2 downsample) ride in the launch of step `a` (its 3x3 /
// stride-2 partner): both read the same parity planes and produce the same output geometry, so they share the
// M x N tiling; b adopts a's N tile and kernel flavour (its weight boxes are re-made to match).
static int try_merge_conv_steps(idn_engine* e, int a, int b, const ConvWeights& wb) {
  Step& sa = e->steps[a];
  Step& sb = e->steps[b];
  if (sa.use_halo || sb.use_halo || sa.cp.mode != MODE_FLAT || sb.cp.mode != MODE_FLAT) return IDN_OK;
  if (sa.cp.img_rows != sb.cp.img_rows || sa.cp.wp != sb.cp.wp || sa.cp.n_total != sb.cp.n_total) return IDN_OK;
  sb.block_n = sa.block_n;
  sb.cp.tc_pair = sa.cp.tc_pair;
  sb.cp.num_n_tiles = sa.cp.num_n_tiles;
  IDN_TRY(make_tmap_2d(&sb.tmB, wb.wt, (uint64_t)wb.ktot, (uint64_t)wb.n_total, (uint64_t)wb.ktot * 2, kBlockK,
                       (uint32_t)(sb.cp.tc_pair ? sb.block_n / 2 : sb.block_n)));
  sa.second = b;
  sb.merged_into = a;
  return IDN_OK;
}

static int run_plan(idn_engine* e, const uint8_t* d_imgs, int n, int32_t* d_idx, float* d_prob,
                    float* d_logits, const PreMeta* d_meta = nullptr) {
  if (d_meta == nullptr) d_meta = e->d_meta;
  cudaStream_t st = e->stream;
  e->last_n = n;
  const bool fuse = e->fuse_stem_step >= 0 && e->impl == IDN_IMPL_TC && !e->stem_simt && !e->no_pool_fusion &&
                    e->stop_after != e->fuse_stem_step &&    // a tap of the convolution itself needs its own output
                    (e->steps[e->fuse_stem_step].cp.mode == MODE_STEM || e->steps[e->fuse_stem_step].use_halo);
  // the conv -> pool pairs of the CTA-pair kernel that run fused in this pass
  auto pair_fused = [&](const idn_engine::FusePair& fp) {
    const Step& c = e->steps[fp.conv_step];
    return e->impl == IDN_IMPL_TC && !e->no_pool_fusion && e->stop_after != fp.conv_step && c.use_halo && c.cp.pair &&
           !c.b_resident && c.cp.num_n_tiles == 1 && ((c.block_n == 192 && c.cp.b_group == 5) || (c.block_n == 256 && c.cp.b_group == 3));
  };
  const Step* deferred_fc = nullptr;
  int step_no = 0;
  for (Step& s : e->steps) {
    if (e->stop_after >= 0 && step_no > e->stop_after) break;
    const int step_idx = step_no;
    ++step_no;
    if (e->timing) {
      if (!s.ev0) { IDN_CUDA_TRY(cudaEventCreate(&s.ev0)); IDN_CUDA_TRY(cudaEventCreate(&s.ev1)); }
      IDN_CUDA_TRY(cudaEventRecord(s.ev0, st));
    }
    switch (s.kind) {
      case STEP_PRE:
        IDN_TRY(launch_preprocess(d_imgs, d_meta, e->d_tables, e->d_lut, n, e->pre_mode, e->xs,
                                  e->stem_left, e->stem_ws, st));
        break;
      case STEP_CONV:
        if (step_idx == e->fuse_stem_step) {
          s.cp.pool_out = fuse ? e->pool_buf : nullptr;
          s.cp.pool_img_rows = e->pool_img_rows; s.cp.pool_wp = e->pool_wp;
          if (e->pool_oh > 0) { s.cp.pool_oh = e->pool_oh; s.cp.pool_ow = e->pool_ow; }
        }
        for (const auto& fp : e->fuse_pairs)
          if (fp.conv_step == step_idx) {
            const bool on = pair_fused(fp);
            s.cp.pool_out = on ? fp.buf : nullptr;
            s.cp.pool_img_rows = fp.img_rows; s.cp.pool_wp = fp.wp; s.cp.pool_oh = fp.oh; s.cp.pool_ow = fp.ow;
            s.cp.pool_ring = e->pool_ring;
          }
        // AlexNet's FC8 (split-K, straight before the softmax step): its partial tiles are added by the softmax kernel
        // instead of by waiting CTAs -- unless this very step's output is wanted (activation tap) or IDN_NO_SPLIT_DEFER=1
        s.cp.split_defer = 0;
        if (s.cp.split_k > 1 && e->impl == IDN_IMPL_TC && step_idx + 1 < (int)e->steps.size() &&
            e->steps[step_idx + 1].kind == STEP_SOFTMAX && e->stop_after != step_idx && !e->no_split_defer) {
          s.cp.split_defer = 1;
          deferred_fc = &s;
        }
        if (s.merged_into >= 0 && ds_fusion_active(e, e->impl)) { e->launches -= 1; break; }   // rode in that step's launch
        IDN_TRY(launch_conv_step(e, s, n, e->impl, st));
        break;
      case STEP_POOL:
        if (fuse && step_idx == e->fuse_pool_step) { e->launches -= 1; break; }   // done by the stem's epilogue
        {
          bool done = false;
          for (const auto& fp : e->fuse_pairs) done = done || (fp.pool_step == step_idx && pair_fused(fp));
          if (done) { e->launches -= 1; break; }                                  // done by the convolution before it
        }
        s.pp.nimg = n;
        IDN_TRY(launch_maxpool(s.pp, st));
        break;
      case STEP_RHEAD:
        if (!e->no_head_fusion) {
          IDN_TRY(launch_resnet_head_fused(e->head_act, n, e->head_g.img_rows(), e->head_g.Wp, e->head_g.H,
                                           e->head_g.W, e->head_C, e->fc_wt, e->fc_b, e->feats, d_logits, d_idx, d_prob, st));
          break;
        }
        IDN_TRY(launch_resnet_head(e->head_act, n, e->head_g.img_rows(), e->head_g.Wp, e->head_g.H,
                                   e->head_g.W, e->head_C, e->fc_w, e->fc_b, e->feats,
                                   d_logits ? d_logits : e->d_logits, d_idx, d_prob, st));
        e->launches += 2;                       // avgpool, fc, softmax/top-1
        break;
      case STEP_SOFTMAX:
        if (deferred_fc != nullptr) {
          const ConvParams& f = deferred_fc->cp;
          IDN_TRY(launch_softmax_top1_splitk(e->splitk_ws, f.split_k, f.num_n_tiles, f.scale, f.shift, n, d_logits, d_idx, d_prob, st));
          break;
        }
        IDN_TRY(launch_softmax_top1(e->logits_pad, 1024, n, d_logits, d_idx, d_prob, st));
        break;
    }
    e->launches += 1;
    if (e->timing) {
      IDN_CUDA_TRY(cudaEventRecord(s.ev1, st));
      if (s.kind == STEP_CONV && e->wait_profile) {   // per-role wait cycles of this kernel alone
        IDN_CUDA_TRY(cudaStreamSynchronize(st));
        IDN_TRY(read_tc_wait_cycles(s.wait_cycles));
      }
    }
  }
  return IDN_OK;
}

static int check_device(int device) {
  int count = 0;
  cudaError_t err = cudaGetDeviceCount(&count);
  if (err != cudaSuccess || count <= 0) {
    set_error(std::string("no CUDA device: ") + cudaGetErrorString(err) +
              " -- idunno_b200 has no CPU fallback");
    return IDN_ENODEV;
  }
  if (device < 0 || device >= count) { set_error("device index out of range"); return IDN_EINVAL; }
  cudaDeviceProp prop;
  IDN_CUDA_TRY(cudaGetDeviceProperties(&prop, device));
  if (prop.major != 10) {
    set_error(std::string("device ") + prop.name + " is sm_" + std::to_string(prop.major) +
              std::to_string(prop.minor) + "; this library is built for sm_100a only");
    return IDN_ENODEV;
  }
  IDN_CUDA_TRY(cudaSetDevice(device));
  return IDN_OK;
}

}  // namespace idn

// ========================================================================================= C ABI
extern "C" {

const char* idn_last_error(void) { return idn::last_error_cstr(); }
const char* idn_version(void) { return "idunno_b200 0.1 sm_100a"; }

int idn_host_alloc(int device, size_t n_bytes, void** out) {
  if (!out || n_bytes == 0) { set_error("idn_host_alloc: bad argument"); return IDN_EINVAL; }
  *out = nullptr;
  if (int rc = check_device(device)) return rc;
  if (cudaSetDevice(device) != cudaSuccess) { set_error("idn_host_alloc: cudaSetDevice failed"); return IDN_ECUDA; }
  cudaError_t err = cudaHostAlloc(out, n_bytes, cudaHostAllocPortable);
  if (err != cudaSuccess) {
    (void)cudaGetLastError();
    set_error(std::string("idn_host_alloc: ") + cudaGetErrorString(err));
    *out = nullptr;
    return IDN_ENOMEM;
  }
  return IDN_OK;
}

int idn_host_free(void* p) {
  if (!p) return IDN_OK;
  if (cudaFreeHost(p) != cudaSuccess) { (void)cudaGetLastError(); set_error("idn_host_free: not a pointer from idn_host_alloc"); return IDN_EINVAL; }
  return IDN_OK;
}

int64_t idn_weight_count(int model_id) {
  if (model_id == IDN_MODEL_ALEXNET) return 61100840LL;
  if (model_id == IDN_MODEL_RESNET18) return 11699112LL;
  return -1;
}

int idn_destroy(idn_handle h) {
  if (!h) return IDN_OK;
  cudaSetDevice(h->device);
  if (h->stream) cudaStreamSynchronize(h->stream);
  for (void* p : h->dev_allocs) cudaFree(p);
  if (h->d_tables) cudaFree(h->d_tables);
  if (h->h_tables) cudaFreeHost(h->h_tables);
  for (int b = 0; b < kSlots; ++b) {
    idn_engine::Slot& sl = h->slot[b];
    if (sl.d_stage) cudaFree(sl.d_stage);
    if (sl.d_logits) cudaFree(sl.d_logits);
    if (sl.d_meta) cudaFree(sl.d_meta);
    if (sl.h_meta) cudaFreeHost(sl.h_meta);
    if (sl.h_idx) cudaFreeHost(sl.h_idx);
    if (sl.h_prob) cudaFreeHost(sl.h_prob);
    if (sl.ev_copied) cudaEventDestroy(sl.ev_copied);
    if (sl.ev_done) cudaEventDestroy(sl.ev_done);
  }
  if (h->copy_stream) cudaStreamDestroy(h->copy_stream);
  if (h->h_meta) cudaFreeHost(h->h_meta);
  for (cudaEvent_t& ev : h->meta_ev) if (ev) { cudaEventDestroy(ev); ev = nullptr; }
  for (Step& s : h->steps) { if (s.ev0) cudaEventDestroy(s.ev0); if (s.ev1) cudaEventDestroy(s.ev1); }
  if (h->own_stream) cudaStreamDestroy(h->own_stream);
  delete h;
  return IDN_OK;
}

int idn_create(int device, int model_id, const void* weights_blob, size_t n_bytes, int max_batch,
               idn_handle* out) {
  if (!out || !weights_blob || max_batch <= 0 || max_batch > 8192 ||
      (model_id != IDN_MODEL_ALEXNET && model_id != IDN_MODEL_RESNET18)) {
    set_error("idn_create: bad argument");
    return IDN_EINVAL;
  }
  *out = nullptr;
  if ((int64_t)n_bytes != idn_weight_count(model_id) * 4) {
    set_error("idn_create: weight blob is " + std::to_string(n_bytes) + " bytes, expected " +
              std::to_string(idn_weight_count(model_id) * 4));
    return IDN_EWEIGHTS;
  }
  IDN_TRY(check_device(device));
  std::unique_ptr<idn_engine> e(new idn_engine());
  e->device = device; e->model = model_id; e->cap = max_batch;
  { const char* v = std::getenv("IDN_STEM_SIMT"); e->stem_simt = v && v[0] == '1'; }
  { const char* v = std::getenv("IDN_NO_POOL_FUSION"); e->no_pool_fusion = v && v[0] == '1'; }
  { const char* v = std::getenv("IDN_FULL_H2D"); e->full_h2d = v && v[0] == '1'; }
  { const char* v = std::getenv("IDN_NO_SPLIT_DEFER"); e->no_split_defer = v && v[0] == '1'; }
  { const char* v = std::getenv("IDN_NO_HEAD_FUSION"); e->no_head_fusion = v && v[0] == '1'; }
  { const char* v = std::getenv("IDN_NO_DS_FUSION"); e->no_ds_fusion = v && v[0] == '1'; }
  int rc = IDN_OK;
  auto fail = [&](int code) { idn_engine* raw = e.release(); idn_destroy(raw); return code; };
  {
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) == cudaSuccess) e->num_sms = prop.multiProcessorCount;
  }
  if (cudaStreamCreateWithFlags(&e->own_stream, cudaStreamNonBlocking) != cudaSuccess) {
    set_error("cudaStreamCreate failed"); return fail(IDN_ECUDA);
  }
  e->stream = e->own_stream;
  if ((rc = conv_tc_configure()) != IDN_OK) return fail(rc);
  if ((rc = dev_alloc(e.get(), &e->d_meta, (size_t)max_batch)) != IDN_OK) return fail(rc);
  e->meta_ring = 4;
  if (cudaMallocHost(&e->h_meta, sizeof(PreMeta) * max_batch * e->meta_ring) != cudaSuccess) {
    set_error("cudaMallocHost failed"); return fail(IDN_ENOMEM);
  }
  uint16_t lut[768];
  build_norm_lut(lut);
  if ((rc = dev_alloc(e.get(), &e->d_lut, 768)) != IDN_OK) return fail(rc);
  if (cudaMemcpy(e->d_lut, lut, sizeof(lut), cudaMemcpyHostToDevice) != cudaSuccess) {
    set_error("lut upload failed"); return fail(IDN_ECUDA);
  }
  if ((rc = dev_alloc(e.get(), &e->d_logits, (size_t)max_batch * IDN_NUM_CLASSES)) != IDN_OK) return fail(rc);
  const float* blob = reinterpret_cast<const float*>(weights_blob);
  rc = model_id == IDN_MODEL_ALEXNET ? build_alexnet(e.get(), blob, n_bytes / 4)
                                     : build_resnet18(e.get(), blob, n_bytes / 4);
  if (rc != IDN_OK) return fail(rc);
  if (cudaDeviceSynchronize() != cudaSuccess) { set_error("sync after build failed"); return fail(IDN_ECUDA); }
  *out = e.release();
  return IDN_OK;
}

int idn_set_stream(idn_handle h, void* cuda_stream) {
  if (!h) return IDN_EINVAL;
  h->stream = cuda_stream ? reinterpret_cast<cudaStream_t>(cuda_stream) : h->own_stream;
  return IDN_OK;
}

int idn_set_impl(idn_handle h, int impl) {
  if (!h || (impl != IDN_IMPL_TC && impl != IDN_IMPL_SIMT)) { set_error("bad impl"); return IDN_EINVAL; }
  h->impl = impl;
  return IDN_OK;
}

int idn_sync(idn_handle h) {
  if (!h) return IDN_EINVAL;
  IDN_CUDA_TRY(cudaSetDevice(h->device));
  IDN_CUDA_TRY(cudaStreamSynchronize(h->stream));
  return read_tc_watchdog();
}

int idn_abort(idn_handle h) {
  if (!h) return IDN_EINVAL;
  IDN_CUDA_TRY(cudaSetDevice(h->device));
  cudaError_t e1 = h->copy_stream ? cudaStreamSynchronize(h->copy_stream) : cudaSuccess;
  cudaError_t e2 = cudaStreamSynchronize(h->stream);
  for (int b = 0; b < kSlots; ++b) h->slot[b].busy = false;
  h->next_slot = 0;
  if (e1 != cudaSuccess || e2 != cudaSuccess) {
    set_error(std::string("idn_abort: ") + cudaGetErrorString(e1 != cudaSuccess ? e1 : e2));
    return IDN_ECUDA;
  }
  return read_tc_watchdog();
}

int idn_infer_u8hwc_device(idn_handle h, const uint8_t* d_imgs, const int64_t* offsets,
                           const int32_t* hw, int n, int32_t* d_idx, float* d_prob, float* d_logits) {
  if (!h || !d_imgs || !offsets || !hw || !d_idx || !d_prob || n < 0 || n > h->cap) {
    set_error("idn_infer_u8hwc_device: bad argument (n must be <= max_batch)");
    return IDN_EINVAL;
  }
  if (n == 0) return IDN_OK;
  IDN_CUDA_TRY(cudaSetDevice(h->device));
  IDN_TRY(fill_meta(h, offsets, hw, n, 0));
  return run_plan(h, d_imgs, n, d_idx, d_prob, d_logits);
}

// ---- asynchronous host-buffer path: submit one tensor batch, collect it later -----------------
static int slot_setup(idn_handle h, idn_engine::Slot& sl) {
  if (!h->copy_stream) IDN_CUDA_TRY(cudaStreamCreateWithFlags(&h->copy_stream, cudaStreamNonBlocking));
  if (sl.ev_done) return IDN_OK;
  IDN_CUDA_TRY(cudaEventCreateWithFlags(&sl.ev_copied, cudaEventDisableTiming));
  IDN_CUDA_TRY(cudaEventCreateWithFlags(&sl.ev_done, cudaEventDisableTiming));
  IDN_CUDA_TRY(cudaMallocHost(&sl.h_meta, sizeof(PreMeta) * (size_t)h->cap));
  IDN_CUDA_TRY(cudaMallocHost(&sl.h_idx, sizeof(int32_t) * (size_t)h->cap));
  IDN_CUDA_TRY(cudaMallocHost(&sl.h_prob, sizeof(float) * (size_t)h->cap));
  IDN_CUDA_TRY(cudaMalloc(&sl.d_logits, sizeof(float) * (size_t)h->cap * IDN_NUM_CLASSES));
  IDN_CUDA_TRY(cudaMalloc(&sl.d_meta, sizeof(PreMeta) * (size_t)h->cap));
  return IDN_OK;
}

int idn_submit_u8hwc(idn_handle h, const uint8_t* imgs, const int64_t* offsets, const int32_t* hw,
                     int n, int want_logits, int* ticket) {
  if (!h || !imgs || !offsets || !hw || !ticket || n <= 0 || n > h->cap) {
    set_error("idn_submit_u8hwc: bad argument (0 < n <= max_batch)");
    return IDN_EINVAL;
  }
  IDN_CUDA_TRY(cudaSetDevice(h->device));
  const int b = h->next_slot;
  idn_engine::Slot& sl = h->slot[b];
  if (sl.busy) { set_error("idn_submit_u8hwc: three submissions are already in flight; collect one first"); return IDN_EINVAL; }
  IDN_TRY(slot_setup(h, sl));
  int64_t lo = offsets[0], hi = 0;
  for (int i = 0; i < n; ++i) {
    lo = std::min(lo, offsets[i]);
    hi = std::max(hi, offsets[i] + (int64_t)hw[2 * i] * hw[2 * i + 1] * 3);
  }
  const size_t bytes = (size_t)(hi - lo);
  if (bytes + 64 > sl.d_stage_cap) {          // the slot is free, so nothing reads its staging buffer
    if (sl.d_stage) cudaFree(sl.d_stage);
    sl.d_stage = nullptr; sl.d_stage_cap = 0;
    IDN_CUDA_TRY(cudaMalloc(&sl.d_stage, bytes + 64));
    sl.d_stage_cap = bytes + 64;
  }
  // copy stream: the image bytes; the handle's stream: everything else, after the copy has landed
  // ... and the per-image records, so that nothing of this batch is left to fetch from the host when
  // the kernels of the previous batch finish
  // A batch of equally sized images at a constant stride (the usual case: one dataset, one packed buffer) is copied
  // as ONE pitched copy of the source rows the centre crop reads -- Resize(256) + CenterCrop(224) never looks at
  // the outer (1 - 224/256) / 2 of the shorter side: 198 of 224 rows for a 224 x 224 image, 11.6 % fewer bytes on a
  // host link that bounds the whole end-to-end rate.  The rows keep their place in the staging buffer, so nothing
  // else changes (the kernel reads at most one row past the last one it needs, inside the buffer).
  bool rows_only = !h->full_h2d && n > 1;
  const int H0 = hw[0], W0 = hw[1];
  const int64_t stride = n > 1 ? offsets[1] - offsets[0] : 0;
  for (int i = 0; i < n && rows_only; ++i)
    rows_only = hw[2 * i] == H0 && hw[2 * i + 1] == W0 && offsets[i] == offsets[0] + (int64_t)i * stride;
  rows_only = rows_only && stride >= (int64_t)H0 * W0 * 3 && H0 > 0 && W0 > 0;
  size_t copied = bytes;
  if (rows_only) {
    int nh, nw, r0, r1;
    resized_size(H0, W0, &nh, &nw);
    needed_source_range(H0, nh, crop_offset(nh), &r0, &r1);
    r1 = std::min(H0, r1 + 1);                                  // the kernel fetches one row ahead
    const size_t row_bytes = (size_t)W0 * 3, first = (size_t)r0 * row_bytes, width = (size_t)(r1 - r0) * row_bytes;
    IDN_CUDA_TRY(cudaMemcpy2DAsync(sl.d_stage + (offsets[0] - lo) + first, (size_t)stride, imgs + offsets[0] + first, (size_t)stride,
                                   width, (size_t)n, cudaMemcpyHostToDevice, h->copy_stream));
    copied = width * (size_t)n;
  } else {
    IDN_CUDA_TRY(cudaMemcpyAsync(sl.d_stage, imgs + lo, bytes, cudaMemcpyHostToDevice, h->copy_stream));
  }
  h->h2d_bytes += (int64_t)copied;
  IDN_TRY(fill_meta(h, offsets, hw, n, lo, sl.h_meta, sl.d_meta, h->copy_stream));
  IDN_CUDA_TRY(cudaEventRecord(sl.ev_copied, h->copy_stream));
  IDN_CUDA_TRY(cudaStreamWaitEvent(h->stream, sl.ev_copied, 0));
  // The softmax/top-1 kernel stores its 8 bytes per image straight into the slot's page-locked host
  // buffers (device-visible under unified addressing): no device-to-host copy to queue anywhere.
  IDN_TRY(run_plan(h, sl.d_stage, n, sl.h_idx, sl.h_prob, want_logits ? sl.d_logits : nullptr, sl.d_meta));
  IDN_CUDA_TRY(cudaEventRecord(sl.ev_done, h->stream));
  sl.n = n; sl.busy = true; sl.logits = want_logits != 0;
  *ticket = b;
  h->next_slot = (b + 1) % kSlots;
  return IDN_OK;
}

int idn_collect(idn_handle h, int ticket, int32_t* top1_idx, float* top1_prob, float* logits) {
  if (!h || ticket < 0 || ticket >= kSlots || !top1_idx || !top1_prob || !h->slot[ticket].busy) {
    set_error("idn_collect: bad ticket or argument");
    return IDN_EINVAL;
  }
  IDN_CUDA_TRY(cudaSetDevice(h->device));
  idn_engine::Slot& sl = h->slot[ticket];
  sl.busy = false;
  IDN_CUDA_TRY(cudaEventSynchronize(sl.ev_done));
  std::memcpy(top1_idx, sl.h_idx, sizeof(int32_t) * (size_t)sl.n);
  std::memcpy(top1_prob, sl.h_prob, sizeof(float) * (size_t)sl.n);
  if (logits) {
    if (!sl.logits) { set_error("idn_collect: the submission did not ask for logits"); return IDN_EINVAL; }
    IDN_CUDA_TRY(cudaMemcpy(logits, sl.d_logits, sizeof(float) * (size_t)sl.n * IDN_NUM_CLASSES, cudaMemcpyDeviceToHost));
  }
  return read_tc_watchdog();
}

int idn_infer_u8hwc(idn_handle h, const uint8_t* imgs, const int64_t* offsets, const int32_t* hw,
                    int n, int32_t* top1_idx, float* top1_prob, float* logits) {
  if (!h || !imgs || !offsets || !hw || !top1_idx || !top1_prob || n < 0) {
    set_error("idn_infer_u8hwc: bad argument");
    return IDN_EINVAL;
  }
  for (int b = 0; b < kSlots; ++b)
    if (h->slot[b].busy) { set_error("idn_infer_u8hwc: collect pending submissions first"); return IDN_EINVAL; }
  // Tensor batches with up to kSlots of them in flight: the copy of the next batch overlaps the kernels of the
  // current one.  An input longer than max_batch is cut into batches of max_batch.  An input that fits ONE
  // tensor batch -- the reference's call shape, one blocking deeplearning() per shard
  // (mp4_machinelearning.py:596) -- would copy all its bytes before the first kernel starts (60 MB = 1.1 ms
  // for 400 images); it is cut into a few sub-batches instead, so that only the first sub-batch's copy and the
  // last one's kernels are exposed.  Results do not depend on the cut (per-image arithmetic is independent of
  // the batch an image falls in; tests/test_gpu_model.py).  Measured on B200 (profiles/README.md, round 2):
  // every extra pass costs its kernels' fill and drain (~0.1 ms AlexNet, ~0.3 ms ResNet18), so few cuts win --
  // thirds for AlexNet (its pass is shorter than its copy), 10/30/60 % for ResNet18 (a small first sub-batch
  // starts the GPU early, the big ones behind it keep their tile efficiency).
  // IDN_SINGLE_CALL_CUTS="a,b,c" (per cent) overrides.
  std::vector<int> cuts;
  if (n <= h->cap && n >= 128) {
    std::vector<int> pct = h->model == IDN_MODEL_ALEXNET ? std::vector<int>{34, 33, 33} : std::vector<int>{10, 30, 60};
    if (const char* v = std::getenv("IDN_SINGLE_CALL_CUTS")) {
      std::vector<int> q;
      int sum = 0;
      for (const char* c = v; *c;) {
        char* end = nullptr;
        const long x = std::strtol(c, &end, 10);
        if (end == c || x <= 0) { q.clear(); break; }
        q.push_back((int)x); sum += (int)x;
        c = *end == ',' ? end + 1 : end;
      }
      if (!q.empty() && sum == 100) pct = q;
    }
    int done = 0, acc = 0;
    for (size_t k = 0; k < pct.size(); ++k) {
      acc += pct[k];
      const int upto = k + 1 == pct.size() ? n : (int)((long long)n * acc / 100);
      if (upto > done) { cuts.push_back(upto - done); done = upto; }
    }
  } else {
    for (int i0 = 0; i0 < n; i0 += h->cap) cuts.push_back(std::min(h->cap, n - i0));
  }
  int tickets[kSlots], first[kSlots], head = 0, count = 0;
  auto collect_oldest = [&]() {
    const int i0 = first[head];
    const int rc = idn_collect(h, tickets[head], top1_idx + i0, top1_prob + i0,
                               logits ? logits + (size_t)i0 * IDN_NUM_CLASSES : nullptr);
    head = (head + 1) % kSlots; --count;
    return rc;
  };
  int rc = IDN_OK;
  int i0 = 0;
  for (size_t k = 0; k < cuts.size() && rc == IDN_OK; ++k) {
    if (count == kSlots) rc = collect_oldest();
    if (rc != IDN_OK) break;
    const int m = cuts[k];
    const int at = (head + count) % kSlots;
    rc = idn_submit_u8hwc(h, imgs, offsets + i0, hw + 2 * i0, m, logits != nullptr, &tickets[at]);
    if (rc == IDN_OK) { first[at] = i0; ++count; i0 += m; }
  }
  while (rc == IDN_OK && count > 0) rc = collect_oldest();
  if (rc != IDN_OK) {                           // drain so that the handle stays usable
    cudaStreamSynchronize(h->stream);
    for (int b = 0; b < kSlots; ++b) h->slot[b].busy = false;
  }
  return rc;
}

int idn_preprocess_u8hwc(int device, const uint8_t* imgs, const int64_t* offsets, const int32_t* hw,
                         int n, uint16_t* out) {
  if (!imgs || !offsets || !hw || !out || n < 0) { set_error("idn_preprocess_u8hwc: bad argument"); return IDN_EINVAL; }
  if (n == 0) return IDN_OK;
  IDN_TRY(check_device(device));
  idn_engine tmp;                       // borrows the table cache / meta plumbing of a handle
  tmp.device = device; tmp.cap = n;
  int rc = IDN_OK;
  uint8_t* d_in = nullptr; uint16_t* d_out = nullptr;
  auto cleanup = [&]() {
    for (void* p : tmp.dev_allocs) cudaFree(p);
    if (tmp.d_tables) cudaFree(tmp.d_tables);
    if (tmp.h_tables) cudaFreeHost(tmp.h_tables);
    if (tmp.h_meta) cudaFreeHost(tmp.h_meta);
    for (cudaEvent_t& ev : tmp.meta_ev) if (ev) { cudaEventDestroy(ev); ev = nullptr; }
    if (d_in) cudaFree(d_in);
    if (d_out) cudaFree(d_out);
  };
  int64_t hi = 0;
  for (int i = 0; i < n; ++i) hi = std::max(hi, offsets[i] + (int64_t)hw[2 * i] * hw[2 * i + 1] * 3);
  uint16_t lut[768];
  build_norm_lut(lut);
  const size_t out_elems = (size_t)n * 3 * IDN_CROP * IDN_CROP;
  if (cudaMalloc(&d_in, hi + 64) != cudaSuccess || cudaMalloc(&d_out, out_elems * 2) != cudaSuccess ||
      cudaMallocHost(&tmp.h_meta, sizeof(PreMeta) * n) != cudaSuccess) {
    set_error("idn_preprocess_u8hwc: allocation failed"); cleanup(); return IDN_ENOMEM;
  }
  if ((rc = dev_alloc(&tmp, &tmp.d_meta, (size_t)n)) != IDN_OK) { cleanup(); return rc; }
  if ((rc = dev_alloc(&tmp, &tmp.d_lut, 768)) != IDN_OK) { cleanup(); return rc; }
  cudaMemcpy(tmp.d_lut, lut, sizeof(lut), cudaMemcpyHostToDevice);
  cudaMemcpy(d_in, imgs, hi, cudaMemcpyHostToDevice);
  rc = fill_meta(&tmp, offsets, hw, n, 0);
  if (rc == IDN_OK) rc = launch_preprocess(d_in, tmp.d_meta, tmp.d_tables, tmp.d_lut, n, PRE_OUT_NCHW, d_out, 0, 0, nullptr);
  if (rc == IDN_OK && cudaDeviceSynchronize() != cudaSuccess) { set_error("preprocess kernel failed"); rc = IDN_ECUDA; }
  if (rc == IDN_OK && cudaMemcpy(out, d_out, out_elems * 2, cudaMemcpyDeviceToHost) != cudaSuccess) {
    set_error("preprocess D2H failed"); rc = IDN_ECUDA;
  }
  cleanup();
  return rc;
}

int idn_num_steps(idn_handle h) { return h ? (int)h->steps.size() : IDN_EINVAL; }

int idn_step_info(idn_handle h, int step, char* name, int name_len, int32_t* chw) {
  if (!h || step < 0 || step >= (int)h->steps.size()) { set_error("bad step"); return IDN_EINVAL; }
  const Step& s = h->steps[step];
  if (name && name_len > 0) { std::snprintf(name, name_len, "%s", s.name.c_str()); }
  if (chw) {
    if (s.out_kind == 0) { chw[0] = chw[1] = chw[2] = 0; }
    else if (s.out_kind == 3) { chw[0] = IDN_NUM_CLASSES; chw[1] = 1; chw[2] = 1; }
    else { chw[0] = s.out_C; chw[1] = s.out_H; chw[2] = s.out_W; }
  }
  return IDN_OK;
}

int idn_step_kernel(idn_handle h, int step, char* name, int name_len) {
  if (!h || !name || name_len <= 0 || step < 0 || step >= (int)h->steps.size()) { set_error("bad step"); return IDN_EINVAL; }
  const Step& s = h->steps[step];
  const ConvParams& cp = s.cp;
  std::string k;
  auto N = [](int v) { return std::to_string(v); };
  const bool fuse = h->fuse_stem_step == step && !h->no_pool_fusion && !h->stem_simt;
  auto pair_fused = [&](const idn_engine::FusePair& fp) {
    const Step& c = h->steps[fp.conv_step];
    return h->impl == IDN_IMPL_TC && !h->no_pool_fusion && c.use_halo && c.cp.pair && !c.b_resident && c.cp.num_n_tiles == 1 &&
           ((c.block_n == 192 && c.cp.b_group == 5) || (c.block_n == 256 && c.cp.b_group == 3));
  };
  bool pair_conv = false, pair_pool = false;
  for (const auto& fp : h->fuse_pairs) {
    pair_conv = pair_conv || (fp.conv_step == step && pair_fused(fp));
    pair_pool = pair_pool || (fp.pool_step == step && pair_fused(fp));
  }
  switch (s.kind) {
    case STEP_PRE: k = "preprocess_kernel<" + N(h->pre_mode) + ">"; break;
    case STEP_POOL: k = pair_pool ? "(folded into the convolution before it)" : (h->fuse_pool_step == step && !h->no_pool_fusion && !h->stem_simt) ? (h->steps[h->fuse_stem_step].cp.mode == MODE_STEM ? "(folded into stem_pool_kernel)" : "(folded into the convolution before it)") : "maxpool3x3s2_kernel"; break;
    case STEP_RHEAD: k = h->no_head_fusion ? "avgpool_kernel + fc_fp32_kernel + softmax_top1_kernel" : "resnet_head_kernel"; break;
    case STEP_SOFTMAX:
      k = (step > 0 && h->steps[step - 1].kind == STEP_CONV && h->steps[step - 1].cp.split_k > 1 && h->impl == IDN_IMPL_TC &&
           !h->no_split_defer) ? "softmax_top1_splitk_kernel" : "softmax_top1_kernel";
      break;
    case STEP_CONV:
      if (s.merged_into >= 0 && h->impl == IDN_IMPL_TC && !h->no_ds_fusion) k = "(rides in the launch of step " + N(s.merged_into) + ")";
      else if (h->impl != IDN_IMPL_TC || (cp.mode == MODE_STEM && h->stem_simt)) k = "conv_simt_kernel";
      else if (cp.mode == MODE_STEM) k = fuse ? "stem_pool_kernel" : "stem_kernel";
      else if (s.use_halo && cp.pair && s.b_resident) k = "conv_pair_res_kernel<" + N(cp.msub) + ", " + N(cp.num_taps == 9 ? 9 : 0) + ">";
      else if (s.use_halo && cp.pair) k = "conv_pair_kernel<" + N(s.block_n) + ", " + N(cp.b_group) + (pair_conv ? ", 1>" : ", 0>");
      else if (s.use_halo) k = "conv_halo_kernel<" + N(s.block_n) + ", " + N(cp.msub) + ", " + N(s.b_resident ? 1 : 0) + ", " +
                               N(s.b_resident ? (cp.num_taps == 9 ? 9 : 0) : (cp.b_group == 3 ? 3 : 0)) + (fuse ? ", 1>" : ", 0>");
      else if (cp.tc_pair) k = "conv_tc_pair_kernel<" + N(s.block_n) + ">";
      else k = "conv_tc_kernel<" + N(s.block_n) + ">";
      break;
  }
  std::snprintf(name, name_len, "%s", k.c_str());
  return IDN_OK;
}

int idn_read_activation(idn_handle h, int step, int n, float* out) {
  if (!h || !out || step < 0 || step >= (int)h->steps.size() || n <= 0 || n > h->cap) { set_error("bad argument"); return IDN_EINVAL; }
  const Step& s = h->steps[step];
  if (s.out_kind == 0) { set_error("step has no image-shaped output"); return IDN_EINVAL; }
  IDN_CUDA_TRY(cudaSetDevice(h->device));
  IDN_CUDA_TRY(cudaStreamSynchronize(h->stream));
  if (s.out_kind == 3) {
    std::vector<float> tmp((size_t)n * s.out_ld);
    IDN_CUDA_TRY(cudaMemcpy(tmp.data(), s.out_ptr, tmp.size() * 4, cudaMemcpyDeviceToHost));
    for (int i = 0; i < n; ++i)
      for (int c = 0; c < IDN_NUM_CLASSES; ++c) out[(size_t)i * IDN_NUM_CLASSES + c] = tmp[(size_t)i * s.out_ld + c];
    return IDN_OK;
  }
  const int planes = s.out_kind == 2 ? 4 : 1;
  const size_t rows = (size_t)s.out_plane_rows * planes;
  std::vector<uint16_t> tmp(rows * s.out_ld);
  IDN_CUDA_TRY(cudaMemcpy(tmp.data(), s.out_ptr, tmp.size() * 2, cudaMemcpyDeviceToHost));
  const int C = s.out_C, H = s.out_H, W = s.out_W;
  for (int i = 0; i < n; ++i)
    for (int y = 0; y < H; ++y)
      for (int x = 0; x < W; ++x) {
        size_t row;
        if (planes == 1) row = (size_t)i * s.out_g.img_rows() + (size_t)y * s.out_g.Wp + x;
        else row = (size_t)((y & 1) * 2 + (x & 1)) * s.out_plane_rows + (size_t)i * s.out_g.img_rows() +
                   (size_t)(y >> 1) * s.out_g.Wp + (x >> 1);
        const uint16_t* src = tmp.data() + row * s.out_ld;
        for (int c = 0; c < C; ++c)
          out[(((size_t)i * C + c) * H + y) * W + x] = bf16_bits_to_f32(src[c]);
      }
  return IDN_OK;
}

int idn_debug_stop_after(idn_handle h, int step) {
  if (!h) return IDN_EINVAL;
  h->stop_after = step;
  return IDN_OK;
}

int64_t idn_kernel_launches(idn_handle h) { return h ? h->launches : -1; }
int64_t idn_h2d_bytes(idn_handle h) { return h ? h->h2d_bytes : -1; }

int idn_debug_needed_rows(int H, int W, int* first_row, int* end_row) {
  if (H <= 0 || W <= 0 || !first_row || !end_row) { set_error("idn_debug_needed_rows: bad argument"); return IDN_EINVAL; }
  int nh, nw;
  resized_size(H, W, &nh, &nw);
  needed_source_range(H, nh, crop_offset(nh), first_row, end_row);
  return IDN_OK;
}

int idn_set_timing(idn_handle h, int on) {
  if (!h) return IDN_EINVAL;
  h->timing = on != 0;
  h->wait_profile = on == 2;
  return IDN_OK;
}

int idn_step_wait_cycles(idn_handle h, int step, uint64_t* out8) {
  if (!h || !out8 || step < 0 || step >= (int)h->steps.size()) { set_error("bad step"); return IDN_EINVAL; }
  for (int i = 0; i < 8; ++i) out8[i] = h->steps[step].wait_cycles[i];
  return IDN_OK;
}

int idn_step_time_ms(idn_handle h, int step, float* ms) {
  if (!h || !ms || step < 0 || step >= (int)h->steps.size()) { set_error("bad step"); return IDN_EINVAL; }
  const Step& s = h->steps[step];
  if (!s.ev0 || !s.ev1) { set_error("timing was not enabled for the last pass"); return IDN_EINVAL; }
  IDN_CUDA_TRY(cudaEventSynchronize(s.ev1));
  IDN_CUDA_TRY(cudaEventElapsedTime(ms, s.ev0, s.ev1));
  return IDN_OK;
}

// ---------------------------------------------------------------- stand-alone layer (tests only)
int idn_conv2d_nchw(int device, int impl, const float* x, int n, int cin, int h, int w,
                    const float* wgt, int cout, int r, int s, int stride, int pad,
                    const float* scale, const float* shift, const float* residual, int relu,
                    float* y) {
  if (!x || !wgt || !y || n <= 0 || r != s) { set_error("idn_conv2d_nchw: bad argument"); return IDN_EINVAL; }
  IDN_TRY(check_device(device));
  std::unique_ptr<idn_engine> e(new idn_engine());
  e->device = device; e->cap = n; e->impl = impl;
  { cudaDeviceProp prop; if (cudaGetDeviceProperties(&prop, device) == cudaSuccess) e->num_sms = prop.multiProcessorCount; }
  int rc = IDN_OK;
  auto done = [&](int code) { idn_engine* raw = e.release(); idn_destroy(raw); return code; };
  if (cudaStreamCreateWithFlags(&e->own_stream, cudaStreamNonBlocking) != cudaSuccess) { set_error("stream"); return done(IDN_ECUDA); }
  e->stream = e->own_stream;
  if ((rc = conv_tc_configure()) != IDN_OK) return done(rc);
  const int oh = (h + 2 * pad - r) / stride + 1, ow = (w + 2 * pad - s) / stride + 1;
  Act dst;
  std::vector<uint16_t> hbuf;
  if (cin == 3) {
    // stem path: x must be [n][3][224][224]
    if (h != IDN_CROP || w != IDN_CROP || cout != 64 || !(r == 7 && stride == 2 && pad == 3)) {
      set_error("idn_conv2d_nchw: cin=3 supports the 7x7 stride-2 stem only"); return done(IDN_EINVAL);
    }
    const int ws = 232, left = pad;
    bf16* xs = nullptr;
    if ((rc = dev_alloc(e.get(), &xs, (size_t)n * IDN_CROP * ws * 4 + 4096)) != IDN_OK) return done(rc);
    hbuf.assign((size_t)n * IDN_CROP * ws * 4, 0);
    for (int i = 0; i < n; ++i) for (int c = 0; c < 3; ++c) for (int yy = 0; yy < h; ++yy) for (int xx = 0; xx < w; ++xx)
      hbuf[(((size_t)i * IDN_CROP + yy) * ws + left + xx) * 4 + c] = f32_to_bf16_bits(x[(((size_t)i * 3 + c) * h + yy) * w + xx]);
    cudaMemcpy(xs, hbuf.data(), hbuf.size() * 2, cudaMemcpyHostToDevice);
    ConvWeights cw;
    if ((rc = upload_stem_weights(e.get(), &cw, wgt, cout, r, s, scale, shift)) != IDN_OK) return done(rc);
    if ((rc = alloc_act(e.get(), &dst, make_geom(oh, ow, 0), 64, n)) != IDN_OK) return done(rc);
    if ((rc = add_stem(e.get(), "stem", xs, ws, left, cw, r, stride, pad, oh, ow, dst, n)) != IDN_OK) return done(rc);
    e->steps.back().cp.relu = relu;
  } else {
    if (cin % 64 != 0 || cout % 64 != 0) { set_error("idn_conv2d_nchw: cin/cout must be multiples of 64"); return done(IDN_EINVAL); }
    Act src;
    if (stride == 1) {
      if (2 * pad != r - 1) { set_error("idn_conv2d_nchw: stride 1 needs 'same' padding"); return done(IDN_EINVAL); }
      if ((rc = alloc_act(e.get(), &src, make_geom(h, w, pad), cin, n)) != IDN_OK) return done(rc);
      hbuf.assign((size_t)src.rows() * cin, 0);
      for (int i = 0; i < n; ++i) for (int c = 0; c < cin; ++c) for (int yy = 0; yy < h; ++yy) for (int xx = 0; xx < w; ++xx)
        hbuf[((size_t)i * src.g.img_rows() + (size_t)yy * src.g.Wp + xx) * cin + c] = f32_to_bf16_bits(x[(((size_t)i * cin + c) * h + yy) * w + xx]);
      if ((rc = alloc_act(e.get(), &dst, src.g, cout, n)) != IDN_OK) return done(rc);
    } else {
      if (stride != 2 || (h & 1) || (w & 1)) { set_error("idn_conv2d_nchw: stride 2 needs even H, W"); return done(IDN_EINVAL); }
      if ((rc = alloc_act(e.get(), &src, make_geom(h / 2, w / 2, 1), cin, n, 4)) != IDN_OK) return done(rc);
      hbuf.assign((size_t)src.rows() * cin, 0);
      for (int i = 0; i < n; ++i) for (int c = 0; c < cin; ++c) for (int yy = 0; yy < h; ++yy) for (int xx = 0; xx < w; ++xx) {
        const size_t row = (size_t)((yy & 1) * 2 + (xx & 1)) * src.plane_rows + (size_t)i * src.g.img_rows() + (size_t)(yy >> 1) * src.g.Wp + (xx >> 1);
        hbuf[row * cin + c] = f32_to_bf16_bits(x[(((size_t)i * cin + c) * h + yy) * w + xx]);
      }
      if ((rc = alloc_act(e.get(), &dst, make_geom(oh, ow, 1), cout, n)) != IDN_OK) return done(rc);
      if (oh != h / 2 || ow != w / 2) { set_error("idn_conv2d_nchw: unsupported stride-2 shape"); return done(IDN_EINVAL); }
    }
    cudaMemcpy(src.ptr, hbuf.data(), hbuf.size() * 2, cudaMemcpyHostToDevice);
    ConvWeights cw;
    if ((rc = upload_conv_weights(e.get(), &cw, wgt, cout, cin, r, s, scale, shift)) != IDN_OK) return done(rc);
    Act res; Act* resp = nullptr;
    if (residual) {
      if ((rc = alloc_act(e.get(), &res, dst.g, cout, n)) != IDN_OK) return done(rc);
      std::vector<uint16_t> rb((size_t)res.rows() * cout, 0);
      for (int i = 0; i < n; ++i) for (int c = 0; c < cout; ++c) for (int yy = 0; yy < oh; ++yy) for (int xx = 0; xx < ow; ++xx)
        rb[((size_t)i * res.g.img_rows() + (size_t)yy * res.g.Wp + xx) * cout + c] = f32_to_bf16_bits(residual[(((size_t)i * cout + c) * oh + yy) * ow + xx]);
      cudaMemcpy(res.ptr, rb.data(), rb.size() * 2, cudaMemcpyHostToDevice);
      resp = &res;
    }
    ConvSpec spec; spec.R = r; spec.S = s; spec.stride = stride; spec.pad = pad; spec.relu = relu != 0;
    if ((rc = add_conv(e.get(), "conv", src, cw, spec, dst, resp, n)) != IDN_OK) return done(rc);
  }
  // uploads above went through the legacy default stream (a pageable cudaMemcpy may return before its
  // DMA has landed); the kernel runs on a non-blocking stream, so drain the device first
  if (cudaDeviceSynchronize() != cudaSuccess) { set_error("sync before conv failed"); return done(IDN_ECUDA); }
  rc = launch_conv_step(e.get(), e->steps.back(), n, impl, e->stream);
  if (rc != IDN_OK) return done(rc);
  if (cudaStreamSynchronize(e->stream) != cudaSuccess) {
    set_error(std::string("conv kernel failed: ") + cudaGetErrorString(cudaGetLastError()));
    return done(IDN_ECUDA);
  }
  if ((rc = read_tc_watchdog()) != IDN_OK) return done(rc);
  rc = idn_read_activation(e.get(), (int)e->steps.size() - 1, n, y);
  return done(rc);
}

}  // extern "C"
