// idn_preprocess_u8hwc_noalloc: the preprocess kernel for callers that must not allocate.
//
// Who that is: code that runs where a device allocation is impossible or forbidden -- a worker
// whose GPU memory has been handed over completely to another framework's pool (a fresh cudaMalloc
// would fail or fragment it), a latency-critical loop that may not take the allocator's implicit
// device-wide synchronisation, or a plug-in that is only given buffers and a stream.  The ordinary
// entry point, idn_preprocess_u8hwc, allocates its staging buffers, its per-image metadata and its
// resample tables with cudaMalloc on every call.  Here the caller brings the image bytes and the
// output buffer (both in device memory), and everything else the kernel needs -- the 768-entry
// normalisation table, one PreMeta record per image and the fixed-point resample tables -- lives in
// ONE file-scope `static __device__` array that the library reserves on every GPU that loads it,
// for the life of the process.  The array has a fixed, modest size, so a batch is limited to
// IDN_NOALLOC_MAX_IMAGES images and to resample tables that fit; larger requests are refused with
// an error instead of being served some other way.
//
// Host code cannot use the array's name as a device address: `g_noalloc_scratch` below is a
// host-side stand-in, and a kernel handed `&g_noalloc_scratch` would fault (or, where the GPU can
// reach host memory, quietly use the wrong memory).  The address is taken once per device, with that
// device current, through cudaGetSymbolAddress, and that pointer is what the kernel receives.
#include <cmath>
#include <mutex>

#include "common.cuh"

namespace idn {
namespace {

static __device__ __align__(256) unsigned char g_noalloc_scratch[IDN_NOALLOC_SCRATCH_BYTES];

// layout of the array
constexpr size_t kLutOff = 0;                                   // 768 x uint16
constexpr size_t kMetaOff = 2048;                               // IDN_NOALLOC_MAX_IMAGES x PreMeta
constexpr size_t kTablesOff = 4096;                             // int32 resample tables
constexpr int kTableIntsCap = (int)((IDN_NOALLOC_SCRATCH_BYTES - kTablesOff) / sizeof(int));
static_assert(sizeof(PreMeta) * IDN_NOALLOC_MAX_IMAGES <= kTablesOff - kMetaOff, "meta region");

constexpr int kMaxDevices = 16;
struct TableKey { int in_size, out_size, crop, off; };
// Everything a call touches on the host is per device and static (this entry point does not use the heap
// either): its own lock, its own staging arrays.  Calls on different GPUs therefore never wait for each
// other; calls on the same GPU are serialised on the host for as long as it takes to fill and hand over
// the staging arrays (they share the device array anyway).  Untouched entries are untouched zero pages.
struct DeviceState {
  std::mutex mu;
  bool ready = false;
  unsigned char* scratch = nullptr;     // device address of g_noalloc_scratch on this device
  int host_tables[kTableIntsCap];
  PreMeta host_meta[IDN_NOALLOC_MAX_IMAGES];
  TableKey keys[2 * IDN_NOALLOC_MAX_IMAGES];
};
DeviceState g_dev[kMaxDevices];

// Pillow's fixed-point BILINEAR coefficients for the 224 cropped outputs of one axis, written
// straight into `dst` ([ksize][xmin[224]][xcnt[224]][k[224][ksize]]).  Same arithmetic as
// build_axis_table in preprocess.cu, without any allocation.  Returns the number of ints written,
// or -1 when `cap` ints are not enough.
int axis_table_into(int in_size, int out_size, int crop_off, int* dst, int cap) {
  const int prec = 22;
  int ksize = 1;
  double scale = 1.0, filterscale = 1.0, support = 1.0;
  if (in_size != out_size) {
    scale = (double)in_size / (double)out_size;
    filterscale = scale < 1.0 ? 1.0 : scale;
    support = 1.0 * filterscale;
    ksize = (int)std::ceil(support) * 2 + 1;
  }
  const long long need = 1 + 2LL * IDN_CROP + (long long)IDN_CROP * ksize;
  if (need > cap) return -1;
  int* xmin = dst + 1;
  int* xcnt = dst + 1 + IDN_CROP;
  int* kk = dst + 1 + 2 * IDN_CROP;
  dst[0] = ksize;
  for (long long i = 0; i < (long long)IDN_CROP * ksize; ++i) kk[i] = 0;
  if (in_size == out_size) {
    for (int i = 0; i < IDN_CROP; ++i) { xmin[i] = i + crop_off; xcnt[i] = 1; kk[i] = 1 << prec; }
    return (int)need;
  }
  const double ss = 1.0 / filterscale;
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
      ww += a < 1.0 ? 1.0 - a : 0.0;
    }
    for (int x = 0; x < n; ++x) {
      double a = (x + x0 - center + 0.5) * ss;
      if (a < 0.0) a = -a;
      double w = a < 1.0 ? 1.0 - a : 0.0;
      if (ww != 0.0) w /= ww;
      const double v = w * (double)(1 << prec);
      kk[(long long)i * ksize + x] = w < 0 ? (int)(v - 0.5) : (int)(v + 0.5);
    }
    xmin[i] = x0;
    xcnt[i] = n;
  }
  return (int)need;
}

}  // namespace
}  // namespace idn

using namespace idn;

extern "C" int idn_preprocess_u8hwc_noalloc(int device, const uint8_t* d_imgs_packed,
                                            const int64_t* offsets, const int32_t* hw, int n,
                                            uint16_t* d_out_nchw_bf16, void* cuda_stream) {
  if (!d_imgs_packed || !offsets || !hw || !d_out_nchw_bf16 || n < 0) {
    set_error("idn_preprocess_u8hwc_noalloc: bad argument");
    return IDN_EINVAL;
  }
  if (n > IDN_NOALLOC_MAX_IMAGES) {
    set_error("idn_preprocess_u8hwc_noalloc: " + std::to_string(n) + " images in one call, the static scratch holds " +
              std::to_string(IDN_NOALLOC_MAX_IMAGES) + " (IDN_NOALLOC_MAX_IMAGES); split the batch");
    return IDN_EINVAL;
  }
  if (device < 0 || device >= kMaxDevices) {
    set_error("idn_preprocess_u8hwc_noalloc: device index out of range (static state exists for devices 0.." +
              std::to_string(kMaxDevices - 1) + ")");
    return IDN_EINVAL;
  }
  if (n == 0) return IDN_OK;
  int count = 0;
  if (cudaGetDeviceCount(&count) != cudaSuccess || device >= count) {
    set_error("idn_preprocess_u8hwc_noalloc: no such CUDA device (there is no CPU fallback)");
    return IDN_ENODEV;
  }
  IDN_CUDA_TRY(cudaSetDevice(device));
  cudaStream_t st = reinterpret_cast<cudaStream_t>(cuda_stream);
  DeviceState& ds = g_dev[device];
  std::lock_guard<std::mutex> lock(ds.mu);
  int* const g_host_tables = ds.host_tables;
  PreMeta* const g_host_meta = ds.host_meta;
  TableKey* const g_keys = ds.keys;

  // ---- set-up, once per device, with that device current
  if (!ds.ready) {
    unsigned char* ptr = nullptr;
    cudaError_t err = cudaGetSymbolAddress(reinterpret_cast<void**>(&ptr), g_noalloc_scratch);
    if (err != cudaSuccess || ptr == nullptr) {
      set_error(std::string("cudaGetSymbolAddress(g_noalloc_scratch) failed: ") + cudaGetErrorString(err));
      return IDN_ECUDA;
    }
    uint16_t lut[768];
    build_norm_lut(lut);
    // on the caller's stream, so that the kernel below is ordered after it whatever kind of stream it is
    IDN_CUDA_TRY(cudaMemcpyAsync(ptr + kLutOff, lut, sizeof(lut), cudaMemcpyHostToDevice, st));
    ds.scratch = ptr;
    ds.ready = true;
  }

  // ---- per-call contents of the array: tables and per-image records are rewritten every call
  //      (nothing from an earlier call is read, so nothing has to be cleared)
  int used = 0, nkeys = 0;
  auto table_for = [&](int in_size, int out_size, int crop) -> int {
    for (int k = 0; k < nkeys; ++k)
      if (g_keys[k].in_size == in_size && g_keys[k].out_size == out_size && g_keys[k].crop == crop)
        return g_keys[k].off;
    const int wrote = axis_table_into(in_size, out_size, crop, g_host_tables + used, kTableIntsCap - used);
    if (wrote < 0) return -1;
    g_keys[nkeys++] = TableKey{in_size, out_size, crop, used};
    used += wrote;
    return used - wrote;
  };
  for (int i = 0; i < n; ++i) {
    const int H = hw[2 * i], W = hw[2 * i + 1];
    if (H <= 0 || W <= 0) { set_error("idn_preprocess_u8hwc_noalloc: image with non-positive size"); return IDN_EINVAL; }
    int nh, nw;
    resized_size(H, W, &nh, &nw);
    const int ht = table_for(W, nw, crop_offset(nw));
    const int vt = ht < 0 ? -1 : table_for(H, nh, crop_offset(nh));
    if (ht < 0 || vt < 0) {
      set_error("idn_preprocess_u8hwc_noalloc: the resample tables of this batch do not fit the static scratch (" +
                std::to_string(IDN_NOALLOC_SCRATCH_BYTES) + " bytes, IDN_NOALLOC_SCRATCH_BYTES); use fewer distinct or "
                "smaller source sizes per call, or idn_preprocess_u8hwc");
      return IDN_EINVAL;
    }
    PreMeta& m = g_host_meta[i];
    m.offset = offsets[i]; m.H = H; m.W = W; m.htab = ht; m.vtab = vt;
  }
  // The sources are ordinary (pageable) static arrays: each of these copies makes the HOST wait until the
  // work already queued on `st` has finished and the bytes have been staged, and returns only then -- so
  // the arrays may be refilled by the next call on this device as soon as the lock is released.
  IDN_CUDA_TRY(cudaMemcpyAsync(ds.scratch + kMetaOff, g_host_meta, sizeof(PreMeta) * n, cudaMemcpyHostToDevice, st));
  IDN_CUDA_TRY(cudaMemcpyAsync(ds.scratch + kTablesOff, g_host_tables, sizeof(int) * used, cudaMemcpyHostToDevice, st));
  return launch_preprocess(d_imgs_packed, reinterpret_cast<const PreMeta*>(ds.scratch + kMetaOff),
                           reinterpret_cast<const int*>(ds.scratch + kTablesOff),
                           reinterpret_cast<const uint16_t*>(ds.scratch + kLutOff), n, PRE_OUT_NCHW,
                           d_out_nchw_bf16, 0, 0, st);
}
