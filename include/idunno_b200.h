/*
 * idunno_b200.h -- C ABI of the B200-native inference hot path of the IDunno cluster.
 *
 * The reference (kentchen831213/-Distributed-Machine-Learning-System) has no FFI: its worker calls
 * the plain Python function  alexnet_resnet.deeplearning(filename, modelname, start, end)
 * (alexnet_resnet.py:12-92; call sites mp4_machinelearning.py:596 and :765).  The drop-in module
 * of the same name shipped in this repo keeps that signature and binds the entry points below
 * through ctypes.  Every function:
 *   - has C linkage, plain pointers and sizes, no C++/torch types;
 *   - returns 0 on success and a negative IDN_E* code otherwise; the message of the last error on
 *     the calling thread is available from idn_last_error();
 *   - never throws across the boundary;
 *   - never falls back to a CPU implementation: without a usable sm_100 device it fails.
 *
 * Threading: calls on one handle must be serialised by the caller (the reference calls
 * deeplearning() from one receiver thread, mp4_machinelearning.py:552-616).  No entry point holds
 * the Python GIL (ctypes releases it around foreign calls).
 */
#ifndef IDUNNO_B200_H
#define IDUNNO_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define IDN_OK            0
#define IDN_EINVAL       -1   /* bad argument                                   */
#define IDN_ECUDA        -2   /* CUDA runtime / driver error                    */
#define IDN_ENODEV       -3   /* no sm_100 device / extension cannot run here   */
#define IDN_EWEIGHTS     -4   /* weight blob has the wrong size for the model   */
#define IDN_ENOMEM       -5

#define IDN_MODEL_ALEXNET  0  /* modelname == "alexnet"  (alexnet_resnet.py:17-19) */
#define IDN_MODEL_RESNET18 1  /* any other modelname     (alexnet_resnet.py:20-22) */

#define IDN_NUM_CLASSES 1000
#define IDN_CROP        224   /* transforms.CenterCrop(224), alexnet_resnet.py:59 */

/* Convolution back-ends.  TC is the product path (tcgen05 / TMEM / TMA implicit GEMM).  SIMT is a
 * slow CUDA-core kernel over the same operand layouts kept as an on-device cross-check for tests
 * and bring-up; it is never selected implicitly. */
#define IDN_IMPL_TC   0
#define IDN_IMPL_SIMT 1

typedef struct idn_engine* idn_handle;

/* Replaces: model construction/load + .eval() at alexnet_resnet.py:17-22, done once per process
 * instead of once per call.  weights_blob = the model's state_dict as fp32 values concatenated in
 * torchvision state_dict order, skipping the integer num_batches_tracked entries (layout in
 * INTEGRATION.md; idunno_b200.weights.pack_state_dict builds it).  BatchNorm is folded to a
 * per-channel fp32 scale/shift, weights are converted to bf16 and re-laid-out for the kernels.
 * max_batch = largest tensor batch one pass processes (activation buffers are sized for it; longer
 * inputs are processed in chunks of max_batch).
 * One handle's work is ordered on its stream; several handles may run at the same time on one GPU (the concurrent
 * AlexNet + ResNet18 jobs of the reference): no CTA of any kernel waits for a CTA outside its own cluster.  (The
 * one kernel that could -- AlexNet's last FC layer with its K loop split over CTAs -- leaves the sum of its partial
 * tiles to the softmax kernel behind it; only with IDN_NO_SPLIT_DEFER=1, or when that layer's own output is read back
 * as an activation tap, do its CTAs wait for one another, and two such launches must then not share a GPU.) */
int idn_create(int device, int model_id, const void* weights_blob, size_t n_bytes, int max_batch,
               idn_handle* out);

/* Frees all device memory owned by the handle. */
int idn_destroy(idn_handle h);

/* Number of fp32 values idn_create expects for model_id (61 100 840 for AlexNet,
 * 11 699 112 for ResNet18 = parameters + BatchNorm running statistics). */
int64_t idn_weight_count(int model_id);

/* Work is enqueued on this CUDA stream (a cudaStream_t); default: a stream owned by the handle. */
int idn_set_stream(idn_handle h, void* cuda_stream);

/* Selects IDN_IMPL_TC (default) or IDN_IMPL_SIMT for the convolution / FC layers. */
int idn_set_impl(idn_handle h, int impl);

/* Replaces the per-image body of the reference loop, alexnet_resnet.py:57-87:
 * Resize(256)/CenterCrop(224)/ToTensor/Normalize, model forward, softmax, top-1 -- for n images at
 * once (eval-mode results are independent per image, so batching is exact).
 *   imgs_packed : uint8 RGB, HWC, images back to back (HOST memory; pinned is faster)
 *   offsets[n]  : byte offset of image i in imgs_packed
 *   hw[2n]      : (H_i, W_i)
 *   top1_idx[n] : argmax class index (lowest index on exact ties)
 *   top1_prob[n]: softmax probability of that class, fp32
 *   logits      : optional [n][1000] fp32 (NULL to skip)
 * Host buffers in and out: the call copies inputs to the device, runs, copies results back and
 * returns after they have arrived.  Inputs longer than max_batch are processed as tensor batches of
 * max_batch images with up to IDN_MAX_IN_FLIGHT of them in flight (see idn_submit_u8hwc).  An input of
 * 128 .. max_batch images -- the reference's call shape, one blocking deeplearning() per shard -- is cut
 * into a few sub-batches the same way, so that its copy overlaps its kernels (results do not depend on
 * the cut). */
int idn_infer_u8hwc(idn_handle h, const uint8_t* imgs_packed, const int64_t* offsets,
                    const int32_t* hw, int n, int32_t* top1_idx, float* top1_prob, float* logits);

/* The same call split in two, for hosts that stage the next query while the current one runs (the
 * reference's worker loop decodes and classifies strictly one image after the other,
 * alexnet_resnet.py:48-90).  idn_submit_u8hwc enqueues the H2D copy of one tensor batch
 * (n <= max_batch) on a copy stream and the kernels on the handle's stream and returns at once;
 * idn_collect blocks until that batch is done and copies the results out.  At most
 * IDN_MAX_IN_FLIGHT submissions may be in flight (each owns a staging buffer of the batch's size on
 * the device), so the copies of the next batches overlap the kernels of the current one and the copy
 * engine never waits for the host to come back from idn_collect; idn_infer_u8hwc uses exactly this
 * for inputs longer than max_batch.  The image buffer must stay valid (and is best
 * pinned) until the batch has been collected.  ticket identifies the submission. */
#define IDN_MAX_IN_FLIGHT 3
int idn_submit_u8hwc(idn_handle h, const uint8_t* imgs_packed, const int64_t* offsets,
                     const int32_t* hw, int n, int want_logits, int* ticket);
int idn_collect(idn_handle h, int ticket, int32_t* top1_idx, float* top1_prob, float* logits);

/* Same computation with the image bytes and the results resident in DEVICE memory (offsets/hw stay
 * host arrays: a few bytes of metadata).  Asynchronous on the handle's stream; n <= max_batch. */
int idn_infer_u8hwc_device(idn_handle h, const uint8_t* d_imgs_packed, const int64_t* offsets,
                           const int32_t* hw, int n, int32_t* d_top1_idx, float* d_top1_prob,
                           float* d_logits);

/* Blocks until everything enqueued on the handle's stream has finished. */
int idn_sync(idn_handle h);

/* Error recovery for the submit/collect pair (no reference counterpart: the reference's worker loop has
 * nothing in flight, alexnet_resnet.py:48-90; its caller drops a failed shard, mp4_machinelearning.py:615).
 * Waits for the copy stream and the handle's stream, then frees every submission slot without returning
 * results: tickets issued before the call are void.  A host that hits an exception between idn_submit_u8hwc
 * and idn_collect (a file that fails to decode for the next batch) calls this so that the handle is usable
 * for the next shard. */
int idn_abort(idn_handle h);

/* Replaces the preprocess Compose at alexnet_resnet.py:57-66 alone (parity tests, staging):
 * host uint8 HWC images -> host bf16 (raw uint16 bit patterns) NCHW [n][3][224][224]. */
int idn_preprocess_u8hwc(int device, const uint8_t* imgs_packed, const int64_t* offsets,
                         const int32_t* hw, int n, uint16_t* out_nchw_bf16);

/* Test/diagnostic taps (no reference counterpart).  After idn_infer_*, copy the activation written
 * by plan step `step` of the last pass to host as fp32 NCHW [n][C][H][W] (buffers are reused: the
 * value is the LAST writer's unless idn_debug_stop_after stopped the pass there); idn_num_steps /
 * idn_step_info describe the plan.  Returns IDN_EINVAL for steps without an image-shaped output. */
int idn_num_steps(idn_handle h);
int idn_step_info(idn_handle h, int step, char* name, int name_len, int32_t* chw /* [3] */);
int idn_read_activation(idn_handle h, int step, int n, float* out_nchw);
/* Name of the CUDA kernel (template instance as a profiler prints it, e.g. "conv_pair_kernel<192, 5>") that
 * plan step `step` launches; bench.py uses it to check that a committed profiler capture it quotes DRAM
 * traffic from really is the kernel it timed. */
int idn_step_kernel(idn_handle h, int step, char* name, int name_len);
/* Later passes run the plan only up to and including `step` (-1: whole plan), so that an
 * activation buffer that a later step reuses can be read back right after its producer. */
int idn_debug_stop_after(idn_handle h, int step);

/* Counters for bench.py: kernels launched by this handle since creation, and the duration in
 * milliseconds of plan step `step` in the last pass when timing was switched on
 * (idn_set_timing(h, 1) records CUDA events around every step; 2 additionally books the pipeline
 * waits idn_step_wait_cycles reports, which slows the kernels a little). */
int64_t idn_kernel_launches(idn_handle h);
/* Image bytes this handle has queued host-to-device since creation (idn_submit_u8hwc / idn_infer_u8hwc).  A batch
 * of equally sized images at a constant stride is copied as one pitched copy of the source rows that
 * Resize(256) + CenterCrop(224) reads (198 of 224 rows for a 224 x 224 image; IDN_FULL_H2D=1 copies whole images). */
int64_t idn_h2d_bytes(idn_handle h);
/* Host-only: the source rows [*first_row, *end_row) of an H x W image that Resize(256) + CenterCrop(224) reads --
 * what the pitched copy sends (plus the one row the kernel fetches ahead).  Needs no GPU; tests/test_host.py checks
 * it against the oracle's Pillow coefficient bounds over a sweep of sizes. */
int idn_debug_needed_rows(int H, int W, int* first_row, int* end_row);
int idn_set_timing(idn_handle h, int on);
int idn_step_time_ms(idn_handle h, int step, float* ms);
/* With timing on: SM cycles the tensor-core kernel of plan step `step` spent blocked in each kind of
 * pipeline wait during the last pass, summed over CTAs -- [0] TMA producer on a free stage, [1] MMA on
 * the epilogue, [2] MMA on activations, [3] epilogue on the MMA, [4] weight producer on a free stage,
 * [5] MMA on resident weights, [6] MMA on streamed weights. */
int idn_step_wait_cycles(idn_handle h, int step, uint64_t* out8);

/* Stand-alone layer entry (tests): y = act(conv(x, w) * scale + shift (+ residual)) with NCHW fp32
 * host tensors, computed through the same operand layouts and kernels as the model plans.
 * x [n][cin][h][w], w [cout][cin][r][s] fp32 (rounded to bf16 inside), scale/shift [cout],
 * residual NULL or [n][cout][oh][ow], y [n][cout][oh][ow].  cin must be a multiple of 64 and cout
 * a multiple of 64; stride 1 or 2 (stride 2 supports r=s=3,pad=1 and r=s=1,pad=0). */
int idn_conv2d_nchw(int device, int impl, const float* x, int n, int cin, int h, int w,
                    const float* wgt, int cout, int r, int s, int stride, int pad,
                    const float* scale, const float* shift, const float* residual, int relu,
                    float* y);

/* Has no counterpart in the reference: an entry point for callers that cannot allocate device
 * memory (all of it already belongs to another framework's pool, an allocator call would stall a
 * latency-critical loop, or the caller is a plug-in that only receives buffers and a stream).
 * Same computation and same NCHW bf16 output as idn_preprocess_u8hwc, but the image bytes and the
 * output are DEVICE buffers the caller owns, and the kernel's working data (normalisation table,
 * per-image records, resample tables) lives in a static device array inside the library instead of
 * cudaMalloc'ed buffers: IDN_NOALLOC_SCRATCH_BYTES per GPU that loads the library, held until the
 * process exits.  Limits: n <= IDN_NOALLOC_MAX_IMAGES per call and resample tables that fit the
 * array (always true for a batch of equally sized images); anything larger returns IDN_EINVAL with
 * the reason in idn_last_error().  The call makes no cudaMalloc/cudaFree/cudaMallocHost and no
 * host heap allocation on its success path.  It is asynchronous: work is enqueued on cuda_stream
 * (NULL = the legacy default stream) and d_out_nchw_bf16 is complete once that stream has been
 * synchronised.  offsets/hw are host arrays as elsewhere.
 * Host-side cost: the per-image records and the resample tables are copied from static (pageable)
 * host arrays with two cudaMemcpyAsync calls; such copies make the calling thread wait until the
 * work already queued on cuda_stream has finished and the bytes have been staged.  The call never
 * waits for the allocator, but it is not free of host stalls: keep cuda_stream short, or give this
 * call a stream of its own.  Host state (lock, staging arrays) is per device, devices 0..15.
 * The first call on a device looks up the array's device address and uploads the normalisation
 * table (an extra copy on cuda_stream); make that first call before starting any stream capture.
 * Every call on a device shares the one array: it is rewritten by each call (nothing from an
 * earlier call is read), so two calls whose GPU work overlaps -- different streams, or different
 * threads -- race on it.  Serialise calls per device, or order them on one stream. */
#define IDN_NOALLOC_MAX_IMAGES    64
#define IDN_NOALLOC_SCRATCH_BYTES (512 * 1024)
int idn_preprocess_u8hwc_noalloc(int device, const uint8_t* d_imgs_packed, const int64_t* offsets,
                                 const int32_t* hw, int n, uint16_t* d_out_nchw_bf16,
                                 void* cuda_stream);

/* Page-locked host memory for image staging (no reference counterpart: the reference keeps decoded
 * images in PIL objects and copies one 602 KB tensor per image, alexnet_resnet.py:67-72).  A batch
 * staged in such a buffer is copied by DMA at the full host-link rate and lets idn_submit_u8hwc
 * return before the copy has finished; pageable memory works too but is staged through a bounce
 * buffer by the driver.  `device` only selects the context the allocation is made in; the memory is
 * usable with every device (portable). */
int idn_host_alloc(int device, size_t n_bytes, void** out);
int idn_host_free(void* p);

const char* idn_last_error(void);

/* Library build identification ("idunno_b200 <version> sm_100a"). */
const char* idn_version(void);

#ifdef __cplusplus
}
#endif
#endif /* IDUNNO_B200_H */
