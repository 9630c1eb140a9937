"""Python face of one C-ABI handle: one (process, GPU, model), weights resident, BN folded."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from .weights import pack_state_dict


def model_id(modelname: str) -> int:
    # any modelname other than "alexnet" selects ResNet18 (reference alexnet_resnet.py:17-22)
    return _lib.IDN_MODEL_ALEXNET if modelname == "alexnet" else _lib.IDN_MODEL_RESNET18


def pack_images(images):
    """list of uint8 HWC arrays -> (packed bytes, offsets int64[n], hw int32[n, 2])."""
    n = len(images)
    hw = np.zeros((n, 2), dtype=np.int32)
    offsets = np.zeros(n, dtype=np.int64)
    total = 0
    for i, im in enumerate(images):
        if im.dtype != np.uint8 or im.ndim != 3 or im.shape[2] != 3:
            raise ValueError("images must be uint8 [H, W, 3]")
        hw[i] = im.shape[:2]
        offsets[i] = total
        total += im.shape[0] * im.shape[1] * 3
        total = (total + 15) & ~15      # 16-byte aligned image starts
    packed = np.zeros(max(total, 16), dtype=np.uint8)
    for i, im in enumerate(images):
        o = int(offsets[i])
        packed[o:o + im.size] = np.ascontiguousarray(im).reshape(-1)
    return packed, offsets, hw


MAX_IN_FLIGHT = 3          # IDN_MAX_IN_FLIGHT in include/idunno_b200.h


class Engine:
    def __init__(self, modelname: str, state_dict, device: int = 0, max_batch: int = 400,
                 impl: str = "tc"):
        self._lib = _lib.load_library()
        self.modelname = "alexnet" if modelname == "alexnet" else "resnet18"
        self.max_batch = int(max_batch)
        self.device = int(device)
        blob = pack_state_dict(state_dict)
        want = self._lib.idn_weight_count(model_id(modelname))
        if blob.size != want:
            raise _lib.IdunnoError(f"state_dict has {blob.size} values, {self.modelname} needs {want}")
        h = C.c_void_p()
        _lib.check(self._lib.idn_create(device, model_id(modelname), blob.ctypes.data, blob.nbytes,
                                        self.max_batch, C.byref(h)), "idn_create")
        self._h = h
        if impl != "tc":
            self.set_impl(impl)

    # -- lifetime ------------------------------------------------------------------------------
    def close(self):
        if getattr(self, "_h", None):
            self._lib.idn_destroy(self._h)
            self._h = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- configuration -------------------------------------------------------------------------
    def set_impl(self, impl: str):
        code = {"tc": _lib.IDN_IMPL_TC, "simt": _lib.IDN_IMPL_SIMT}[impl]
        _lib.check(self._lib.idn_set_impl(self._h, code), "idn_set_impl")

    def set_stream(self, cuda_stream_ptr: int):
        _lib.check(self._lib.idn_set_stream(self._h, C.c_void_p(cuda_stream_ptr)), "idn_set_stream")

    def set_timing(self, on):
        """0/False off, 1/True CUDA events per plan step, 2 events + per-role pipeline-wait cycles."""
        _lib.check(self._lib.idn_set_timing(self._h, int(on)), "idn_set_timing")

    # -- inference -----------------------------------------------------------------------------
    def infer_packed(self, packed: np.ndarray, offsets: np.ndarray, hw: np.ndarray,
                     want_logits: bool = False, packed_ptr: int | None = None):
        """Host buffers in, host results out (H2D and D2H inside the call)."""
        n = int(offsets.shape[0])
        idx = np.zeros(n, dtype=np.int32)
        prob = np.zeros(n, dtype=np.float32)
        logits = np.zeros((n, 1000), dtype=np.float32) if want_logits else None
        ptr = packed_ptr if packed_ptr is not None else packed.ctypes.data
        _lib.check(self._lib.idn_infer_u8hwc(
            self._h, ptr, offsets.ctypes.data, np.ascontiguousarray(hw, dtype=np.int32).ctypes.data,
            n, idx.ctypes.data, prob.ctypes.data,
            logits.ctypes.data if logits is not None else None), "idn_infer_u8hwc")
        return idx, prob, logits

    def infer(self, images, want_logits: bool = False):
        packed, offsets, hw = pack_images(images)
        return self.infer_packed(packed, offsets, hw, want_logits)

    def submit_packed(self, offsets: np.ndarray, hw: np.ndarray, packed_ptr: int, want_logits: bool = False):
        """Enqueue one tensor batch (host buffer at ``packed_ptr``, best pinned; it must stay alive until
        ``collect``).  Returns a ticket; at most ``MAX_IN_FLIGHT`` (3) submissions may be in flight."""
        n = int(offsets.shape[0])
        t = C.c_int(-1)
        hw32 = np.ascontiguousarray(hw, dtype=np.int32)
        _lib.check(self._lib.idn_submit_u8hwc(self._h, C.c_void_p(packed_ptr), offsets.ctypes.data,
                                              hw32.ctypes.data, n, int(want_logits), C.byref(t)),
                   "idn_submit_u8hwc")
        return (t.value, n, want_logits)

    def collect(self, ticket):
        t, n, want_logits = ticket
        idx = np.zeros(n, dtype=np.int32)
        prob = np.zeros(n, dtype=np.float32)
        logits = np.zeros((n, 1000), dtype=np.float32) if want_logits else None
        _lib.check(self._lib.idn_collect(self._h, t, idx.ctypes.data, prob.ctypes.data,
                                         logits.ctypes.data if logits is not None else None), "idn_collect")
        return idx, prob, logits

    def infer_device(self, d_imgs_ptr: int, offsets: np.ndarray, hw: np.ndarray, d_idx_ptr: int,
                     d_prob_ptr: int, d_logits_ptr: int | None = None):
        """Device-resident images and results; asynchronous on the handle's stream."""
        n = int(offsets.shape[0])
        _lib.check(self._lib.idn_infer_u8hwc_device(
            self._h, C.c_void_p(d_imgs_ptr), offsets.ctypes.data,
            np.ascontiguousarray(hw, dtype=np.int32).ctypes.data, n, C.c_void_p(d_idx_ptr),
            C.c_void_p(d_prob_ptr), C.c_void_p(d_logits_ptr) if d_logits_ptr else None),
            "idn_infer_u8hwc_device")

    def sync(self):
        _lib.check(self._lib.idn_sync(self._h), "idn_sync")

    def abort(self):
        """Waits for everything in flight and frees every submission slot without reading results
        (error recovery: tickets whose results nobody will collect)."""
        _lib.check(self._lib.idn_abort(self._h), "idn_abort")

    # -- diagnostics ---------------------------------------------------------------------------
    def steps(self):
        out = []
        for i in range(self._lib.idn_num_steps(self._h)):
            name = C.create_string_buffer(96)
            chw = (C.c_int32 * 3)()
            _lib.check(self._lib.idn_step_info(self._h, i, name, 96, chw), "idn_step_info")
            out.append((name.value.decode(), tuple(chw)))
        return out

    def step_kernels(self):
        """[(step name, CUDA kernel the step launches)]"""
        out = []
        for i, (name, _) in enumerate(self.steps()):
            buf = C.create_string_buffer(160)
            _lib.check(self._lib.idn_step_kernel(self._h, i, buf, 160), "idn_step_kernel")
            out.append((name, buf.value.decode()))
        return out

    def read_activation(self, step: int, n: int) -> np.ndarray:
        name, (c, h, w) = self.steps()[step]
        out = np.zeros((n, c, h, w), dtype=np.float32)
        _lib.check(self._lib.idn_read_activation(self._h, step, n, out.ctypes.data), "idn_read_activation")
        return out

    def stop_after(self, step: int):
        _lib.check(self._lib.idn_debug_stop_after(self._h, int(step)), "idn_debug_stop_after")

    def step_times_ms(self):
        out = []
        for i, (name, _) in enumerate(self.steps()):
            ms = C.c_float()
            _lib.check(self._lib.idn_step_time_ms(self._h, i, C.byref(ms)), "idn_step_time_ms")
            out.append((name, ms.value))
        return out

    def step_wait_cycles(self):
        out = []
        for i, (name, _) in enumerate(self.steps()):
            w = np.zeros(8, dtype=np.uint64)
            _lib.check(self._lib.idn_step_wait_cycles(self._h, i, w.ctypes.data), "idn_step_wait_cycles")
            out.append((name, w))
        return out

    @property
    def kernel_launches(self) -> int:
        return int(self._lib.idn_kernel_launches(self._h))

    @property
    def h2d_bytes(self) -> int:
        """Image bytes queued host-to-device so far (only the source rows the centre crop reads are copied)."""
        return int(self._lib.idn_h2d_bytes(self._h))


def preprocess(images, device: int = 0) -> np.ndarray:
    """uint8 HWC images -> bf16 bit patterns (uint16) NCHW [n, 3, 224, 224] via the CUDA kernel."""
    lib = _lib.load_library()
    packed, offsets, hw = pack_images(images)
    out = np.zeros((len(images), 3, 224, 224), dtype=np.uint16)
    _lib.check(lib.idn_preprocess_u8hwc(device, packed.ctypes.data, offsets.ctypes.data,
                                        hw.ctypes.data, len(images), out.ctypes.data),
               "idn_preprocess_u8hwc")
    return out


def conv2d_nchw(x, w, scale=None, shift=None, residual=None, stride=1, pad=0, relu=False,
                impl="tc", device=0) -> np.ndarray:
    """Stand-alone convolution layer through the library (tests)."""
    lib = _lib.load_library()
    x = np.ascontiguousarray(x, dtype=np.float32)
    w = np.ascontiguousarray(w, dtype=np.float32)
    n, cin, h, wd = x.shape
    cout, _, r, s = w.shape
    oh = (h + 2 * pad - r) // stride + 1
    ow = (wd + 2 * pad - s) // stride + 1
    y = np.zeros((n, cout, oh, ow), dtype=np.float32)
    sc = np.ascontiguousarray(scale, dtype=np.float32) if scale is not None else None
    sh = np.ascontiguousarray(shift, dtype=np.float32) if shift is not None else None
    rs = np.ascontiguousarray(residual, dtype=np.float32) if residual is not None else None
    code = {"tc": _lib.IDN_IMPL_TC, "simt": _lib.IDN_IMPL_SIMT}[impl]
    _lib.check(lib.idn_conv2d_nchw(
        device, code, x.ctypes.data, n, cin, h, wd, w.ctypes.data, cout, r, s, stride, pad,
        sc.ctypes.data if sc is not None else None, sh.ctypes.data if sh is not None else None,
        rs.ctypes.data if rs is not None else None, int(relu), y.ctypes.data), "idn_conv2d_nchw")
    return y


class PinnedBuffer:
    """Page-locked host bytes from idn_host_alloc, exposed as a numpy uint8 array (``.array``).
    Keep the object alive while a submission that reads it is in flight."""

    def __init__(self, n_bytes: int, device: int = 0):
        self._lib = _lib.load_library()
        p = C.c_void_p()
        _lib.check(self._lib.idn_host_alloc(device, int(n_bytes), C.byref(p)), "idn_host_alloc")
        self.ptr = p.value
        self.nbytes = int(n_bytes)
        self.array = np.ctypeslib.as_array((C.c_uint8 * self.nbytes).from_address(self.ptr))

    def close(self):
        if getattr(self, "ptr", None):
            self.array = None
            self._lib.idn_host_free(C.c_void_p(self.ptr))
            self.ptr = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def bind_host_thread_to_gpu(device: int) -> list[int]:
    """Pins the calling process to the CPU cores next to ``device`` (NVML's affinity mask for the GPU,
    intersected with what the process may use), so that page-locked staging buffers allocated afterwards
    are first-touched on the GPU's own NUMA node.  With one process per GPU on a two-socket host, image
    batches otherwise cross the socket interconnect on their way to half of the GPUs.  Returns the cores
    chosen ([] = left alone: no NVML, or the mask does not intersect the allowed set)."""
    import os
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(device)
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (os.cpu_count() + 63) // 64)
    except Exception:
        return []
    near = {64 * w + b for w, word in enumerate(words) for b in range(64) if (int(word) >> b) & 1}
    allowed = os.sched_getaffinity(0)
    pick = sorted(near & allowed)
    if not pick or len(pick) == len(allowed):
        return []
    os.sched_setaffinity(0, pick)
    return pick
