"""GPU: idn_preprocess_u8hwc_noalloc, the entry point for hosts that cannot allocate device memory
(their GPU memory already belongs to another pool, or an allocator stall is not acceptable).

The caller owns the input and output device buffers -- torch allocates them here, standing in for
that host -- and the library's working data sits in its static device array.  The output must be
bit-identical to idn_preprocess_u8hwc (the bar every preprocess test in this repo uses), on two
calls in a row: the second call reuses the array with different images, different sizes and fewer
of them, so stale contents of the first call would show.  Collected last on purpose (file name)."""
import ctypes as C

import numpy as np
import pytest
import torch

from idunno_b200 import _lib, pack_images, preprocess
from idunno_b200.synth import synth_image

pytestmark = pytest.mark.gpu


def _noalloc(lib, images, stream):
    packed, offsets, hw = pack_images(images)
    d_in = torch.from_numpy(packed).cuda()                       # caller-owned device buffers
    d_out = torch.zeros(len(images) * 3 * 224 * 224, dtype=torch.int16, device="cuda")
    torch.cuda.synchronize()
    rc = lib.idn_preprocess_u8hwc_noalloc(0, C.c_void_p(d_in.data_ptr()), offsets.ctypes.data,
                                          hw.ctypes.data, len(images), C.c_void_p(d_out.data_ptr()),
                                          C.c_void_p(stream.cuda_stream))
    _lib.check(rc, "idn_preprocess_u8hwc_noalloc")
    stream.synchronize()
    return d_out.cpu().numpy().view(np.uint16).reshape(len(images), 3, 224, 224)


def test_noalloc_matches_allocating_entry_point_twice(built_library):
    """Two calls in a row, different images, sizes and counts; the second on ANOTHER stream than the one
    the first call did its one-time set-up on (the set-up copy has completed by then: _noalloc
    synchronises the stream it used)."""
    lib = built_library
    first = [synth_image(i, mixed=True) for i in range(10)]     # ten different source sizes
    second = [synth_image(200 + i) for i in range(3)] + [synth_image(7, size=(300, 224))]
    for images, stream in ((first, torch.cuda.Stream()), (second, torch.cuda.Stream())):
        got = _noalloc(lib, images, stream)
        want = preprocess(images)                               # idn_preprocess_u8hwc
        assert got.dtype == want.dtype and got.shape == want.shape
        assert np.array_equal(got, want), f"{int((got != want).sum())} bf16 values differ"


def test_noalloc_allocates_nothing_and_runs_with_the_device_full(built_library):
    """What the entry point is for: the caller's pool owns the GPU.  A filler tensor takes all free device
    memory but a small reserve (what the driver itself needs to launch a kernel), then the call must still
    succeed, give the right answer, and leave cudaMemGetInfo exactly where it was."""
    lib = built_library
    images = [synth_image(300 + i, mixed=True) for i in range(6)]
    want = preprocess(images)
    stream = torch.cuda.Stream()
    _noalloc(lib, images[:1], stream)                            # one-time set-up outside the measured window
    packed, offsets, hw = pack_images(images)
    d_in = torch.from_numpy(packed).cuda()
    d_out = torch.zeros(len(images) * 3 * 224 * 224, dtype=torch.int16, device="cuda")
    torch.cuda.synchronize()
    torch.cuda.empty_cache()
    reserve = 192 << 20
    free0, _total = torch.cuda.mem_get_info()
    filler = torch.empty(max(free0 - reserve, 1 << 20), dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 <= reserve + (64 << 20), "the filler did not take the device's memory"
    for _ in range(2):
        rc = lib.idn_preprocess_u8hwc_noalloc(0, C.c_void_p(d_in.data_ptr()), offsets.ctypes.data,
                                              hw.ctypes.data, len(images), C.c_void_p(d_out.data_ptr()),
                                              C.c_void_p(stream.cuda_stream))
        _lib.check(rc, "idn_preprocess_u8hwc_noalloc")
        stream.synchronize()
        free2, _ = torch.cuda.mem_get_info()
        assert free2 == free1, f"device memory in use changed by {free1 - free2} bytes across the call"
    del filler
    torch.cuda.empty_cache()
    got = d_out.cpu().numpy().view(np.uint16).reshape(len(images), 3, 224, 224)
    assert np.array_equal(got, want)


def test_noalloc_rejects_what_does_not_fit(built_library):
    lib = built_library
    packed, offsets, hw = pack_images([synth_image(0)])
    d_in = torch.from_numpy(packed).cuda()
    d_out = torch.zeros(3 * 224 * 224, dtype=torch.int16, device="cuda")
    too_many = 65
    rc = lib.idn_preprocess_u8hwc_noalloc(0, C.c_void_p(d_in.data_ptr()),
                                          np.zeros(too_many, np.int64).ctypes.data,
                                          np.tile(hw, (too_many, 1)).ctypes.data, too_many,
                                          C.c_void_p(d_out.data_ptr()), None)
    assert rc == -1 and b"IDN_NOALLOC_MAX_IMAGES" in lib.idn_last_error()
    # one image whose resample tables alone exceed the static array (a 2000x downscale)
    huge_hw = np.asarray([[512000, 512000]], dtype=np.int32)
    rc = lib.idn_preprocess_u8hwc_noalloc(0, C.c_void_p(d_in.data_ptr()), offsets.ctypes.data,
                                          huge_hw.ctypes.data, 1, C.c_void_p(d_out.data_ptr()), None)
    assert rc == -1 and b"do not fit the static scratch" in lib.idn_last_error()
