"""idunno_b200 -- B200-native worker-side inference for the IDunno cluster (hot path only).

Host side of the C ABI in include/idunno_b200.h.  The drop-in module ``alexnet_resnet`` next to
this package keeps the reference's ``deeplearning(filename, modelname, start, end)`` signature.
"""
from ._lib import IdunnoError, load_library, LIB_PATH  # noqa: F401
from .engine import Engine, bind_host_thread_to_gpu, conv2d_nchw, pack_images, preprocess  # noqa: F401
from . import sharding, synth, weights  # noqa: F401
