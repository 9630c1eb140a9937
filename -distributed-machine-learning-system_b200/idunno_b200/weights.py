"""state_dict -> the fp32 weight blob idn_create() takes (see include/idunno_b200.h)."""
from __future__ import annotations

import numpy as np


def pack_state_dict(sd) -> np.ndarray:
    """Concatenate every floating-point entry of a torchvision AlexNet/ResNet18 state_dict, in
    state_dict order, as fp32; the integer ``num_batches_tracked`` entries are skipped."""
    parts = []
    for k, v in sd.items():
        if k.endswith("num_batches_tracked"):
            continue
        a = v.detach().cpu().numpy() if hasattr(v, "detach") else np.asarray(v)
        parts.append(np.ascontiguousarray(a, dtype=np.float32).reshape(-1))
    return np.concatenate(parts)
