"""`labels_to_contours()` -- the step in front of the path (ultrack/utils/edge.py:18-96).

Merges a sequence of label timelapses into ultrack's inputs: foreground = OR of `labels > 0`,
contours = mean of `skimage.segmentation.find_boundaries(mode="outer")`.  One CUDA kernel per
frame (`ulk_labels_to_contours`).  The reference writes zarr arrays; zarr is optional here, so
the result is a pair of numpy arrays unless a torch CUDA output is requested by the caller of
`labels_to_contours_device`.  `sigma` (Gaussian smoothing) is not implemented and raises.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence, Tuple, Union

import numpy as np

from ultrack_b200 import _lib


def labels_to_contours_device(lib, labels_dev_ptr: int, n_sets: int, shape3, fg_ptr: int, ct_ptr: int,
                              stream: int = 0) -> None:
    """labels int32 [n_sets, D, H, W] on the device -> foreground uint8, contours float32."""
    shp = (C.c_int64 * 3)(*[int(s) for s in shape3])
    _lib.check(lib, lib.ulk_labels_to_contours(C.c_void_p(labels_dev_ptr), int(n_sets), shp,
                                               C.c_void_p(fg_ptr), C.c_void_p(ct_ptr), C.c_void_p(stream)))


def labels_to_contours(
    labels: Union[np.ndarray, Sequence[np.ndarray]],
    sigma: Optional[Union[Sequence[float], float]] = None,
    foreground_store_or_path=None,
    contours_store_or_path=None,
    overwrite: bool = False,
) -> Tuple[np.ndarray, np.ndarray]:
    """Converts and merges a sequence of labels into ultrack input format (foreground and contours)."""
    import torch

    if sigma is not None:
        raise NotImplementedError("`sigma` smoothing is not implemented by ultrack_b200.labels_to_contours")
    if foreground_store_or_path is not None or contours_store_or_path is not None:
        raise NotImplementedError("zarr stores are not supported; the result is returned as numpy arrays")
    if not isinstance(labels, Sequence):
        labels = [labels]
    shape = tuple(labels[0].shape)
    for lb in labels:
        if shape != tuple(lb.shape):
            raise ValueError(f"All labels must have the same shape. Found {shape} and {lb.shape}")
    lib = _lib.get_lib()
    frame_shape = shape[1:]
    shape3 = frame_shape if len(frame_shape) == 3 else (1,) + frame_shape
    foreground = np.zeros(shape, dtype=bool)
    contours = np.zeros(shape, dtype=np.float32)
    dev = torch.device("cuda", torch.cuda.current_device())
    fg_d = torch.empty(shape3, dtype=torch.uint8, device=dev)
    ct_d = torch.empty(shape3, dtype=torch.float32, device=dev)
    for t in range(shape[0]):
        stack = np.stack([np.asarray(lb[t]) for lb in labels])
        if stack.min() < 0 or stack.max() > np.iinfo(np.int32).max:
            raise ValueError("labels must fit int32 and be non-negative")
        lab_d = torch.from_numpy(np.ascontiguousarray(stack, dtype=np.int32)).to(dev)
        labels_to_contours_device(lib, lab_d.data_ptr(), len(labels), shape3, fg_d.data_ptr(), ct_d.data_ptr(),
                                  torch.cuda.current_stream().cuda_stream)
        foreground[t] = fg_d.cpu().numpy().reshape(frame_shape).astype(bool)
        contours[t] = ct_d.cpu().numpy().reshape(frame_shape)
    return foreground, contours
