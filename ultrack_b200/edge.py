"""`labels_to_contours()` -- the step in front of the path (ultrack/utils/edge.py:18-96).

Merges a sequence of label timelapses into ultrack's inputs: foreground = OR of `labels > 0`,
contours = mean of `skimage.segmentation.find_boundaries(mode="outer")`, optionally Gaussian
smoothed and max-normalised (`sigma`).  Per frame: one stencil kernel (`ulk_labels_to_contours`)
and, with `sigma`, a separable Gaussian + global max (`ulk_smooth_contours`).

Outputs.  The reference returns zarr arrays.  Here:
* `device=True`  -> a pair of torch CUDA tensors (bool, float32) that `segment()` consumes in
  place: BASELINE configs 2 and 4 enter the path through this function and never touch the host;
* store / path arguments -> zarr arrays in those stores (needs zarr; raises ImportError otherwise);
* default -> numpy arrays (pinned staging, one asynchronous copy per frame, one sync at the end).
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence, Tuple, Union

import numpy as np

from ultrack_b200 import _lib


def labels_to_contours_device(lib, labels_dev_ptr: int, n_sets: int, shape3, fg_ptr: int, ct_ptr: int,
                              stream: int = 0, sigma3: Optional[Sequence[float]] = None) -> None:
    """labels int32 [n_sets, D, H, W] on the device -> foreground uint8, contours float32."""
    shp = (C.c_int64 * 3)(*[int(s) for s in shape3])
    _lib.check(lib, lib.ulk_labels_to_contours(C.c_void_p(labels_dev_ptr), int(n_sets), shp,
                                               C.c_void_p(fg_ptr), C.c_void_p(ct_ptr), C.c_void_p(stream)))
    if sigma3 is not None:
        sg = (C.c_double * 3)(*[float(v) for v in sigma3])
        _lib.check(lib, lib.ulk_smooth_contours(C.c_void_p(ct_ptr), shp, sg, C.c_void_p(stream)))


def _sigma3(sigma, ndim: int):
    if sigma is None:
        return None
    if np.isscalar(sigma):
        sg = [float(sigma)] * ndim
    else:
        sg = [float(v) for v in sigma]
        if len(sg) != ndim:
            raise RuntimeError("sequence argument must have length equal to input rank")     # scipy's error
    return [0.0] * (3 - ndim) + sg


def labels_to_contours(
    labels: Union[np.ndarray, Sequence[np.ndarray]],
    sigma: Optional[Union[Sequence[float], float]] = None,
    foreground_store_or_path=None,
    contours_store_or_path=None,
    overwrite: bool = False,
    device: bool = False,
) -> Tuple[np.ndarray, np.ndarray]:
    """Converts and merges a sequence of labels into ultrack input format (foreground and contours)."""
    import torch

    if not isinstance(labels, Sequence):
        labels = [labels]
    shape = tuple(labels[0].shape)
    for lb in labels:
        if shape != tuple(lb.shape):
            raise ValueError(f"All labels must have the same shape. Found {shape} and {lb.shape}")
    lib = _lib.get_lib()
    frame_shape = shape[1:]
    shape3 = frame_shape if len(frame_shape) == 3 else (1,) + frame_shape
    sg3 = _sigma3(sigma, len(frame_shape))
    dev = torch.device("cuda", torch.cuda.current_device())
    stream = torch.cuda.current_stream().cuda_stream
    to_store = foreground_store_or_path is not None or contours_store_or_path is not None
    if to_store:
        try:
            import zarr
        except ImportError as e:  # pragma: no cover - zarr is not in the build image
            raise ImportError("zarr is needed for `foreground_store_or_path` / `contours_store_or_path`") from e
        mode = "w" if overwrite else "w-"
        fg_out = zarr.open(foreground_store_or_path, mode=mode, shape=shape, dtype=bool, chunks=(1,) + frame_shape)
        ct_out = zarr.open(contours_store_or_path, mode=mode, shape=shape, dtype=np.float32, chunks=(1,) + frame_shape)
    if device:
        fg_all = torch.empty(shape, dtype=torch.bool, device=dev)
        ct_all = torch.empty(shape, dtype=torch.float32, device=dev)
    elif not to_store:
        fg_all = torch.empty(shape, dtype=torch.bool).pin_memory()
        ct_all = torch.empty(shape, dtype=torch.float32).pin_memory()
    fg_d = torch.empty(shape3, dtype=torch.uint8, device=dev)
    for t in range(shape[0]):
        frames = [lb[t] for lb in labels]
        if all(type(f).__module__.startswith("torch") and f.is_cuda for f in frames):
            lab_d = torch.stack([f.to(torch.int32) for f in frames]).contiguous()
        else:
            stack = np.stack([np.asarray(f) for f in frames])
            if stack.min() < 0 or stack.max() > np.iinfo(np.int32).max:
                raise ValueError("labels must fit int32 and be non-negative")
            lab_d = torch.from_numpy(np.ascontiguousarray(stack, dtype=np.int32)).to(dev, non_blocking=True)
        ct_d = ct_all[t].view(shape3) if device else torch.empty(shape3, dtype=torch.float32, device=dev)
        labels_to_contours_device(lib, lab_d.data_ptr(), len(labels), shape3, fg_d.data_ptr(), ct_d.data_ptr(), stream, sg3)
        if device:
            fg_all[t] = fg_d.view(frame_shape).bool()
        elif to_store:
            fg_out[t] = fg_d.view(frame_shape).bool().cpu().numpy()
            ct_out[t] = ct_d.view(frame_shape).cpu().numpy()
        else:
            fg_all[t].copy_(fg_d.view(frame_shape).bool(), non_blocking=True)
            ct_all[t].copy_(ct_d.view(frame_shape), non_blocking=True)
    torch.cuda.current_stream().synchronize()
    if to_store:
        return fg_out, ct_out
    if device:
        return fg_all, ct_all
    return fg_all.numpy(), ct_all.numpy()


labels_to_edges = labels_to_contours      # utils/edge.py:99-101
