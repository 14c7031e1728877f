"""add_flow: vector-field shifts of the candidate nodes.

Mirror of `ultrack.imgproc.flow.add_flow` (imgproc/flow.py:639-751): for every timepoint the
vector field is sampled at each node's centroid mapped into field coordinates and the result,
scaled to voxels, is written to the `z_shift / y_shift / x_shift` columns that `link()` reads
(core/linking/processing.py:140-148,211-213).  The per-timepoint gather runs on the device
(`ulk_node_shifts`) on the node table `segment()` left resident; only the field of that
timepoint is uploaded (fields are usually coarser than the image).
"""
from __future__ import annotations

import ctypes as C
import logging
from typing import Sequence, Union

import numpy as np

from . import _lib, database, frames

LOG = logging.getLogger(__name__)


def _field_at(vector_field, t: int, is_sequence: bool, nvecs: int) -> np.ndarray:
    """[n_vec, ...spatial] array of timepoint t (np.asarray because of lazy formats, flow.py:728-731)."""
    if is_sequence:
        return np.stack([np.asarray(vector_field[v][t]) for v in range(nvecs)])
    return np.asarray(vector_field[t])


def add_flow(config, vector_field: Union[np.ndarray, Sequence[np.ndarray]]) -> None:
    """Adds vector field (coordinate shift) data into nodes.

    Same arguments, checks and error behaviour as the reference (flow.py:639-700): fewer
    vector fields than dimensions update the last (z, y, x) axes; the field's spatial shape may
    differ from the image's; values must be normalised to [-1, 1].
    """
    LOG.info("Adding shift (vector field) to nodes")
    if isinstance(vector_field, Sequence):
        assert all(v.shape == vector_field[0].shape for v in vector_field)
        ndim = vector_field[0].ndim
        nvecs = len(vector_field)
        vec_shape = np.asarray(vector_field[0].shape)
        is_sequence = True
    else:
        ndim = vector_field.ndim - 1
        nvecs = vector_field.shape[1]
        vec_shape = np.asarray([vector_field.shape[0], *vector_field.shape[2:]])
        is_sequence = False

    shape = np.asarray(config.data_config.metadata["shape"])
    if len(shape) != ndim:
        raise ValueError(
            "Original data shape and vector field must have same number of dimensions (ignoring channels)."
            f" Found {len(shape)} and {ndim}."
        )
    first = np.asarray(vector_field[0])
    if np.any(np.abs(first) > 1):
        raise ValueError(
            "Vector field values must be normalized. "
            f"Found {first.min()} and {first.max()}."
        )

    lib = _lib.get_lib()
    import torch

    db_path = config.data_config.database_path
    nd = len(shape) - 1
    shape3 = tuple(int(s) for s in shape[1:]) if nd == 3 else (1,) + tuple(int(s) for s in shape[1:])
    fshape3 = tuple(int(s) for s in vec_shape[1:]) if nd == 3 else (1,) + tuple(int(s) for s in vec_shape[1:])
    fs = (C.c_int64 * 3)(*fshape3)

    for t in range(int(shape[0])):
        frame = frames.STORE.get(db_path, t)
        if frame is None:
            frame, _ = frames.frame_from_database(lib, db_path, t, shape3, nd)
            if frame is not None:
                frames.STORE.put(db_path, t, frame)
        if frame is None or frame.n_nodes == 0:
            LOG.warning(f"No node found at time point {t}.")
            continue
        field = np.ascontiguousarray(_field_at(vector_field, t, is_sequence, nvecs))
        code = _lib.dtype_code(field.dtype)
        dev = torch.from_numpy(field.view(np.uint8).reshape(-1)).cuda()
        out = np.zeros((frame.n_nodes, 3), np.float64)
        _lib.check(lib, lib.ulk_node_shifts(frame.handle, C.c_void_p(dev.data_ptr()), code, nvecs, fs,
                                            out.ctypes.data_as(C.c_void_p), None))
        database.update_shifts(db_path, frame.ids, out)
