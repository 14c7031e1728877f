"""CPU restatement of ultrack.imgproc.flow.add_flow's per-timepoint arithmetic (TEST
INFRASTRUCTURE ONLY; imgproc/flow.py:700-737).  Parity unpinned: the reference function needs
sqlalchemy/pandas sessions that are not installable here; the arithmetic is restated line by line.
"""
from __future__ import annotations

import numpy as np


def node_shifts(centroids_zyx: np.ndarray, field: np.ndarray, image_shape) -> np.ndarray:
    """centroids_zyx [n, ndim] node coordinates (the z/y/x columns), field [n_vec, *spatial] of one
    timepoint, image_shape = spatial shape of the image.  Returns [n, 3] (z_shift, y_shift, x_shift)."""
    image_shape = np.asarray(image_shape)
    nvecs = field.shape[0]
    vec_spatial = np.asarray(field.shape[1:])
    coords_scaling = (vec_spatial - 1) / (image_shape - 1)                       # flow.py:708
    vec_scaling = np.asarray(image_shape[len(coords_scaling) - nvecs:])[::-1]    # flow.py:713 (x, y, z)
    coords = np.round(np.asarray(centroids_zyx, float) * coords_scaling).astype(int)   # flow.py:717
    coords = np.minimum(np.maximum(0, coords), vec_spatial - 1)                  # flow.py:718-720
    idx = tuple(coords.T)
    out = np.zeros((len(coords), 3), np.float64)                                  # columns x, y, z -> stored z, y, x
    for col, v in enumerate(reversed(range(nvecs))):                              # flow.py:711,726
        out[:, 2 - col] = np.asarray(field[v])[idx].astype(np.float64) * float(vec_scaling[col])   # flow.py:737
    return out
