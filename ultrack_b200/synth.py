"""Seeded synthetic timelapses (numpy, host side) for tests and small benches.

The reference's own fixtures (`ultrack/utils/data.py:15-32`,
`ultrack/utils/test_utils.py:75-111`) need scikit-image, which is not available
here, so the build ships its own generator:

* cells are balls on a jittered lattice, radius ~ N(mu, 0.15 mu) clipped to
  [0.6 mu, 1.4 mu], lattice pitch chosen for a target foreground fraction;
* variant "T" (tie-heavy): bool foreground, bool contours = boundary map of the
  generating labels -> edge weights in {1e-8, 0.5, 1.0}, like the mock data;
* variant "S" (smooth): float32 foreground and float32 contours that are low in
  the cell centre and 1 on cell borders / between touching cells, plus a small
  uniform noise that breaks ties.

The CUDA library has a device-side generator with the same recipe
(`ulk_synth_frame`) for frames that must never exist on the host.
"""
from __future__ import annotations

import itertools
from typing import Tuple

import numpy as np


def _hash_u01(seed: int, *keys: np.ndarray) -> np.ndarray:
    """Counter-based uniform(0,1) from integer keys (splitmix64 finaliser)."""
    with np.errstate(over="ignore"):
        x = np.uint64(seed) * np.uint64(0x9E3779B97F4A7C15)
        for k in keys:
            x = (x ^ np.asarray(k).astype(np.uint64)) * np.uint64(0xBF58476D1CE4E5B9)
            x ^= x >> np.uint64(31)
            x = x * np.uint64(0x94D049BB133111EB)
            x ^= x >> np.uint64(29)
    return (x >> np.uint64(11)).astype(np.float64) * (1.0 / (1 << 53))


def cell_fields(shape: Tuple[int, ...], t: int = 0, seed: int = 0, radius: float = 8.0,
                fill: float = 0.3, motion: float = 0.0) -> Tuple[np.ndarray, np.ndarray, np.ndarray]:
    """Returns (labels int32, s1, s2): nearest / second nearest normalised
    distance to a cell centre (d / r) for every voxel."""
    ndim = len(shape)
    unit = np.pi if ndim == 2 else 4.0 / 3.0 * np.pi
    pitch = radius * (unit / fill) ** (1.0 / ndim)
    ncell = [int(np.ceil(s / pitch)) + 2 for s in shape]
    # per-cell centre and radius on the (small) cell lattice, gathered per voxel below
    cell_idx = np.meshgrid(*[np.arange(-1, n - 1, dtype=np.int64) for n in ncell], indexing="ij")
    cid_cell = np.zeros(cell_idx[0].shape, dtype=np.int64)
    for a in range(ndim):
        cid_cell = cid_cell * ncell[a] + (cell_idx[a] + 1)
    centre = []
    for a in range(ndim):
        jit = (_hash_u01(seed, cid_cell, 17 + a) - 0.5) * 0.5 * pitch
        mov = (_hash_u01(seed, cid_cell, 101 + a, t) - 0.5) * 2.0 * motion if motion > 0 else 0.0
        centre.append(((cell_idx[a] + 0.5) * pitch + jit + mov).ravel())
    u1 = _hash_u01(seed, cid_cell, 7)
    u2 = _hash_u01(seed, cid_cell, 8)
    g = np.sqrt(-2.0 * np.log(np.maximum(u1, 1e-12))) * np.cos(2 * np.pi * u2)
    rad = np.clip(radius * (1.0 + 0.15 * g), 0.6 * radius, 1.4 * radius).ravel()

    axes = [np.arange(s, dtype=np.float64) for s in shape]
    cell = [np.floor(ax / pitch).astype(np.int64) for ax in axes]
    s1 = np.full(shape, np.inf)
    s2 = np.full(shape, np.inf)
    lab = np.zeros(shape, dtype=np.int64)
    for off in itertools.product((-1, 0, 1), repeat=ndim):
        cid = np.zeros(shape, dtype=np.int64)
        for a in range(ndim):
            sl = [None] * ndim
            sl[a] = slice(None)
            cid = cid * ncell[a] + (cell[a] + off[a] + 1)[tuple(sl)]
        d2 = np.zeros(shape)
        for a in range(ndim):
            sl = [None] * ndim
            sl[a] = slice(None)
            d2 += (axes[a][tuple(sl)] - centre[a][cid]) ** 2
        s = np.sqrt(d2) / rad[cid]
        closer = s < s1
        s2 = np.where(closer, s1, np.minimum(s2, s))
        lab = np.where(closer, cid + 1, lab)
        s1 = np.where(closer, s, s1)
    lab = np.where(s1 < 1.0, lab, 0)
    return lab.astype(np.int64), s1, s2


def boundary_map(labels: np.ndarray) -> np.ndarray:
    """Bool map of voxels that touch (4/6-connectivity) a different label."""
    out = np.zeros(labels.shape, dtype=bool)
    for a in range(labels.ndim):
        sl_a = [slice(None)] * labels.ndim
        sl_b = [slice(None)] * labels.ndim
        sl_a[a] = slice(0, -1)
        sl_b[a] = slice(1, None)
        diff = labels[tuple(sl_a)] != labels[tuple(sl_b)]
        out[tuple(sl_a)] |= diff
        out[tuple(sl_b)] |= diff
    return out


def make_frame(shape: Tuple[int, ...], t: int = 0, seed: int = 0, variant: str = "S",
               radius: float = 8.0, fill: float = 0.3, motion: float = 0.0,
               noise: float = 1e-3) -> Tuple[np.ndarray, np.ndarray, np.ndarray]:
    """(foreground, contours, labels) for one frame."""
    lab, s1, s2 = cell_fields(shape, t=t, seed=seed, radius=radius, fill=fill, motion=motion)
    if variant == "T":
        fg = lab > 0
        ct = boundary_map(lab) & fg
        return fg, ct, lab.astype(np.int32)
    idx = np.arange(lab.size, dtype=np.int64).reshape(shape)
    nz = _hash_u01(seed, idx, 1000 + t) * noise
    ct = np.clip(np.maximum(s1, 1.0 - 3.0 * (s2 - s1)), 0.0, 1.0) + nz
    fg = np.clip(4.0 * (1.0 - s1), 0.0, 1.0)
    return fg.astype(np.float32), ct.astype(np.float32), lab.astype(np.int32)


def make_timelapse(shape: Tuple[int, ...], length: int, seed: int = 0, variant: str = "S",
                   repeat: bool = False, **kw) -> Tuple[np.ndarray, np.ndarray]:
    """(foreground, contours) of shape (T, *shape).  `repeat=True` repeats frame 0,
    as ultrack's `timelapse_mock_data` does (utils/test_utils.py:106-109)."""
    frames = []
    for t in range(length):
        if repeat and t > 0:
            frames.append(frames[0])
        else:
            fg, ct, _ = make_frame(shape, t=t, seed=seed, variant=variant, **kw)
            frames.append((fg, ct))
    return np.stack([f[0] for f in frames]), np.stack([f[1] for f in frames])
