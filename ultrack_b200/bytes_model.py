"""Algorithmic (compulsory) byte model of the path -- what bench.py's `roofline` divides by time.

Whole pipeline, per frame:
    B_alg = V*(s_fg + s_ct) + 4*n_fg + 64*N + mask_bytes + 8*P
every input byte read once, one int32 finest-region label per foreground voxel, a 64-byte
record per emitted node, the bit-packed masks, 8 bytes per overlap pair.
Linking, per frame pair:  20 bytes per emitted link + per evaluated pair 2*36 bytes of
bbox/centroid records + the mask words of both bbox-overlap windows (bounded above by the two
masks themselves; the model charges min(mask_i, mask_j) words twice).

Per kernel: every array the kernel must read or write once per voxel (its own voxel's
element; neighbour elements belong to other threads and are counted once there).  Values are
bytes per voxel of the frame for float32 inputs; kernels that run on lists (basins, RAG edges,
candidates) are not voxel-bound and carry no entry.
"""
from __future__ import annotations

from typing import Dict, Optional

# kernel name (as reported by ulk_profile_report) -> bytes per frame voxel per launch
KERNEL_BYTES_PER_VOXEL: Dict[str, float] = {
    "K_Threshold": 4 + 4 + 4 + 4,        # fg, ct in; staged ct, parent out
    "K_CclUnion": 4 + 4,                 # parent in / out
    "K_Flatten": 4 + 4,
    "K_Compress": 4 + 4,
    "K_CountByRoot": 4,
    "K_CcStats": 4,
    "K_RemoveSmall": 4 + 1,
    "K_ApplyDrop": 1 + 4,
    "K_FlagRoots": 4 + 4,
    "K_FlagBigRoots": 4 + 4,
    "K_LabelBigCc": 4 + 4,
    "K_LabelSmallOnly": 4 + 4,
    "K_WriteLabels": 4 + 4 + 4,
    "K_MinEdge": 4 + 4 + 4 + 4,          # label, ct in; min weight, min key out
    "K_Hook1": 4 + 4 + 4 + 4 + 1 + 4 + 4 + 4,
    "K_CandInit": 4 + 4 + 1,
    "K_BasinRootInfo": 4,
    "K_AssignBasin": 4 + 4 + 4,
    "K_BasinAccum": 4 + 4 + 4 + 4 + 4 + 4,
    "K_JoinKey": 4 + 4 + 4 + 4,
    "K_Attach": 4 + 4 + 4,
    "K_RagInsert": 4 + 4 + 4,
    "K_MinKeyInBasin": 4,
    "K_K3Init": 4 + 4,
    "K_K3Union": 4 + 4,
    "K_K3CountRoots": 4 + 4,
    "K_MaskPaint": 4,
}


def frame_bytes(voxels: int, fg_itemsize: int, ct_itemsize: int, n_foreground: int, n_nodes: int,
                mask_words: int, n_overlaps: int) -> int:
    return (voxels * (fg_itemsize + ct_itemsize) + 4 * n_foreground + 64 * n_nodes + 4 * mask_words
            + 8 * n_overlaps)


def link_bytes(n_links: int, n_pairs: int, mean_mask_words: float) -> int:
    return int(20 * n_links + n_pairs * (2 * 36 + 2 * 4 * mean_mask_words))


def kernel_bytes(name: str, voxels: int) -> Optional[float]:
    base = name.split("<")[0]
    bpv = KERNEL_BYTES_PER_VOXEL.get(base)
    return None if bpv is None else bpv * voxels
