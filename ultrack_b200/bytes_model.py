"""Algorithmic (compulsory) byte model of the path -- what bench.py's `roofline` divides by time.

Whole pipeline, per frame:
    B_alg = V*(s_fg + s_ct) + 4*n_fg + 64*N + mask_bytes + 8*P
every input byte read once, one int32 finest-region label per foreground voxel, a 64-byte
record per emitted node, the bit-packed masks, 8 bytes per overlap pair.
Linking, per frame pair:  20 bytes per emitted link + per evaluated pair 2*36 bytes of
bbox/centroid records + the mask words of both bbox-overlap windows (bounded above by the two
masks themselves; the model charges min(mask_i, mask_j) words twice).

Per kernel: every array element the kernel must read or write once for its own voxel
(neighbour elements belong to other threads and are counted once there).  Dense kernels run
over all V voxels of the frame; list kernels run over the n_fg foreground voxels only and also
read their 4-byte list entry.  Kernels that run on small lists (basins, RAG edges, candidates,
roots) are not voxel-bound and carry no entry.
"""
from __future__ import annotations

from typing import Dict, Optional

# kernel name (as reported by ulk_profile_report) -> bytes per frame voxel per launch (float32 foreground)
DENSE_BYTES_PER_VOXEL: Dict[str, float] = {
    "K_Threshold": 4 + 0.125,            # foreground in; one foreground bit out (contours are not touched here)
}
# the tile-resident front end reads the dense contour volume once (TMA bricks; the halo is re-read
# from L2, not counted) and writes six rank-space arrays per foreground voxel
FRONT_TILE = {"dense": 4 + 0.25, "fg": 6 * 4}    # contours + words / offsets in; fgl, ct, forest, mw, pk, po out
# kernel name -> bytes per FOREGROUND voxel per launch.  Every array is rank-indexed (n_fg long);
# an entry is what the kernel must read or write once for its own voxel.
LIST_BYTES_PER_FG_VOXEL: Dict[str, float] = {
    "K_Expand": 4 + 4 + 4 + 4,           # contour in (from the volume); list, rank-space contour, forest out
    "K_CclWords": 0.5 + 4 + 4,           # segment records in, forest in / out
    "K_FlattenCount": 4 + 4,             # forest in / out
    "K_LabelCc": 4 + 4,                  # forest in, label out (in place)
    "K_WriteLabels": 4 + 4 + 4,          # label, pass-A basin in; label out
    "K_MinEdge": 4 + 4 + 4 + 4 + 4 + 4,  # list, label, contour in; min weight, min key, other endpoint out (pass B: the pass-A voxels only)
    "K_Hook1": 4 + 4 + 4,                # endpoint, the endpoint's endpoint in; forest out
    "K_CompressCand": 4 + 4 + 4,         # forest in / out, weight
    "K_AssignBasin": 4 + 4 + 4,          # forest (two hops) in; basin out
    "K_BasinAccum": 4 + 4 + 4 + 4 + 4 + 4,   # list, basin, weight, key, endpoint in; join key out
    "K_Attach": 4 + 4 + 4,               # basin, weight, join key
    "K_MaskPaint": 4 + 4 + 0.125,        # list, basin in; one mask bit out (per emitted ancestor)
}


def frame_bytes(voxels: int, fg_itemsize: int, ct_itemsize: int, n_foreground: int, n_nodes: int,
                mask_words: int, n_overlaps: int) -> int:
    return (voxels * (fg_itemsize + ct_itemsize) + 4 * n_foreground + 64 * n_nodes + 4 * mask_words
            + 8 * n_overlaps)


def link_bytes(n_links: int, n_pairs: int, mean_mask_words: float) -> int:
    return int(20 * n_links + n_pairs * (2 * 36 + 2 * 4 * mean_mask_words))


# list kernels of the voxel stage: launched once on the voxels of the components above max_area
# (oversegmentation pass A) and once on the whole foreground (pass B) per frame
TWO_PASS = {"K_Hook1", "K_CompressCand", "K_AssignBasin", "K_BasinAccum", "K_Attach"}


def kernel_bytes(name: str, voxels: int, n_foreground: int, pass_a_voxels: Optional[int] = None) -> Optional[float]:
    """Algorithmic bytes of one launch of kernel `name` on a frame of `voxels` voxels of which
    `n_foreground` are foreground.  For the two-pass kernels, when `pass_a_voxels` is given, the
    figure is the AVERAGE over the frame's two launches (pass A on `pass_a_voxels`, pass B on
    `n_foreground`), matching an average launch duration."""
    base = name.split("<")[0]
    if base == "K_FrontTile":
        return FRONT_TILE["dense"] * voxels + FRONT_TILE["fg"] * n_foreground
    if base == "K_MinEdge":       # one launch per frame since the front end: the voxels of the oversegmented components
        return LIST_BYTES_PER_FG_VOXEL[base] * (pass_a_voxels if pass_a_voxels else n_foreground)
    if base in DENSE_BYTES_PER_VOXEL:
        return DENSE_BYTES_PER_VOXEL[base] * voxels
    if base in LIST_BYTES_PER_FG_VOXEL:
        n = n_foreground
        if base in TWO_PASS and pass_a_voxels is not None and pass_a_voxels > 0:
            n = (n_foreground + pass_a_voxels) / 2.0
        return LIST_BYTES_PER_FG_VOXEL[base] * n
    return None
