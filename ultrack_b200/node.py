"""Node records and the `nodes.pickle` blob format.

The reference stores `pickle.dumps(Node)` per candidate segment
(ultrack/core/segmentation/processing.py:254-267); `Node.__getstate__` replaces the bool
mask by `blosc2.pack_array(mask)` (ultrack/core/segmentation/node.py:270-278).  Every
consumer after the solver (export, UltrackArray, linking in the reference) unpickles these
blobs, so a drop-in has to emit bytes that `pickle.loads` turns into an
`ultrack.core.segmentation.node.Node` when ultrack is installed.

* `dumps_node` writes that pickle (protocol 4: STACK_GLOBAL of the reference's class,
  NEWOBJ, state dict, BUILD) without importing ultrack.
* `pack_array` / `unpack_array` follow blosc2's `pack_array` container -- a Blosc chunk
  whose payload is `pickle.dumps(ndarray)`.  With blosc2 installed the real codec is used;
  without it the chunk is written in Blosc's "memcpyed" (stored, flag 0x02) form, which is
  a valid chunk for any Blosc1/Blosc2 decoder, and only stored chunks can be read back.
  blosc2 is not installable in the build container, so byte-compatibility of the stored
  chunk with c-blosc2's decoder is by the format specification, not by test.
* `loads_node` reads either kind into this module's `Node` (or the reference's own class
  when ultrack is importable).
"""
from __future__ import annotations

import io
import pickle
import struct
from typing import Any, Dict, Optional

import numpy as np

try:  # pragma: no cover - not available in the build container
    import blosc2 as _blosc2
except Exception:  # noqa: BLE001
    _blosc2 = None

REF_NODE_MODULE = "ultrack.core.segmentation.node"
REF_NODE_NAME = "Node"

_BLOSC_MEMCPYED = 0x02
_BLOSC1_VERSION_FORMAT = 2
_BLOSC_HEADER = 16


def pack_array(arr: np.ndarray) -> bytes:
    """blosc2.pack_array(arr): Blosc chunk of pickle.dumps(arr, HIGHEST_PROTOCOL)."""
    if _blosc2 is not None:
        return _blosc2.pack_array(arr)
    payload = pickle.dumps(np.ascontiguousarray(arr), pickle.HIGHEST_PROTOCOL)
    n = len(payload)
    if n > 2**31 - 1 - _BLOSC_HEADER:
        raise ValueError("mask too large for a single Blosc chunk")
    # version, versionlz, flags, typesize, nbytes, blocksize, cbytes (little endian)
    header = struct.pack("<BBBBIII", _BLOSC1_VERSION_FORMAT, 1, _BLOSC_MEMCPYED, 1, n, n, n + _BLOSC_HEADER)
    return header + payload


def unpack_array(packed: bytes) -> np.ndarray:
    """blosc2.unpack_array(packed)."""
    if _blosc2 is not None:
        return _blosc2.unpack_array(packed)
    packed = bytes(packed)
    if len(packed) < _BLOSC_HEADER:
        raise ValueError("not a Blosc chunk")
    _, _, flags, _, nbytes, _, cbytes = struct.unpack("<BBBBIII", packed[:_BLOSC_HEADER])
    if not flags & _BLOSC_MEMCPYED:
        raise RuntimeError(
            "this mask blob is Blosc-compressed; install blosc2 to read databases written by "
            "the reference (ultrack_b200 itself writes stored chunks when blosc2 is absent)")
    # an extended (Blosc2) header carries 16 more bytes; flags 0x01|0x04 both set mark it
    overhead = 32 if (flags & 0x01) and (flags & 0x04) else _BLOSC_HEADER
    if cbytes != nbytes + overhead or len(packed) < cbytes:
        raise ValueError("corrupt stored Blosc chunk")
    return pickle.loads(packed[overhead:overhead + nbytes])


class Node:
    """Attribute-compatible stand-in for `ultrack.core.segmentation.node.Node`
    (state keys of vendored/node.py:13-25 and node.py:121-131)."""

    def __init__(self, h_node_index: int = -1, id: int = -1, time: int = -1, **kwargs):
        self.area = None
        for key, value in kwargs.items():
            setattr(self, key, value)
        self._h_node_index = h_node_index
        self._parent = None
        self.bbox = kwargs.get("bbox")
        self.mask = kwargs.get("mask")
        self.id = id
        self.time = time
        self.centroid = kwargs.get("centroid")

    def __getstate__(self) -> Dict:
        d = self.__dict__.copy()
        d["_parent"] = None
        d["mask"] = pack_array(self.mask)
        return d

    def __setstate__(self, d: Dict) -> None:
        d["mask"] = unpack_array(d["mask"])
        self.__dict__ = d

    def IoU(self, other: "Node") -> float:
        raise RuntimeError("Node.IoU runs on the GPU (ulk_link_frames); ultrack_b200 has no CPU path")


def _state(h_node_index: int, id: int, time: int, area: int, frontier: float, height: float,
           bbox: np.ndarray, mask: np.ndarray, centroid: np.ndarray) -> Dict[str, Any]:
    # key order of the reference instance __dict__ (node.py:121-131, vendored/node.py:13-25)
    return {
        "area": int(area), "frontier": float(frontier), "height": float(height),
        "_h_node_index": int(h_node_index), "_parent": None,
        "bbox": bbox, "mask": pack_array(mask), "id": int(id), "time": int(time), "centroid": centroid,
    }


def _short_str(s: str) -> bytes:
    b = s.encode("utf-8")
    return b"\x8c" + bytes([len(b)]) + b        # SHORT_BINUNICODE


_NODE_PREFIX = b"\x80\x04" + _short_str(REF_NODE_MODULE) + _short_str(REF_NODE_NAME) + b"\x93)\x81"


def dumps_node(h_node_index: int, id: int, time: int, area: int, frontier: float, height: float,
               bbox: np.ndarray, mask: np.ndarray, centroid: np.ndarray) -> bytes:
    """pickle.dumps(ultrack Node) -- protocol 4, same opcode structure as copyreg.__reduce_ex__:
    PROTO 4 | STACK_GLOBAL(module, name) | EMPTY_TUPLE | NEWOBJ | <state dict> | BUILD | STOP."""
    body = pickle.dumps(_state(h_node_index, id, time, area, frontier, height, bbox, mask, centroid), protocol=4)
    # strip PROTO (2 bytes), the FRAME opcode (1 + 8 bytes) when present, and the final STOP
    start = 2
    if body[start:start + 1] == b"\x95":
        start += 9
    return _NODE_PREFIX + body[start:-1] + b"b."


class _NodeUnpickler(pickle.Unpickler):
    def find_class(self, module: str, name: str):
        if module == REF_NODE_MODULE and name == REF_NODE_NAME:
            try:
                return super().find_class(module, name)
            except Exception:  # noqa: BLE001  (ultrack not installed)
                return Node
        return super().find_class(module, name)


def loads_node(blob: bytes) -> Node:
    return _NodeUnpickler(io.BytesIO(bytes(blob))).load()
