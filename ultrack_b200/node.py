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


# ---------------------------------------------------------------------------------------------
# Bulk blob writer.  `pickle.dumps(Node)` per node costs tens of
# microseconds of Python; at 2e5 nodes per frame that dwarfs the kernels.  The pickle stream of
# a Node is fully determined by its fields, so it is written for ALL nodes of a frame by one
# CUDA kernel (`ulk_frame_blobs_to_host`): a constant HEAD template, the raw bool mask bytes
# expanded from the bit-packed device mask, a constant TAIL template, with fixed-width holes
# patched per node.  To keep every hole at a fixed offset the stream uses the widest encoding
# of each value (BININT for ints, LONG1/8 for the id, BINBYTES with a 32-bit length) and no
# memo opcodes -- still a plain protocol-4 pickle that loads into the same object.
def _op_str(x: str) -> bytes:
    return _short_str(x)


def _op_global(module: str, name: str) -> bytes:
    return _op_str(module) + _op_str(name) + b"\x93"


def _np_reconstruct_module() -> str:
    return np.zeros(1).__reduce__()[0].__module__      # numpy._core.multiarray (numpy 2) / numpy.core.multiarray


def _op_dtype(code: str, order: str) -> bytes:
    return (_op_global("numpy", "dtype") + _op_str(code) + b"\x89\x88\x87R"
            + b"(K\x03" + _op_str(order) + b"NNNJ\xff\xff\xff\xffJ\xff\xff\xff\xffK\x00tb")


def _op_array_open(shape_ops: bytes, dtype_ops: bytes) -> bytes:
    """Everything of an ndarray pickle up to (excluding) the raw-data bytes opcode."""
    return (_op_global(_np_reconstruct_module(), "_reconstruct") + _op_global("numpy", "ndarray")
            + b"K\x00\x85" + b"C\x01b" + b"\x87R" + b"(K\x01" + shape_ops + dtype_ops + b"\x89")


_ARRAY_CLOSE = b"tb"


class BlobTemplate:
    """HEAD / TAIL byte templates and hole offsets of the Node pickle for `ndim`-D frames."""

    def __init__(self, ndim: int):
        self.ndim = ndim
        d = ndim
        head = bytearray(_NODE_PREFIX + b"}(")
        off = {}

        def hole(name: str, width: int):
            off[name] = len(head)
            head.extend(b"\x00" * width)

        head += _op_str("area") + b"J"; hole("area", 4)
        head += _op_str("frontier") + b"G"; hole("frontier", 8)
        head += _op_str("height") + b"G"; hole("height", 8)
        head += _op_str("_h_node_index") + b"J"; hole("hidx", 4)
        head += _op_str("_parent") + b"N"
        head += _op_str("bbox") + _op_array_open(b"K" + bytes([2 * d]) + b"\x85", _op_dtype("i8", "<")) + b"C" + bytes([16 * d])
        hole("bbox", 16 * d)
        head += _ARRAY_CLOSE
        head += _op_str("mask") + b"B"; hole("chunk_len", 4)
        # Blosc1 stored chunk header: version 2, versionlz 1, flags MEMCPYED, typesize 1, nbytes, blocksize, cbytes
        head += bytes([_BLOSC1_VERSION_FORMAT, 1, _BLOSC_MEMCPYED, 1]); hole("blosc_nbytes", 4); hole("blosc_blocksize", 4)
        hole("blosc_cbytes", 4)
        self.payload_start = len(head)
        head += b"\x80\x04"
        pre = (_op_global(_np_reconstruct_module(), "_reconstruct") + _op_global("numpy", "ndarray")
               + b"K\x00\x85" + b"C\x01b" + b"\x87R" + b"(K\x01")
        head += pre
        off["shape"] = len(head) + 1
        for _ in range(d):
            head += b"J\x00\x00\x00\x00"
        head += b"\x86" if d == 2 else b"\x87"
        head += _op_dtype("b1", "|") + b"\x89" + b"B"; hole("raw_len", 4)
        tail = bytearray(_ARRAY_CLOSE + b".")       # end of the mask array pickle (payload ends here)
        self.payload_tail = len(tail)
        toff = {}

        def thole(name: str, width: int):
            toff[name] = len(tail)
            tail.extend(b"\x00" * width)

        tail += _op_str("id") + b"\x8a\x08"; thole("id", 8)
        tail += _op_str("time") + b"J"; thole("time", 4)
        tail += _op_str("centroid") + _op_array_open(b"K" + bytes([d]) + b"\x85", _op_dtype("i8", "<")) + b"C" + bytes([8 * d])
        thole("centroid", 8 * d)
        tail += _ARRAY_CLOSE + b"ub."
        self.head, self.tail, self.off, self.toff = bytes(head), bytes(tail), off, toff

    def blob_sizes(self, volumes: np.ndarray) -> np.ndarray:
        return volumes.astype(np.int64) + (len(self.head) + len(self.tail))

    def spec(self):
        from ultrack_b200 import _lib
        o, t = self.off, self.toff
        return _lib.BlobSpec(len(self.head), len(self.tail), self.ndim, self.payload_start, self.payload_tail,
                             o["area"], o["frontier"], o["height"], o["hidx"], o["bbox"], o["chunk_len"],
                             o["blosc_nbytes"], o["blosc_blocksize"], o["blosc_cbytes"], o["shape"], o["raw_len"],
                             t["id"], t["time"], t["centroid"])


_TEMPLATES: Dict[int, BlobTemplate] = {}


def blob_template(ndim: int) -> BlobTemplate:
    if ndim not in _TEMPLATES:
        _TEMPLATES[ndim] = BlobTemplate(ndim)
    return _TEMPLATES[ndim]


def frame_blobs(lib, handle, nodes: Dict[str, np.ndarray], time: int, ndim: int, id_base: int):
    """All `nodes.pickle` blobs of a frame from the device node table: (uint8 buffer, offsets[n+1]).
    blob i = buffer[offsets[i]:offsets[i+1]] loads as the reference's Node (`ulk_frame_blobs_to_host`)."""
    import ctypes as C

    from ultrack_b200 import _lib
    tpl = blob_template(ndim)
    bb = nodes["bbox"].astype(np.int64)
    vol = np.prod(bb[:, 3:] - bb[:, :3], axis=1) if len(bb) else np.zeros(0, np.int64)
    off = np.zeros(len(bb) + 1, np.int64)
    np.cumsum(tpl.blob_sizes(vol), out=off[1:])
    buf = _pinned_bytes(int(off[-1]), lib)
    if len(bb):
        spec = tpl.spec()
        _lib.check(lib, lib.ulk_frame_blobs_to_host(
            handle, C.c_char_p(tpl.head), C.c_char_p(tpl.tail), C.byref(spec), int(id_base), int(time),
            off.ctypes.data_as(C.c_void_p), C.c_void_p(buf.ctypes.data), None))
    return buf, off


def _pinned_bytes(n: int, lib) -> np.ndarray:
    """uint8[n] in pinned host memory when a CUDA device is there (the D2H copy of the blobs runs
    at PCIe speed only from pinned pages); plain numpy under the host-emulation harness."""
    if n > 0 and not getattr(lib, "_ulk_is_emu", False):
        try:
            import torch
            return torch.empty(n, dtype=torch.uint8, pin_memory=True).numpy()
        except Exception:  # noqa: BLE001
            pass
    return np.empty(max(n, 1), np.uint8)[:n]
