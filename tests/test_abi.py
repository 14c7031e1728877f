"""The C-ABI library loads and exports every symbol include/ultrack_b200.h declares
(no compute calls: this tier has no GPU)."""
import ctypes as C
import re
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent


def _declared():
    text = (ROOT / "include" / "ultrack_b200.h").read_text()
    return sorted(set(re.findall(r"ULK_API\s+[\w\s\*]+?\b(ulk_\w+)\s*\(", text)))


def test_header_declares_the_path():
    names = _declared()
    for must in ("ulk_segment_frame", "ulk_segment_frame_host", "ulk_link_frames", "ulk_labels_to_contours",
                 "ulk_frame_nodes_to_host", "ulk_frame_nodes_from_device", "ulk_last_error"):
        assert must in names


def test_cuda_library_exports_every_declared_symbol():
    from ultrack_b200.csrc.build import build
    lib = C.CDLL(str(build()))
    for name in _declared():
        assert hasattr(lib, name), f"{name} declared in include/ultrack_b200.h but not exported"


def test_python_binding_lists_every_declared_symbol():
    from ultrack_b200 import _lib
    assert sorted(_lib.EXPORTS) == _declared()


def test_struct_layouts_match_header(tmp_path):
    """sizeof / offsetof of every ctypes mirror against the header itself, compiled by gcc."""
    import subprocess
    from ultrack_b200 import _lib
    structs = {"ulk_seg_params": _lib.SegParams, "ulk_link_params": _lib.LinkParams, "ulk_links": _lib.Links,
               "ulk_seg_stats": _lib.SegStats, "ulk_frame_nodes": _lib.FrameNodes}
    lines = ['#include <stdio.h>', '#include <stddef.h>', f'#include "{ROOT / "include" / "ultrack_b200.h"}"', "int main(void){"]
    for cname, cls in structs.items():
        lines.append(f'printf("{cname} %zu\\n", sizeof({cname}));')
        for fname, _ in cls._fields_:
            lines.append(f'printf("{cname}.{fname} %zu\\n", offsetof({cname}, {fname}));')
    lines.append("return 0;}")
    src = tmp_path / "layout.c"
    src.write_text("\n".join(lines))
    exe = tmp_path / "layout"
    subprocess.run(["gcc", "-o", str(exe), str(src)], check=True)
    out = dict(l.split() for l in subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout.splitlines())
    for cname, cls in structs.items():
        assert int(out[cname]) == C.sizeof(cls), cname
        for fname, _ in cls._fields_:
            assert int(out[f"{cname}.{fname}"]) == getattr(cls, fname).offset, (cname, fname)
    assert C.sizeof(_lib.LinkParams) == 96


def test_product_path_fails_loudly_without_gpu(monkeypatch):
    """No CPU fallback: without a CUDA device get_lib() raises."""
    import torch
    from ultrack_b200 import _lib
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    monkeypatch.setattr(_lib, "_LIB", None)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        _lib.get_lib()


def test_package_never_imports_the_oracle():
    for path in (ROOT / "ultrack_b200").rglob("*.py"):
        text = path.read_text()
        assert "import oracle" not in text and "from oracle" not in text, path
    for path in (ROOT / "ultrack_b200" / "csrc").glob("*.*"):
        if path.suffix in (".cu", ".h"):
            assert "ws_oracle" not in path.read_text(), path
