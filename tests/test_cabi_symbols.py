"""CPU: libbff.so loads and exports every symbol include/bff.h declares (no compute call is made),
and the Python struct mirrors sizeof(bff_lift_desc)."""
import ctypes
import os
import re
import subprocess

from conftest import ROOT


def _declared():
    text = open(os.path.join(ROOT, "include", "bff.h")).read()
    return sorted(set(re.findall(r"BFF_API\s+[\w\s\*]+?\b(bff_\w+)\s*\(", text)))


def test_every_declared_symbol_is_exported_and_bound():
    from bff_b200 import _cabi
    names = _declared()
    assert len(names) >= 19 and "bff_lift_scene" in names and "bff_mask_iou" in names
    h = _cabi.lib()                        # raises if the library is missing: there is no fallback
    for n in names:
        assert hasattr(h, n), f"{n} declared in bff.h but not exported by libbff.so"
        assert n in _cabi.SYMBOLS, f"{n} has no ctypes prototype in _cabi.SYMBOLS"
    assert sorted(_cabi.SYMBOLS) == names
    assert h.bff_version() == 100


def test_struct_layout_matches_c(tmp_path):
    from bff_b200 import _cabi
    src = tmp_path / "sz.c"
    src.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "bff.h"\nint main(){printf("%zu %zu %zu %zu",'
                   'sizeof(bff_lift_desc),offsetof(bff_lift_desc,depth_thresh),offsetof(bff_lift_desc,rows),'
                   'offsetof(bff_lift_desc,frames_per_chunk));return 0;}\n')
    exe = tmp_path / "sz"
    subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)])
    size, o1, o2, o3 = (int(x) for x in subprocess.check_output([str(exe)]).split())
    D = _cabi.LiftDesc
    assert (ctypes.sizeof(D), D.depth_thresh.offset, D.rows.offset, D.frames_per_chunk.offset) == (size, o1, o2, o3)


def test_missing_library_fails_loudly(monkeypatch):
    from bff_b200 import _cabi
    monkeypatch.setattr(_cabi, "_lib", None)
    monkeypatch.setattr(_cabi, "LIB_PATH", "/nonexistent/libbff.so")
    try:
        _cabi.lib()
    except _cabi.BffError as e:
        assert "no CPU" in str(e)
    else:
        raise AssertionError("expected BffError")


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "beyond-fixed-forms_b200")
    for dp, _, fns in os.walk(pkg):
        for fn in fns:
            if fn.endswith((".py", ".cu", ".cuh")):
                text = open(os.path.join(dp, fn)).read()
                assert "reference_port" not in text and "bff_oracle" not in text and "import oracle" not in text, fn
    for fn in ("tools/projection_2d_to_3d.py", "tools/refinement.py"):
        assert "oracle" not in open(os.path.join(ROOT, fn)).read()
