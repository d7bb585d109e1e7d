"""CPU: the C-ABI library loads and exports every symbol include/fisher_b200.h declares (no compute)."""
import ctypes
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "fisher_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(fb_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_are_exported_and_bound():
    from fisher_nerf_customized_b200 import _lib, build
    build.build()
    L = ctypes.CDLL(_lib.LIB_PATH)
    names = _declared()
    assert len(names) >= 19
    for n in names:
        assert hasattr(L, n), f"{n} declared in include/fisher_b200.h but not exported"
        assert n in _lib.SIGNATURES, f"{n} has no ctypes signature"
    assert set(_lib.SIGNATURES) == set(names)


def test_host_only_entry_points():
    from fisher_nerf_customized_b200 import _lib
    L = _lib.lib()
    assert L.fb_version() >= 100
    assert L.fb_error_string(0) == b"ok"
    assert L.fb_error_string(_lib.FB_E_CAPACITY) == b"instance capacity exceeded"
    assert L.fb_geom_bytes(1000) >= 1000 * 32 and L.fb_geom_bytes(0) > 0
    assert L.fb_image_bytes(256, 256) >= 256 * 256 * 8
    assert L.fb_inst_bytes(10) >= 80
    assert L.fb_score_workspace_bytes(1000, 256, 256, 4, 1 << 16, 0) > 4 * 1000 * 112
    # records are sized by the visible capacity, not by the map
    assert L.fb_score_workspace_bytes(100000, 256, 256, 4, 1 << 16, 25000) < L.fb_score_workspace_bytes(100000, 256, 256, 4, 1 << 16, 0) // 2
    assert L.fb_hessian_workspace_bytes(1000, 256, 256, 4, 1 << 16) > 4 * 1000 * 112
    assert L.fb_launch_count() == 0
    assert L.fb_knn_workspace_bytes(1000) > 0


def test_struct_layout_matches_header():
    from fisher_nerf_customized_b200 import _lib
    assert ctypes.sizeof(_lib.fb_camera) == 4 * (8 + 3 + 3 + 16 + 16)
    assert ctypes.sizeof(_lib.fb_map) == 8 + 5 * 8
    assert ctypes.sizeof(_lib.fb_gaussians) == 8 + 7 * 8


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "fisher_nerf_customized_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".inl", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle|#include\s+\".*oracle|liboracle|oracle\.", text, re.M), \
                    f"{f} uses the oracle"


def test_missing_library_fails_loudly(monkeypatch, tmp_path):
    from fisher_nerf_customized_b200 import _lib
    import pytest
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", str(tmp_path / "nope.so"))
    with pytest.raises(RuntimeError):
        _lib.lib()
