"""CPU-side checks of the drop-in boundary: libimgtx.so loads without a GPU, exports every symbol include/imgtx.h
declares (and the ctypes table binds exactly that set), reports errors through the C ABI instead of raising/aborting,
and the product package never imports the oracle."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "imgtx.h")


def declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(imgtx_[a-z0-9_]+)\s*\(", src)))


def test_header_declares_entry_points():
    names = declared_symbols()
    assert len(names) >= 30
    for must in ("imgtx_pointwise_u8", "imgtx_geometry_u8", "imgtx_augmix_f32", "imgtx_views_f32",
                 "imgtx_sign_step_f32", "imgtx_cc_pointwise_u8", "imgtx_cc_jpeg_u8", "imgtx_to_tensor_f32"):
        assert must in names


def test_library_exports_every_declared_symbol():
    from imagetransformations_b200 import _lib
    assert os.path.exists(_lib.LIB_PATH), "libimgtx.so has not been built (python __graft_entry__.py)"
    lib = ctypes.CDLL(_lib.LIB_PATH)
    missing = [n for n in declared_symbols() if not hasattr(lib, n)]
    assert not missing, f"declared in imgtx.h but not exported: {missing}"


def test_ctypes_table_matches_header():
    from imagetransformations_b200 import _lib
    assert sorted(_lib.exported_names()) == declared_symbols()


def test_version_and_error_reporting_without_gpu():
    from imagetransformations_b200 import _lib
    lib = _lib.load()
    assert lib.imgtx_version() >= 100
    # invalid arguments are reported by status + thread-local message; nothing is launched
    rc = lib.imgtx_pointwise_u8(None, None, 4, 3072, 99, 1.0, None, None)
    assert rc < 0
    msg = lib.imgtx_last_error()
    assert msg and len(msg) > 0
    with pytest.raises(_lib.ImgtxError):
        _lib.check(rc)
    # n == 0 is a no-op that succeeds
    assert lib.imgtx_pointwise_u8(None, None, 0, 3072, 0, 1.0, None, None) == 0
    # host helpers run without a device
    n = lib.imgtx_resample_table_words(32, 36, _lib.FILTER_BILINEAR)
    assert n > 0
    buf = (ctypes.c_int32 * n)()
    assert lib.imgtx_resample_table_host(32, 36, _lib.FILTER_BILINEAR, buf) == 0
    assert any(buf)


def test_product_never_imports_oracle():
    """The oracle is test infrastructure: no module of the product package may reference it."""
    pkg = os.path.join(ROOT, "imagetransformations_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", text, flags=re.M), f
                assert "/root/reference" not in text, f
