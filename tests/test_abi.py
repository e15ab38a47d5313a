"""The C-ABI library loads without a GPU and exports every symbol include/evr.h declares (no compute calls)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "evr.h")


def declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(evr_[a-z0-9_]+)\s*\(", src)))


def test_header_declares_the_hot_path():
    syms = declared_symbols()
    for s in ("evr_store_prepare", "evr_topk_search", "evr_topk_merge", "evr_rescore", "evr_vit_create",
              "evr_vit_encode", "evr_vit_destroy", "evr_last_error", "evr_launch_count"):
        assert s in syms


def test_library_exports_every_declared_symbol(pkg):
    lib = ctypes.CDLL(pkg._lib.LIB_PATH)
    missing = [s for s in declared_symbols() if not hasattr(lib, s)]
    assert not missing, missing


def test_python_binding_covers_header(pkg):
    assert sorted(pkg._lib.SIGNATURES.keys()) == declared_symbols()


def test_argument_validation_needs_no_gpu(pkg):
    lib = pkg._lib.lib
    assert lib.evr_version() >= 100
    n = ctypes.c_size_t(0)
    # D not a multiple of 64 -> status 2 + message, no CUDA call made
    assert lib.evr_topk_workspace_bytes(1000, 100, 1, 10, ctypes.byref(n)) != 0
    assert b"D=" in lib.evr_last_error()
    assert lib.evr_topk_workspace_bytes(1000, 512, 1, 0, ctypes.byref(n)) != 0
    with pytest.raises(pkg._lib.EvrError):
        pkg._lib.check(lib.evr_store_prepare(None, 0, 10, 512, None, None), "evr_store_prepare")


def test_no_oracle_import_in_product():
    pkg_dir = os.path.join(ROOT, "event-retrival-in-video-learning-transferable-visual-model-from-supervised-"
                                 "natural-language_b200")
    for dirpath, _, files in os.walk(pkg_dir):
        for f in files:
            if f.endswith(".py"):
                text = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", text, flags=re.M), f
