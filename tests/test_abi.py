"""The C-ABI library loads and exports every symbol include/fsb200.h declares (no compute
calls: this runs without a GPU), and argument errors are reported through the ABI."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    with open(os.path.join(ROOT, "include", "fsb200.h")) as fh:
        src = fh.read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(fsb_[a-z0-9_]+)\s*\(", src)))


@pytest.fixture(scope="module")
def lib():
    from foundationstereo_b200 import build, _lib
    build.build()
    return _lib.lib()


def test_every_declared_symbol_is_exported(lib):
    names = declared_symbols()
    assert "fsb_volume_gwc_concat" in names and "fsb_geo_lookup" in names and len(names) >= 7
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/fsb200.h but not exported by libfsb200.so"


def test_version_and_error_channel(lib):
    assert lib.fsb_version() >= 100
    # bad arguments are rejected before any CUDA call, with a message
    rc = lib.fsb_volume_gwc_concat(None, None, None, None, None, None, 1, 8, 4, 8, 2, 2, 2, 0, None)
    assert rc == -1
    assert b"NULL" in lib.fsb_last_error()
    rc = lib.fsb_geo_lookup(None, None, None, None, None, 1, 28, 8, 2, 2, 2, 2, 4, 0, None)
    assert rc == -1
    assert lib.fsb_disparity_regression(None, None, 1, 8, 2, 2, 0, 0, None) == -1
    assert lib.fsb_context_upsample(None, None, None, 1, 2, 2, 0, 1.0, 0, None) == -1


def test_python_ops_refuse_cpu_tensors():
    import torch
    import foundationstereo_b200 as fsb
    x = torch.zeros(1, 16, 2, 4)
    with pytest.raises(Exception):
        fsb.build_gwc_volume(x, x, 2, 8)
    with pytest.raises(Exception):
        fsb.SelectiveUpdateBlock({}, "cpu")


def test_patch_rebinds_reference_names():
    """install() swaps the three names in a module that holds them (stand-in for
    core.foundation_stereo, which is not on the GPU box) and uninstall() restores them."""
    import types
    import foundationstereo_b200 as fsb
    mod = types.ModuleType("core.foundation_stereo")
    sentinel = object()
    mod.build_gwc_volume = mod.build_concat_volume = mod.Combined_Geo_Encoding_Volume = sentinel
    mod.disparity_regression = mod.context_upsample = sentinel
    fsb.install(modules=[mod])
    assert mod.build_gwc_volume is fsb.build_gwc_volume
    assert mod.build_concat_volume is fsb.build_concat_volume
    assert mod.Combined_Geo_Encoding_Volume is fsb.Combined_Geo_Encoding_Volume
    assert mod.disparity_regression is fsb.disparity_regression
    assert mod.context_upsample is fsb.context_upsample
