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


def test_convc1_weight_layout_for_the_fused_lookup():
    """ops.pack_convc1_weight: every pyramid level's K*(Cg+1) columns of convc1.weight start at a multiple of
    Kl = K*(Cg+1) rounded up to 64, zeros in between (the layout fsb_lookup_convc1 documents in include/fsb200.h)."""
    import torch
    from foundationstereo_b200 import ops
    L, K, Cg = 2, 9, 28
    n = K * (Cg + 1)
    w = torch.arange(256 * L * n, dtype=torch.float32).reshape(256, L * n, 1, 1) / 1000.0
    p = ops.pack_convc1_weight(w, L, K, Cg, torch.float32)
    kl = 320
    assert p.shape == (256, L * kl) and p.is_contiguous()
    for l in range(L):
        assert torch.equal(p[:, l * kl:l * kl + n], w[:, l * n:(l + 1) * n, 0, 0])
        assert (p[:, l * kl + n:(l + 1) * kl] == 0).all()
    assert ops.pack_convc1_weight(w[:, :n], 1, K, Cg, torch.float16).shape == (256, kl)


def test_fused_stem_patch_keeps_the_module_and_restores_it():
    """install(model, fused_stem=True): corr_stem[0] stays in the module tree (state_dict keys for
    checkpoints) with an identity forward, the deferred build_gwc_volume returns a zero-channel tensor that torch.cat
    passes over, and uninstall() puts the convolution's forward back.  (The fused launch itself is a GPU test.)"""
    import types
    import torch
    from foundationstereo_b200 import patch

    class Stub(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.corr_stem = torch.nn.Sequential(torch.nn.Conv3d(32, 28, kernel_size=1))

    model = Stub()
    keys = sorted(model.state_dict().keys())
    x = torch.randn(1, 32, 2, 3, 4)
    want = model.corr_stem(x)
    gwc_fn, cat_fn = patch.install_stem(model)
    try:
        assert sorted(model.state_dict().keys()) == keys
        assert model.corr_stem(x) is x                              # identity: the fused kernel already applied it
        placeholder = gwc_fn(torch.zeros(2, 16, 3, 4), torch.zeros(2, 16, 3, 4), 5, 8)
        assert placeholder.shape == (2, 0, 5, 3, 4)
        y = torch.zeros(2, 28, 5, 3, 4)
        assert torch.cat([placeholder, y], dim=1).shape == y.shape   # what core/foundation_stereo.py:212 does with it
    finally:
        patch.uninstall(model)
    assert torch.equal(model.corr_stem(x), want)
