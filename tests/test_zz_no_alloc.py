"""fsb_edgenext_block_noalloc (csrc/noalloc.cu): the EdgeNeXt block with its depthwise scratch in a static device array
of the library, for callers that cannot call cudaMalloc.  No counterpart in the reference,
so the check is against the library's EXISTING entry points on the same inputs: fsb_edgenext_block
and the pair fsb_dwconv7x7_nhwc + fsb_edgenext_mlp with a caller-owned buffer, all three documented as bit-identical.

Named test_zz_* so that pytest collects it last; it passes alone in a fresh process
and after the other tests (the scratch array keeps whatever an earlier call left in it; nothing has to start at zero).
"""
import pytest
import torch

pytestmark = pytest.mark.gpu

C_, F_ = 128, 512


def _rand(*shape, seed=0):
    g = torch.Generator().manual_seed(seed)
    return torch.randn(*shape, generator=g)


def _params(dtype, seed):
    dw_w, dw_b = (_rand(C_, 49, seed=seed + 2) / 7).cuda(), (_rand(C_, seed=seed + 3) * 0.1).cuda()
    w1 = (_rand(F_, C_, seed=seed + 4) / C_ ** 0.5).to(dtype).cuda()
    w2 = (_rand(C_, F_, seed=seed + 5) / F_ ** 0.5).to(dtype).cuda()
    b1, b2, gamma = (_rand(F_, seed=seed + 6) * 0.1).cuda(), (_rand(C_, seed=seed + 7) * 0.1).cuda(), _rand(C_, seed=seed + 8).cuda()
    return dw_w, dw_b, w1, b1, w2, b2, gamma


def _call(lib, name, x, prm, out):
    """Straight through the C ABI (ctypes): name(x, dw_w, dw_b, w1, b1, w2, b2, gamma, out, B, H, W, C, F, dtype, stream)."""
    dw_w, dw_b, w1, b1, w2, b2, gamma = prm
    B, H, W, C = x.shape
    dt = 1 if x.dtype == torch.float16 else 2
    return getattr(lib, name)(x.data_ptr(), dw_w.data_ptr(), dw_b.data_ptr(), w1.data_ptr(), b1.data_ptr(), w2.data_ptr(),
                              b2.data_ptr(), gamma.data_ptr(), out.data_ptr(), B, H, W, C, F_, dt,
                              torch.cuda.current_stream().cuda_stream)


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
def test_noalloc_block_matches_existing_entry_points_twice_in_a_row(dtype):
    from foundationstereo_b200 import _lib
    lib = _lib.lib()
    assert lib.fsb_noalloc_max_pixels() == 32768
    # two calls in a row on DIFFERENT inputs and shapes: the second (smaller) call finds the first call's depthwise
    # output in the shared array and must not be affected by it
    cases = [((2, 96, 160), 1), ((1, 13, 37), 11)]
    data = []
    for shape, seed in cases:
        B, H, W = shape
        x = _rand(B, H, W, C_, seed=seed).to(dtype).cuda()
        data.append((x, _params(dtype, seed), torch.full_like(x, 7.0), torch.full_like(x, -3.0), torch.empty_like(x),
                     torch.full_like(x, 5.0)))
    import gc
    gc.collect()   # garbage of earlier tests must not be released between the two readings below
    torch.cuda.synchronize()
    for i, (x, prm, got, want_block, tmp, want_pair) in enumerate(data):
        before = torch.cuda.memory_allocated(), torch.cuda.mem_get_info()[0]
        rc = _call(lib, "fsb_edgenext_block_noalloc", x, prm, got)
        assert rc == 0, lib.fsb_last_error()
        torch.cuda.synchronize()
        after = torch.cuda.memory_allocated(), torch.cuda.mem_get_info()[0]
        assert after[0] <= before[0], "the call went through torch's allocator"
        if i > 0:  # (the first call may load the kernels' module images lazily)
            assert after[1] >= before[1], f"device memory shrank by {before[1] - after[1]} bytes across a no-alloc call"
    for x, prm, got, want_block, tmp, want_pair in data:
        dw_w, dw_b, w1, b1, w2, b2, gamma = prm
        B, H, W, _ = x.shape
        dt = 1 if dtype == torch.float16 else 2
        s = torch.cuda.current_stream().cuda_stream
        assert _call(lib, "fsb_edgenext_block", x, prm, want_block) == 0, lib.fsb_last_error()
        assert lib.fsb_dwconv7x7_nhwc(x.data_ptr(), dw_w.data_ptr(), dw_b.data_ptr(), tmp.data_ptr(), B, C_, H, W, dt, s) == 0
        assert lib.fsb_edgenext_mlp(tmp.data_ptr(), w1.data_ptr(), b1.data_ptr(), w2.data_ptr(), b2.data_ptr(),
                                    gamma.data_ptr(), x.data_ptr(), want_pair.data_ptr(), B * H * W, C_, F_, dt, s) == 0
    torch.cuda.synchronize()
    for x, prm, got, want_block, tmp, want_pair in data:
        assert torch.isfinite(got.float()).all()
        # tolerance: the three forms run the same arithmetic in the same order -> bit-identical (the tests of
        # fsb_edgenext_block in test_gpu_conv.py use torch.equal as well)
        assert torch.equal(got, want_pair), float((got.float() - want_pair.float()).abs().max())
        assert torch.equal(got, want_block), float((got.float() - want_block.float()).abs().max())


def test_noalloc_block_rejects_sizes_beyond_the_static_scratch():
    from foundationstereo_b200 import _lib
    lib = _lib.lib()
    dtype = torch.float16
    x = torch.zeros(1, 1, 8, C_, dtype=dtype, device="cuda")  # the check precedes every access: sizes are only claimed
    out = torch.empty_like(x)
    prm = _params(dtype, 3)
    dw_w, dw_b, w1, b1, w2, b2, gamma = prm
    s = torch.cuda.current_stream().cuda_stream
    args = (x.data_ptr(), dw_w.data_ptr(), dw_b.data_ptr(), w1.data_ptr(), b1.data_ptr(), w2.data_ptr(), b2.data_ptr(),
            gamma.data_ptr(), out.data_ptr())
    rc = lib.fsb_edgenext_block_noalloc(*args, 2, 128, 129, C_, F_, 1, s)  # 33024 pixels > 32768
    assert rc == -2
    msg = lib.fsb_last_error().decode()
    assert "32768" in msg and "exceed" in msg, msg
    rc = lib.fsb_edgenext_block_noalloc(*args, 1, 1, 8, 64, F_, 1, s)      # channel count the scratch is not laid out for
    assert rc == -2 and "C=64" in lib.fsb_last_error().decode()
    torch.cuda.synchronize()
