"""The allocation-free BatchNorm-statistics entry point, called through the C ABI.

``tfop_bn_fwd_stats_noalloc`` exists for callers that are not allowed to allocate device
memory while they run (the GPU belongs to the host application's pool allocator, or a
cudaMalloc-induced device synchronisation would break a latency budget), so it carries its
scratch in a built-in ``__device__`` array instead of taking a workspace. This test drives it
the way such a caller would — plain ``ctypes`` on the shared library, raw device pointers —
and checks, on two calls in a row (the array is shared and keeps the previous call's
contents), that it produces what the workspace-taking entry point ``tfop_bn_fwd_stats``
produces. The file name keeps it last in collection order.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _lib():
    from tf_operator_b200.ops import _lib
    path = _lib.lib_path()
    assert path.exists(), f"{path} is missing: run __graft_entry__.build()"
    lib = C.CDLL(str(path))
    vp, f32 = C.c_void_p, C.c_float
    common = [vp, C.c_size_t, C.c_int, vp, vp, C.c_int, vp, vp, f32, f32, vp, vp, vp]
    lib.tfop_bn_fwd_stats.restype = C.c_int
    lib.tfop_bn_fwd_stats.argtypes = common + [vp, vp]
    lib.tfop_bn_fwd_stats_noalloc.restype = C.c_int
    lib.tfop_bn_fwd_stats_noalloc.argtypes = common + [vp]
    lib.tfop_bn_workspace_bytes.restype = C.c_size_t
    lib.tfop_last_error.restype = C.c_char_p
    return lib


def _run(lib, fn_name, x, gamma, beta, workspace=None):
    m = x.shape[0] * x.shape[2] * x.shape[3]
    c = x.shape[1]
    dev = x.device
    rm = torch.zeros(c, device=dev)
    rv = torch.ones(c, device=dev)
    ss = torch.empty(2 * c, device=dev)
    mean = torch.empty(c, device=dev)
    invstd = torch.empty(c, device=dev)
    stream = torch.cuda.current_stream().cuda_stream
    args = [x.data_ptr(), m, c, gamma.data_ptr(), beta.data_ptr(), 0, rm.data_ptr(), rv.data_ptr(),
            0.1, 1e-5, ss.data_ptr(), mean.data_ptr(), invstd.data_ptr()]
    if workspace is not None:
        args.append(workspace.data_ptr())
    rc = getattr(lib, fn_name)(*args, stream)
    assert rc == 0, lib.tfop_last_error()
    torch.cuda.synchronize()
    return ss, mean, invstd, rm, rv


def test_noalloc_entry_point_matches_workspace_entry_point_twice():
    lib = _lib()
    dev = torch.device("cuda", 0)
    torch.manual_seed(11)
    ws = torch.zeros(lib.tfop_bn_workspace_bytes(), dtype=torch.uint8, device=dev)
    # two calls in a row, different shapes: the second one sees the first one's leftovers
    for shape in [(8, 256, 14, 14), (4, 64, 28, 28)]:
        x = (torch.randn(shape, device=dev) * 2 + 0.5).bfloat16().contiguous(
            memory_format=torch.channels_last)
        gamma = torch.rand(shape[1], device=dev) + 0.5
        beta = torch.randn(shape[1], device=dev)
        ref = _run(lib, "tfop_bn_fwd_stats", x, gamma, beta, ws)
        got = _run(lib, "tfop_bn_fwd_stats_noalloc", x, gamma, beta)
        # same kernel behind both entry points: only the order of the fp32 atomics differs
        for r, g in zip(ref, got):
            torch.testing.assert_close(g, r, rtol=1e-5, atol=1e-5)
        # and every output agrees with the fp32 definition of batch-norm statistics
        xf = x.float()
        mean = xf.mean(dim=(0, 2, 3))
        var = xf.var(dim=(0, 2, 3), unbiased=False)
        invstd = torch.rsqrt(var + 1e-5)
        ss, got_mean, got_invstd, rm, rv = got
        c = shape[1]
        torch.testing.assert_close(got_mean, mean, rtol=1e-4, atol=1e-4)
        torch.testing.assert_close(got_invstd, invstd, rtol=1e-3, atol=1e-4)
        torch.testing.assert_close(ss[:c], gamma * invstd, rtol=1e-3, atol=1e-4)
        torch.testing.assert_close(ss[c:], beta - mean * gamma * invstd, rtol=1e-3, atol=1e-3)
        n = xf.numel() // c
        torch.testing.assert_close(rm, 0.1 * mean, rtol=1e-4, atol=1e-4)
        torch.testing.assert_close(rv, 0.9 + 0.1 * var * n / (n - 1), rtol=1e-3, atol=1e-4)


def test_noalloc_entry_point_rejects_sizes_beyond_its_scratch():
    lib = _lib()
    dev = torch.device("cuda", 0)
    c = 4096 + 64
    x = torch.zeros((1, c, 1, 1), device=dev, dtype=torch.bfloat16)
    g = torch.ones(c, device=dev)
    out = torch.empty(2 * c, device=dev)
    rc = lib.tfop_bn_fwd_stats_noalloc(x.data_ptr(), 1, c, g.data_ptr(), g.data_ptr(), 0, None, None,
                                       0.1, 1e-5, out.data_ptr(), out.data_ptr(), out.data_ptr(),
                                       torch.cuda.current_stream().cuda_stream)
    assert rc != 0
    assert b"4096" in lib.tfop_last_error()
