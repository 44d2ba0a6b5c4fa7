"""dlion_sumsq_static: the sum-of-squares entry point for callers that cannot allocate.

Who needs it: code that wants a gradient norm but has no way to get a scratch buffer — a
plugin inside a host application that handed its device-memory budget out at start-up and
allows no cudaMalloc afterwards, a monitoring hook that must leave the training framework's
caching allocator alone, a bare C caller with no allocator layer.  The library therefore
carries a small fixed scratch array of its own (16,640 bytes per GPU) and this entry point
uses it; inputs and outputs still belong to the caller.

Checked here, through the C ABI only: the result equals dlion_sumsq's (the entry point with a
caller-provided scratch) on two calls in a row — the second call sees whatever the first left in
the shared array — and an input larger than the array can serve is refused with a message.
This file sorts last on purpose: a run that stops at the first failure has then already run
every other test.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _lib():
    from distributed_lion_pytorch_b200.ops import _cabi
    L = _cabi.lib()
    L.dlion_sumsq_static.argtypes = [C.c_void_p, C.c_int64, C.c_int, C.c_void_p, C.c_void_p]
    L.dlion_sumsq_static.restype = C.c_int
    L.dlion_sumsq_static_max_n.restype = C.c_int64
    return L, _cabi


def _reference(L, _cabi, x):
    """dlion_sumsq with a scratch buffer the caller owns."""
    scratch = torch.zeros(int(L.dlion_sumsq_scratch_bytes()) // 4 + 1, dtype=torch.int32, device=x.device)
    out = torch.zeros(1, dtype=torch.float32, device=x.device)
    rc = L.dlion_sumsq(x.data_ptr(), x.numel(), _cabi.dtype_code(x.dtype), scratch.data_ptr(), out.data_ptr(),
                       torch.cuda.current_stream().cuda_stream)
    assert rc == 0, L.dlion_last_error()
    return out


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
def test_static_scratch_matches_caller_scratch_twice_in_a_row(dtype):
    L, _cabi = _lib()
    gen = torch.Generator(device="cuda").manual_seed(5)
    stream = torch.cuda.current_stream().cuda_stream
    sizes = [(1 << 22) + 11, 70001]          # the second, smaller call reuses slots the first one filled
    for n in sizes:
        x = torch.randn(n, device="cuda", generator=gen).to(dtype)
        out = torch.full((1,), -1.0, dtype=torch.float32, device="cuda")
        before = torch.cuda.memory_allocated()
        rc = L.dlion_sumsq_static(x.data_ptr(), n, _cabi.dtype_code(dtype), out.data_ptr(), stream)
        assert rc == 0, L.dlion_last_error()
        assert torch.cuda.memory_allocated() == before
        want = _reference(L, _cabi, x)
        torch.cuda.synchronize()
        assert torch.allclose(out, want, rtol=1e-5), (float(out), float(want))
        assert torch.allclose(out, x.double().pow(2).sum().float().view(1), rtol=1e-5)


def test_static_scratch_refuses_oversized_input():
    L, _cabi = _lib()
    limit = int(L.dlion_sumsq_static_max_n())
    assert limit == 4096 * 65536
    x = torch.zeros(8, device="cuda")
    out = torch.zeros(1, device="cuda")
    rc = L.dlion_sumsq_static(x.data_ptr(), limit + 1, _cabi.dtype_code(x.dtype), out.data_ptr(),
                              torch.cuda.current_stream().cuda_stream)          # rejected before anything is read
    assert rc == -3
    msg = L.dlion_last_error().decode()
    assert "exceeds the built-in scratch" in msg and str(limit) in msg


def test_static_scratch_empty_input_and_largest_input():
    """The two ends of the size range: n = 0 gives 0, and n = the stated maximum (every one of the 4,096 slots in
    use) still equals the caller-scratch entry point."""
    L, _cabi = _lib()
    stream = torch.cuda.current_stream().cuda_stream
    out = torch.full((1,), -1.0, dtype=torch.float32, device="cuda")
    x = torch.ones(8, device="cuda", dtype=torch.bfloat16)
    rc = L.dlion_sumsq_static(x.data_ptr(), 0, _cabi.dtype_code(x.dtype), out.data_ptr(), stream)
    assert rc == 0, L.dlion_last_error()
    torch.cuda.synchronize()
    assert float(out) == 0.0
    limit = int(L.dlion_sumsq_static_max_n())
    try:
        big = torch.empty(limit, device="cuda", dtype=torch.bfloat16)      # 512 MiB
    except torch.OutOfMemoryError:
        pytest.skip("not enough free device memory for the largest input")
    big.fill_(0.5)
    big[::4099] = 2.0
    for _ in range(2):
        rc = L.dlion_sumsq_static(big.data_ptr(), limit, _cabi.dtype_code(big.dtype), out.data_ptr(), stream)
        assert rc == 0, L.dlion_last_error()
        want = _reference(L, _cabi, big)
        torch.cuda.synchronize()
        assert torch.allclose(out, want, rtol=1e-5), (float(out), float(want))


def test_static_scratch_on_a_side_stream_and_on_every_device():
    """A caller's own (non-default) stream, and — where the box has more than one GPU — each device's own copy of
    the array (the address is looked up per device, with that device current)."""
    L, _cabi = _lib()
    for d in range(torch.cuda.device_count()):
        dev = torch.device("cuda", d)
        with torch.cuda.device(dev):
            side = torch.cuda.Stream(device=dev)
            gen = torch.Generator(device=dev).manual_seed(11 + d)
            x = torch.randn((1 << 20) + 3, device=dev, generator=gen).bfloat16()
            out = torch.full((1,), -1.0, dtype=torch.float32, device=dev)
            torch.cuda.synchronize(dev)
            for _ in range(2):
                rc = L.dlion_sumsq_static(x.data_ptr(), x.numel(), _cabi.dtype_code(x.dtype), out.data_ptr(),
                                          side.cuda_stream)
                assert rc == 0, L.dlion_last_error()
                side.synchronize()
                want = _reference(L, _cabi, x)
                torch.cuda.synchronize(dev)
                assert torch.allclose(out, want, rtol=1e-5), (d, float(out), float(want))
