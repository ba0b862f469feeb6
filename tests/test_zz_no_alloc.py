"""b2_sumsq_static: the sum-of-squares entry point that brings its own scratch (a fixed array inside the library)
for callers that cannot allocate device memory — a host holding only borrowed pointers, a process whose framework
allocator owns the whole GPU, a hook where an allocation's device-wide sync is unwelcome.

Collected last on purpose (the file name sorts after every other test file), so a `-x` run has already exercised
everything else when it gets here. Works in a fresh process and after the other tests alike: it needs no state from
them and leaves none behind.
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _both(lib, g, accumulate_into=None):
    """(static-scratch result, caller-scratch result) for the same input, through the C ABI."""
    dev = g.device
    stream = torch.cuda.current_stream().cuda_stream
    code = lib.dtype_code(g.dtype)
    out_static = torch.full((1,), 123.0, device=dev) if accumulate_into is None else accumulate_into.clone()
    out_plain = out_static.clone()
    acc = 0 if accumulate_into is None else 1
    lib.call("b2_sumsq_static", g.data_ptr(), g.numel(), code, out_static.data_ptr(), acc, stream)
    scratch = torch.zeros(lib.lib.b2_sumsq_scratch_bytes() // 4, dtype=torch.float32, device=dev)
    lib.call("b2_sumsq", g.data_ptr(), g.numel(), code, scratch.data_ptr(), out_plain.data_ptr(), acc, stream)
    torch.cuda.synchronize()
    return float(out_static.item()), float(out_plain.item())


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
def test_static_scratch_matches_caller_scratch_twice_in_a_row(dtype):
    from b200_train import _lib
    torch.cuda.set_device(0)
    gen = torch.Generator(device="cpu").manual_seed(17)
    # two calls back to back on different data and sizes: the second must not see anything of the first
    for n in ((1 << 22) + 40, 1000):
        g = (torch.randn(n, generator=gen) * 0.05).to("cuda", dtype)
        got, want = _both(_lib, g)
        np.testing.assert_allclose(got, want, rtol=1e-5)            # same tolerance as the sumsq parity tests
        np.testing.assert_allclose(got, g.float().double().pow(2).sum().item(), rtol=1e-5)
    # accumulate on top of an existing value
    g = (torch.randn(4096, generator=gen)).to("cuda", dtype)
    base = torch.tensor([7.5], device="cuda")
    got, want = _both(_lib, g, accumulate_into=base)
    np.testing.assert_allclose(got, want, rtol=1e-5)
    assert got > 7.5


def test_static_scratch_call_leaves_device_memory_untouched():
    """The one property the entry point exists for: a call takes no device memory. After a warm-up call (the first call on
    a device resolves the scratch array's address), the driver's free-memory figure and torch's own allocator counters
    are the same before and after further calls."""
    from b200_train import _lib
    torch.cuda.set_device(0)
    g = (torch.randn(1 << 20, device="cuda") * 0.1).to(torch.bfloat16)
    out = torch.zeros(1, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    _lib.call("b2_sumsq_static", g.data_ptr(), g.numel(), _lib.BF16, out.data_ptr(), 0, stream)
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    reserved0, allocated0 = torch.cuda.memory_reserved(), torch.cuda.memory_allocated()
    for _ in range(3):
        _lib.call("b2_sumsq_static", g.data_ptr(), g.numel(), _lib.BF16, out.data_ptr(), 0, stream)
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 == free0, (free0, free1)
    assert (torch.cuda.memory_reserved(), torch.cuda.memory_allocated()) == (reserved0, allocated0)
    np.testing.assert_allclose(float(out.item()), g.float().double().pow(2).sum().item(), rtol=1e-5)


def test_static_scratch_rejects_oversized_input_and_bad_arguments():
    from b200_train import _lib
    L = _lib.lib
    limit = L.b2_sumsq_static_max_elems()
    assert limit == 1 << 31
    out = torch.zeros(1, device="cuda")
    g = torch.zeros(64, device="cuda")
    rc = L.b2_sumsq_static(g.data_ptr(), limit + 1, 0, out.data_ptr(), 0, None)
    assert rc == -3 and b"exceed the limit" in L.b2_last_error()
    assert L.b2_sumsq_static(g.data_ptr(), 64, 0, None, 0, None) == -1
    assert L.b2_sumsq_static(g.data_ptr(), 64, 5, out.data_ptr(), 0, None) == -1
    # an empty input is legal and gives 0
    _lib.call("b2_sumsq_static", None, 0, 0, out.data_ptr(), 0, torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    assert float(out.item()) == 0.0
