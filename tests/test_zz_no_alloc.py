"""gmt_daxpy_sum_noalloc_f64 through the C ABI.

That entry point exists for callers that cannot allocate device memory at call time: programs
whose arrays or memory pool already fill the GPU (a late cudaMalloc may fail) and
time-stepping loops that cannot afford an allocator call, which may synchronise the device.
Its scratch is a fixed array inside the library, shared by all calls on a device and never
cleared between them except for the counter each call resets -- so the thing to test is that
two calls in a row (the second one finding the first one's leftovers) both agree with the
caller-scratch entry point gmt_daxpy_sum_f64, and that sizes beyond the fixed capacity are
refused instead of overrunning the array.

The file name sorts last on purpose: a run that stops at the first failure has then already
run every other test.  It passes in a fresh process and after the other tests alike (it
touches no state of theirs).
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu
F64 = torch.float64
RTOL = 1e-12  # the tolerance the other kernel tests use


def _ptr(t):
    return C.c_void_p(t.data_ptr())


@pytest.mark.parametrize("n", [1, 1000, 4096, 4097, (1 << 20) + 6, 5 * (1 << 20), 33_554_432])
def test_noalloc_matches_caller_scratch_entry_point_twice_in_a_row(native, dev, n):
    noalloc = native.raw("gmt_daxpy_sum_noalloc_f64")
    withscratch = native.raw("gmt_daxpy_sum_f64")
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    scratch = torch.zeros(int(native.gmt_reduce_scratch_bytes()), dtype=torch.uint8, device=dev)
    g = torch.Generator(device="cuda").manual_seed(n)
    x = torch.rand(n, dtype=F64, device=dev, generator=g)
    y0 = torch.rand(n, dtype=F64, device=dev, generator=g)
    for call, a in enumerate((2.0, -0.75)):   # two calls in a row, different scale each time
        y_ref, y_new = y0.clone(), y0.clone()
        s_ref = torch.zeros(1, dtype=F64, device=dev)
        s_new = torch.full((1,), float("nan"), dtype=F64, device=dev)
        assert withscratch(n, a, _ptr(x), _ptr(y_ref), _ptr(s_ref), _ptr(scratch), stream) == 0
        assert noalloc(n, a, _ptr(x), _ptr(y_new), _ptr(s_new), stream) == 0, native.last_error()
        torch.cuda.synchronize()
        assert torch.equal(y_new, y_ref), f"call {call}: y differs"
        ref = s_ref.item()
        assert abs(s_new.item() - ref) <= RTOL * abs(ref) + 1e-300, f"call {call}: {s_new.item()} vs {ref}"
        # and against plain PyTorch fp64 arithmetic, not only against the sibling kernel
        y_torch = a * x + y0
        assert float((y_new - y_torch).abs().max()) <= 1e-15 * (1.0 + float(y_torch.abs().max()))
        t_sum = y_torch.sum().item()
        assert abs(s_new.item() - t_sum) <= RTOL * max(abs(t_sum), float(y_torch.abs().sum().item()) * 1e-3) + 1e-300
        y0 = y_ref  # the second call starts from the first call's result


def test_noalloc_unaligned_views_and_empty(native, dev):
    noalloc = native.raw("gmt_daxpy_sum_noalloc_f64")
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    x = torch.rand(10001, dtype=F64, device=dev)
    y = torch.rand(10001, dtype=F64, device=dev)
    xs, ys = x[1:], y[1:]          # 8-byte aligned only: scalar path
    ref = 3.0 * xs + ys
    s = torch.zeros(1, dtype=F64, device=dev)
    for _ in range(2):
        yy = ys.clone()
        assert noalloc(yy.numel(), 3.0, _ptr(xs), _ptr(yy), _ptr(s), stream) == 0
        torch.cuda.synchronize()
        assert float((yy - ref).abs().max()) < 1e-14
        assert abs(s.item() - ref.sum().item()) <= RTOL * abs(ref.sum().item())
    s.fill_(7.0)
    assert noalloc(0, 1.0, None, None, _ptr(s), stream) == 0
    torch.cuda.synchronize()
    assert s.item() == 0.0


def test_noalloc_rejects_sizes_beyond_the_fixed_scratch(native, dev):
    noalloc = native.raw("gmt_daxpy_sum_noalloc_f64")
    max_n = int(native.gmt_noalloc_max_n())
    assert max_n == 8192 * 4096
    assert int(native.gmt_noalloc_scratch_bytes()) == 256 + 8 * 8192
    s = torch.zeros(1, dtype=F64, device=dev)
    x = torch.zeros(8, dtype=F64, device=dev)  # never dereferenced: the size check comes first
    rc = noalloc(max_n + 1, 1.0, _ptr(x), _ptr(x), _ptr(s), None)
    assert rc == -2  # GMT_ETOOBIG
    msg = native.last_error()
    assert "exceeds the fixed scratch capacity" in msg and str(max_n) in msg


def test_noalloc_on_a_second_device(native):
    """Every GPU holds its own copy of the array; the address is looked up per device."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    noalloc = native.raw("gmt_daxpy_sum_noalloc_f64")
    for d in (1, 0, 1):
        dev = torch.device("cuda", d)
        with torch.cuda.device(dev):
            x = torch.rand(100_003, dtype=F64, device=dev)
            y = torch.rand(100_003, dtype=F64, device=dev)
            ref = 1.5 * x + y
            s = torch.zeros(1, dtype=F64, device=dev)
            stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
            assert noalloc(x.numel(), 1.5, _ptr(x), _ptr(y), _ptr(s), stream) == 0, native.last_error()
            torch.cuda.synchronize(dev)
            assert float((y - ref).abs().max()) < 1e-14
            assert abs(s.item() - ref.sum().item()) <= RTOL * abs(ref.sum().item())
