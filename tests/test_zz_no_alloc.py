"""GPU (-m gpu): the allocation-free entry point asf_channel_stats_noalloc (scratch in a file-scope device array inside
the library) against asf_channel_stats with a caller-owned workspace, through the C ABI.  Collected last on purpose;
passes alone in a fresh process and after the other tests."""
import pytest
import torch

pytestmark = pytest.mark.gpu

from asf_b200 import _lib as L  # noqa: E402

MAX_CHANNELS = 256  # ASF_NOALLOC_MAX_CHANNELS (include/asf_b200.h)


def _reference(x, rows, Cn, ld):
    stats = torch.full((2 * Cn,), -7.0, dtype=torch.float64, device="cuda")
    ws = torch.empty(L.stats_ws_floats(Cn), device="cuda")
    tk = torch.zeros(L.STATS_TICKETS, dtype=torch.int32, device="cuda")
    L.call("asf_channel_stats", L.ptr(x), L.dtype_code(x.dtype), rows, Cn, ld, L.ptr(stats), L.ptr(ws), L.ptr(tk),
           L.stream_ptr())
    return stats


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
def test_channel_stats_noalloc_matches_workspace_version_twice_in_a_row(dtype):
    g = torch.Generator().manual_seed(11)
    # two calls in a row with DIFFERENT shapes: the second starts from the scratch contents the first left behind
    cases = [(6000, 256, 256), (3001, 64, 96), (6000, 256, 256)]
    for rows, Cn, ld in cases:
        x = (torch.randn(rows, ld, generator=g) * 2 + 0.5).cuda().to(dtype)
        ref = _reference(x, rows, Cn, ld)
        out = torch.full((2 * Cn,), 99.0, dtype=torch.float64, device="cuda")
        out2 = torch.full((2 * Cn,), 98.0, dtype=torch.float64, device="cuda")
        L.call("asf_channel_stats_noalloc", L.ptr(x), L.dtype_code(dtype), rows, Cn, ld, L.ptr(out), L.stream_ptr())
        L.call("asf_channel_stats_noalloc", L.ptr(x), L.dtype_code(dtype), rows, Cn, ld, L.ptr(out2), L.stream_ptr())
        torch.cuda.synchronize()
        # same kernels, same ordered reduction: bit-identical, which is also the tolerance of the existing statistics tests
        assert torch.equal(out, ref), float((out - ref).abs().max())
        assert torch.equal(out2, ref), float((out2 - ref).abs().max())
        xd = x[:, :Cn].double()
        tol = 1e-6 if dtype == torch.bfloat16 else 1e-12
        assert torch.allclose(out[:Cn], xd.sum(0), rtol=tol) and torch.allclose(out[Cn:], (xd * xd).sum(0), rtol=tol)


def test_channel_stats_noalloc_allocates_nothing_and_rejects_oversize():
    rows, Cn = 4096, 128
    x = torch.randn(rows, Cn, device="cuda").to(torch.bfloat16)
    out = torch.empty(2 * Cn, dtype=torch.float64, device="cuda")
    L.call("asf_channel_stats_noalloc", L.ptr(x), L.BF16, rows, Cn, Cn, L.ptr(out), L.stream_ptr())  # set-up call
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    for _ in range(5):
        L.call("asf_channel_stats_noalloc", L.ptr(x), L.BF16, rows, Cn, Cn, L.ptr(out), L.stream_ptr())
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 >= free0, (free0, free1)  # the driver handed out no device memory
    # beyond the built-in scratch: a clear error, nothing launched
    big = MAX_CHANNELS + 8
    xb = torch.randn(64, big, device="cuda").to(torch.bfloat16)
    ob = torch.zeros(2 * big, dtype=torch.float64, device="cuda")
    with pytest.raises(RuntimeError, match="ASF_NOALLOC_MAX_CHANNELS"):
        L.call("asf_channel_stats_noalloc", L.ptr(xb), L.BF16, 64, big, big, L.ptr(ob), L.stream_ptr())
    assert float(ob.abs().sum()) == 0.0
    with pytest.raises(RuntimeError):
        L.call("asf_channel_stats_noalloc", None, L.BF16, 64, 8, 8, L.ptr(ob), L.stream_ptr())
