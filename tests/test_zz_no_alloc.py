"""``b200_bn_stats_noalloc`` — the allocation-free variant of the BN statistics pass.

Some users of the C ABI cannot allocate device memory at call time: a process whose own pool already owns
the whole GPU, a serving loop that cannot afford ``cudaMalloc``'s device-wide synchronisation, a plain-C
runtime with no caching allocator to borrow scratch from.  For them the reduction scratch sits in a fixed
1 MiB array inside the library.  This test drives the entry point through ctypes exactly as such a caller
would and checks it, twice in a row, against the ordinary entry point that takes caller-provided scratch;
it also checks that an over-sized problem is refused with a readable error.

The file name sorts last on purpose: a run that stops at the first failure has then already run every
other test.
"""
import ctypes
from ctypes import c_float, c_int, c_size_t, c_void_p

import pytest
import torch

pytestmark = pytest.mark.gpu


def _args(x, rows, C, gamma, beta, stat, rm, rv):
    return [c_void_p(x.data_ptr()), c_size_t(rows), c_int(C), c_void_p(gamma.data_ptr()), c_void_p(beta.data_ptr()),
            c_float(1e-5), c_float(0.1), c_void_p(stat.data_ptr()), c_void_p(rm.data_ptr()), c_void_p(rv.data_ptr())]


@pytest.mark.parametrize("shape", [(8, 64, 8, 8), (32, 512, 2, 2), (4, 256, 28, 28)])
def test_bn_stats_noalloc_matches_allocating_entry_point(shape):
    if not torch.cuda.is_available():
        pytest.skip("no GPU")
    from b200ddp.ops import _lib as L
    lib = L.lib()
    lib.b200_bn_stats_noalloc.restype = c_int
    lib.b200_bn_stats.restype = c_int
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    torch.manual_seed(11)
    N, C, H, W = shape
    rows = N * H * W
    stream = c_void_p(torch.cuda.current_stream().cuda_stream)
    gamma = torch.rand(C, device=dev) + 0.5
    beta = torch.randn(C, device=dev)
    need = lib.b200_bn_scratch_bytes(c_size_t(rows), c_int(C))
    scratch = torch.zeros(need, dtype=torch.uint8, device=dev)
    for call in range(2):                         # two calls in a row share the built-in array
        x = (torch.randn(N, C, H, W, device=dev) * (1.5 + call)).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
        stat_ref = torch.empty(4, C, device=dev)
        stat_new = torch.full((4, C), float("nan"), device=dev)
        rm_ref, rv_ref = torch.zeros(C, device=dev), torch.ones(C, device=dev)
        rm_new, rv_new = torch.zeros(C, device=dev), torch.ones(C, device=dev)
        rc = lib.b200_bn_stats(*_args(x, rows, C, gamma, beta, stat_ref, rm_ref, rv_ref), c_void_p(scratch.data_ptr()),
                               c_size_t(need), stream)
        assert rc == 0, lib.b200_last_error()
        rc = lib.b200_bn_stats_noalloc(*_args(x, rows, C, gamma, beta, stat_new, rm_new, rv_new), stream)
        assert rc == 0, lib.b200_last_error()
        torch.cuda.synchronize()
        assert torch.allclose(stat_new, stat_ref, rtol=1e-5, atol=1e-6), "call %d" % call
        assert torch.allclose(rm_new, rm_ref, rtol=1e-5, atol=1e-6) and torch.allclose(rv_new, rv_ref, rtol=1e-5, atol=1e-6)
        # and both agree with plain PyTorch fp32 statistics
        xf = x.float()
        assert torch.allclose(stat_new[0], xf.mean((0, 2, 3)), atol=1e-3, rtol=1e-3)


def test_bn_stats_noalloc_rejects_oversized_problem():
    if not torch.cuda.is_available():
        pytest.skip("no GPU")
    from b200ddp.ops import _lib as L
    lib = L.lib()
    lib.b200_bn_stats_noalloc.restype = c_int
    dev = torch.device("cuda", 0)
    C, rows = 8192, 128 * 256                     # 128 channel groups x 128 row blocks: 8 MiB of partials
    assert lib.b200_bn_scratch_bytes(c_size_t(rows), c_int(C)) > (1 << 20)
    x = torch.zeros(8, dtype=torch.bfloat16, device=dev)      # never read: the call must fail before launching
    g = torch.ones(C, device=dev)
    stat = torch.zeros(4, C, device=dev)
    rc = lib.b200_bn_stats_noalloc(*_args(x, rows, C, g, g, stat, g, g), c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc != 0
    msg = lib.b200_last_error().decode()
    assert "needs" in msg and "built-in array" in msg
    # the documented boundary: 1 MiB = 4 KiB of counters + 2040 tiles of 512 B.  C = 1024 is 16 channel groups, so 127 row blocks
    # (rows < 32768) fit and 128 do not; C = 960 (15 groups) fits at any row count
    limit = 1 << 20
    assert lib.b200_bn_scratch_bytes(c_size_t(32767), c_int(1024)) <= limit
    assert lib.b200_bn_scratch_bytes(c_size_t(32768), c_int(1024)) > limit
    assert lib.b200_bn_scratch_bytes(c_size_t(1 << 24), c_int(960)) <= limit
    g1 = torch.ones(1024, device=dev)
    stat1 = torch.zeros(4, 1024, device=dev)
    rc = lib.b200_bn_stats_noalloc(*_args(x, 32768, 1024, g1, g1, stat1, g1, g1), c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc != 0 and "built-in array" in lib.b200_last_error().decode()
    torch.cuda.synchronize()
