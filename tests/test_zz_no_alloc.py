"""The allocation-free Gated CRF entry point (fedicra_gcrf_fwd_bwd_noalloc), called straight through the C ABI.

It serves callers that cannot provide a scratch buffer because they cannot allocate device memory (their framework owns
the GPU's memory, or a cudaMalloc's device-wide synchronisation is not acceptable where they run): the scratch is a fixed
128 KiB array inside the library.  This file checks that it returns what the ordinary entry point returns, twice in a row
(the array keeps the previous call's contents, so the second call proves nothing stale leaks through), that oversized
problems are refused, and that no device memory is allocated by the call.  It is named to be collected last.
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu
TOL = 1e-4


def _rel(a, b):
    return float((a.double() - b.double()).abs().max() / b.double().abs().max().clamp_min(1e-30))


def _call_noalloc(y, img, radius, loss, grad):
    from fedicra_b200._lib import GcrfDesc, lib
    N, C, H, W = y.shape
    descs = (GcrfDesc * 1)(GcrfDesc(1.0, 6.0, 0.1))
    return lib.fedicra_gcrf_fwd_bwd_noalloc(y.data_ptr(), img.data_ptr(), N, C, img.shape[1], H, W, radius, descs, 1,
                                            loss.data_ptr(), grad.data_ptr(), torch.cuda.current_stream().cuda_stream)


def test_noalloc_matches_the_workspace_entry_point_twice_in_a_row():
    from fedicra_b200 import ops
    from fedicra_b200._lib import lib
    assert lib.fedicra_gcrf_noalloc_max_bytes() == 64 + 4 * 32768
    g = torch.Generator().manual_seed(42)
    cases = [(2, 2, 1, 96, 80), (3, 3, 3, 50, 71)]   # second call: other shape, other kernel instantiation
    for N, C, Cin, H, W in cases:
        y = torch.softmax(torch.randn(N, C, H, W, generator=g), 1).cuda()
        img = torch.rand(N, Cin, H, W, generator=g).cuda()
        ref_loss, ref_grad = ops.gcrf_fwd_bwd(y, img, 5, [(1.0, 6.0, 0.1)])
        loss = torch.full((1,), float("nan"), device="cuda")
        grad = torch.full_like(y, float("nan"))
        rc = _call_noalloc(y, img, 5, loss, grad)          # also the set-up call: loads the module, looks the array up
        assert rc == 0, lib.fedicra_last_error().decode()
        torch.cuda.synchronize()
        assert _rel(loss, ref_loss) < TOL and _rel(grad, ref_grad) < TOL
        # the driver's own count of free device memory (cudaMemGetInfo) sees every allocation, whoever made it
        loss.fill_(float("nan"))
        grad.fill_(float("nan"))
        torch.cuda.synchronize()
        free_before, _ = torch.cuda.mem_get_info()
        rc = _call_noalloc(y, img, 5, loss, grad)
        assert rc == 0, lib.fedicra_last_error().decode()
        torch.cuda.synchronize()
        free_after, _ = torch.cuda.mem_get_info()
        assert free_after == free_before, "the call changed the amount of free device memory"
        assert _rel(loss, ref_loss) < TOL and _rel(grad, ref_grad) < TOL


def test_noalloc_after_a_refused_call_and_on_a_second_device():
    """A refused call must leave the shared array usable, and every GPU has its own copy of it (the address is looked up
    per device)."""
    from fedicra_b200 import ops
    from fedicra_b200._lib import lib
    assert lib.fedicra_gcrf_fwd_bwd_noalloc(None, None, 20000, 2, 1, 1, 33, 5, None, 1, None, None, None) < 0
    g = torch.Generator().manual_seed(7)
    for dev in range(min(torch.cuda.device_count(), 2)):
        with torch.cuda.device(dev):
            y = torch.softmax(torch.randn(2, 2, 40, 64, generator=g), 1).cuda()
            img = torch.rand(2, 1, 40, 64, generator=g).cuda()
            ref_loss, ref_grad = ops.gcrf_fwd_bwd(y, img, 5, [(1.0, 6.0, 0.1)])
            loss = torch.full((1,), float("nan"), device="cuda")
            grad = torch.full_like(y, float("nan"))
            rc = _call_noalloc(y, img, 5, loss, grad)
            assert rc == 0, lib.fedicra_last_error().decode()
            torch.cuda.synchronize()
            assert _rel(loss, ref_loss) < TOL and _rel(grad, ref_grad) < TOL


def test_noalloc_refuses_problems_beyond_its_built_in_scratch():
    from fedicra_b200._lib import lib
    # 40 x 512 x 512 = 10.5 M pixels > 4 194 304; rejected before any pointer is touched
    rc = lib.fedicra_gcrf_fwd_bwd_noalloc(None, None, 40, 2, 1, 512, 512, 5, None, 1, None, None, None)
    assert rc < 0
    msg = lib.fedicra_last_error().decode()
    assert "built-in array" in msg and "32768" in msg
    # thin images are limited by the tile count, not by the pixel count: 660 000 pixels, 40 000 tiles
    rc = lib.fedicra_gcrf_fwd_bwd_noalloc(None, None, 20000, 2, 1, 1, 33, 5, None, 1, None, None, None)
    assert rc < 0 and "40000 thread blocks" in lib.fedicra_last_error().decode()
