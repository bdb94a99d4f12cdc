"""GPU: rac_assemble_input_static — input assembly for hosts that must not allocate device memory.

Who needs it: an embedder whose process is not allowed to call cudaMalloc once it is running (a serving or control
loop with an allocation freeze after start-up, a plug-in inside a framework that owns the GPU heap, a process with a
fixed, pre-audited memory budget).  rac_assemble_input asks such a caller for a scratch buffer it has no way to obtain;
the _static variant keeps that scratch in a fixed 64 KiB array inside the library instead (include/rac.h).

What is checked, all through the C ABI (ctypes), nothing through the Python convenience layer:
  * the new entry point writes exactly what rac_assemble_input writes, on two calls in a row with DIFFERENT inputs and
    batch sizes — the second call runs on an array still holding the first call's depth sums, so a missing clear shows;
  * a batch above RAC_ASSEMBLE_STATIC_MAX_B is refused with a message naming the limit, and nothing is written.

The file name sorts after every other test file on purpose: `pytest -x` stops at the first failure, and a failure here
must not hide the parity tests.  The test holds no state from other tests and passes alone in a fresh process too.
"""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

MAX_B = 1024        # RAC_ASSEMBLE_STATIC_MAX_B (include/rac.h)


def _inputs(B, H, W, seed):
    rng = np.random.default_rng(seed)
    init = rng.uniform(0, 1, (B, H, W, 3)).astype(np.float32)
    gt = (rng.integers(0, 256, (B, H, W, 3)) / 255.0).astype(np.float32)
    depth = np.where(rng.uniform(size=(B, H, W)) > 0.6, rng.uniform(2.2, 2.8, (B, H, W)), 0).astype(np.float32)
    depth[0] = 0.0                                              # normalize_depth early return: max <= 0.1
    if B > 2:
        depth[2] = np.where(depth[2] > 0, 2.5, 0)               # early return: std == 0
    return [torch.from_numpy(a).cuda() for a in (init, gt, depth)]


def _both(L, _lib, init, gt, depth, use_depth):
    """x from the existing entry point (caller-allocated scratch) and from the static one, same inputs, same stream."""
    B, H, W, _ = init.shape
    C = 7 if use_depth else 6
    s = _lib.stream_ptr()
    ws_bytes = L.rac_assemble_workspace_bytes(B)
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device="cuda")
    x_ref = torch.full((B, C, H, W), float("nan"), device="cuda")
    x_new = torch.full((B, C, H, W), float("nan"), device="cuda")
    _lib.check(L.rac_assemble_input(_lib.ptr(init), _lib.ptr(gt), _lib.ptr(depth), B, W, H, int(use_depth),
                                    _lib.ptr(x_ref), _lib.ptr(ws), ws_bytes, s))
    _lib.check(L.rac_assemble_input_static(_lib.ptr(init), _lib.ptr(gt), _lib.ptr(depth), B, W, H, int(use_depth),
                                           _lib.ptr(x_new), s))
    torch.cuda.synchronize()
    return x_ref.cpu().numpy(), x_new.cpu().numpy()


def test_static_scratch_entry_point_matches_the_allocating_one():
    from rac import _lib
    _lib.require_cuda()
    L = _lib.lib()
    # first call on this device does the one-time address look-up; two more follow on the now-dirty array
    for (B, H, W, use_depth, seed) in [(5, 60, 80, True, 11), (3, 48, 64, True, 12), (9, 60, 80, False, 13), (MAX_B, 8, 8, True, 14)]:
        init, gt, depth = _inputs(B, H, W, seed)
        x_ref, x_new = _both(L, _lib, init, gt, depth, use_depth)
        assert np.isfinite(x_ref).all() and np.isfinite(x_new).all()
        # same kernels, same order of operations per image -> the tolerance of test_assemble_input_only is met with 0 to spare
        np.testing.assert_allclose(x_new, x_ref, rtol=1e-5, atol=2e-5)
        assert np.array_equal(x_new, x_ref), "static-scratch variant differs bitwise from rac_assemble_input"


def test_static_scratch_rejects_batches_beyond_its_limit():
    from rac import _lib
    _lib.require_cuda()
    L = _lib.lib()
    B, H, W = MAX_B + 1, 4, 4
    init = torch.zeros((B, H, W, 3), device="cuda")
    depth = torch.zeros((B, H, W), device="cuda")
    x = torch.full((B, 7, H, W), -7.0, device="cuda")
    n0 = _lib.launch_count()
    rc = L.rac_assemble_input_static(_lib.ptr(init), _lib.ptr(init), _lib.ptr(depth), B, W, H, 1, _lib.ptr(x), _lib.stream_ptr())
    assert rc == -1
    msg = _lib.last_error()
    assert "RAC_ASSEMBLE_STATIC_MAX_B" in msg and str(MAX_B) in msg, msg
    torch.cuda.synchronize()
    assert _lib.launch_count() == n0 and bool((x == -7.0).all())
    assert L.rac_assemble_input_static(_lib.ptr(init), _lib.ptr(init), _lib.ptr(depth), 0, W, H, 1, _lib.ptr(x), _lib.stream_ptr()) == -1
