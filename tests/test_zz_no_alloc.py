"""GPU: hipe_keypoint_measures_noalloc (scratch in a file-scope __device__ array, no allocation inside
the call) against hipe_keypoint_measures (caller-owned workspace), both through the C ABI, on two
calls in a row that share the array.  The entry point has no counterpart in the reference (see
include/hipe_b200.h); the tolerance is the one tests/test_chain.py holds the measures to (1e-12).

Collected last on purpose (file name)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

TOL = 1e-12


def _case(N, K, steps, seed, vis_frac):
    rng = np.random.default_rng(seed)
    dist = np.abs(rng.normal(0.0, 0.02, size=(N, K)))
    vis = (rng.random((N, K)) < vis_frac).astype(np.uint8)
    vis[0] = 1                       # every keypoint fed at least once
    return torch.as_tensor(dist).cuda(), torch.as_tensor(vis).cuda()


def _general(lib, dist, vis, N, K, vmin, vmax, steps):
    need = lib.hipe_keypoint_measures_workspace_bytes(K, steps)
    ws = torch.empty((need // 8,), dtype=torch.float64, device="cuda")
    summary = torch.empty(3, dtype=torch.float64, device="cuda")
    curve = torch.empty(steps, dtype=torch.float64, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    rc = lib.hipe_keypoint_measures(dist.data_ptr(), vis.data_ptr(), N, K, vmin, vmax, steps, summary.data_ptr(),
                                    curve.data_ptr(), ws.data_ptr(), need, st)
    assert rc == 0, rc
    return summary, curve


def test_noalloc_matches_the_general_entry_point_twice_in_a_row():
    from hipe_b200 import _lib
    lib = _lib.load()
    st = torch.cuda.current_stream().cuda_stream
    # second call smaller than the first: whatever the first left in the shared array lies beyond and
    # inside what the second uses
    cases = [(3960, 21, 100, 1, 0.9), (500, 5, 37, 2, 0.6)]
    inputs = [(c, _case(c[0], c[1], c[2], c[3], c[4])) for c in cases]
    want = [_general(lib, d, v, c[0], c[1], 0.0, 0.05, c[2]) for c, (d, v) in inputs]
    # every output of the calls under test exists before the first of them: nothing is allocated between
    outs = [(torch.full((3,), float("nan"), dtype=torch.float64, device="cuda"),
             torch.full((c[2],), float("nan"), dtype=torch.float64, device="cuda")) for c, _ in inputs]
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    for (c, (d, v)), (summary, curve) in zip(inputs, outs):
        rc = lib.hipe_keypoint_measures_noalloc(d.data_ptr(), v.data_ptr(), c[0], c[1], 0.0, 0.05, c[2],
                                                summary.data_ptr(), curve.data_ptr(), st)
        assert rc == 0, rc
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before
    for (ws, wc), (summary, curve) in zip(want, outs):
        assert torch.isfinite(summary).all() and torch.isfinite(curve).all()
        assert (summary - ws).abs().max().item() <= TOL * max(1.0, ws.abs().max().item())
        assert (curve - wc).abs().max().item() <= TOL


def test_noalloc_without_a_mask_and_repeated():
    from hipe_b200 import _lib
    lib = _lib.load()
    st = torch.cuda.current_stream().cuda_stream
    d, v = _case(1000, 21, 64, 3, 1.0)
    ones = torch.ones_like(v)
    ws, wc = _general(lib, d, ones, 1000, 21, 0.0, 0.08, 64)
    for _ in range(2):
        summary = torch.empty(3, dtype=torch.float64, device="cuda")
        curve = torch.empty(64, dtype=torch.float64, device="cuda")
        rc = lib.hipe_keypoint_measures_noalloc(d.data_ptr(), None, 1000, 21, 0.0, 0.08, 64, summary.data_ptr(),
                                                curve.data_ptr(), st)
        assert rc == 0, rc
        torch.cuda.synchronize()
        assert (summary - ws).abs().max().item() <= TOL and (curve - wc).abs().max().item() <= TOL


def test_noalloc_rejects_sizes_beyond_its_scratch():
    from hipe_b200 import _lib
    lib = _lib.load()
    d, v = _case(10, 65, 10, 4, 1.0)
    summary = torch.zeros(3, dtype=torch.float64, device="cuda")
    curve = torch.zeros(10, dtype=torch.float64, device="cuda")
    rc = lib.hipe_keypoint_measures_noalloc(d.data_ptr(), v.data_ptr(), 10, 65, 0.0, 0.05, 10, summary.data_ptr(),
                                            curve.data_ptr(), torch.cuda.current_stream().cuda_stream)
    assert rc == _lib.E_LIMIT
    assert "fixed scratch" in lib.hipe_strerror(rc).decode()
    with pytest.raises(_lib.HipeError, match="fixed scratch"):
        _lib.check(rc)
    torch.cuda.synchronize()
    assert summary.abs().sum().item() == 0.0          # nothing was launched
