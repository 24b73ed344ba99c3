"""GPU parity: the CUDA path (through the C ABI / ctypes) against the oracle on identical inputs.

Gates (BASELINE.json north_star): coords |ours-ref64| <= 1e-5*max(|ref64|,0.5);
loss relative 1e-5; fp32 grads <= 1e-5*max|grad_ref64| per volume (+1e-12); Procrustes
PA-MPJPE within 1e-6 m and Z within 1e-9 m (fp64)."""
import numpy as np
import pytest
import torch

from oracle import integral_np as O
from oracle import procrustes_np as P
from tests import synth

pytestmark = pytest.mark.gpu

COORD_RTOL = 1e-5
GEOMS = [(8, 8, 8), (16, 16, 16), (56, 56, 56), (64, 64, 64), (12, 20, 24), (5, 6, 7), (3, 9, 100), (96, 96, 96)]


def _cuda(a, dtype=None):
    t = torch.as_tensor(a).cuda()
    return t if dtype is None else t.to(dtype)


def _coord_ok(ours, ref):
    tol = COORD_RTOL * np.maximum(np.abs(ref), 0.5)
    err = np.abs(ours.astype(np.float64) - ref)
    return err, (err <= tol).all()


@pytest.mark.parametrize("geom", GEOMS)
@pytest.mark.parametrize("regime", ["noise", "peaked", "sharp"])
@pytest.mark.parametrize("variant", [1, 2])
def test_forward_coords(geom, regime, variant):
    from hipe_b200 import ops
    D, H, W = geom
    J, B = 21, 2 if D * H * W > 100000 else 3
    hm = synth.heatmaps(B, J, D, H, W, regime, seed=11)
    ref = O.integral_coords(hm.astype(np.float64), J, W, H, D)
    for cluster in ([0, 1, 2, 4, 8] if D * H >= 8 else [0, 1]):
        c, e, s = ops.softargmax_forward(_cuda(hm), B * J, D, H, W, cluster=cluster, variant=variant)
        err, ok = _coord_ok(c.view(B, 3 * J).cpu().numpy(), ref)
        assert ok, "geom %s regime %s cluster %d variant %d: max err %.3e" % (geom, regime, cluster, variant, err.max())
        # stats = (max, sum exp(s - max))
        s64 = hm.astype(np.float64).reshape(B * J, -1)
        m = s64.max(axis=1)
        l = np.exp(s64 - m[:, None]).sum(axis=1)
        np.testing.assert_array_equal(s[:, 0].cpu().numpy(), m.astype(np.float32))
        np.testing.assert_allclose(s[:, 1].cpu().numpy(), l, rtol=1e-5)


@pytest.mark.parametrize("geom", GEOMS)
@pytest.mark.parametrize("variant", [1, 2])
def test_backward_grads(geom, variant):
    from hipe_b200 import ops
    D, H, W = geom
    J, B = 21, 2
    hm = synth.heatmaps(B, J, D, H, W, "peaked", seed=5)
    rng = np.random.default_rng(3)
    g = rng.standard_normal((B, 3 * J))
    ref = O.integral_coords_vjp(hm.astype(np.float64), J, W, H, D, g).reshape(B * J, -1)
    x = _cuda(hm)
    for cluster in [0, 1, 4]:
        c, e, s = ops.softargmax_forward(x, B * J, D, H, W, cluster=cluster, variant=variant)
        d = ops.softargmax_backward(x, B * J, D, H, W, s, e, _cuda(g.reshape(-1, 3), torch.float32),
                                    cluster=cluster, variant=variant)
        d = d.cpu().numpy().reshape(B * J, -1).astype(np.float64)
        scale = np.abs(ref).max(axis=1, keepdims=True)
        assert (np.abs(d - ref) <= 1e-5 * scale + 1e-12).all(), \
            "geom %s cluster %d: %.3e" % (geom, cluster, (np.abs(d - ref) / scale).max())


def _near_one_hot(B, J, D, H, W, seed):
    """Volumes whose mass sits in ONE voxel (p ~ 0.9-0.999) with E a few 1e-3 voxel off its centre:
    the gradient at the peak is p * (idx - E) with idx - E ~ 1e-3, so any absolute error in E shows
    up undiluted (the sharp-regime case that first missed the gate: scripts/diag_resident.py)."""
    rng = np.random.default_rng(seed)
    hm = rng.standard_normal((B * J, D, H, W)).astype(np.float32) * 0.5
    for v in range(B * J):
        z, y, x = rng.integers(1, D - 1), rng.integers(1, H - 1), rng.integers(1, W - 1)
        peak = rng.uniform(14.0, 24.0)
        hm[v, z, y, x] += peak
        for dz, dy, dx in ((0, 0, 1), (0, 1, 0), (1, 0, 0), (0, 0, -1), (0, -1, 0), (-1, 0, 0)):
            hm[v, z + dz, y + dy, x + dx] += peak - rng.uniform(4.0, 9.0)
    return hm.reshape(B, J * D, H, W)


@pytest.mark.parametrize("case", ["sharp5", "onehot"])
@pytest.mark.parametrize("geom", [(64, 64, 64), (16, 16, 16), (12, 20, 24)])
def test_backward_grads_when_E_sits_on_the_peak(geom, case):
    from hipe_b200 import ops
    D, H, W = geom
    J, B = 21, 3
    hm = synth.heatmaps(B, J, D, H, W, "sharp", seed=5) if case == "sharp5" else _near_one_hot(B, J, D, H, W, 17)
    gt, vis = synth.targets(B, J, seed=5)
    up = hm.astype(np.float64)
    _, ref_c = O.joint_location_loss(up, gt, vis, return_coords=True)
    ref = O.joint_location_loss_grad(up, gt, vis).reshape(B * J, -1)
    scale = np.abs(ref).max(axis=1, keepdims=True)
    x = _cuda(hm)
    for cluster in [0, 1, 2, 4, 8]:
        for variant in (1, 2):
            c, e, s = ops.softargmax_forward(x, B * J, D, H, W, cluster=cluster, variant=variant)
            err, ok = _coord_ok(c.view(B, 3 * J).cpu().numpy(), ref_c)
            assert ok
            _, g = ops.l1_loss(c.view(B, 3 * J), _cuda(gt), _cuda(vis), 1.0 / B)
            d = ops.softargmax_backward(x, B * J, D, H, W, s, e, g.view(-1, 3), cluster=cluster, variant=variant)
            d = d.cpu().numpy().reshape(B * J, -1).astype(np.float64)
            rel = np.abs(d - ref) / scale
            assert (np.abs(d - ref) <= 1e-5 * scale + 1e-12).all(), \
                "geom %s %s cluster %d variant %d: %.3e" % (geom, case, cluster, variant, rel.max())


@pytest.mark.parametrize("fused", [False, True, None])
@pytest.mark.parametrize("tgt", [np.float64, np.float32])
@pytest.mark.parametrize("geom", [(16, 16, 16), (64, 64, 64), (5, 6, 7)])
def test_joint_location_loss_module(fused, tgt, geom):
    from nets import loss as L
    D, H, W = geom
    J, B = 21, 4
    hm = synth.heatmaps(B, J, D, H, W, "peaked", seed=2)
    gt, vis = synth.targets(B, J, seed=2, zero_vis_frac=0.1)
    gt, vis = gt.astype(tgt), vis.astype(tgt)
    ref_loss, ref_c = O.joint_location_loss(hm.astype(np.float64), gt.astype(np.float64), vis.astype(np.float64),
                                            return_coords=True)
    ref_g = O.joint_location_loss_grad(hm.astype(np.float64), gt.astype(np.float64), vis.astype(np.float64))
    x = _cuda(hm).requires_grad_(True)
    crit = L.JointLocationLoss(fused=fused)
    loss = crit(x, _cuda(gt), _cuda(vis))
    assert loss.dtype == (torch.float64 if tgt == np.float64 else torch.float32)
    assert abs(loss.item() - ref_loss) <= 1e-5 * abs(ref_loss)
    loss.backward()
    d = x.grad.cpu().numpy().astype(np.float64).reshape(B * J, -1)
    # only an axis whose sign(c - gt) is undecidable takes our sign; the rest of the volume is checked
    from tests import parity
    with torch.no_grad():
        ours_c = L.softmax_integral_tensor(x.detach(), J, W, H, D).view(B * J, 3).cpu().numpy()
    parity.check_volumes(hm.astype(np.float64).reshape(B * J, D, H, W), gt.reshape(B * J, 3), vis.reshape(B * J, 3),
                         1.0 / B, ours_c, None, d)


def test_upstream_gradient_scaling_fused_and_unfused():
    from nets import loss as L
    D = H = W = 16
    J, B = 21, 2
    hm = synth.heatmaps(B, J, D, H, W, "peaked", seed=9)
    gt, vis = synth.targets(B, J, seed=9)
    grads = []
    for fused in (False, True):
        x = _cuda(hm).requires_grad_(True)
        loss = L.JointLocationLoss(fused=fused)(x, _cuda(gt), _cuda(vis))
        (loss * 0.25).backward()
        grads.append(x.grad.clone())
    x = _cuda(hm).requires_grad_(True)
    L.JointLocationLoss()(x, _cuda(gt), _cuda(vis)).backward()
    torch.testing.assert_close(grads[0], 0.25 * x.grad, rtol=1e-6, atol=0)
    # the fused launch is a different kernel (its own summation order): compare against the
    # per-volume gradient scale, the way the parity gate is defined, not element-relative
    want = (0.25 * x.grad).view(B * J, -1)
    scale = want.abs().amax(dim=1, keepdim=True)
    assert ((grads[1].view(B * J, -1) - want).abs() <= 5e-6 * scale + 1e-12).all()


def test_bf16_forward_backward():
    from hipe_b200 import ops
    D = H = W = 64
    J, B = 21, 2
    hm32 = synth.heatmaps(B, J, D, H, W, "peaked", seed=4)
    xb = _cuda(hm32).to(torch.bfloat16)
    up = xb.float().cpu().numpy().astype(np.float64)       # the rounded tensor is the oracle's input
    ref = O.integral_coords(up, J, W, H, D)
    for variant in (1, 2):
        c, e, s = ops.softargmax_forward(xb, B * J, D, H, W, variant=variant)
        err, ok = _coord_ok(c.view(B, 3 * J).cpu().numpy(), ref)
        assert ok, err.max()
        g = np.random.default_rng(1).standard_normal((B, 3 * J))
        refg = O.integral_coords_vjp(up, J, W, H, D, g)
        d = ops.softargmax_backward(xb, B * J, D, H, W, s, e, _cuda(g.reshape(-1, 3), torch.float32), variant=variant)
        want = torch.as_tensor(refg).to(torch.bfloat16).float().numpy()
        got = d.float().cpu().numpy()
        # within 1 bf16 ulp of the rounded fp64 reference
        ulp = np.maximum(np.abs(want), 1e-30) * 2.0 ** -7
        assert (np.abs(got - want) <= ulp + 1e-12).all()


@pytest.mark.parametrize("geom", [(64, 64, 64), (56, 56, 56), (16, 16, 16)])
def test_bf16_module_default_plan_vs_oracle(geom):
    """bf16 heatmaps through the drop-in module with its own launch plan (the streaming single
    launch for volumes >= 128 KB, forward + recompute backward below): loss to 1e-5, gradient
    within 1 bf16 ulp of the rounded fp64 reference, on the rounded input."""
    from hipe_b200 import ops
    from nets import loss as L
    D, H, W = geom
    J, B = 21, 2
    hm32 = synth.heatmaps(B, J, D, H, W, "peaked", seed=14)
    gt, vis = synth.targets(B, J, seed=14, zero_vis_frac=0.1)
    xb = _cuda(hm32).to(torch.bfloat16)
    up = xb.float().cpu().numpy().astype(np.float64)
    plan = ops.fused_plan(xb, B * J, D, H, W)
    assert (plan is None) == (D * H * W * 2 < (1 << 16))
    x = xb.clone().requires_grad_(True)
    ops.KERNEL_EVENTS = {}
    try:
        loss = L.JointLocationLoss()(x, _cuda(gt), _cuda(vis))
        loss.backward()
        assert ("joint_loss_fused" in ops.KERNEL_EVENTS) == (plan == "stream")
        assert ("joint_loss_cluster" in ops.KERNEL_EVENTS) == (plan == "cluster")
    finally:
        ops.KERNEL_EVENTS = None
    ref_loss, ref_c = O.joint_location_loss(up, gt, vis, return_coords=True)
    assert abs(loss.item() - ref_loss) <= 1e-5 * abs(ref_loss)
    refg = O.joint_location_loss_grad(up, gt, vis)
    want = torch.as_tensor(refg).to(torch.bfloat16).float().numpy().reshape(B * J, -1)
    assert x.grad.dtype == torch.bfloat16
    got = x.grad.float().cpu().numpy().reshape(B * J, -1)
    from tests import parity
    with torch.no_grad():
        ours_c = L.softmax_integral_tensor(xb, J, W, H, D).view(B * J, 3).cpu().numpy()
    parity.check_volumes(up.reshape(B * J, D, H, W), gt.reshape(B * J, 3), vis.reshape(B * J, 3), 1.0 / B,
                         ours_c, None, got, bf16=True)


def test_plain_expectations_and_volume_softmax():
    from nets import loss as L
    D, H, W = 16, 12, 24
    J, B = 21, 3
    hm = synth.heatmaps(B, J, D, H, W, "peaked", seed=6)
    p_ref = O.volume_softmax(hm.astype(np.float64), J)
    x = _cuda(hm)
    p = L.softmax_integral_tensor2(x, J, W, H, D)
    assert p.shape == (B, J, D * H * W)
    np.testing.assert_allclose(p.cpu().numpy(), p_ref, rtol=3e-6, atol=1e-30)
    ex, ey, ez = L.generate_3d_integral_preds_tensor(p, J, W, H, D)
    rx, ry, rz = O.marginal_expectations(p.cpu().numpy().astype(np.float64), J, W, H, D)
    assert ex.shape == (B, J, 1)
    np.testing.assert_allclose(ex.cpu().numpy(), rx, rtol=1e-5)
    np.testing.assert_allclose(ey.cpu().numpy(), ry, rtol=1e-5)
    np.testing.assert_allclose(ez.cpu().numpy(), rz, rtol=1e-5)


def test_analytic_properties():
    from hipe_b200 import ops
    D, H, W = 8, 16, 32
    n = 4
    # one-hot-like volume -> idx/dim - 0.5 ; uniform -> ((dim-1)/2)/dim - 0.5 ; shift invariance
    x = torch.full((n, D, H, W), -50.0, device="cuda")
    x[0, 3, 5, 7] = 50.0
    x[1] = 0.0
    x[2] = torch.randn(D, H, W, device="cuda")
    x[3] = x[2] + 123.0
    c, _, _ = ops.softargmax_forward(x, n, D, H, W)
    c = c.cpu().numpy()
    np.testing.assert_allclose(c[0], [7 / W - 0.5, 5 / H - 0.5, 3 / D - 0.5], atol=1e-6)
    np.testing.assert_allclose(c[1], [(W - 1) / 2 / W - 0.5, (H - 1) / 2 / H - 0.5, (D - 1) / 2 / D - 0.5], atol=1e-6)
    np.testing.assert_allclose(c[2], c[3], atol=2e-6)
    # gradient sums to zero over each volume
    _, e, s = ops.softargmax_forward(x, n, D, H, W)
    g = torch.randn(n, 3, device="cuda")
    d = ops.softargmax_backward(x, n, D, H, W, s, e, g)
    assert d.double().sum(dim=(1, 2, 3)).abs().max().item() < 1e-6


def test_empty_batch_and_errors():
    from hipe_b200 import ops, _lib
    x = torch.empty((0, 21 * 8, 8, 8), device="cuda")
    c, e, s = ops.softargmax_forward(x, 0, 8, 8, 8)
    assert c.shape == (0, 3)
    with pytest.raises(TypeError):
        ops.softargmax_forward(torch.zeros(21 * 8, 8, 8, device="cuda", dtype=torch.float16), 21, 8, 8, 8)
    with pytest.raises(_lib.HipeError):
        ops.softargmax_forward(torch.zeros(21 * 8, 8, 8, device="cuda"), 21, 8, 8, 8, cluster=3)
    with pytest.raises(RuntimeError):
        ops.softargmax_forward(torch.zeros(21 * 8, 8, 8), 21, 8, 8, 8)


@pytest.mark.parametrize("optimal_scale", [True, False])
def test_procrustes_vs_oracle(optimal_scale):
    from hipe_b200 import evaluate as E
    X, Y = synth.procrustes_pairs(3960, seed=0, mirror_frac=0.05)
    d, Z, T, b, c = E.compute_similarity_transform_batched(X, Y, optimal_scale)
    rd, rZ, rT, rb, rc = P.similarity_transform(X, Y, optimal_scale)
    assert np.abs(Z.cpu().numpy() - rZ).max() <= 1e-9
    np.testing.assert_allclose(T.cpu().numpy(), rT, atol=1e-9)
    np.testing.assert_allclose(d.cpu().numpy(), rd, atol=1e-10)
    np.testing.assert_allclose(b.cpu().numpy(), rb, rtol=1e-10)
    np.testing.assert_allclose(c.cpu().numpy(), rc, atol=1e-9)
    assert np.allclose(np.linalg.det(T.cpu().numpy()), 1.0, atol=1e-9)
    pa, raw, _ = E.pa_mpjpe(X, Y)
    rpa, rraw = P.pa_mpjpe(X, Y)
    assert abs(pa.item() - rpa) <= 1e-6 and abs(raw.item() - rraw) <= 1e-6
    assert abs(pa.item() - rpa) <= 1e-12


def test_align_w_scale_vs_oracle():
    from hipe_b200 import evaluate as E
    X, Y = synth.procrustes_pairs(512, seed=3, mirror_frac=0.2)
    Z = E.align_w_scale_batched(X, Y).cpu().numpy()
    np.testing.assert_allclose(Z, P.align_w_scale(X, Y), atol=1e-9)


def test_procrustes_degenerate_inputs():
    from hipe_b200 import evaluate as E
    rng = np.random.default_rng(0)
    # planar hands (rank-2 correlation), exact similarity (d ~ 0), single sample, J = 40 (> one warp)
    Y = synth.hand_like_joints(64, seed=5)
    Y[:, :, 2] = 0.7
    R = synth.random_rotations(64, rng)
    X = 0.9 * np.einsum("njk,nkl->njl", Y, R) + 0.1
    d, Z, T, b, c = E.compute_similarity_transform_batched(X, Y, True)
    assert np.abs(Z.cpu().numpy() - X).max() < 1e-9
    assert np.abs(d.cpu().numpy()).max() < 1e-9
    np.testing.assert_allclose(b.cpu().numpy(), 0.9, rtol=1e-9)
    X, Y = synth.procrustes_pairs(5, J=40, seed=8)
    _, Z, _, _, _ = E.compute_similarity_transform_batched(X, Y, True)
    np.testing.assert_allclose(Z.cpu().numpy(), P.similarity_transform(X, Y, True)[1], atol=1e-9)


def test_host_entry_points():
    import ctypes
    from hipe_b200 import _lib
    lib = _lib.load()
    D = H = W = 16
    J, B = 21, 5
    hm = synth.heatmaps(B, J, D, H, W, "peaked", seed=12)
    gt, vis = synth.targets(B, J, seed=12)
    loss = ctypes.c_double()
    coords = np.empty((B, 3 * J), np.float32)
    dl = np.empty_like(hm)
    rc = lib.hipe_joint_location_loss_host(hm.ctypes.data, _lib.F32, B, J, D, H, W, gt.ctypes.data, vis.ctypes.data,
                                           _lib.F64, 1, ctypes.byref(loss), coords.ctypes.data, dl.ctypes.data, None, 2)
    assert rc == 0
    ref_loss, ref_c = O.joint_location_loss(hm.astype(np.float64), gt, vis, return_coords=True)
    assert abs(loss.value - ref_loss) <= 1e-5 * ref_loss
    _, ok = _coord_ok(coords, ref_c)
    assert ok
    ref_g = O.joint_location_loss_grad(hm.astype(np.float64), gt, vis).reshape(B * J, -1)
    scale = np.abs(ref_g).max(axis=1, keepdims=True)
    assert (np.abs(dl.reshape(B * J, -1) - ref_g) <= 1e-5 * scale + 1e-12).all()
    X, Y = synth.procrustes_pairs(100, seed=1)
    Z = np.empty_like(X)
    pa, raw = ctypes.c_double(), ctypes.c_double()
    assert lib.hipe_procrustes_host(X.ctypes.data, Y.ctypes.data, 100, 21, 0, 1, Z.ctypes.data,
                                    ctypes.byref(pa), ctypes.byref(raw)) == 0
    rpa, rraw = P.pa_mpjpe(X, Y)
    assert abs(pa.value - rpa) < 1e-12 and abs(raw.value - rraw) < 1e-12
