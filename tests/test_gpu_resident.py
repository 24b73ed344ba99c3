"""GPU parity of the persistent on-chip fused launch (hipe_joint_loss_resident) against the oracle,
and its agreement with the two-launch path.  Same gates as tests/test_gpu_parity.py."""
import numpy as np
import pytest
import torch

from oracle import integral_np as O
from tests import synth

pytestmark = pytest.mark.gpu

# geometry -> items per volume the tiling gives (P); None = not tileable, must report unsupported
RES_GEOMS = {
    (16, 16, 16): 1,     # stage = 8 z-slices, half the consumers idle
    (32, 32, 32): 1,
    (64, 64, 64): 8,     # headline: stage == one z-slice
    (56, 56, 56): 7,     # reference default (main/config.py:23-25): 14 rows per visit
    (96, 96, 96): 36,    # BASELINE configs[3]: stage = a third of a z-slice
    (16, 128, 64): 4,    # stage = half a z-slice (SR | H, non-zero y base)
    (12, 20, 24): 1,     # 30 rows per visit, 180 of 256 consumers active
}
UNSUPPORTED = [(8, 8, 8), (5, 6, 7), (3, 9, 100)]


def _cuda(a, dtype=None):
    t = torch.as_tensor(a).cuda()
    return t if dtype is None else t.to(dtype)


def _check(hm, gt, vis, D, H, W, B, J, out, tgt=np.float64):
    """Every volume against the fp64 oracle (tests/parity.py: coordinates, loss terms, gradients;
    only an axis whose sign(c - gt) is undecidable takes our sign).  Returns the worst gradient
    error as a fraction of the 1e-5 gate."""
    from tests import parity
    coords, expect, stats, terms, dlogits = out
    h64 = hm.astype(np.float64)
    n = B * J
    worst = parity.check_volumes(h64.reshape(n, D, H, W), gt.reshape(n, 3), vis.reshape(n, 3), 1.0 / B,
                                 coords.cpu().numpy(), terms.view(n, 3).cpu().numpy(),
                                 dlogits.float().cpu().numpy().reshape(n, -1))
    ref_loss = O.joint_location_loss(h64, gt.astype(np.float64), vis.astype(np.float64))
    loss = terms.sum().item() / B
    assert abs(loss - ref_loss) <= 1e-5 * abs(ref_loss), (loss, ref_loss)
    # stats: (max, sum exp) per volume
    v = h64.reshape(n, -1)
    m = v.max(axis=1)
    np.testing.assert_allclose(stats[:, 0].cpu().numpy(), m, rtol=0, atol=0)
    np.testing.assert_allclose(stats[:, 1].cpu().numpy(), np.exp(v - m[:, None]).sum(axis=1), rtol=2e-6)
    return worst


@pytest.mark.parametrize("regime", ["noise", "peaked", "sharp"])
@pytest.mark.parametrize("geom", list(RES_GEOMS))
def test_resident_vs_oracle(geom, regime):
    from hipe_b200 import ops
    D, H, W = geom
    J, B = 21, 2
    hm = synth.heatmaps(B, J, D, H, W, regime, seed=11)
    gt, vis = synth.targets(B, J, seed=11, zero_vis_frac=0.1)
    x = _cuda(hm)
    assert ops.resident_supported(x, B * J, D, H, W)
    out = ops.joint_loss_resident(x, B * J, D, H, W, _cuda(gt), _cuda(vis), 1.0 / B)
    torch.cuda.synchronize()
    assert not ops.resident_timed_out(x.device, B * J)
    _check(hm, gt, vis, D, H, W, B, J, out)


@pytest.mark.parametrize("geom", UNSUPPORTED)
def test_unsupported_geometries_fall_back(geom):
    from hipe_b200 import _lib, ops
    D, H, W = geom
    J, B = 21, 2
    hm = synth.heatmaps(B, J, D, H, W, "peaked", seed=3)
    gt, vis = synth.targets(B, J, seed=3)
    x = _cuda(hm)
    assert not ops.resident_supported(x, B * J, D, H, W)
    with pytest.raises(_lib.HipeError):
        ops.joint_loss_resident(x, B * J, D, H, W, _cuda(gt), _cuda(vis), 1.0 / B)
    # "auto" silently takes the L2 re-read launch (still a CUDA kernel of this library)
    out = ops.joint_loss_fused(x, B * J, D, H, W, _cuda(gt), _cuda(vis), 1.0 / B)
    _check(hm, gt, vis, D, H, W, B, J, out)


def test_many_items_per_cta_and_straddling_rounds():
    """More items than CTAs (each CTA loops, the pipeline's head/tail split is exercised) with a
    volume count that makes item rounds straddle volumes (148 is not a multiple of 8)."""
    from hipe_b200 import ops
    D = H = W = 64
    J, B = 21, 5             # 105 volumes x 8 items = 840 items over 148 CTAs
    hm = synth.heatmaps(B, J, D, H, W, "peaked", seed=21)
    gt, vis = synth.targets(B, J, seed=21)
    x = _cuda(hm)
    out = ops.joint_loss_resident(x, B * J, D, H, W, _cuda(gt), _cuda(vis), 1.0 / B)
    torch.cuda.synchronize()
    assert not ops.resident_timed_out(x.device, B * J)
    _check(hm, gt, vis, D, H, W, B, J, out)
    # bit-reproducible: a second launch gives identical bytes (fixed merge order)
    out2 = ops.joint_loss_resident(x, B * J, D, H, W, _cuda(gt), _cuda(vis), 1.0 / B)
    assert torch.equal(out[4], out2[4]) and torch.equal(out[0], out2[0]) and torch.equal(out[3], out2[3])


def test_resident_agrees_with_two_launch_path():
    from hipe_b200 import ops
    D = H = W = 64
    J, B = 21, 3
    hm = synth.heatmaps(B, J, D, H, W, "sharp", seed=5)
    gt, vis = synth.targets(B, J, seed=5)
    x = _cuda(hm)
    c1, e1, s1, t1, d1 = ops.joint_loss_resident(x, B * J, D, H, W, _cuda(gt), _cuda(vis), 1.0 / B)
    c0, e0, s0 = ops.softargmax_forward(x, B * J, D, H, W)
    _, g = ops.l1_loss(c0.view(B, 3 * J), _cuda(gt), _cuda(vis), 1.0 / B)
    d0 = ops.softargmax_backward(x, B * J, D, H, W, s0, e0, g.view(-1, 3))
    assert (c1 - c0).abs().max().item() <= 2e-7
    assert (e1 - e0).abs().max().item() <= 1e-5          # voxel units
    same_sign = ((c1 - _cuda(gt).view(-1, 3)).sign() == (c0 - _cuda(gt).view(-1, 3)).sign()).all(dim=1)
    scale = d0.view(B * J, -1).abs().amax(dim=1, keepdim=True)
    rel = ((d1 - d0).view(B * J, -1).abs() / scale)[same_sign]
    assert rel.max().item() <= 2e-5


def test_resident_bf16():
    from hipe_b200 import ops
    D = H = W = 64
    J, B = 21, 2
    hm32 = synth.heatmaps(B, J, D, H, W, "peaked", seed=4)
    gt, vis = synth.targets(B, J, seed=4)
    xb = _cuda(hm32).to(torch.bfloat16)
    up = xb.float().cpu().numpy().astype(np.float64)       # the rounded tensor is the oracle's input
    assert ops.resident_supported(xb, B * J, D, H, W)
    coords, expect, stats, terms, d = ops.joint_loss_resident(xb, B * J, D, H, W, _cuda(gt), _cuda(vis), 1.0 / B)
    ref_loss, ref_c = O.joint_location_loss(up, gt, vis, return_coords=True)
    c = coords.view(B, 3 * J).cpu().numpy().astype(np.float64)
    assert (np.abs(c - ref_c) <= 1e-5 * np.maximum(np.abs(ref_c), 0.5)).all()
    assert abs(terms.sum().item() / B - ref_loss) <= 1e-5 * abs(ref_loss)
    refg = O.joint_location_loss_grad(up, gt, vis)
    want = torch.as_tensor(refg).to(torch.bfloat16).float().numpy().reshape(B * J, -1)
    got = d.float().cpu().numpy().reshape(B * J, -1)
    from tests import parity
    parity.check_volumes(up.reshape(B * J, D, H, W), gt.reshape(B * J, 3), vis.reshape(B * J, 3), 1.0 / B,
                         coords.cpu().numpy(), terms.view(B * J, 3).cpu().numpy(), got, bf16=True)


def test_module_uses_resident_launch_and_scales_upstream_gradient():
    from hipe_b200 import ops
    from nets import loss as L
    D = H = W = 32
    J, B = 21, 2
    hm = synth.heatmaps(B, J, D, H, W, "peaked", seed=9)
    gt, vis = synth.targets(B, J, seed=9)
    ops.KERNEL_EVENTS = {}
    saved = ops.FUSED_MODE
    ops.FUSED_MODE = "resident"          # 32^3 is not a geometry fused_plan gives to this launch
    try:
        x = _cuda(hm).requires_grad_(True)
        loss = L.JointLocationLoss(fused=True)(x, _cuda(gt), _cuda(vis))
        (loss * 0.25).backward()
        assert "joint_loss_resident" in ops.KERNEL_EVENTS
    finally:
        ops.KERNEL_EVENTS = None
        ops.FUSED_MODE = saved
    y = _cuda(hm).requires_grad_(True)
    L.JointLocationLoss()(y, _cuda(gt), _cuda(vis)).backward()
    scale = y.grad.view(B * J, -1).abs().amax(dim=1, keepdim=True)
    assert ((x.grad - 0.25 * y.grad).view(B * J, -1).abs() <= 2e-5 * 0.25 * scale + 1e-12).all()


def test_default_module_picks_the_measured_fastest_launch():
    """fused=None: 64^3 / 56^3 / 32^3 fp32 and bf16 64^3 / 56^3 batches go to the cluster-resident
    launch, other volumes >= 128 KB to the streaming fused launch, smaller volumes to forward +
    recompute backward."""
    from hipe_b200 import ops
    from nets import loss as L
    x64 = torch.empty((16, 21 * 64, 64, 64), device="cuda")
    assert ops.fused_plan(x64, 16 * 21, 64, 64, 64) == "cluster"
    assert ops.fused_plan(x64[:2], 42, 64, 64, 64) == "cluster"
    assert ops.fused_plan(x64.bfloat16(), 16 * 21, 64, 64, 64) == "cluster"
    assert ops.fused_plan(x64.bfloat16()[:, :21 * 32, :32, :32], 16 * 21, 32, 32, 32) == "cluster"
    x56 = torch.empty((2, 21 * 56, 56, 56), device="cuda")
    assert ops.fused_plan(x56, 42, 56, 56, 56) == "cluster"
    x48 = torch.empty((2, 21 * 48, 48, 48), device="cuda")
    assert ops.fused_plan(x48, 42, 48, 48, 48) == "stream"
    x96 = torch.empty((1, 21 * 96, 96, 96), device="cuda")
    assert ops.fused_plan(x96, 21, 96, 96, 96) == "cluster"      # streamed shares on 16 CTAs
    assert ops.fused_plan(x96.bfloat16(), 21, 96, 96, 96) == "stream"
    x16 = torch.empty((2, 21 * 16, 16, 16), device="cuda")
    assert ops.fused_plan(x16, 42, 16, 16, 16) is None
    hm = synth.heatmaps(16, 21, 64, 64, 64, "peaked", seed=11)
    gt, vis = synth.targets(16, 21, seed=11)
    ops.KERNEL_EVENTS = {}
    try:
        x = _cuda(hm).requires_grad_(True)
        loss = L.JointLocationLoss()(x, _cuda(gt), _cuda(vis))
        loss.backward()
        assert "joint_loss_cluster" in ops.KERNEL_EVENTS and "softargmax_bwd" not in ops.KERNEL_EVENTS
        assert "sum_f64" not in ops.KERNEL_EVENTS            # the launch writes the loss scalar itself
    finally:
        ops.KERNEL_EVENTS = None
    y = _cuda(hm).requires_grad_(True)
    l2 = L.JointLocationLoss(fused=False)(y, _cuda(gt), _cuda(vis))
    l2.backward()
    assert abs(loss.item() - l2.item()) <= 1e-6 * abs(l2.item())   # two kernels, fp32 coordinates
    scale = y.grad.view(16 * 21, -1).abs().amax(dim=1, keepdim=True)
    assert ((x.grad - y.grad).view(16 * 21, -1).abs() <= 2e-5 * scale + 1e-12).all()


def test_resident_launch_writes_the_loss_scalar_itself():
    """loss_dev of hipe_joint_loss_resident: the last CTA's fixed-order sum of the loss terms equals
    hipe_sum_f64 of the same terms to fp64 round-off, repeats bit-identically, and the module's
    step needs no separate reduction launch."""
    from hipe_b200 import ops
    from nets import loss as L
    D = H = W = 64
    J, B = 21, 16
    hm = synth.heatmaps(B, J, D, H, W, "peaked", seed=13)
    gt, vis = synth.targets(B, J, seed=13, zero_vis_frac=0.1)
    x = _cuda(hm)
    vals = []
    for _ in range(3):
        out = torch.full((), float("nan"), dtype=torch.float64, device="cuda")
        c, e, s, terms, d = ops.joint_loss_resident(x, B * J, D, H, W, _cuda(gt), _cuda(vis), 1.0 / B, loss_out=out)
        ref = ops.sum_f64(terms, 1.0 / B)
        assert abs(out.item() - ref.item()) <= 1e-14 * abs(ref.item())
        vals.append(out.item())
    assert vals[0] == vals[1] == vals[2]
    ref_loss = O.joint_location_loss(hm.astype(np.float64), gt, vis)
    assert abs(vals[0] - ref_loss) <= 1e-5 * abs(ref_loss)
    ops.KERNEL_EVENTS = {}
    saved = ops.FUSED_MODE
    ops.FUSED_MODE = "resident"
    try:
        xg = _cuda(hm).requires_grad_(True)
        L.JointLocationLoss(fused=True)(xg, _cuda(gt), _cuda(vis)).backward()
        assert "joint_loss_resident" in ops.KERNEL_EVENTS and "sum_f64" not in ops.KERNEL_EVENTS
    finally:
        ops.KERNEL_EVENTS = None
        ops.FUSED_MODE = saved
