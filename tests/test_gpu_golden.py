"""GPU parity against the committed golden vectors (frozen from the live reference by
tests/golden/make_golden.py): the CUDA path, called through the drop-in modules / C ABI, must
reproduce the reference's fp64 results within the north-star tolerances -- 1e-5 relative on
coordinates (on the natural scale max(|c|, 0.5)) and loss, 1e-5 of each volume's largest
gradient, 1e-3 mm on PA-MPJPE."""
import os

import numpy as np
import pytest
import torch

from oracle import integral_np as O
from tests import synth
from tests.test_oracle_golden import J, SMALL, case_input, gold

pytestmark = pytest.mark.gpu


def _cu(a, dtype=None):
    t = torch.as_tensor(np.ascontiguousarray(a)).cuda()
    return t if dtype is None else t.to(dtype)


@pytest.mark.parametrize("name", SMALL + ["cube64"])
def test_soft_argmax_and_loss_vs_golden(name):
    from nets import loss as L
    g = gold("integral_" + name)
    hm, (B, D, H, W) = case_input(g)
    x = _cu(hm).requires_grad_(True)
    c = L.softmax_integral_tensor(x, J, W, H, D)
    ref = g["coords64"]
    err = np.abs(c.detach().cpu().numpy().astype(np.float64) - ref)
    assert (err <= 1e-5 * np.maximum(np.abs(ref), 0.5)).all(), err.max()
    # ours is at least as close to fp64 as the reference's own fp32 evaluation (plus 2 ulp of 0.5)
    assert err.max() <= np.abs(g["coords32"] - ref).max() + 1.2e-7
    for fused in (False, True):
        x = _cu(hm).requires_grad_(True)
        loss = L.JointLocationLoss(fused=fused)(x, _cu(g["gt"]), _cu(g["vis"]))
        assert loss.dtype == torch.float64
        assert abs(loss.item() - g["loss64"]) <= 1e-5 * abs(g["loss64"])
        loss.backward()
        d = x.grad.cpu().numpy().astype(np.float64).reshape(B * J, -1)
        close = (np.abs(ref - g["gt"]) < 1e-5).reshape(B * J, 3).any(axis=1)   # sign may flip there
        if "loss_grad64" in g.files:
            r = g["loss_grad64"]
            scale = np.abs(r).max(axis=1, keepdims=True)
            ok = (np.abs(d - r) <= 1e-5 * scale + 1e-12).all(axis=1) | close
        else:
            r = g["loss_grad64_sampled"]
            scale = g["loss_grad64_abs_max"][:, None]
            ok = (np.abs(d[:, g["grad_idx"]] - r) <= 1e-5 * scale + 1e-12).all(axis=1) | close
            np.testing.assert_allclose(np.abs(d).sum(axis=1)[~close], g["loss_grad64_abs_sum"][~close], rtol=1e-5)
        assert ok.all()


@pytest.mark.parametrize("name", SMALL + ["cube64"])
def test_vjp_vs_golden(name):
    from nets import loss as L
    g = gold("integral_" + name)
    hm, (B, D, H, W) = case_input(g)
    x = _cu(hm).requires_grad_(True)
    (L.softmax_integral_tensor(x, J, W, H, D) * _cu(g["upstream"], torch.float32)).sum().backward()
    d = x.grad.cpu().numpy().astype(np.float64).reshape(B * J, -1)
    if "vjp64" in g.files:
        r = g["vjp64"]
        scale = np.abs(r).max(axis=1, keepdims=True)
    else:
        r, d = g["vjp64_sampled"], d[:, g["grad_idx"]]
        scale = g["vjp64_abs_max"][:, None]
    assert (np.abs(d - r) <= 1e-5 * scale + 1e-12).all()


@pytest.mark.parametrize("name", SMALL)
def test_probs_expectations_mpjpe_vs_golden(name):
    from nets import loss as L
    g = gold("integral_" + name)
    hm, (B, D, H, W) = case_input(g)
    p = L.softmax_integral_tensor2(_cu(hm), J, W, H, D)
    assert p.shape == (B, J, D * H * W)
    np.testing.assert_allclose(p.cpu().numpy(), g["probs64"], rtol=1e-5, atol=1e-30)
    ex, ey, ez = L.generate_3d_integral_preds_tensor(_cu(g["probs64"], torch.float32), J, W, H, D)
    got = torch.cat((ex, ey, ez), dim=2).cpu().numpy()
    np.testing.assert_allclose(got, g["expect64"], rtol=1e-5)
    c = L.softmax_integral_tensor(_cu(hm), J, W, H, D)
    assert abs(L.computeMPJPE(c.double(), _cu(g["gt"])).item() - g["mpjpe64"]) <= 1e-5 * g["mpjpe64"]


def test_combined_loss_vs_golden():
    from nets import loss as L
    g = gold("combined_cube8")
    x = _cu(g["logits"]).requires_grad_(True)
    loss, smp, tmp, sup, uns = L.CombinedLoss()(x, _cu(g["teacher"]), _cu(g["gt"]), _cu(g["vis"]),
                                                _cu(g["labelled"]))
    assert abs(loss.item() - g["loss64"]) <= 1e-5 * g["loss64"]
    assert abs(smp.item() - g["student_mpjpe"]) <= 1e-5 * g["student_mpjpe"]
    assert abs(tmp.item() - g["teacher_mpjpe"]) <= 1e-12
    assert abs(float(sup) - g["sup"]) <= 1e-5 * g["sup"] and abs(float(uns) - g["unsup"]) <= 1e-5 * g["unsup"]
    loss.backward()
    r = g["grad64"].reshape(-1, 512)
    d = x.grad.cpu().numpy().astype(np.float64).reshape(-1, 512)
    assert (np.abs(d - r) <= 1e-5 * np.abs(r).max(axis=1, keepdims=True) + 1e-12).all()


@pytest.mark.parametrize("name", ["combined_cube16", "combined_cube64"])
def test_combined_loss_single_launch_vs_golden(name):
    """lambda = 0 on volumes the single-launch plans take (64^3: the production shape): one launch for
    loss + gradient, no host synchronisation in forward, the reference's five outputs."""
    from hipe_b200 import ops
    from nets import loss as L
    g = gold(name)
    B, D, H, W = (int(v) for v in g["shape"])
    hm = synth.heatmaps(B, 21, D, H, W, "peaked", seed=int(g["seed"]))
    x = _cu(hm).requires_grad_(True)
    ops.KERNEL_EVENTS = {}
    try:
        loss, smp, tmp, sup, uns = L.CombinedLoss()(x, _cu(g["teacher"]), _cu(g["gt"]), _cu(g["vis"]),
                                                    _cu(g["labelled"]))
        loss.backward()
        names = set(ops.KERNEL_EVENTS)
    finally:
        ops.KERNEL_EVENTS = None
    if ops.fused_plan(x, B * 21, D, H, W) is not None:
        assert "softargmax_fwd" not in names and "softargmax_bwd" not in names, names
    assert loss.dtype == torch.float64
    assert abs(loss.item() - g["loss64"]) <= 1e-5 * g["loss64"]
    assert abs(smp.item() - g["student_mpjpe"]) <= 1e-5 * g["student_mpjpe"]
    assert abs(tmp.item() - g["teacher_mpjpe"]) <= 1e-12
    assert abs(float(sup) - g["sup"]) <= 1e-5 * g["sup"] and abs(float(uns) - g["unsup"]) <= 1e-5 * g["unsup"]
    d = x.grad.cpu().numpy().astype(np.float64).reshape(B * 21, -1)
    r = g["grad64_sampled"]
    assert (np.abs(d[:, g["grad_idx"]] - r) <= 1e-5 * g["grad64_abs_max"][:, None] + 1e-12).all()
    np.testing.assert_allclose(np.abs(d).max(axis=1), g["grad64_abs_max"], rtol=2e-5)
    np.testing.assert_allclose(np.abs(d).sum(axis=1), g["grad64_abs_sum"], rtol=2e-5)


def test_combined_loss_all_labelled_and_none_labelled():
    """Degenerate masks: the reference returns a Python 0 for the empty part (loss.py:129-130);
    here a 0-valued tensor (no host synchronisation to find out)."""
    from nets import loss as L
    g = gold("combined_cube16")
    B, D, H, W = (int(v) for v in g["shape"])
    hm = synth.heatmaps(B, 21, D, H, W, "peaked", seed=int(g["seed"]))
    for lab in (np.ones(B, bool), np.zeros(B, bool)):
        x = _cu(hm).requires_grad_(True)
        loss, _, _, sup, uns = L.CombinedLoss()(x, _cu(g["teacher"]), _cu(g["gt"]), _cu(g["vis"]), _cu(lab))
        ref = O.combined_loss(hm.astype(np.float64), g["teacher"], g["gt"], g["vis"], lab)
        assert abs(loss.item() - ref[0]) <= 1e-5 * ref[0]
        assert abs(float(sup) - ref[3]) <= 1e-5 * max(ref[3], 1e-30) and abs(float(uns) - ref[4]) <= 1e-5 * max(ref[4], 1e-30)
        assert (float(uns) == 0.0) if lab.all() else (float(sup) == 0.0)


def test_joint_location_loss2_vs_golden(capsys):
    """loss.py:207-277 on the device: soft-argmax -> hipe_patch_to_camera -> fixed-order fp64 sum."""
    from nets import loss as L
    g = gold("loss2")
    args = [_cu(g[k]) for k in ("logits", "gt_label")] + [_cu(np.ones_like(g["gt_label"]))] + \
        [_cu(g[k]) for k in ("joint_cam", "joint_cam_normalized", "bbox", "scale", "R", "trans", "K", "tprime")]
    crit = L.JointLocationLoss2()
    v = crit(*args)
    assert v.dtype == torch.float64 and v.dim() == 0 and v.is_cuda and not v.requires_grad
    assert abs(v.item() - g["loss64"]) <= 1e-5 * g["loss64"]
    assert crit.size_average is False                    # loss.py:270
    assert "Warning" not in capsys.readouterr().out      # the labels reproduce joint_cam_normalized
    # an inconsistent joint_cam_normalized trips the reference's warning (loss.py:253-256)
    bad = list(args)
    bad[4] = args[4] + 1e-3
    L.JointLocationLoss2()(*bad)
    assert "label_3d_kpt and joint_cam_normalized are not equal" in capsys.readouterr().out
    ref = O.joint_location_loss2(g["logits"].astype(np.float64), g["gt_label"], g["joint_cam_normalized"], g["R"],
                                 g["trans"], g["K"], g["tprime"])
    assert abs(v.item() - ref) <= 1e-5 * ref


@pytest.mark.parametrize("tag", ["freihand", "panet"])
def test_procrustes_vs_golden(tag):
    import augment as A
    from hipe_b200 import evaluate as E
    g = gold("procrustes")
    X, Y = g[tag + "_X"], g[tag + "_Y"]
    for scale in (True, False):
        d, Z, T, b, c = E.compute_similarity_transform_batched(X, Y, scale)
        k = "%s_s%d_" % (tag, int(scale))
        assert np.abs(Z.cpu().numpy() - g[k + "Z"]).max() <= 1e-9
        np.testing.assert_allclose(T.cpu().numpy(), g[k + "T"], rtol=0, atol=1e-9)
        np.testing.assert_allclose(d.cpu().numpy(), g[k + "d"], rtol=0, atol=1e-10)
        np.testing.assert_allclose(b.cpu().numpy(), g[k + "b"], rtol=1e-10)
        np.testing.assert_allclose(c.cpu().numpy(), g[k + "c"], rtol=0, atol=1e-9)
    # the per-sample numpy-in / numpy-out form the reference's callers use (FreiHand.py:742)
    d0, Z0, T0, b0, c0 = A.compute_similarity_transform(X[3], Y[3], compute_optimal_scale=True)
    np.testing.assert_allclose(Z0, g[tag + "_s1_Z"][3], rtol=0, atol=1e-9)
    assert isinstance(d0, float) and Z0.shape == (21, 3) and T0.shape == (3, 3) and c0.shape == (3,)
    np.testing.assert_allclose(E.align_w_scale_batched(X, Y).cpu().numpy(), g[tag + "_aws_Z"], rtol=0, atol=1e-9)
    pa, raw, _ = E.pa_mpjpe(X, Y)
    assert abs(pa.item() - g[tag + "_pa_mpjpe"]) <= 1e-6       # 1e-3 mm
    assert abs(raw.item() - g[tag + "_mpjpe"]) <= 1e-6
    assert abs(pa.item() - g[tag + "_pa_mpjpe"]) <= 1e-12      # and in fact to fp64 round-off
    for nm, pred, vmax in (("raw", Y, 1.0), ("aligned", g[tag + "_aws_Z"], 0.05)):
        dist = torch.linalg.norm(_cu(X) - _cu(pred), dim=2)
        em, emed, auc, pck, thr = E.keypoint_measures(dist, 0.0, vmax, 100)
        np.testing.assert_allclose([em.item(), emed.item(), auc.item()], g["%s_%s_measures" % (tag, nm)], rtol=1e-9)
        np.testing.assert_allclose(pck.cpu().numpy(), g["%s_%s_pck" % (tag, nm)], rtol=1e-9, atol=1e-12)
