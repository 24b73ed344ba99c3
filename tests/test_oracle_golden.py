"""CPU: the oracle (numpy restatement, and the torch restatement bench.py times) against the
golden vectors frozen from the live reference, and against the live reference itself when
/root/reference is mounted (build container only).  This is what pins the oracle."""
import hashlib
import os

import numpy as np
import pytest
import torch

from oracle import integral_np as O
from oracle import integral_torch as OT
from oracle import procrustes_np as P
from tests import synth
from tests.golden import ref_loader

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
J = 21
SMALL = ["cube8", "ragged567", "box6x10x12"]


def gold(name):
    return np.load(os.path.join(GOLD, name + ".npz"))


def case_input(g):
    B, D, H, W = (int(v) for v in g["shape"])
    if "logits" in g.files:
        hm = g["logits"]
    else:
        hm = synth.heatmaps(B, J, D, H, W, str(g["regime"]), seed=int(g["seed"]))
        if hashlib.sha256(hm.tobytes()).hexdigest() != str(g["input_sha256"]):
            pytest.skip("numpy's Generator stream differs from the one the golden was made with")
    return hm, (B, D, H, W)


@pytest.mark.parametrize("name", SMALL + ["cube64"])
def test_numpy_oracle_coords_and_loss(name):
    g = gold("integral_" + name)
    hm, (B, D, H, W) = case_input(g)
    c = O.integral_coords(hm.astype(np.float64), J, W, H, D)
    np.testing.assert_allclose(c, g["coords64"], rtol=0, atol=5e-15)
    loss = O.joint_location_loss(hm.astype(np.float64), g["gt"], g["vis"])
    assert abs(loss - g["loss64"]) <= 1e-13 * abs(g["loss64"])
    # the reference's own fp32 arithmetic stays within its documented distance of fp64
    assert np.abs(g["coords32"] - g["coords64"]).max() < 2e-4


@pytest.mark.parametrize("name", SMALL)
def test_numpy_oracle_gradients_probs_expectations(name):
    g = gold("integral_" + name)
    hm, (B, D, H, W) = case_input(g)
    x = hm.astype(np.float64)
    grad = O.joint_location_loss_grad(x, g["gt"], g["vis"]).reshape(B * J, -1)
    np.testing.assert_allclose(grad, g["loss_grad64"], rtol=0, atol=1e-15)
    vjp = O.integral_coords_vjp(x, J, W, H, D, g["upstream"]).reshape(B * J, -1)
    np.testing.assert_allclose(vjp, g["vjp64"], rtol=0, atol=2e-15)
    p = O.volume_softmax(x, J)
    np.testing.assert_allclose(p, g["probs64"], rtol=1e-13, atol=0)
    ex, ey, ez = O.marginal_expectations(g["probs64"], J, W, H, D)
    np.testing.assert_allclose(np.concatenate([ex, ey, ez], axis=2), g["expect64"], rtol=1e-13)
    assert abs(O.compute_mpjpe(g["coords64"], g["gt"]) - g["mpjpe64"]) < 1e-14
    r = O.joint_location_result(224, 224, x)
    c = g["coords64"].reshape(B, J, 3)
    np.testing.assert_allclose(r[..., 0], (c[..., 0] + 0.5) * 224, atol=1e-11)
    np.testing.assert_allclose(r[..., 2], c[..., 2] * 224, atol=1e-11)
    assert (r[..., 3] == 1).all()


def test_numpy_oracle_gradient_cube64_sampled():
    g = gold("integral_cube64")
    hm, (B, D, H, W) = case_input(g)
    x = hm.astype(np.float64)
    idx = g["grad_idx"]
    grad = O.joint_location_loss_grad(x, g["gt"], g["vis"]).reshape(B * J, -1)
    np.testing.assert_allclose(grad[:, idx], g["loss_grad64_sampled"], rtol=0, atol=1e-16)
    np.testing.assert_allclose(np.abs(grad).sum(axis=1), g["loss_grad64_abs_sum"], rtol=1e-12)
    vjp = O.integral_coords_vjp(x, J, W, H, D, g["upstream"]).reshape(B * J, -1)
    np.testing.assert_allclose(vjp[:, idx], g["vjp64_sampled"], rtol=0, atol=1e-15)
    np.testing.assert_allclose(np.abs(vjp).max(axis=1), g["vjp64_abs_max"], rtol=1e-12)


@pytest.mark.parametrize("name", SMALL)
def test_torch_port_matches_golden(name):
    g = gold("integral_" + name)
    hm, (B, D, H, W) = case_input(g)
    x = torch.as_tensor(hm).double()
    c = OT.integral_coords(x, J, W, H, D).numpy()
    np.testing.assert_array_equal(c, g["coords64"])     # same ATen ops -> same bits
    loss, grad = OT.fwd_bwd_step(x, torch.as_tensor(g["gt"]), torch.as_tensor(g["vis"]))
    assert loss.item() == float(g["loss64"])
    np.testing.assert_allclose(grad.numpy().reshape(B * J, -1), g["loss_grad64"], rtol=0, atol=1e-16)
    c32 = OT.integral_coords(torch.as_tensor(hm), J, W, H, D).numpy()
    np.testing.assert_allclose(c32, g["coords32"], rtol=0, atol=1e-6)


def test_combined_loss_oracle():
    g = gold("combined_cube8")
    loss, smp, tmp, sup, uns = O.combined_loss(g["logits"].astype(np.float64), g["teacher"], g["gt"], g["vis"],
                                               g["labelled"])
    assert abs(loss - g["loss64"]) < 1e-13
    assert abs(smp - g["student_mpjpe"]) < 1e-14 and abs(tmp - g["teacher_mpjpe"]) < 1e-14
    assert abs(sup - g["sup"]) < 1e-12 and abs(uns - g["unsup"]) < 1e-12


@pytest.mark.parametrize("name", ["combined_cube16", "combined_cube64"])
def test_combined_loss_oracle_larger_volumes(name):
    """16^3 and 64^3 (the production volume): input regenerated from the seed, hash-checked."""
    import hashlib
    g = gold(name)
    B, D, H, W = (int(v) for v in g["shape"])
    hm = synth.heatmaps(B, 21, D, H, W, "peaked", seed=int(g["seed"]))
    assert hashlib.sha256(np.ascontiguousarray(hm).tobytes()).hexdigest() == str(g["input_sha256"])
    h64 = hm.astype(np.float64)
    loss, smp, tmp, sup, uns = O.combined_loss(h64, g["teacher"], g["gt"], g["vis"], g["labelled"])
    assert abs(loss - g["loss64"]) < 1e-12
    assert abs(smp - g["student_mpjpe"]) < 1e-13 and abs(tmp - g["teacher_mpjpe"]) < 1e-14
    assert abs(sup - g["sup"]) < 1e-11 and abs(uns - g["unsup"]) < 1e-11
    # the gradient of the lambda = 0 total = JointLocationLoss against where(labelled, gt, teacher)
    tgt = np.where(g["labelled"][:, None], g["gt"], g["teacher"][0])
    grad = O.joint_location_loss_grad(h64, tgt, g["vis"]).reshape(B * 21, -1)
    np.testing.assert_allclose(grad[:, g["grad_idx"]], g["grad64_sampled"], rtol=0, atol=1e-15)
    np.testing.assert_allclose(np.abs(grad).max(axis=1), g["grad64_abs_max"], rtol=1e-10)


def test_joint_location_loss2_oracle():
    g = gold("loss2")
    v = O.joint_location_loss2(g["logits"].astype(np.float64), g["gt_label"], g["joint_cam_normalized"], g["R"],
                               g["trans"], g["K"], g["tprime"])
    assert abs(v - g["loss64"]) <= 1e-12 * g["loss64"]


@pytest.mark.parametrize("tag", ["freihand", "panet"])
def test_procrustes_oracle_golden(tag):
    g = gold("procrustes")
    X, Y = g[tag + "_X"], g[tag + "_Y"]
    for scale in (True, False):
        d, Z, T, b, c = P.similarity_transform(X, Y, scale)
        k = "%s_s%d_" % (tag, int(scale))
        np.testing.assert_allclose(Z, g[k + "Z"], rtol=0, atol=1e-13)
        np.testing.assert_allclose(T, g[k + "T"], rtol=0, atol=1e-12)
        np.testing.assert_allclose(d, g[k + "d"], rtol=0, atol=1e-13)
        np.testing.assert_allclose(b, g[k + "b"], rtol=1e-13)
        np.testing.assert_allclose(c, g[k + "c"], rtol=0, atol=1e-12)
        np.testing.assert_allclose(np.linalg.det(T), 1.0, atol=1e-12)
        np.testing.assert_allclose(Z, b[:, None, None] * (Y @ T) + c[:, None, :], atol=1e-12)
    np.testing.assert_allclose(P.align_w_scale(X, Y), g[tag + "_aws_Z"], rtol=0, atol=1e-13)
    pa, raw = P.pa_mpjpe(X, Y)
    assert abs(pa - g[tag + "_pa_mpjpe"]) < 1e-14 and abs(raw - g[tag + "_mpjpe"]) < 1e-14
    for nm, pred, vmax in (("raw", Y, 1.0), ("aligned", g[tag + "_aws_Z"], 0.05)):
        em, emed, auc, pck, thr = P.keypoint_measures(X, pred, 0.0, vmax, 100)
        np.testing.assert_allclose([em, emed, auc], g["%s_%s_measures" % (tag, nm)], rtol=1e-12)
        np.testing.assert_allclose(pck, g["%s_%s_pck" % (tag, nm)], rtol=1e-12, atol=1e-15)
        np.testing.assert_allclose(thr, g["%s_%s_thr" % (tag, nm)], rtol=0, atol=0)


def test_mirrored_targets_differ_between_the_two_alignments():
    """align_w_scale has no reflection repair, so on mirrored targets it fits
    better than compute_similarity_transform (which must keep det T = +1)."""
    g = gold("procrustes")
    X, Y = g["freihand_X"][::16], g["freihand_Y"][::16]          # the mirrored rows
    _, Z, T, _, _ = P.similarity_transform(X, Y, True)
    Za = P.align_w_scale(X, Y)
    e_rep = np.sqrt(((Z - X) ** 2).sum(axis=2)).mean()
    e_aws = np.sqrt(((Za - X) ** 2).sum(axis=2)).mean()
    assert e_aws < e_rep
    np.testing.assert_allclose(np.linalg.det(T), 1.0, atol=1e-12)


# ---- against the live reference (build container only) --------------------------------
live = pytest.mark.skipif(not ref_loader.available(), reason="/root/reference is not mounted here")


@live
def test_oracles_vs_live_reference_random_shapes():
    L = ref_loader.load_loss_module()
    rng = np.random.default_rng(77)
    for trial in range(4):
        D, H, W = (int(v) for v in rng.integers(3, 14, size=3))
        B = int(rng.integers(1, 4))
        hm = synth.heatmaps(B, J, D, H, W, ["noise", "peaked", "sharp"][trial % 3], seed=300 + trial)
        gt, vis = synth.targets(B, J, seed=trial, zero_vis_frac=0.2)
        x = torch.as_tensor(hm).double().requires_grad_(True)
        loss = L.JointLocationLoss()(x, torch.as_tensor(gt), torch.as_tensor(vis))
        loss.backward()
        assert abs(O.joint_location_loss(hm.astype(np.float64), gt, vis) - loss.item()) < 1e-13 * abs(loss.item())
        np.testing.assert_allclose(O.joint_location_loss_grad(hm.astype(np.float64), gt, vis), x.grad.numpy(),
                                   rtol=0, atol=1e-15)
        l2, g2 = OT.fwd_bwd_step(torch.as_tensor(hm).double(), torch.as_tensor(gt), torch.as_tensor(vis))
        assert l2.item() == loss.item()
        np.testing.assert_array_equal(g2.numpy(), x.grad.numpy())


@live
def test_procrustes_oracle_vs_live_reference_full_fixture():
    cst = ref_loader.load_compute_similarity_transform()
    preds, _ = ref_loader.fixture_arrays()
    rng = np.random.default_rng(5)
    Y = preds[:400]
    X = 1.1 * np.einsum("njk,nkl->njl", Y, synth.random_rotations(len(Y), rng)) + 0.3 \
        + rng.standard_normal(Y.shape) * 0.005
    _, Z, T, b, c = P.similarity_transform(X, Y, True)
    for i in range(0, len(Y), 7):
        d0, Z0, T0, b0, c0 = cst(X[i], Y[i], compute_optimal_scale=True)
        np.testing.assert_allclose(Z[i], Z0, rtol=0, atol=1e-13)
        np.testing.assert_allclose(T[i], T0, rtol=0, atol=1e-12)
