"""The patch -> camera coordinate chain: oracle against the golden vectors
frozen from the live reference (CPU) and the CUDA kernels against both (GPU).  Tolerances:
1e-9 m on camera-space joints and aligned joints, 1e-9 px on image-space coordinates relative
to a 224 px patch, 1e-6 m (1e-3 mm) on PA-MPJPE -- all fp64 work."""
import numpy as np
import pytest
import torch

from oracle import chain_np as C
from oracle import procrustes_np as P
from tests import synth
from tests.golden import ref_loader
from tests.test_oracle_golden import gold


def _oracle_all(g):
    img, cam = C.camera_chain(g["patch"], g["trans"], g["tprime"], g["K"], R=g["R"], bone_length=g["bone_len"])
    _, cam_norot = C.camera_chain(g["patch"], g["trans"], g["tprime"], g["K"], R=None, bone_length=g["bone_len"])
    return img, cam, cam_norot


def test_chain_oracle_golden():
    g = gold("chain")
    img, cam, cam_norot = _oracle_all(g)
    assert np.abs(img - g["img"][..., 0:3]).max() < 1e-10
    assert np.abs(cam - g["cam"]).max() < 1e-12
    assert np.abs(cam_norot - g["cam_norot"]).max() < 1e-12
    T = C.gen_trans_from_patch(*g["bbox"].T, 224, 224, 1.0, inv=True)
    assert np.abs(T - g["trans"]).max() < 1e-12
    _, Z, _, _, _ = P.similarity_transform(g["gt_cam"], cam, True)
    assert np.abs(Z - g["Z"]).max() < 1e-12
    pa = np.mean(np.sqrt(((Z - g["gt_cam"]) ** 2).sum(axis=2)))
    assert abs(pa - g["pa_mpjpe"]) < 1e-12


def test_joint_location_result_oracle():
    rng = np.random.default_rng(5)
    c = rng.uniform(-0.5, 0.5, size=(3, 63)).astype(np.float32)
    o = C.joint_location_result(c, 224, 192)
    assert o.shape == (3, 21, 4) and np.all(o[..., 3] == 1)
    assert np.allclose(o[0, 4, 0], (float(c[0, 12]) + 0.5) * 224)
    assert np.allclose(o[1, 2, 1], (float(c[1, 7]) + 0.5) * 192)
    assert np.allclose(o[2, 20, 2], float(c[2, 62]) * 224)


@pytest.mark.skipif(not ref_loader.available(), reason="/root/reference is not mounted here")
def test_chain_oracle_vs_live_reference():
    ch = ref_loader.load_chain_functions()
    c = synth.chain_inputs(12, seed=77)
    for inv in (True, False):
        T = np.array([ch.gen_trans_from_patch_cv(*c["bbox"][i], 224, 224, 1.25, inv=inv) for i in range(12)])
        assert np.abs(T - C.gen_trans_from_patch(*c["bbox"].T, 224, 224, 1.25, inv=inv)).max() < 1e-9
    T = C.gen_trans_from_patch(*c["bbox"].T, 224, 224, 1.0, inv=True)
    img, cam = C.camera_chain(c["patch"], T, c["tprime"], c["K"], R=c["R"], bone_length=c["bone_len"])
    for i in range(12):
        pim = ch.trans_coords_from_patch_to_org_3d(c["patch"][i].copy(), 0, 0, 0, 0, 224, 224, 1.0, T[i], c["tprime"][i])
        assert np.abs(pim[:, 0:3] - img[i]).max() < 1e-10
        p3 = np.matmul(c["R"][i].T, ch.pixel2cam(pim[:, 0:3].copy(), c["K"][i]).T).T
        p3, _ = ch.scale_result(None, p3, method="scale", bone_length=float(c["bone_len"][i]))
        assert np.abs(p3 - cam[i]).max() < 1e-12


# ------------------------------------------------------------------------------------------
# GPU
# ------------------------------------------------------------------------------------------
def _cu(a):
    return torch.as_tensor(np.ascontiguousarray(a)).cuda()


@pytest.mark.gpu
def test_chain_kernel_vs_golden():
    from hipe_b200 import evaluate as E
    g = gold("chain")
    o = E.camera_joints(_cu(g["patch"]), g["tprime"], g["K"], trans=g["trans"], R=g["R"], bone_len=g["bone_len"],
                        want=("img", "cam"))
    assert np.abs(o["img"].cpu().numpy() - g["img"][..., 0:3]).max() < 1e-9 * 224
    assert np.abs(o["cam"].cpu().numpy() - g["cam"]).max() < 1e-9
    # evaluation split: trans from the crop box, no rotation (FreiHand.py:853-883)
    o2 = E.camera_joints(_cu(g["patch"]), g["tprime"], g["K"], bbox=g["bbox"], bone_len=g["bone_len"])
    assert np.abs(o2["cam"].cpu().numpy() - g["cam_norot"]).max() < 1e-9


@pytest.mark.gpu
def test_fused_chain_procrustes_vs_golden():
    from hipe_b200 import evaluate as E
    g = gold("chain")
    pa, raw, o = E.evaluate_chain(_cu(g["patch"]), g["gt_cam"], g["tprime"], g["K"], trans=g["trans"], R=g["R"],
                                  bone_len=g["bone_len"])
    assert np.abs(o["cam"].cpu().numpy() - g["cam"]).max() < 1e-9
    assert np.abs(o["Z"].cpu().numpy() - g["Z"]).max() < 1e-9
    assert abs(pa.item() - float(g["pa_mpjpe"])) < 1e-6
    assert abs(raw.item() - float(g["mpjpe"])) < 1e-6


@pytest.mark.gpu
@pytest.mark.parametrize("n,J", [(1, 21), (37, 21), (5, 40), (3960, 21)])
def test_chain_from_soft_argmax_output_vs_oracle(n, J):
    """coords_kind 0: the float32 (N, 3J) soft-argmax output goes straight in (no host round trip)."""
    from hipe_b200 import evaluate as E
    c = synth.chain_inputs(n, J=J, seed=n + J) if J == 21 else None
    if c is None:
        c = synth.chain_inputs(n, J=21, seed=n + J)
        reps = (J + 20) // 21
        c["patch"] = np.tile(c["patch"], (1, reps, 1))[:, :J]
        c["gt_cam"] = np.tile(c["gt_cam"], (1, reps, 1))[:, :J] + np.arange(J)[None, :, None] * 1e-3
    coords = np.empty((n, J, 3))
    coords[..., 0] = c["patch"][..., 0] / 224 - 0.5
    coords[..., 1] = c["patch"][..., 1] / 224 - 0.5
    coords[..., 2] = c["patch"][..., 2] / 224
    coords32 = coords.reshape(n, 3 * J).astype(np.float32)
    patch = C.joint_location_result(coords32, 224, 224)
    T = C.gen_trans_from_patch(*c["bbox"].T, 224, 224, 1.0, inv=True)
    _, cam = C.camera_chain(patch, T, c["tprime"], c["K"], R=c["R"], bone_length=c["bone_len"])
    _, Z, _, _, _ = P.similarity_transform(c["gt_cam"], cam, True)
    pa, raw, o = E.evaluate_chain(_cu(coords32), c["gt_cam"], c["tprime"], c["K"], bbox=c["bbox"], R=c["R"],
                                  bone_len=c["bone_len"], num_joints=J)
    assert np.abs(o["cam"].cpu().numpy() - cam).max() < 1e-9
    assert np.abs(o["Z"].cpu().numpy() - Z).max() < 1e-9
    assert abs(pa.item() - np.mean(np.sqrt(((Z - c["gt_cam"]) ** 2).sum(axis=2)))) < 1e-6


@pytest.mark.gpu
def test_per_sample_drop_ins():
    import augment as A
    g = gold("chain")
    pim = A.trans_coords_from_patch_to_org_3d(g["patch"][3], 0, 0, 0, 0, 224, 224, 1.0, g["trans"][3], g["tprime"][3])
    assert pim.shape == g["patch"][3].shape and np.abs(pim - g["img"][3]).max() < 1e-9 * 224
    px = g["img"][3][:, 0:3].copy()
    cam = A.pixel2cam(px, g["K"][3])
    ref = C.pixel2cam(g["img"][3:4, :, 0:3], g["K"][3:4])[0]
    assert np.abs(cam - ref).max() < 1e-9
    assert np.allclose(px[:, 2], g["img"][3][:, 2] / 1000)     # the reference's in-place side effect


@pytest.mark.gpu
def test_chain_argument_errors():
    from hipe_b200 import _lib, ops
    c = synth.chain_inputs(2, seed=1)
    with pytest.raises(ValueError):
        ops.patch_to_camera(_cu(c["patch"]), 2, 21, c["tprime"], c["K"])           # neither trans nor bbox
    with pytest.raises(_lib.HipeError):
        ops.patch_to_camera(_cu(c["patch"]), 2, 21, c["tprime"], c["K"], bbox=c["bbox"], bone_len=c["bone_len"],
                            bone=(9, 99))
    with pytest.raises(RuntimeError):
        ops.patch_to_camera(torch.as_tensor(c["patch"]), 2, 21, c["tprime"], c["K"], bbox=c["bbox"])  # CPU tensor


# ------------------------------------------------------------------------------------------
# keypoint measures with a visibility mask
# ------------------------------------------------------------------------------------------
def _masked_case(n, K=21, seed=0):
    rng = np.random.default_rng(seed)
    dist = np.abs(rng.standard_normal((n, K))) * 0.02
    dist[rng.uniform(size=dist.shape) < 0.05] = 0.0           # ties at the first threshold
    dist[0, 1] = 0.05                                         # exactly the last threshold
    vis = rng.uniform(size=(n, K)) < 0.8
    vis[:, 3] = False                                         # a keypoint that is never fed
    vis[:, 5] = False
    vis[: min(n, 4), 5] = True                                # an even, tiny count
    return dist, vis


@pytest.mark.skipif(not ref_loader.available(), reason="/root/reference is not mounted here")
def test_masked_measures_oracle_vs_live_evalutil():
    ev = ref_loader.load_eval_module()
    dist, vis = _masked_case(257, seed=3)
    u = ev.EvalUtil(num_kp=21)
    for i in range(len(dist)):                                # feed() takes points; make them 1-D offsets
        gt = np.zeros((21, 3))
        pred = np.zeros((21, 3))
        pred[:, 0] = dist[i]
        u.feed(gt, vis[i], pred)
    m = u.get_measures(0.0, 0.05, 100)
    o = P.keypoint_measures_masked(dist, vis, 0.0, 0.05, 100)
    np.testing.assert_allclose([o[0], o[1], o[2]], [m[0], m[1], m[2]], rtol=1e-13)
    np.testing.assert_allclose(o[3], m[3], rtol=0, atol=1e-15)


@pytest.mark.gpu
@pytest.mark.parametrize("n", [1, 4, 257, 3960])
def test_masked_measures_kernel_vs_oracle(n):
    from hipe_b200 import evaluate as E
    dist, vis = _masked_case(n, seed=n)
    o = P.keypoint_measures_masked(dist, vis, 0.0, 0.05, 100)
    em, emed, auc, pck, thr, per = E.keypoint_measures(_cu(dist), 0.0, 0.05, 100, vis=vis, per_keypoint=True)
    np.testing.assert_allclose([em.item(), emed.item(), auc.item()], [o[0], o[1], o[2]], rtol=1e-12)
    np.testing.assert_allclose(pck.cpu().numpy(), o[3], rtol=0, atol=1e-14)
    assert np.array_equal(thr.cpu().numpy(), o[4])
    cnt = per["count"].cpu().numpy()
    assert np.array_equal(cnt, vis.sum(axis=0).astype(np.float64)) and cnt[3] == 0
    k = int(np.flatnonzero(cnt > 0)[-1])
    assert per["median"][k].item() == np.median(dist[vis[:, k], k])          # selection is exact
    # all-visible form equals the unmasked one
    a = E.keypoint_measures(_cu(dist), 0.0, 0.05, 100)
    b = P.keypoint_measures_masked(dist, np.ones_like(vis), 0.0, 0.05, 100)
    np.testing.assert_allclose([a[0].item(), a[1].item(), a[2].item()], [b[0], b[1], b[2]], rtol=1e-12)


# ------------------------------------------------------------------------------------------
# per-joint softmax entropy
# ------------------------------------------------------------------------------------------
@pytest.mark.skipif(not ref_loader.available(), reason="/root/reference is not mounted here")
def test_entropy_oracle_vs_live_reference_and_scipy():
    from scipy import stats
    from oracle import integral_np as O
    Lm = ref_loader.load_loss_module()
    hm = synth.heatmaps(2, 21, 6, 7, 8, "peaked", seed=4)
    p = Lm.softmax_integral_tensor2(torch.as_tensor(hm).double(), 21, 8, 7, 6).numpy()
    ref = np.array([[stats.entropy(p[b, j]) for j in range(21)] for b in range(2)])
    assert np.abs(O.softmax_entropy(hm, 21, 6, 7, 8) - ref).max() < 1e-12


@pytest.mark.gpu
@pytest.mark.parametrize("geom,regime,dtype", [((16, 16, 16), "peaked", "f32"), ((64, 64, 64), "sharp", "f32"),
                                               ((5, 6, 7), "noise", "f32"), ((32, 32, 32), "peaked", "bf16")])
def test_entropy_kernel_vs_oracle(geom, regime, dtype):
    """Tolerance 1e-5 relative (+1e-6 nats): fp32 exponentials, fp64 accumulation."""
    from nets import loss as L
    from oracle import integral_np as O
    D, H, W = geom
    hm = synth.heatmaps(2, 21, D, H, W, regime, seed=8)
    x = torch.as_tensor(hm).cuda()
    if dtype == "bf16":
        x = x.bfloat16()
        hm = x.float().cpu().numpy()           # the oracle sees the rounded logits
    ref = O.softmax_entropy(hm, 21, D, H, W)
    got = L.softmax_entropy(x, 21, W, H, D).cpu().numpy()
    assert got.shape == (2, 21)
    assert (np.abs(got - ref) <= 1e-5 * np.abs(ref) + 1e-6).all(), np.abs(got - ref).max()
    hm2 = hm.copy().reshape(2, 21, -1)
    hm2[0, 0, :5] = -np.inf                    # -inf logits contribute 0, not NaN
    y = torch.as_tensor(hm2.reshape(hm.shape)).cuda()
    if dtype == "bf16":
        y = y.bfloat16()
    g2 = L.softmax_entropy(y, 21, W, H, D).cpu().numpy()
    assert np.isfinite(g2).all() and abs(g2[0, 0] - O.softmax_entropy(hm2.reshape(hm.shape), 21, D, H, W)[0, 0]) < 1e-4
