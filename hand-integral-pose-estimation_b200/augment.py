"""Drop-in for the reference's common/augment.py.

When the reference's own common/augment.py is further down sys.path (the INTEGRATION.md set-up),
its source is executed in this module first, so every name the rest of the reference imports
from `augment` (get_aug_config, generate_patch_image, trans_point2d, prepare_panet_input,
test_get_joint_loc_res, projectPoints, gen_trans_from_patch_cv, ...: data/dataset.py:112-243,
data/FreiHand/FreiHand.py:745,865,888, common/nets/loss.py:141,169,228-229) is the reference's
own; then the hot-path functions below replace theirs with the sm_100a kernels:

    compute_similarity_transform          augment.py:18-77    (+ a batched form)
    get_joint_location_result             augment.py:160-184
    trans_coords_from_patch_to_org_3d     augment.py:196-217  numpy callers; tensors go to the reference's
    pixel2cam                             augment.py:89-103

Without a reference tree on the path only these functions exist (`REFERENCE_AUGMENT` is None)."""
import numpy as np
import torch

from hipe_b200 import _overlay

REFERENCE_AUGMENT, REFERENCE_AUGMENT_ERROR = _overlay.exec_shadowed(globals(), __file__, "augment.py")
_ref_trans3d = globals().get("trans_coords_from_patch_to_org_3d") if REFERENCE_AUGMENT_ERROR is None else None
_ref_pixel2cam = globals().get("pixel2cam") if REFERENCE_AUGMENT_ERROR is None else None

from hipe_b200 import evaluate as _ev                                   # noqa: E402
from nets.loss import FreiHandConfig, softmax_integral_tensor           # noqa: E402


def compute_similarity_transform(X, Y, compute_optimal_scale=False):
    """augment.py:18-77 for one (J,3) pair, numpy in / numpy out: (d, Z, T, b, c)."""
    X = np.asarray(X, dtype=np.float64)
    Y = np.asarray(Y, dtype=np.float64)
    d, Z, T, b, c = _ev.compute_similarity_transform_batched(X[None], Y[None], compute_optimal_scale)
    bb = float(b[0].item())
    return float(d[0].item()), Z[0].cpu().numpy(), T[0].cpu().numpy(), (bb if compute_optimal_scale else 1), \
        c[0].cpu().numpy()


compute_similarity_transform_batched = _ev.compute_similarity_transform_batched


def get_joint_location_result(patch_width, patch_height, preds):
    """augment.py:160-184: soft-argmax, rescale to patch pixels, append a score column of ones.
    Returns a (B, J, 4) float64 numpy array."""
    hm_width = preds.shape[-1]
    hm_height = preds.shape[-2]
    hm_depth = preds.shape[-3] // FreiHandConfig.num_joints
    num_joints = preds.shape[1] // hm_depth
    with torch.no_grad():
        pred_jts = softmax_integral_tensor(preds, num_joints, hm_width, hm_height, hm_depth)
    coords = pred_jts.detach().cpu().numpy().astype(float)
    coords = coords.reshape((coords.shape[0], int(coords.shape[1] / 3), 3))
    coords[:, :, 0] = (coords[:, :, 0] + 0.5) * patch_width
    coords[:, :, 1] = (coords[:, :, 1] + 0.5) * patch_height
    coords[:, :, 2] = coords[:, :, 2] * patch_width
    scores = np.ones((coords.shape[0], coords.shape[1], 1), dtype=float)
    return np.concatenate((coords, scores), axis=2)


def _one(a, shape):
    return torch.as_tensor(np.ascontiguousarray(np.asarray(a, dtype=np.float64).reshape(shape))).cuda()


def trans_coords_from_patch_to_org_3d(coords_in_patch, c_x, c_y, bb_width, bb_height, patch_width, patch_height,
                                      scale, trans, tprime, p=False):
    """augment.py:196-217 for one (J, C>=3) sample, numpy in / numpy out (same positional
    signature; like the reference only `trans` and `tprime` enter the arithmetic).  Tensor inputs
    (prepare_panet_input's differentiable use, augment.py:556-574) take the reference's own code."""
    if isinstance(coords_in_patch, torch.Tensor):
        if _ref_trans3d is None:
            raise TypeError("tensor inputs need the reference's common/augment.py on sys.path")
        return _ref_trans3d(coords_in_patch, c_x, c_y, bb_width, bb_height, patch_width, patch_height, scale,
                            trans, tprime, p)
    c = np.asarray(coords_in_patch, dtype=np.float64)
    J = c.shape[0]
    K = np.eye(3)[None]
    o = _ev.camera_joints(_one(c, (1, J, c.shape[1])), _one(tprime, (1,)), _one(K, (1, 3, 3)),
                          trans=_one(trans, (1, 3, 3)), num_joints=J, want=("img",))
    out = c.copy()
    out[:, 0:3] = o["img"][0].cpu().numpy()
    return out


def pixel2cam(pixel_coord, K, pp=False):
    """augment.py:89-103 for one (J, 3) sample: K^-1 (u, v, 1) z / 1000.  (The reference also
    divides the caller's z column by 1000 in place; so does this.)"""
    if isinstance(pixel_coord, torch.Tensor) and _ref_pixel2cam is not None:
        return _ref_pixel2cam(pixel_coord, K, pp)
    pc = np.asarray(pixel_coord, dtype=np.float64)
    J = pc.shape[0]
    ident = np.array([[1.0, 0.0, 0.0], [0.0, 1.0, 0.0], [0.0, 0.0, 1.0]])
    o = _ev.camera_joints(_one(pc, (1, J, 3)), _one(0.0, (1,)), _one(K, (1, 3, 3)), trans=_one(ident, (1, 3, 3)),
                          num_joints=J, want=("cam",))
    if isinstance(pixel_coord, np.ndarray) and pixel_coord.dtype.kind == "f":
        pixel_coord[..., 2] = pixel_coord[..., 2] / 1000
    return o["cam"][0].cpu().numpy()
