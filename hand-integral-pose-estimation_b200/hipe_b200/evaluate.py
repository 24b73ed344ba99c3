"""Evaluation-side operators: batched Procrustes alignment and the MPJPE reductions.

Replaces the per-sample Python loop of data/FreiHand/FreiHand.py:707-767,804-807 around
augment.compute_similarity_transform (common/augment.py:18-77) and main/eval.py:142-165's
align_w_scale, with one kernel launch over all samples (one warp per sample, 3x3 SVD in
fp64 registers) plus fixed-order fp64 reductions.  CUDA only.
"""
import numpy as np
import torch

from . import ops

MODE_SIMILARITY = 0   # common/augment.py:18-77 (reflection repaired)
MODE_ALIGN_W_SCALE = 1  # main/eval.py:142-165 (no repair, +1e-8 on the norms)


def _dev(device=None):
    if not torch.cuda.is_available():
        raise RuntimeError("hipe_b200 has no CPU path: no CUDA device is available")
    return torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)


def _as_cuda64(a, device=None):
    if isinstance(a, torch.Tensor):
        if not a.is_cuda:
            a = a.to(_dev(device))
        return a.to(torch.float64)
    return torch.as_tensor(np.ascontiguousarray(a, dtype=np.float64)).to(_dev(device))


def compute_similarity_transform_batched(X, Y, compute_optimal_scale=False, device=None):
    """(N,J,3) targets X and inputs Y -> (d (N,), Z (N,J,3), T (N,3,3), b (N,), c (N,3)) CUDA fp64
    tensors with Z = b * Y @ T + c, exactly the per-sample outputs of augment.py:18-77."""
    Xc, Yc = _as_cuda64(X, device), _as_cuda64(Y, device)
    o = ops.procrustes(Xc, Yc, MODE_SIMILARITY, compute_optimal_scale, want=("d", "Z", "T", "b", "c"))
    return o["d"], o["Z"], o["T"], o["b"], o["c"]


def align_w_scale_batched(gt, pred, device=None):
    """main/eval.py:142-165 for a batch: the aligned predictions (N,J,3), CUDA fp64."""
    Xc, Yc = _as_cuda64(gt, device), _as_cuda64(pred, device)
    return ops.procrustes(Xc, Yc, MODE_ALIGN_W_SCALE, True, want=("Z",))["Z"]


def pa_mpjpe(gt, pred, device=None, mode=MODE_SIMILARITY):
    """data/FreiHand/FreiHand.py:742,766-767,804-807: (PA-MPJPE, MPJPE) over all samples and joints,
    as 0-dim CUDA fp64 tensors, plus the aligned predictions."""
    Xc, Yc = _as_cuda64(gt, device), _as_cuda64(pred, device)
    o = ops.procrustes(Xc, Yc, mode, True, want=("Z", "pa_dist", "raw_dist"))
    n = o["pa_dist"].numel()
    return ops.sum_f64(o["pa_dist"], 1.0 / n), ops.sum_f64(o["raw_dist"], 1.0 / n), o["Z"]


def camera_joints(coords, tprime, K, trans=None, bbox=None, R=None, bone_len=None, num_joints=21,
                  patch=(224.0, 224.0), dst=(224.0, 224.0), want=("cam",)):
    """Soft-argmax output (B, 3J) float32 -- or patch pixels (B, J, 3|4) -- to camera-space joints in
    metres: get_joint_location_result's rescale, trans_coords_from_patch_to_org_3d, pixel2cam, R^T
    and the bone-length rescale (common/augment.py:89-103,176-217; data/FreiHand/FreiHand.py:632-653,
    722), one launch, nothing leaves the device.  Returns a dict with the tensors named in `want`
    ("img" = image-space (u, v, z_mm), "cam")."""
    if not isinstance(coords, torch.Tensor):
        coords = torch.as_tensor(np.ascontiguousarray(coords, dtype=np.float64)).to(_dev())
    N = coords.shape[0]
    return ops.patch_to_camera(coords, N, num_joints, tprime, K, trans=trans, bbox=bbox, R=R, bone_len=bone_len,
                               patch=patch, dst=dst, want=want)


def evaluate_chain(coords, gt_cam, tprime, K, trans=None, bbox=None, R=None, bone_len=None, num_joints=21,
                   patch=(224.0, 224.0), dst=(224.0, 224.0), mode=MODE_SIMILARITY):
    """The per-sample body of FreiHand.evaluate (data/FreiHand/FreiHand.py:703-767,804-807) for a whole
    batch in one launch: chain -> Procrustes -> per-joint distances.  Returns (PA-MPJPE, MPJPE) as
    0-dim CUDA fp64 tensors and a dict with cam, Z, pa_dist, raw_dist."""
    if not isinstance(coords, torch.Tensor):
        coords = torch.as_tensor(np.ascontiguousarray(coords, dtype=np.float64)).to(_dev())
    N = coords.shape[0]
    o = ops.patch_to_camera(coords, N, num_joints, tprime, K, trans=trans, bbox=bbox, R=R, bone_len=bone_len,
                            patch=patch, dst=dst, gt=gt_cam, mode=mode, optimal_scale=True,
                            want=("cam", "Z", "pa_dist", "raw_dist"))
    n = o["pa_dist"].numel()
    return ops.sum_f64(o["pa_dist"], 1.0 / n), ops.sum_f64(o["raw_dist"], 1.0 / n), o


def keypoint_measures(dist, val_min=0.0, val_max=0.05, steps=100, vis=None, per_keypoint=False):
    """main/eval.py:25-106 (EvalUtil.feed + get_measures) from per-keypoint distances (N, K) on the
    device, one launch pair (hipe_keypoint_measures): mean / median EPE per keypoint, PCK over the
    linspace thresholds, trapezoid AUC, then the means over the keypoints that were fed.
    `vis` (N, K), optional: only entries with vis != 0 are fed (main/eval.py:45-47).
    Returns (epe_mean, epe_median, auc, pck_curve, thresholds) as CUDA fp64 tensors; with
    per_keypoint=True also a dict of the per-keypoint count / mean / median / auc / pck."""
    if not isinstance(dist, torch.Tensor):
        dist = torch.as_tensor(np.ascontiguousarray(dist, dtype=np.float64)).to(_dev())
    out = ops.keypoint_measures(dist, val_min, val_max, steps, vis)
    # np.linspace's own roundings (torch.linspace differs in the last bit); the kernel forms the same values
    thr = torch.as_tensor(np.linspace(val_min, val_max, steps)).to(dist.device)
    res = (out["summary"][0], out["summary"][1], out["summary"][2], out["pck_curve"], thr)
    return res + (out["per_keypoint"],) if per_keypoint else res
