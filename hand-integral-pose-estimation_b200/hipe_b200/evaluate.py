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


def keypoint_measures(dist, val_min=0.0, val_max=0.05, steps=100):
    """main/eval.py:25-106 (EvalUtil.feed + get_measures) from per-keypoint distances (N,K) on the
    device: mean/median EPE per keypoint, PCK over linspace thresholds, trapezoid AUC, then the
    means over keypoints.  Returns (epe_mean, epe_median, auc, pck_curve, thresholds)."""
    d = dist.to(torch.float64)
    thr = torch.linspace(val_min, val_max, steps, dtype=torch.float64, device=d.device)
    pck = (d.unsqueeze(-1) <= thr).to(torch.float64).mean(dim=0)            # (K, steps)
    span = torch.trapezoid(torch.ones_like(thr), thr)
    auc = torch.trapezoid(pck, thr, dim=1) / span
    return d.mean(dim=0).mean(), d.median(dim=0).values.mean() if d.shape[0] % 2 else _median_even(d).mean(), \
        auc.mean(), pck.mean(dim=0), thr


def _median_even(d):
    # np.median averages the two middle values for even counts; torch.median takes the lower
    s, _ = torch.sort(d, dim=0)
    n = d.shape[0]
    return 0.5 * (s[n // 2 - 1] + s[n // 2])
