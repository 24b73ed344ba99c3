"""ORACLE (test infrastructure, not product code): numpy restatement of the reference's
Procrustes alignment and evaluation metrics.  See `oracle/integral_np.py` for the import
rules and the parity-pinning statement (goldens: tests/golden/procrustes_*.npz).

All functions are batched over a leading sample axis N and evaluate in float64, the dtype the
reference runs them in.  Citations are into /root/reference.
"""
import numpy as np


def similarity_transform(X, Y, compute_optimal_scale=False):
    """common/augment.py:18-77 for a batch: X, Y are (N, J, 3) targets / inputs.
    Returns (d (N,), Z (N,J,3), T (N,3,3), b (N,), c (N,3)) with Z = b * Y @ T + c.

    Steps and the lines they restate: centroids (:35-36), centring (:38-39), squared
    Frobenius norms and their roots (:41-46), unit scaling (:49-50), the 3x3 correlation
    A = X0^T Y0 (:53), its SVD and T = V U^T (:54-56), the reflection repair that flips the
    last right-singular vector and last singular value by sign(det T) (:59-62), trace (:64),
    the scaled (:66-69) or unscaled (:70-73) branch, and the translation (:75)."""
    X = np.asarray(X, dtype=np.float64)
    Y = np.asarray(Y, dtype=np.float64)
    mu_x = X.mean(axis=1, keepdims=True)
    mu_y = Y.mean(axis=1, keepdims=True)
    xc, yc = X - mu_x, Y - mu_y
    ss_x = (xc ** 2).sum(axis=(1, 2))
    ss_y = (yc ** 2).sum(axis=(1, 2))
    n_x, n_y = np.sqrt(ss_x), np.sqrt(ss_y)
    xc = xc / n_x[:, None, None]
    yc = yc / n_y[:, None, None]
    A = np.einsum("nji,njk->nik", xc, yc)
    U, s, Vt = np.linalg.svd(A, full_matrices=False)
    V = np.swapaxes(Vt, 1, 2).copy()
    T0 = V @ np.swapaxes(U, 1, 2)
    sgn = np.sign(np.linalg.det(T0))
    V[:, :, -1] *= sgn[:, None]
    s = s.copy()
    s[:, -1] *= sgn
    T = V @ np.swapaxes(U, 1, 2)
    tr = s.sum(axis=1)
    if compute_optimal_scale:
        b = tr * n_x / n_y
        d = 1.0 - tr ** 2
        Z = (n_x * tr)[:, None, None] * (yc @ T) + mu_x
    else:
        b = np.ones_like(tr)
        d = 1.0 + ss_y / ss_x - 2.0 * tr * n_y / n_x
        Z = n_y[:, None, None] * (yc @ T) + mu_x
    c = mu_x[:, 0, :] - b[:, None] * np.einsum("ni,nik->nk", mu_y[:, 0, :], T)
    return d, Z, T, b, c


def align_w_scale(gt, pred):
    """main/eval.py:142-165 for a batch (the FreiHAND toolkit alignment): centre (:145-148),
    scale by Frobenius norm + 1e-8 (:151-154), orthogonal Procrustes WITHOUT a reflection
    repair (:157; scipy: R = U Vt of svd(m1^T m2 ^T...), scale = sum of singular values),
    then map the prediction back (:160-161).  Returns the aligned prediction (N, J, 3)."""
    gt = np.asarray(gt, dtype=np.float64)
    pred = np.asarray(pred, dtype=np.float64)
    t1 = gt.mean(axis=1, keepdims=True)
    t2 = pred.mean(axis=1, keepdims=True)
    a, b = gt - t1, pred - t2
    s1 = np.sqrt((a ** 2).sum(axis=(1, 2))) + 1e-8
    s2 = np.sqrt((b ** 2).sum(axis=(1, 2))) + 1e-8
    a = a / s1[:, None, None]
    b = b / s2[:, None, None]
    # scipy.linalg.orthogonal_procrustes(A, B): u, w, vt = svd(B^T A ^T...) -> R = u @ vt
    # with M = (B^T A)^T = A^T B; here A = gt_n, B = pred_n.
    M = np.swapaxes(np.einsum("nji,njk->nik", b, a), 1, 2)
    u, w, vt = np.linalg.svd(M)
    R = u @ vt
    scale = w.sum(axis=1)
    out = (b @ np.swapaxes(R, 1, 2)) * scale[:, None, None]
    return out * s1[:, None, None] + t1


def mpjpe_pair(gt, pred, aligned):
    """data/FreiHand/FreiHand.py:766-767,804-807: squared per-coordinate errors, summed over
    xyz, square-rooted, averaged over all samples and joints.  Returns (PA-MPJPE, MPJPE)."""
    p1 = np.sqrt(((aligned - gt) ** 2).sum(axis=2)).mean()
    p2 = np.sqrt(((pred - gt) ** 2).sum(axis=2)).mean()
    return p1, p2


def pa_mpjpe(gt, pred):
    """data/FreiHand/FreiHand.py:742 + :766-807: align with optimal scale, then `mpjpe_pair`."""
    _, Z, _, _, _ = similarity_transform(gt, pred, True)
    return mpjpe_pair(np.asarray(gt, np.float64), np.asarray(pred, np.float64), Z)


def keypoint_measures(gt, pred, val_min=0.0, val_max=0.05, steps=100):
    """main/eval.py:15-106 (`EvalUtil.feed` + `get_measures`) for all-visible keypoints:
    per-keypoint Euclidean distances (:38-39), mean/median EPE per keypoint (:55-63), PCK
    `mean(dist <= t)` over `linspace(val_min, val_max, steps)` (:46-53,67,90-93), AUC by the
    trapezoid rule normalised by the threshold span (:69,97-98), then means over keypoints
    (:101-104).  Returns (epe_mean, epe_median, auc, pck_curve (steps,), thresholds)."""
    dist = np.sqrt(((np.asarray(gt, np.float64) - np.asarray(pred, np.float64)) ** 2).sum(axis=2))  # (N, K)
    thr = np.linspace(val_min, val_max, steps)
    span = np.trapezoid(np.ones_like(thr), thr)
    pck = (dist[:, :, None] <= thr[None, None, :]).mean(axis=0)      # (K, steps)
    auc = np.trapezoid(pck, thr, axis=1) / span
    return dist.mean(axis=0).mean(), np.median(dist, axis=0).mean(), auc.mean(), pck.mean(axis=0), thr


def keypoint_measures_masked(dist, vis, val_min=0.0, val_max=0.05, steps=100):
    """main/eval.py:25-106 with the visibility gate of `feed` (:45-47): keypoint k only collects the
    distances of samples with vis[n, k] true; keypoints that collected nothing are skipped in
    the final means (:76-78).  dist, vis: (N, K)."""
    dist = np.asarray(dist, np.float64)
    vis = np.asarray(vis).astype(bool)
    thr = np.linspace(val_min, val_max, steps)
    span = np.trapezoid(np.ones_like(thr), thr)
    means, medians, aucs, curves = [], [], [], []
    for k in range(dist.shape[1]):
        d = dist[vis[:, k], k]
        if d.size == 0:
            continue
        means.append(d.mean())
        medians.append(np.median(d))
        pck = np.array([np.mean((d <= t).astype(float)) for t in thr])
        curves.append(pck)
        aucs.append(np.trapezoid(pck, thr) / span)
    return np.mean(means), np.mean(medians), np.mean(aucs), np.mean(np.array(curves), 0), thr
