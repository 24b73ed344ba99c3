"""ORACLE (test infrastructure, not product code): numpy restatement of the reference's
integral-regression path.  Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s
CPU-baseline / `--impl reference` legs may import this package.

Parity status: PINNED against the live reference.  The reference has no golden vectors of
its own for this path, so `tests/golden/make_golden.py` imports the real
`common/nets/loss.py`, `common/augment.py:18-77` and `main/eval.py:15-165` in the build
container and freezes their outputs under `tests/golden/*.npz`;
`tests/test_oracle_golden.py` holds this file to those vectors.

Every function evaluates in the dtype of its input (pass float64 for ground truth, float32
for "the reference's own rounding").  Citations are into /root/reference.
"""
import numpy as np

NUM_JOINTS = 21  # data/FreiHand/FreiHand_config.py:60


# ----------------------------------------------------------------------------------------
# soft-argmax
# ----------------------------------------------------------------------------------------
def volume_softmax(logits, num_joints):
    """common/nets/loss.py:48-49 (and :61-67, `softmax_integral_tensor2`): the logits are
    viewed as (B, J, V) and normalised over the whole D*H*W volume of each joint."""
    s = np.asarray(logits)
    s = s.reshape(s.shape[0], num_joints, -1)
    e = np.exp(s - s.max(axis=2, keepdims=True))
    return e / e.sum(axis=2, keepdims=True)


def marginal_expectations(probs, num_joints, x_dim, y_dim, z_dim):
    """common/nets/loss.py:23-43.  `probs` is any array reshapeable to (B, J, z, y, x)
    (:25).  The x marginal sums out z then y (:27-29), y sums out z then x (:31-32),
    z sums out y then x (:33-34); each marginal is weighted by the 0-based voxel index
    (:35-37) and summed (:39-41).  Returns three (B, J, 1) arrays in voxel units."""
    p = np.asarray(probs)
    p = p.reshape(p.shape[0], num_joints, z_dim, y_dim, x_dim)
    mx = p.sum(axis=(2, 3))
    my = p.sum(axis=(2, 4))
    mz = p.sum(axis=(3, 4))
    ex = (mx * np.arange(x_dim)).sum(axis=2, keepdims=True)
    ey = (my * np.arange(y_dim)).sum(axis=2, keepdims=True)
    ez = (mz * np.arange(z_dim)).sum(axis=2, keepdims=True)
    return ex.astype(p.dtype), ey.astype(p.dtype), ez.astype(p.dtype)


def integral_coords(logits, num_joints, hm_width, hm_height, hm_depth):
    """common/nets/loss.py:46-59: softmax, marginal expectations, `/dim - 0.5`, interleave
    as [x0, y0, z0, x1, ...] -> (B, 3J).  Note the (W, H, D) argument order (:46,51)."""
    p = volume_softmax(logits, num_joints)
    ex, ey, ez = marginal_expectations(p, num_joints, hm_width, hm_height, hm_depth)
    out = np.concatenate([ex / float(hm_width) - 0.5,
                          ey / float(hm_height) - 0.5,
                          ez / float(hm_depth) - 0.5], axis=2)
    return out.reshape(out.shape[0], num_joints * 3)


def _dims_of(heatmap_out):
    """common/nets/loss.py:78-80: W, H from the last two axes, D = channels // 21."""
    return heatmap_out.shape[-1], heatmap_out.shape[-2], heatmap_out.shape[-3] // NUM_JOINTS


def joint_location_loss(heatmap_out, gt_coord, gt_vis, size_average=True, return_coords=False):
    """common/nets/loss.py:74-91: L1 between soft-argmax coordinates and targets, masked by
    visibility, summed, divided by the batch length when `size_average` (:88-89).  The result
    carries the promoted dtype of (coords, targets), as torch's type promotion does."""
    joint_num = int(gt_coord.shape[1] / 3)
    w, h, d = _dims_of(heatmap_out)
    c = integral_coords(heatmap_out, joint_num, w, h, d)
    terms = np.abs(c - gt_coord) * gt_vis
    total = terms.sum()
    if size_average:
        total = total / len(c)
    return (total, c) if return_coords else total


def integral_coords_vjp(logits, num_joints, hm_width, hm_height, hm_depth, grad_coords):
    """What autograd produces for `loss.backward()` through common/nets/loss.py:46-59
    (main/train.py:113), written in closed form:
        dL/ds_i = p_i * sum_a g_a * (idx_a(i) - E_a) / dim_a .
    `grad_coords` is (B, 3J).  Returns an array shaped like `logits`."""
    s = np.asarray(logits)
    B = s.shape[0]
    p = volume_softmax(s, num_joints).reshape(B, num_joints, hm_depth, hm_height, hm_width)
    ex, ey, ez = marginal_expectations(p, num_joints, hm_width, hm_height, hm_depth)
    g = np.asarray(grad_coords).reshape(B, num_joints, 3).astype(p.dtype)
    ix = np.arange(hm_width, dtype=p.dtype).reshape(1, 1, 1, 1, -1)
    iy = np.arange(hm_height, dtype=p.dtype).reshape(1, 1, 1, -1, 1)
    iz = np.arange(hm_depth, dtype=p.dtype).reshape(1, 1, -1, 1, 1)
    e = lambda a: a.reshape(B, num_joints, 1, 1, 1)  # noqa: E731
    w = (e(g[..., 0]) * (ix - e(ex)) / hm_width
         + e(g[..., 1]) * (iy - e(ey)) / hm_height
         + e(g[..., 2]) * (iz - e(ez)) / hm_depth)
    return (p * w).reshape(s.shape)


def joint_location_loss_grad(heatmap_out, gt_coord, gt_vis, size_average=True, upstream=1.0):
    """Gradient of `joint_location_loss` w.r.t. the heatmap logits: d|c-t|/dc = sign(c-t)
    (torch's `abs` backward uses sgn, so 0 at c == t), times vis, times 1/B (:86-89)."""
    joint_num = int(gt_coord.shape[1] / 3)
    w, h, d = _dims_of(heatmap_out)
    c = integral_coords(heatmap_out, joint_num, w, h, d)
    g = np.sign(c - gt_coord) * gt_vis * upstream
    if size_average:
        g = g / len(c)
    return integral_coords_vjp(heatmap_out, joint_num, w, h, d, g)


def compute_mpjpe(pred, gt):
    """common/nets/loss.py:93-96: mean over batch of the mean over joints of ||pred-gt||_2."""
    p = np.asarray(pred).reshape(len(pred), NUM_JOINTS, 3)
    g = np.asarray(gt).reshape(len(gt), NUM_JOINTS, 3)
    return np.sqrt(((p - g) ** 2).sum(axis=2)).mean(axis=-1).mean()


def joint_location_result(patch_width, patch_height, heatmaps):
    """common/augment.py:160-184: soft-argmax, then x,y -> (v+0.5)*patch, z -> v*patch_width,
    and a constant score column.  Returns (B, J, 4) float64."""
    w, h = heatmaps.shape[-1], heatmaps.shape[-2]
    d = heatmaps.shape[-3] // NUM_JOINTS
    j = heatmaps.shape[1] // d
    c = integral_coords(heatmaps, j, w, h, d).astype(np.float64).reshape(len(heatmaps), j, 3)
    out = np.ones((len(heatmaps), j, 4))
    out[..., 0] = (c[..., 0] + 0.5) * patch_width
    out[..., 1] = (c[..., 1] + 0.5) * patch_height
    out[..., 2] = c[..., 2] * patch_width
    return out


def combined_loss(heatmap_out, coord_teacher, gt_coord, gt_vis, labelled):
    """common/nets/loss.py:117-203 with `cfg._lambda == 0` (main/config.py:50), i.e. without
    the PANet term: supervised L1 on labelled rows (:174), teacher L1 on the rest (:154),
    both summed and divided by the full batch length (:197).  Returns
    (loss, student_mpjpe, teacher_mpjpe, supervised_sum, unsupervised_sum)."""
    joint_num = int(gt_coord.shape[1] / 3)
    w, h, d = _dims_of(heatmap_out)
    c = integral_coords(heatmap_out, joint_num, w, h, d)
    teacher = np.asarray(coord_teacher)
    if teacher.ndim == 3:
        teacher = teacher[0]                     # torch.squeeze(dim=0), :125
    lab = np.asarray(labelled, dtype=bool)
    sup = (np.abs(c[lab] - gt_coord[lab]) * gt_vis[lab]).sum() if lab.any() else 0.0
    uns = (np.abs(c[~lab] - teacher[~lab]) * gt_vis[~lab]).sum() if (~lab).any() else 0.0
    return (sup + uns) / len(c), compute_mpjpe(c, gt_coord), compute_mpjpe(teacher, gt_coord), sup, uns


def joint_location_loss2(heatmap_out, gt_label, joint_cam_normalized, R, trans, K, tprime, patch_width=224.0,
                         patch_height=224.0):
    """common/nets/loss.py:207-277: soft-argmax (:219), rescale to patch pixels (:229 =
    common/augment.py:186-194), trans inverted (:232), per sample patch -> image with z + tprime
    (:236-239), pixel2cam (:245), R^T (:249); the loss is the SUM over samples and joints of the
    squared distance to joint_cam_normalized (:262-276; size_average is forced off at :270)."""
    from oracle import chain_np
    joint_num = int(np.asarray(gt_label).shape[1] / 3)
    w, h, d = _dims_of(heatmap_out)
    c = integral_coords(heatmap_out, joint_num, w, h, d)
    patch = chain_np.joint_location_result(c, patch_width, patch_height)[..., 0:3]
    tinv = np.linalg.inv(np.asarray(trans, dtype=np.float64))
    _, cam = chain_np.camera_chain(patch, tinv, tprime, K, R=R)
    diff = np.asarray(joint_cam_normalized, dtype=np.float64) - cam
    return (diff ** 2).sum()


def softmax_entropy(logits, J, D, H, W):
    """scipy.stats.entropy of each row of softmax_integral_tensor2's output
    (main/generate_correlation_graph.py:116-141 on common/nets/loss.py:61-67): -sum p ln p per
    (sample, joint) volume, float64.  (B, J)."""
    x = np.asarray(logits, dtype=np.float64).reshape(logits.shape[0], J, D * H * W)
    x = x - x.max(axis=2, keepdims=True)
    e = np.exp(x)
    p = e / e.sum(axis=2, keepdims=True)
    with np.errstate(divide="ignore", invalid="ignore"):
        t = np.where(p > 0, p * np.log(p), 0.0)
    return -t.sum(axis=2)
