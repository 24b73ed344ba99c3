"""Drop-in replacement for the reference's common/nets/loss.py on the integral-regression path.

Same names, positional signatures, return shapes and dtypes; every function runs the
hand-written sm_100a kernels of libhipe_b200.so (through hipe_b200.ops) instead of the
ATen op chain.  CUDA tensors only -- a CPU tensor raises.

    reference (file:line)                                   here
    generate_3d_integral_preds_tensor  loss.py:23-43        one streaming pass, no exp
    softmax_integral_tensor            loss.py:46-59        SoftArgmaxFunction (fwd 1 pass, bwd 1R+1W)
    softmax_integral_tensor2           loss.py:61-67        stats pass + write pass
    JointLocationLoss                  loss.py:69-91        soft-argmax + fp64 L1 kernel (or fused)
    computeMPJPE                       loss.py:93-96        (B,63) arithmetic, torch
    CombinedLoss                       loss.py:99-203       lambda = 0: ONE loss+gradient launch against
                                                            where(labelled, gt, teacher); else the
                                                            reference's masking logic on top of the op
    JointLocationLoss2                 loss.py:207-277      soft-argmax -> patch->camera chain -> sum of
                                                            squared distances, all on the device

`nets` is a namespace package here as in the reference (no __init__.py): with this directory in
front of the reference's common/ on sys.path, `nets.loss` is this file while `nets.resnet` and
`nets.balanced_parallel` stay the reference's (main/model.py:2, common/base.py:15).
"""
import torch
import torch.nn as nn

from hipe_b200 import ops

NUM_JOINTS = 21  # data/FreiHand/FreiHand_config.py:60 (FreiHandConfig.num_joints)


class FreiHandConfig:  # the two attributes the reference's loss.py reads
    num_joints = NUM_JOINTS
    root_idx = 9


def generate_3d_integral_preds_tensor(heatmaps, num_joints, x_dim, y_dim, z_dim):
    """loss.py:23-43: index-weighted marginal sums of (already normalised) heatmaps.
    Returns three (B, J, 1) tensors in voxel units."""
    assert isinstance(heatmaps, torch.Tensor)
    B = heatmaps.shape[0]
    _, expect, _ = ops.softargmax_forward(heatmaps, B * num_joints, z_dim, y_dim, x_dim, apply_softmax=False)
    e = expect.view(B, num_joints, 3).to(torch.float32)
    return e[:, :, 0:1], e[:, :, 1:2], e[:, :, 2:3]


def softmax_integral_tensor(preds, num_joints, hm_width, hm_height, hm_depth):
    """loss.py:46-59: (B, J*D, H, W) logits -> (B, 3J) coordinates in [-0.5, 0.5), differentiable."""
    return ops.SoftArgmaxFunction.apply(preds, int(num_joints), int(hm_width), int(hm_height), int(hm_depth))


def softmax_integral_tensor2(preds, num_joints, hm_width, hm_height, hm_depth):
    """loss.py:61-67: the softmax probabilities themselves, shaped (B, J, D*H*W)."""
    B = preds.shape[0]
    p = ops.volume_softmax(preds.contiguous(), B * num_joints, hm_depth, hm_height, hm_width)
    return p.view(B, num_joints, -1)


def softmax_entropy(preds, num_joints, hm_width, hm_height, hm_depth):
    """Per-joint entropy (nats) of the volume softmax, shape (B, J) fp64: what
    main/generate_correlation_graph.py:116-141 computes as scipy.stats.entropy over the rows of
    softmax_integral_tensor2(...).cpu() -- here from the logits in one extra streaming read, no
    probability tensor, no host copy.  (An addition next to the reference's names, not one of them.)"""
    B = preds.shape[0]
    h = ops.volume_entropy(preds.contiguous(), B * num_joints, hm_depth, hm_height, hm_width)
    return h.view(B, num_joints)


class JointLocationLoss(nn.Module):
    """loss.py:69-91.  When the heatmaps require a gradient the loss and `dlogits` can come out
    of ONE launch (one read + one write of each volume from HBM instead of two reads + one
    write); backward then only hands the stored gradient over.  `fused=None` (default) takes
    that path for the geometries where it is measured faster (ops.fused_plan), `True` / `False`
    force / forbid it."""

    def __init__(self, fused=None):
        super(JointLocationLoss, self).__init__()
        self.size_average = True
        self.fused = fused

    def forward(self, heatmap_out, gt_coord, gt_vis):
        joint_num = int(gt_coord.shape[1] / 3)
        hm_width = heatmap_out.shape[-1]
        hm_height = heatmap_out.shape[-2]
        hm_depth = heatmap_out.shape[-3] // FreiHandConfig.num_joints
        B = heatmap_out.shape[0]
        norm = 1.0 / B if self.size_average else 1.0
        fused = self.fused
        if fused is None and torch.is_grad_enabled() and heatmap_out.requires_grad and heatmap_out.is_cuda:
            fused = ops.fused_plan(heatmap_out, B * joint_num, hm_depth, hm_height, hm_width) is not None
        if fused and torch.is_grad_enabled() and heatmap_out.requires_grad:
            loss, _, _ = ops.FusedJointLossFunction.apply(heatmap_out, gt_coord, gt_vis, joint_num, hm_depth,
                                                       hm_height, hm_width, norm)
            return loss
        coord_out = softmax_integral_tensor(heatmap_out, joint_num, hm_width, hm_height, hm_depth)
        return ops.IntegralL1Function.apply(coord_out, gt_coord, gt_vis, norm)


def computeMPJPE(pred, gt):
    """loss.py:93-96."""
    pred = pred.reshape((pred.shape[0], FreiHandConfig.num_joints, 3))
    gt = gt.reshape((gt.shape[0], FreiHandConfig.num_joints, 3))
    return (pred - gt).norm(dim=2).mean(-1).mean()


def _cfg():
    """The reference's `config.cfg` when it is importable (main/config.py), else None."""
    try:
        from config import cfg
        return cfg
    except Exception:
        return None


class CombinedLoss(nn.Module):
    """loss.py:99-203.  `lambda_panet` is the weight of the PANet term: cfg._lambda when the
    reference's config is importable (0 in the shipped config, main/config.py:50), else 0.

    lambda == 0 (the shipped training configuration, main/config.py:49-50): the total
        (sum_lab |c - gt| vis + sum_unlab |c - teacher| vis) / B          (loss.py:154,174,197)
    is JointLocationLoss against target = where(labelled, gt, teacher), so the loss and the heatmap
    gradient come out of ONE launch (ops.FusedJointLossFunction); the supervised / unsupervised
    parts are masked sums of the kernel's per-coordinate loss terms.  No boolean gathers and no
    host synchronisation (the reference's `coord_out[~labelled]` and its print of the two counts
    synchronise every step); the two partial sums are returned as 0-dim tensors, 0 when empty.

    lambda != 0: the reference's own sequence on top of the differentiable soft-argmax; the caller
    passes `nrsfm_tester`, and `prepare_panet_input` comes from `augment` (the reference's)."""

    def __init__(self, lambda_panet=None, prepare_panet_input=None):
        super(CombinedLoss, self).__init__()
        self.size_average = True
        if lambda_panet is None:
            cfg = _cfg()
            lambda_panet = getattr(cfg, "_lambda", 0.0) if cfg is not None else 0.0
        self.lambda_panet = lambda_panet
        self.prepare_panet_input = prepare_panet_input

    def _panet_input_fn(self):
        if self.prepare_panet_input is None:
            import augment
            self.prepare_panet_input = augment.prepare_panet_input
        return self.prepare_panet_input

    def forward(self, heatmap_out, coord_out_teacher, gt_coord, gt_vis, labelled, tprime=None, trans=None,
                bbox=None, K=None, R=None, scale=None, joint_cam_normalized=None, nrsfm_tester=None):
        joint_num = int(gt_coord.shape[1] / 3)
        hm_width = heatmap_out.shape[-1]
        hm_height = heatmap_out.shape[-2]
        hm_depth = heatmap_out.shape[-3] // FreiHandConfig.num_joints
        B = heatmap_out.shape[0]
        dev = heatmap_out.device
        gt_coord = gt_coord.to(dev)
        gt_vis = gt_vis.to(dev)
        labelled = labelled.to(dev)
        coord_out_teacher = torch.squeeze(coord_out_teacher, dim=0).to(dev)
        if (self.lambda_panet == 0 and self.size_average and heatmap_out.is_cuda and torch.is_grad_enabled()
                and heatmap_out.requires_grad
                and ops.fused_plan(heatmap_out, B * joint_num, hm_depth, hm_height, hm_width) is not None):
            lab = labelled.to(torch.bool).view(B, 1)
            target = torch.where(lab, gt_coord, coord_out_teacher.to(gt_coord.dtype))
            loss, coord_out, terms = ops.FusedJointLossFunction.apply(heatmap_out, target, gt_vis, joint_num,
                                                                      hm_depth, hm_height, hm_width, 1.0 / B)
            with torch.no_grad():
                per_sample = terms.view(B, 3 * joint_num).sum(dim=1)
                w = lab.view(B).to(per_sample.dtype)
                sup = (per_sample * w).sum()
                unsup = (per_sample * (1.0 - w)).sum()
                student_mpjpe = computeMPJPE(coord_out, gt_coord)
                teacher_mpjpe = computeMPJPE(coord_out_teacher, gt_coord)
            return loss, student_mpjpe, teacher_mpjpe, sup, unsup
        coord_out = softmax_integral_tensor(heatmap_out, joint_num, hm_width, hm_height, hm_depth)
        labelled = labelled.to(torch.bool)
        n_unsup = int((~labelled).sum())
        n_sup = int(labelled.sum())
        with torch.no_grad():
            student_mpjpe = computeMPJPE(coord_out, gt_coord)
            teacher_mpjpe = computeMPJPE(coord_out_teacher, gt_coord)
        loss_unsup = 0
        loss_sup = 0
        if n_unsup > 0:
            l_teacher = (torch.abs(coord_out[~labelled] - coord_out_teacher[~labelled]) * gt_vis[~labelled]).sum()
            loss_unsup = l_teacher
            if self.lambda_panet != 0:
                pin = coord_out[~labelled].reshape(n_unsup, FreiHandConfig.num_joints, 3)
                pin = self._panet_input_fn()(pin, tprime[~labelled], trans[~labelled], bbox[~labelled],
                                             K[~labelled], R[~labelled], scale[~labelled])
                pout = nrsfm_tester.forward(pin)[0].reshape(n_unsup, FreiHandConfig.num_joints * 3)
                pin = pin.reshape(n_unsup, FreiHandConfig.num_joints * 3)
                loss_unsup = loss_unsup + ((self.lambda_panet * torch.abs(pin - pout)) * gt_vis[~labelled]).sum()
        if n_sup > 0:
            loss_sup = (torch.abs(coord_out[labelled] - gt_coord[labelled]) * gt_vis[labelled]).sum()
        loss = loss_sup + loss_unsup
        sup_tmp = loss_sup.detach() if torch.is_tensor(loss_sup) else 0
        unsup_tmp = loss_unsup.detach() if torch.is_tensor(loss_unsup) else 0
        loss = loss / len(coord_out)
        return loss, student_mpjpe, teacher_mpjpe, sup_tmp, unsup_tmp


class JointLocationLoss2(nn.Module):
    """loss.py:207-277: the evaluation-time camera-space error
        sum_{b,j} || joint_cam_normalized[b,j] - R_b^T pixel2cam(patch->image(soft-argmax), K_b) ||^2
    with trans inverted first (:232) and size_average forced off (:270-274).  The reference copies
    everything to the host and loops over samples in numpy; here the soft-argmax output goes
    straight into hipe_patch_to_camera (rescale :229 = augment.test_get_joint_loc_res, homography +
    tprime :236-239, pixel2cam :245, R^T :249) and a fixed-order fp64 sum, on the device.  Returns a
    0-dim fp64 CUDA tensor that does not require grad, like the reference's.

    `verify_labels` keeps the reference's consistency check (:240-256: the ground-truth label pushed
    through the same chain must reproduce joint_cam_normalized, else a warning is printed); it costs
    one host synchronisation per call and can be switched off."""

    def __init__(self, patch_width=None, patch_height=None, verify_labels=True):
        super(JointLocationLoss2, self).__init__()
        self.size_average = True
        cfg = _cfg()
        self.patch_width = patch_width if patch_width is not None else getattr(cfg, "patch_width", 224)
        self.patch_height = patch_height if patch_height is not None else getattr(cfg, "patch_height", 224)
        self.verify_labels = verify_labels

    def forward(self, heatmap_out, gt_label, gt_vis, joint_cam, joint_cam_normalized, bbox, scale, R, trans, K,
                tprime):
        joint_num = int(gt_label.shape[1] / 3)
        hm_width = heatmap_out.shape[-1]
        hm_height = heatmap_out.shape[-2]
        hm_depth = heatmap_out.shape[-3] // FreiHandConfig.num_joints
        B = heatmap_out.shape[0]
        dev = heatmap_out.device
        patch = (float(self.patch_width), float(self.patch_height))
        with torch.no_grad():
            coord_out = softmax_integral_tensor(heatmap_out, joint_num, hm_width, hm_height, hm_depth)
            f64 = dict(device=dev, dtype=torch.float64)
            trans_inv = torch.linalg.inv(trans.to(**f64).reshape(B, 3, 3))                      # :232
            Kd, Rd, tp = K.to(**f64).reshape(B, 3, 3), R.to(**f64).reshape(B, 3, 3), tprime.to(**f64).reshape(B)
            target = joint_cam_normalized.to(**f64).reshape(B, joint_num, 3)
            pre = ops.patch_to_camera(coord_out.detach(), B, joint_num, tp, Kd, trans=trans_inv, R=Rd,
                                      patch=patch, want=("cam",))["cam"]
            diff = target - pre
            loss = ops.sum_f64(diff * diff)                                                      # :262-276
            if self.verify_labels:
                lab = ops.patch_to_camera(gt_label.to(device=dev, dtype=torch.float32).contiguous(), B, joint_num,
                                          tp, Kd, trans=trans_inv, R=Rd, patch=patch, want=("cam",))["cam"]
                if not torch.allclose(lab, target, rtol=1e-6, atol=1e-6):
                    print("(loss::JointLocationLoss2) Warning: label_3d_kpt and joint_cam_normalized are not equal.")
        self.size_average = False                                                                # :270
        return loss
