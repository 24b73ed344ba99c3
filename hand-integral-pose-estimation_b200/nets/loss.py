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
    CombinedLoss                       loss.py:99-203       same masking logic on top of the op
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


class CombinedLoss(nn.Module):
    """loss.py:99-203 with the PANet term weighted by `lambda_panet` (cfg._lambda, 0 in the shipped
    config, main/config.py:50).  With a non-zero weight the caller passes `nrsfm_tester` and
    `prepare_panet_input` exactly as the reference does; with 0 they are never touched."""

    def __init__(self, lambda_panet=0.0, prepare_panet_input=None):
        super(CombinedLoss, self).__init__()
        self.size_average = True
        self.lambda_panet = lambda_panet
        self.prepare_panet_input = prepare_panet_input

    def forward(self, heatmap_out, coord_out_teacher, gt_coord, gt_vis, labelled, tprime=None, trans=None,
                bbox=None, K=None, R=None, scale=None, joint_cam_normalized=None, nrsfm_tester=None):
        joint_num = int(gt_coord.shape[1] / 3)
        hm_width = heatmap_out.shape[-1]
        hm_height = heatmap_out.shape[-2]
        hm_depth = heatmap_out.shape[-3] // FreiHandConfig.num_joints
        coord_out = softmax_integral_tensor(heatmap_out, joint_num, hm_width, hm_height, hm_depth)
        gt_coord = gt_coord.to(coord_out.device)
        gt_vis = gt_vis.to(coord_out.device)
        labelled = labelled.to(coord_out.device)
        coord_out_teacher = torch.squeeze(coord_out_teacher, dim=0).to(coord_out.device)
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
                pin = self.prepare_panet_input(pin, tprime[~labelled], trans[~labelled], bbox[~labelled],
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
