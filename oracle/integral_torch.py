"""ORACLE (test infrastructure, not product code): the reference's integral-regression path
restated with the same ATen operator classes it executes (softmax, axis sums, index-weighted
sums, abs, autograd), on CPU tensors with all host threads.

This is what `bench.py` times as `cpu_baseline` (kind "port") and under `--impl reference`:
the reference is pure PyTorch-eager (common/nets/loss.py), `/root/reference` does not travel to
the GPU box, so its op sequence is restated here.  Only `tests/`, `__graft_entry__.smoke()` and
those two bench legs may import it.

Parity status: PINNED -- `tests/test_oracle_golden.py` holds it to the vectors that
`tests/golden/make_golden.py` froze from the live `common/nets/loss.py`, and (in the build
container) to the live module itself, bit for bit in fp64.
Citations are into /root/reference.
"""
import torch

NUM_JOINTS = 21  # data/FreiHand/FreiHand_config.py:60


def _axis_indices(n, like):
    # common/nets/loss.py:35-37 multiplies by an int64 arange (promoted by the multiply)
    return torch.arange(n, device=like.device)


def marginal_expectations(heatmaps, num_joints, x_dim, y_dim, z_dim):
    """common/nets/loss.py:23-43: (B, J, z, y, x) view, three marginals, index-weighted sums."""
    h = heatmaps.reshape(heatmaps.shape[0], num_joints, z_dim, y_dim, x_dim)
    over_z = h.sum(dim=2)                      # (B, J, y, x)   :27 and :31 (same reduction twice)
    mx = over_z.sum(dim=2)                     # (B, J, x)      :29
    my = h.sum(dim=2).sum(dim=3)               # (B, J, y)      :31-32
    mz = h.sum(dim=3).sum(dim=3)               # (B, J, z)      :33-34
    ex = (mx * _axis_indices(x_dim, h)).sum(dim=2, keepdim=True)   # :35,39
    ey = (my * _axis_indices(y_dim, h)).sum(dim=2, keepdim=True)   # :36,40
    ez = (mz * _axis_indices(z_dim, h)).sum(dim=2, keepdim=True)   # :37,41
    return ex, ey, ez


def integral_coords(preds, num_joints, hm_width, hm_height, hm_depth):
    """common/nets/loss.py:46-59."""
    rows = preds.reshape(preds.shape[0], num_joints, -1)
    p = torch.softmax(rows, dim=2)                                  # :49
    ex, ey, ez = marginal_expectations(p, num_joints, hm_width, hm_height, hm_depth)
    out = torch.cat((ex / float(hm_width) - 0.5,
                     ey / float(hm_height) - 0.5,
                     ez / float(hm_depth) - 0.5), dim=2)           # :54-57
    return out.reshape(out.shape[0], num_joints * 3)               # :58


def joint_location_loss(heatmap_out, gt_coord, gt_vis, size_average=True):
    """common/nets/loss.py:74-91."""
    joint_num = int(gt_coord.shape[1] / 3)
    w, h = heatmap_out.shape[-1], heatmap_out.shape[-2]
    d = heatmap_out.shape[-3] // NUM_JOINTS
    c = integral_coords(heatmap_out, joint_num, w, h, d)
    total = (torch.abs(c - gt_coord) * gt_vis).sum()
    return total / len(c) if size_average else total


def fwd_bwd_step(heatmap_out, gt_coord, gt_vis):
    """One training-step pass of the path: loss forward + autograd backward to the heatmaps
    (main/train.py:101-113).  Returns (loss, dlogits)."""
    x = heatmap_out.detach().requires_grad_(True)
    loss = joint_location_loss(x, gt_coord, gt_vis)
    loss.backward()
    return loss.detach(), x.grad
