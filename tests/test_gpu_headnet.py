"""Integration with the heat-map PRODUCER (main/model.py:14-20,40-44): a 1x1
convolution head feeds the drop-in JointLocationLoss and the gradient flows back into the conv
weights -- a non-leaf producer, channels_last (non-contiguous) and autocast-bf16 outputs, a scaled
upstream gradient (GradScaler style) and repeated backward through one graph."""
import numpy as np
import pytest
import torch

from oracle import integral_torch as OT
from tests import parity

pytestmark = pytest.mark.gpu
J = 21


def _head(D, seed=0):
    torch.manual_seed(seed)
    conv = torch.nn.Conv2d(256, J * D, kernel_size=1, stride=1, padding=0)      # HeadNet.final_layer, model.py:14-20
    with torch.no_grad():
        conv.weight.mul_(4.0)                                                    # peaked enough to matter
    return conv


def _batch(B, D, seed=1):
    g = torch.Generator().manual_seed(seed)
    feat = torch.randn(B, 256, D, D, generator=g)
    gt = torch.rand(B, 3 * J, generator=g, dtype=torch.float64) - 0.5
    vis = (torch.rand(B, 3 * J, generator=g) > 0.1).double()
    return feat, gt, vis


def _cpu_reference(conv, feat, gt, vis):
    c64 = torch.nn.Conv2d(256, conv.out_channels, 1).double()
    c64.load_state_dict({k: v.detach().cpu().double() for k, v in conv.state_dict().items()})
    hm = c64(feat.double())
    hm.retain_grad()
    loss = OT.joint_location_loss(hm, gt, vis)
    loss.backward()
    return loss.item(), c64.weight.grad, c64.bias.grad, hm.detach(), hm.grad


@pytest.mark.parametrize("D", [16, 32], ids=["two-launch", "single-launch"])
@pytest.mark.parametrize("layout", ["contiguous", "channels_last"])
def test_conv_head_gets_the_oracle_gradient(D, layout):
    from hipe_b200 import ops
    from nets import loss as L
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    B = 2
    conv = _head(D).cuda()
    feat, gt, vis = _batch(B, D)
    ref_loss, ref_w, ref_b, ref_hm, ref_hg = _cpu_reference(conv, feat, gt, vis)
    x = feat.cuda()
    if layout == "channels_last":
        conv = conv.to(memory_format=torch.channels_last)
        x = x.contiguous(memory_format=torch.channels_last)
    hm = conv(x)                                                                # non-leaf, requires grad
    assert hm.grad_fn is not None
    if layout == "channels_last":
        assert not hm.is_contiguous()
    hm.retain_grad()
    ops.KERNEL_EVENTS = {}
    try:
        loss = L.JointLocationLoss()(hm, gt.cuda(), vis.cuda())
        loss.backward()
        launched = set(ops.KERNEL_EVENTS)
    finally:
        ops.KERNEL_EVENTS = None
    if D == 32:
        assert launched & {"joint_loss_cluster", "joint_loss_resident", "joint_loss_fused"}, launched
    assert abs(loss.item() - ref_loss) <= 2e-5 * abs(ref_loss)
    # heat-map gradient per volume (the conv output is fp32-rounded: compare on ITS values)
    n = B * J
    with torch.no_grad():
        c = L.softmax_integral_tensor(hm.detach(), J, D, D, D).view(n, 3).cpu().numpy()
    parity.check_volumes(hm.detach().contiguous().view(n, D, D, D).cpu().numpy(), gt.view(n, 3).numpy(),
                         vis.view(n, 3).numpy(), 1.0 / B, c, None, hm.grad.contiguous().view(n, -1).cpu().numpy())
    # and what arrives in the weights
    w, b = conv.weight.grad.detach().cpu().double(), conv.bias.grad.detach().cpu().double()
    assert (w - ref_w).abs().max().item() <= 2e-4 * ref_w.abs().max().item()
    assert (b - ref_b).abs().max().item() <= 2e-4 * ref_b.abs().max().item()


def test_autocast_bf16_head():
    """bf16 heat maps out of an autocast region: loss and heat-map gradient against the oracle on the
    ROUNDED heat maps (1 bf16 ulp), weight gradient close to the fp32 run."""
    from nets import loss as L
    B, D = 2, 32
    conv = _head(D).cuda()
    feat, gt, vis = _batch(B, D)
    with torch.autocast("cuda", dtype=torch.bfloat16):
        hm = conv(feat.cuda())
    assert hm.dtype == torch.bfloat16
    hm.retain_grad()
    loss = L.JointLocationLoss()(hm, gt.cuda(), vis.cuda())
    loss.backward()
    assert hm.grad.dtype == torch.bfloat16 and conv.weight.grad.dtype == torch.float32
    n = B * J
    up = hm.detach().float().contiguous().view(n, D, D, D).cpu().numpy()
    with torch.no_grad():
        c = L.softmax_integral_tensor(hm.detach(), J, D, D, D).view(n, 3).cpu().numpy()
    parity.check_volumes(up, gt.view(n, 3).numpy(), vis.view(n, 3).numpy(), 1.0 / B, c, None,
                         hm.grad.float().contiguous().view(n, -1).cpu().numpy(), bf16=True)
    ref = OT.joint_location_loss(torch.as_tensor(up).double().view(B, J * D, D, D), gt, vis).item()
    assert abs(loss.item() - ref) <= 1e-5 * abs(ref)
    conv32 = _head(D).cuda()
    hm32 = conv32(feat.cuda())
    L.JointLocationLoss()(hm32, gt.cuda(), vis.cuda()).backward()
    assert (conv.weight.grad - conv32.weight.grad).abs().max().item() <= 5e-2 * conv32.weight.grad.abs().max().item()


def test_scaled_loss_and_repeated_backward():
    """(loss * s).backward() scales the stored gradient once; a second backward through the
    single-launch graph raises (its gradient was consumed) instead of compounding the scale;
    fused=False supports retain_graph and accumulates."""
    from nets import loss as L
    B, D = 2, 32
    feat, gt, vis = _batch(B, D)

    def run(fused, scale, twice):
        conv = _head(D).cuda()
        hm = conv(feat.cuda())
        loss = L.JointLocationLoss(fused=fused)(hm, gt.cuda(), vis.cuda())
        (loss * scale).backward(retain_graph=twice)
        if twice:
            (loss * scale).backward()
        return conv.weight.grad.detach().clone()

    base = run(False, 1.0, False)
    for fused in (True, None):
        g = run(fused, 1024.0, False)                                # GradScaler style
        assert (g / 1024.0 - base).abs().max().item() <= 2e-5 * base.abs().max().item()
    g2 = run(False, 0.5, True)                                       # two-launch path: accumulates 2 x 0.5
    assert (g2 - base).abs().max().item() <= 2e-5 * base.abs().max().item()
    with pytest.raises(RuntimeError, match="second backward"):
        run(True, 0.5, True)
    # the coordinates returned next to the loss by the single-launch function carry no graph
    from hipe_b200 import ops
    conv = _head(D).cuda()
    hm = conv(feat.cuda())
    loss, coords, terms = ops.FusedJointLossFunction.apply(hm, gt.cuda(), vis.cuda(), J, D, D, D, 1.0 / B)
    assert loss.requires_grad and not coords.requires_grad and not terms.requires_grad
    assert coords.shape == (B, 3 * J)


def test_graphed_step_matches_eager():
    """hipe_b200.graphed: the captured step replays bit-identically to the eager one, on new inputs
    copied into its static buffers, for a two-launch and a single-launch geometry."""
    import hipe_b200
    from nets import loss as L
    from tests import synth
    for D in (16, 32):
        B = 4
        hm = [torch.as_tensor(synth.heatmaps(B, J, D, D, D, "peaked", seed=s)).cuda() for s in (1, 2)]
        gt, vis = (torch.as_tensor(a).cuda() for a in synth.targets(B, J, seed=3, zero_vis_frac=0.1))
        crit = L.JointLocationLoss()
        step = hipe_b200.graphed(crit, hm[0], gt, vis)
        for x in (hm[1], hm[0], hm[1]):
            xe = x.clone().requires_grad_(True)
            le = crit(xe, gt, vis)
            le.backward()
            lg, gg = step(x)
            assert lg.item() == le.item()
            assert torch.equal(gg, xe.grad)
