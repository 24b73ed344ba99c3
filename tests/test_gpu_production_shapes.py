"""Oracle parity AT THE SHAPES bench.py and training launch (cluster size, launch plan and item
dealing all depend on the volume count): BASELINE configs[1] fp32 64^3 x 64 samples, configs[2]
bf16 64^3 x 512 and x 64 samples, configs[3] fp32 96^3 x 32 samples, and the reference's own
training shape fp32 56^3 x 32 samples (main/config.py:23-25,39).  The default-plan module call is
run at full size; 42 sampled volumes (first, last, and a seeded spread -- the oracle never sees
the multi-GB tensor) are compared with oracle.integral_np in fp64: coordinates, loss terms, full
gradients; all coordinates are also held to the independent forward kernel."""
import numpy as np
import pytest
import torch

from tests import parity

pytestmark = pytest.mark.gpu
J = 21

SHAPES = [
    ("C2 fp32 64^3 x64", torch.float32, 64, 64),
    ("C3 bf16 64^3 x512", torch.bfloat16, 64, 512),
    ("C3/8 bf16 64^3 x64", torch.bfloat16, 64, 64),
    ("C4 fp32 96^3 x32", torch.float32, 96, 32),
    ("train fp32 56^3 x32", torch.float32, 56, 32),
    ("train/2 fp32 56^3 x16", torch.float32, 56, 16),
]


def _inputs(dt, D, B, seed, regime_mix=True):
    import bench
    x = bench.synth_logits(torch, B, D, torch.float32, torch.device("cuda"), seed)
    if regime_mix:
        # every third sample sharp (A=20-like: scale the bump up), every third pure noise
        v = x.view(B, J, D, D, D)
        v[1::3] *= 1.6
        v[2::3] = torch.randn(v[2::3].shape, device="cuda", generator=torch.Generator(device="cuda").manual_seed(seed + 1))
    return x.to(dt).contiguous()


@pytest.mark.parametrize("name,dt,D,B", SHAPES, ids=[s[0] for s in SHAPES])
def test_default_plan_at_production_shape_vs_oracle(name, dt, D, B):
    from hipe_b200 import ops
    from nets import loss as L
    n_vol = B * J
    x = _inputs(dt, D, B, 31)
    g = torch.Generator(device="cuda").manual_seed(32)
    gt = torch.rand(B, 3 * J, generator=g, device="cuda", dtype=torch.float64) - 0.5
    vis = (torch.rand(B, 3 * J, generator=g, device="cuda") > 0.1).double()
    plan = ops.fused_plan(x, n_vol, D, D, D)
    # the module call (what bench.py and a training step execute)
    xr = x.clone().requires_grad_(True)
    ops.KERNEL_EVENTS = {}
    try:
        loss = L.JointLocationLoss()(xr, gt, vis)
        loss.backward()
        launched = set(ops.KERNEL_EVENTS)
    finally:
        ops.KERNEL_EVENTS = None
    # the same launch through ops, for the per-volume outputs the module does not return
    if plan is not None:
        coords, expect, stats, terms, dlog = ops.joint_loss_fused(x, n_vol, D, D, D, gt, vis, 1.0 / B)
        assert torch.equal(dlog.view_as(xr.grad), xr.grad), "the launch is not bit-reproducible"
        assert {"cluster": "joint_loss_cluster", "resident": "joint_loss_resident", "stream": "joint_loss_fused"}[plan] in launched
    else:
        coords, expect, stats = ops.softargmax_forward(x, n_vol, D, D, D)
        terms = ((coords.view(B, 3 * J).double() - gt).abs() * vis).view(-1)
        dlog = xr.grad
    # the loss scalar is the fixed-order fp64 sum of the terms
    assert abs(loss.item() - terms.sum().item() / B) <= 1e-12 * abs(loss.item())
    # all coordinates against the independent forward kernel
    c_fwd, _, _ = ops.softargmax_forward(x, n_vol, D, D, D)
    assert (coords - c_fwd).abs().max().item() <= 3e-7
    # sampled volumes against the fp64 oracle
    rng = np.random.default_rng(5)
    idx = np.unique(np.concatenate([[0, 1, n_vol - 2, n_vol - 1], rng.integers(0, n_vol, size=38)]))
    it = torch.as_tensor(idx, device="cuda")
    vols = x.view(n_vol, D, D, D)[it].float().cpu().numpy()
    worst = parity.check_volumes(
        vols, gt.view(n_vol, 3)[it].cpu().numpy(), vis.view(n_vol, 3)[it].cpu().numpy(), 1.0 / B,
        coords[it].cpu().numpy(), terms.view(n_vol, 3)[it].cpu().numpy(),
        dlog.view(n_vol, -1)[it].float().cpu().numpy(), bf16=(dt == torch.bfloat16), label=name)
    print("%s: plan %s, %d volumes sampled, worst gradient error = %.3f of the gate" % (name, plan, len(idx), worst))


def test_sharp_regime_margin_is_reported():
    """SURVEY 8d regime (iii), A=20 sigma=1.5: the gradient gate with its margin, per launch plan."""
    from hipe_b200 import ops
    from tests import synth
    D, B = 64, 16
    n_vol = B * J
    hm = synth.heatmaps(B, J, D, D, D, "sharp", seed=3)
    gt, vis = synth.targets(B, J, seed=3)
    x = torch.as_tensor(hm).cuda()
    tg, tv = torch.as_tensor(gt).cuda(), torch.as_tensor(vis).cuda()
    idx = np.arange(0, n_vol, 8)
    for mode in ("cluster", "resident", "stream"):
        if mode == "cluster" and not ops.cluster_supported(x, n_vol, D, D, D):
            continue
        c, e, s, t, d = ops.joint_loss_fused(x, n_vol, D, D, D, tg, tv, 1.0 / B, mode=mode)
        it = torch.as_tensor(idx, device="cuda")
        worst = parity.check_volumes(hm.reshape(n_vol, D, D, D)[idx], gt.reshape(n_vol, 3)[idx], vis.reshape(n_vol, 3)[idx],
                                     1.0 / B, c[it].cpu().numpy(), t.view(n_vol, 3)[it].cpu().numpy(),
                                     d.view(n_vol, -1)[it].cpu().numpy(), label="sharp/" + mode)
        print("sharp regime, %s launch: worst gradient error = %.3f of the 1e-5 gate" % (mode, worst))
