"""GPU parity of the cluster single launch (hipe_joint_loss_cluster) against the fp64 oracle, in both
of its forms: every volume held in the distributed shared memory of one thread-block cluster
between the reduction and the gradient write, or streamed shares (two rings per CTA, the second
read through L2 a bounded lead behind the first).  Same gates as tests/test_gpu_parity.py."""
import os

import numpy as np
import pytest
import torch

from oracle import integral_np as O
from tests import synth
from tests.test_gpu_resident import _check, _cuda

pytestmark = pytest.mark.gpu

# geometries the launch takes (the cluster size is the library's choice: hipe_cluster_plan)
CLU_GEOMS = [
    (16, 16, 16),
    (32, 32, 32),
    (64, 64, 64),        # headline: 1 MiB, streamed shares on 8 CTAs (lead 21 stages)
    (96, 96, 96),        # BASELINE configs[3]: fits no cluster; wrapped 12 KiB stages, streamed on 16 CTAs
    (32, 96, 96),        # same rows, but fits a cluster: whole-row visits, 192 of 256 team threads
    (80, 80, 80),        # fits no cluster either, general loops: streamed on 8 CTAs
    (56, 56, 56),        # reference default (main/config.py:23-25): 12.25 KB stages
    (48, 48, 48),
    (16, 128, 64),       # stage = half a z-slice (SR | H, non-zero y base: general row terms)
    (12, 20, 24),        # 30 rows per visit, 180 of 256 team threads active
]
UNSUPPORTED = [(8, 8, 8), (5, 6, 7), (3, 9, 100)]


@pytest.mark.parametrize("regime", ["noise", "peaked", "sharp"])
@pytest.mark.parametrize("geom", CLU_GEOMS)
def test_cluster_vs_oracle(geom, regime):
    from hipe_b200 import ops
    D, H, W = geom
    J, B = 21, 2
    hm = synth.heatmaps(B, J, D, H, W, regime, seed=11)
    gt, vis = synth.targets(B, J, seed=11, zero_vis_frac=0.1)
    x = _cuda(hm)
    plan = ops.cluster_plan(x, B * J, D, H, W)
    assert plan is not None and 1 <= plan[0] <= 16, plan
    # resident form: the share fits the ring with look-ahead; streamed form: the lead exceeds the share
    assert (plan[4] == 0 and plan[2] + 3 <= plan[1]) or plan[4] > plan[2], plan
    out = ops.joint_loss_cluster(x, B * J, D, H, W, _cuda(gt), _cuda(vis), 1.0 / B)
    torch.cuda.synchronize()
    worst = _check(hm, gt, vis, D, H, W, B, J, out)
    print("cluster %s %s: worst gradient error = %.3f of the 1e-5 gate" % (geom, regime, worst))


@pytest.mark.parametrize("geom", UNSUPPORTED)
def test_cluster_unsupported_geometries(geom):
    from hipe_b200 import _lib, ops
    D, H, W = geom
    x = _cuda(synth.heatmaps(1, 21, D, H, W, "peaked", seed=3))
    gt, vis = synth.targets(1, 21, seed=3)
    assert not ops.cluster_supported(x, 21, D, H, W)
    assert ops.cluster_plan(x, 21, D, H, W) is None
    with pytest.raises(_lib.HipeError):
        ops.joint_loss_cluster(x, 21, D, H, W, _cuda(gt), _cuda(vis), 1.0)


@pytest.mark.parametrize("csize", [1, 2, 3, 4, 5, 6, 7, 8])
def test_every_cluster_size(csize):
    """Forced cluster sizes, powers of two or not (uneven shares): 32^3 fp32 has 8 stages."""
    from hipe_b200 import ops
    D = H = W = 32
    J, B = 21, 3
    hm = synth.heatmaps(B, J, D, H, W, "peaked", seed=17)
    gt, vis = synth.targets(B, J, seed=17)
    x = _cuda(hm)
    os.environ["HIPE_CLUSTER_C"] = str(csize)
    try:
        assert ops.cluster_plan(x, B * J, D, H, W)[0] == csize
        out = ops.joint_loss_cluster(x, B * J, D, H, W, _cuda(gt), _cuda(vis), 1.0 / B)
        torch.cuda.synchronize()
    finally:
        del os.environ["HIPE_CLUSTER_C"]
    _check(hm, gt, vis, D, H, W, B, J, out)


def test_many_volumes_per_cluster_and_reproducible(monkeypatch):
    """More volumes than clusters (every role loops, ring laps and both exchange buffers are
    reused many times), a count that is not a multiple of the cluster count, bit-reproducible."""
    from hipe_b200 import ops
    D = H = W = 64
    J, B = 21, 9             # 189 volumes over <= 18 clusters of 8
    hm = synth.heatmaps(B, J, D, H, W, "peaked", seed=21)
    gt, vis = synth.targets(B, J, seed=21, zero_vis_frac=0.1)
    x = _cuda(hm)
    loss = torch.full((), float("nan"), dtype=torch.float64, device="cuda")
    monkeypatch.setenv("HIPE_CLUSTER_C", "8")      # one cluster size for every variant below: same merge order, same bits
    out = ops.joint_loss_cluster(x, B * J, D, H, W, _cuda(gt), _cuda(vis), 1.0 / B, loss_out=loss)
    torch.cuda.synchronize()
    _check(hm, gt, vis, D, H, W, B, J, out)
    ref = ops.sum_f64(out[3], 1.0 / B)
    assert abs(loss.item() - ref.item()) <= 1e-14 * abs(ref.item())
    out2 = ops.joint_loss_cluster(x, B * J, D, H, W, _cuda(gt), _cuda(vis), 1.0 / B)
    assert torch.equal(out[4], out2[4]) and torch.equal(out[0], out2[0]) and torch.equal(out[3], out2[3])
    # a small ring (3 slots of look-ahead only) and a tiny grid (one cluster walks everything)
    # and the plain-store exchange; the general (non-specialised) loops are another compilation of the
    # same arithmetic: held to the oracle, not to the bits
    for env in ({"HIPE_CLUSTER_SLOTS": "9"}, {"HIPE_CLUSTER_GRID": "1"}, {"HIPE_CLUSTER_GRID": "3"},
                {"HIPE_CLUSTER_STASYNC": "0"}, {"HIPE_CLUSTER_GENERIC": "1"}):
        os.environ.update(env)
        try:
            out3 = ops.joint_loss_cluster(x, B * J, D, H, W, _cuda(gt), _cuda(vis), 1.0 / B)
            torch.cuda.synchronize()
        finally:
            for k_ in env:
                del os.environ[k_]
        if "HIPE_CLUSTER_GENERIC" in env:
            _check(hm, gt, vis, D, H, W, B, J, out3)
        else:
            assert torch.equal(out[0], out3[0]) and torch.equal(out[3], out3[3]), env
            assert torch.equal(out[4], out3[4]), (env, (out[4] - out3[4]).abs().max().item())


STREAM_CASES = [
    # (dtype, geometry, env): the streamed form forced onto geometries that default to the resident one,
    # the resident form forced onto the headline, odd cluster sizes, tiny rings, leads of share + 1
    ("f32", (64, 64, 64), {"HIPE_CLUSTER_STREAM": "0"}),
    ("f32", (64, 64, 64), {"HIPE_CLUSTER_STREAM": "1", "HIPE_CLUSTER_C": "5"}),
    ("f32", (64, 64, 64), {"HIPE_CLUSTER_STREAM": "1", "HIPE_CLUSTER_C": "1", "HIPE_CLUSTER_LEAD": "65"}),
    ("f32", (64, 64, 64), {"HIPE_CLUSTER_LEAD": "9", "HIPE_CLUSTER_RSLOTS": "2"}),
    ("f32", (64, 64, 64), {"HIPE_CLUSTER_SLOTS": "6", "HIPE_CLUSTER_RSLOTS": "4"}),
    ("bf16", (64, 64, 64), {"HIPE_CLUSTER_STREAM": "1"}),
    ("bf16", (32, 32, 32), {"HIPE_CLUSTER_STREAM": "1", "HIPE_CLUSTER_C": "2"}),
    ("f32", (56, 56, 56), {"HIPE_CLUSTER_STREAM": "1"}),
    ("f32", (32, 32, 32), {"HIPE_CLUSTER_STREAM": "1", "HIPE_CLUSTER_C": "3"}),
    ("f32", (12, 20, 24), {"HIPE_CLUSTER_STREAM": "1"}),
    ("f32", (96, 96, 96), {"HIPE_CLUSTER_C": "8"}),
    ("f32", (96, 96, 96), {"HIPE_CLUSTER_C": "7", "HIPE_CLUSTER_LEAD": "43"}),
    ("f32", (96, 96, 96), {"HIPE_CLUSTER_GENERIC": "1"}),
    ("bf16", (96, 96, 96), {}),
]


@pytest.mark.parametrize("dt,geom,env", STREAM_CASES, ids=["%s-%s-%s" % (c[0], "x".join(map(str, c[1])), ",".join(
    "%s=%s" % (k[13:], v) for k, v in c[2].items()) or "default") for c in STREAM_CASES])
def test_streamed_and_forced_forms_vs_oracle(dt, geom, env, monkeypatch):
    from hipe_b200 import ops
    from tests import parity
    D, H, W = geom
    J, B = 21, 2
    hm = synth.heatmaps(B, J, D, H, W, "peaked", seed=23)
    gt, vis = synth.targets(B, J, seed=23, zero_vis_frac=0.1)
    x = _cuda(hm)
    for k, v in env.items():
        monkeypatch.setenv(k, v)
    if dt == "bf16":
        x = x.to(torch.bfloat16)
        hm = x.float().cpu().numpy().astype(np.float64)
    plan = ops.cluster_plan(x, B * J, D, H, W)
    assert plan is not None
    assert (plan[4] > 0) == (env.get("HIPE_CLUSTER_STREAM") != "0" and (env.get("HIPE_CLUSTER_STREAM") == "1" or geom in (
        (64, 64, 64), (96, 96, 96)) and not (dt == "bf16" and geom == (64, 64, 64)))), plan
    out = ops.joint_loss_cluster(x, B * J, D, H, W, _cuda(gt), _cuda(vis), 1.0 / B)
    torch.cuda.synchronize()
    if dt == "bf16":
        coords, expect, stats, terms, d = out
        parity.check_volumes(hm.reshape(B * J, D, H, W), gt.reshape(B * J, 3), vis.reshape(B * J, 3), 1.0 / B,
                             coords.cpu().numpy(), terms.view(B * J, 3).cpu().numpy(),
                             d.float().cpu().numpy().reshape(B * J, -1), bf16=True)
    else:
        _check(hm, gt, vis, D, H, W, B, J, out)


@pytest.mark.parametrize("D", [64, 96])
def test_streamed_many_volumes_reproducible(D):
    """The streamed form over more volumes than clusters (both rings lap many times, the lead throttle
    and the second producer's hand-over are exercised on every volume), bit-reproducible, and the
    same bits whatever the ring split, the lead or the grid (they change timing, not arithmetic)."""
    from hipe_b200 import ops
    H = W = D
    J, B = 21, 7 if D == 64 else 4        # 147 / 84 volumes over 15 / 7 clusters
    hm = synth.heatmaps(B, J, D, H, W, "peaked", seed=29)
    gt, vis = synth.targets(B, J, seed=29, zero_vis_frac=0.1)
    x = _cuda(hm)
    loss = torch.full((), float("nan"), dtype=torch.float64, device="cuda")
    plan = ops.cluster_plan(x, B * J, D, H, W)
    assert plan[4] > plan[2]
    out = ops.joint_loss_cluster(x, B * J, D, H, W, _cuda(gt), _cuda(vis), 1.0 / B, loss_out=loss)
    torch.cuda.synchronize()
    _check(hm, gt, vis, D, H, W, B, J, out)
    ref = ops.sum_f64(out[3], 1.0 / B)
    assert abs(loss.item() - ref.item()) <= 1e-14 * abs(ref.item())
    # (the last two: a write team starved by a 2-slot ring behind an unthrottled reduce team that runs
    # volumes ahead -- the pacing counters must not lose a volume)
    for env in ({}, {"HIPE_CLUSTER_RSLOTS": "3"}, {"HIPE_CLUSTER_LEAD": str(plan[2] + 1)}, {"HIPE_CLUSTER_LEAD": "200"},
                {"HIPE_CLUSTER_LEAD": "100000", "HIPE_CLUSTER_RSLOTS": str(plan[1] - 2), "HIPE_CLUSTER_GRID": "1"},
                {"HIPE_CLUSTER_LEAD": "100000", "HIPE_CLUSTER_RSLOTS": str(plan[1] - 2)},
                {"HIPE_CLUSTER_GRID": "1"}, {"HIPE_CLUSTER_GRID": "2"}, {"HIPE_CLUSTER_SLOTS": "7"}, {"HIPE_CLUSTER_POL1": "0"},
                {"HIPE_CLUSTER_STASYNC": "0"}):
        os.environ.update(env)
        try:
            out3 = ops.joint_loss_cluster(x, B * J, D, H, W, _cuda(gt), _cuda(vis), 1.0 / B)
            torch.cuda.synchronize()
        finally:
            for k_ in env:
                del os.environ[k_]
        assert torch.equal(out[0], out3[0]) and torch.equal(out[3], out3[3]), env
        assert torch.equal(out[4], out3[4]), (env, (out[4] - out3[4]).abs().max().item())


def test_cluster_agrees_with_two_launch_path():
    from hipe_b200 import ops
    D = H = W = 64
    J, B = 21, 3
    hm = synth.heatmaps(B, J, D, H, W, "sharp", seed=5)
    gt, vis = synth.targets(B, J, seed=5)
    x = _cuda(hm)
    c1, e1, s1, t1, d1 = ops.joint_loss_cluster(x, B * J, D, H, W, _cuda(gt), _cuda(vis), 1.0 / B)
    c0, e0, s0 = ops.softargmax_forward(x, B * J, D, H, W)
    _, g = ops.l1_loss(c0.view(B, 3 * J), _cuda(gt), _cuda(vis), 1.0 / B)
    d0 = ops.softargmax_backward(x, B * J, D, H, W, s0, e0, g.view(-1, 3))
    assert (c1 - c0).abs().max().item() <= 2e-7
    assert (e1 - e0).abs().max().item() <= 1e-5          # voxel units
    same_sign = ((c1 - _cuda(gt).view(-1, 3)).sign() == (c0 - _cuda(gt).view(-1, 3)).sign()).all(dim=1)
    scale = d0.view(B * J, -1).abs().amax(dim=1, keepdim=True)
    rel = ((d1 - d0).view(B * J, -1).abs() / scale)[same_sign]
    assert rel.max().item() <= 2e-5


@pytest.mark.parametrize("D", [64, 56, 48, 32])
def test_cluster_bf16(D):
    from hipe_b200 import ops
    H = W = D
    J, B = 21, 3
    hm32 = synth.heatmaps(B, J, D, H, W, "peaked", seed=4)
    gt, vis = synth.targets(B, J, seed=4)
    xb = _cuda(hm32).to(torch.bfloat16)
    up = xb.float().cpu().numpy().astype(np.float64)       # the rounded tensor is the oracle's input
    assert ops.cluster_supported(xb, B * J, D, H, W)
    coords, expect, stats, terms, d = ops.joint_loss_cluster(xb, B * J, D, H, W, _cuda(gt), _cuda(vis), 1.0 / B)
    ref_loss, ref_c = O.joint_location_loss(up, gt, vis, return_coords=True)
    c = coords.view(B, 3 * J).cpu().numpy().astype(np.float64)
    assert (np.abs(c - ref_c) <= 1e-5 * np.maximum(np.abs(ref_c), 0.5)).all()
    assert abs(terms.sum().item() / B - ref_loss) <= 1e-5 * abs(ref_loss)
    refg = O.joint_location_loss_grad(up, gt, vis)
    want = torch.as_tensor(refg).to(torch.bfloat16).float().numpy().reshape(B * J, -1)
    got = d.float().cpu().numpy().reshape(B * J, -1)
    from tests import parity
    parity.check_volumes(up.reshape(B * J, D, H, W), gt.reshape(B * J, 3), vis.reshape(B * J, 3), 1.0 / B,
                         coords.cpu().numpy(), terms.view(B * J, 3).cpu().numpy(), got, bf16=True)
