"""GPU, world_size 2 over NCCL (skipped on a 1-GPU box): the batch-sharded criterion with the real
kernels on each rank against the single-process oracle semantics."""
import os
import socket

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
J = 21


def _worker(rank, world, port, B, out_dir, geom):
    import torch.distributed as dist
    from hipe_b200 import parallel as PL
    from tests import synth
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        D, H, W = geom
        hm = synth.heatmaps(B, J, D, H, W, "peaked", seed=42)
        gt, vis = synth.targets(B, J, seed=42, zero_vis_frac=0.1)
        lo, hi = PL.shard_bounds(B, world, rank)
        x = torch.as_tensor(hm[lo:hi]).cuda().requires_grad_(True)
        crit = PL.ShardedJointLocationLoss(global_batch=B)
        loss = crit(x, torch.as_tensor(gt[lo:hi]).cuda(), torch.as_tensor(vis[lo:hi]).cuda())
        loss.backward()
        # several more steps through the same exchange object (sequence numbers, both parities), then
        # the NCCL form of the same step for comparison
        vals = []
        for _ in range(5):
            x2 = torch.as_tensor(hm[lo:hi]).cuda().requires_grad_(True)
            l2 = crit(x2, torch.as_tensor(gt[lo:hi]).cuda(), torch.as_tensor(vis[lo:hi]).cuda())
            l2.backward()
            vals.append(l2.item())
        peer_used = crit.exchange
        peer_err = bool(crit._peer.timed_out()) if crit._peer else False
        # the one-kernel, in-stream form of the same exchange (defer=False)
        crit_s = PL.ShardedJointLocationLoss(global_batch=B, defer=False)
        x4 = torch.as_tensor(hm[lo:hi]).cuda().requires_grad_(True)
        l4 = crit_s(x4, torch.as_tensor(gt[lo:hi]).cuda(), torch.as_tensor(vis[lo:hi]).cuda())
        l4.backward()
        deferred_type = type(loss).__name__
        os.environ["HIPE_EXCHANGE"] = "nccl"
        crit_n = PL.ShardedJointLocationLoss(global_batch=B)
        x3 = torch.as_tensor(hm[lo:hi]).cuda().requires_grad_(True)
        l3 = crit_n(x3, torch.as_tensor(gt[lo:hi]).cuda(), torch.as_tensor(vis[lo:hi]).cuda())
        l3.backward()
        os.environ.pop("HIPE_EXCHANGE")
        X, Y = synth.procrustes_pairs(101, seed=4)
        l2, h2 = PL.shard_bounds(101, world, rank)
        pa, raw = PL.sharded_pa_mpjpe(X[l2:h2], Y[l2:h2])
        np.savez(os.path.join(out_dir, "r%d.npz" % rank), loss=loss.item(), coords=crit.coords.cpu().numpy(),
                 grad=x.grad.cpu().numpy(), pa=pa.item(), raw=raw.item(), repeats=np.array(vals),
                 exchange=np.array(peer_used), exchange_n=np.array(crit_n.exchange), peer_err=np.array(peer_err),
                 loss_n=l3.item(), coords_n=crit_n.coords.cpu().numpy(), grad_n=x3.grad.cpu().numpy(),
                 loss_s=l4.item(), coords_s=crit_s.coords.cpu().numpy(), grad_s=x4.grad.cpu().numpy(),
                 exchange_s=np.array(crit_s.exchange), loss_type=np.array(deferred_type), loss_s_type=np.array(type(l4).__name__),
                 peer_error=np.array(getattr(crit, "_peer_error", "")))
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
@pytest.mark.parametrize("geom", [(16, 16, 16), (32, 64, 64)], ids=["two-launch", "single-launch"])
def test_sharded_criterion_nccl_world2(tmp_path, geom):
    import torch.multiprocessing as mp
    from oracle import integral_np as O
    from oracle import procrustes_np as P
    from tests import synth
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    B, world = 6, 2
    D, H, W = geom
    mp.spawn(_worker, args=(world, port, B, str(tmp_path), geom), nprocs=world, join=True)
    hm = synth.heatmaps(B, J, D, H, W, "peaked", seed=42).astype(np.float64)
    gt, vis = synth.targets(B, J, seed=42, zero_vis_frac=0.1)
    ref_loss, ref_c = O.joint_location_loss(hm, gt, vis, return_coords=True)      # equal shards: global sum / B
    ref_g = O.joint_location_loss_grad(hm, gt, vis)
    outs = [np.load(os.path.join(str(tmp_path), "r%d.npz" % r)) for r in range(world)]
    assert outs[0]["loss"] == outs[1]["loss"]
    for o in outs:
        # the NVLink peer-memory exchange is the one that ran, it never timed out, repeated steps give
        # the same bits, and the NCCL form of the step agrees
        assert str(o["exchange"]) == "peer", "peer exchange unavailable: %s" % o["peer_error"]
        assert str(o["exchange_n"]) == "nccl" and not bool(o["peer_err"])
        assert (o["repeats"] == o["loss"]).all()
        assert abs(o["loss_n"] - o["loss"]) <= 1e-14 * abs(o["loss"])
        # deferred (publish + side-stream collect) and in-stream forms of the peer exchange: same bits
        if str(o["exchange"]) == "peer":
            assert str(o["loss_type"]) == "DeferredLoss" and str(o["loss_s_type"]) == "Tensor" and str(o["exchange_s"]) == "peer"
        assert o["loss_s"] == o["loss"]
        np.testing.assert_array_equal(o["coords_s"], o["coords"])
        np.testing.assert_array_equal(o["grad_s"], o["grad"])
        np.testing.assert_array_equal(o["coords_n"], o["coords"])
        np.testing.assert_array_equal(o["grad_n"], o["grad"])
    assert abs(outs[0]["loss"] - ref_loss) <= 1e-5 * ref_loss
    for r in range(world):
        np.testing.assert_array_equal(outs[r]["coords"], outs[0]["coords"])
        err = np.abs(outs[r]["coords"] - ref_c)
        assert (err <= 1e-5 * np.maximum(np.abs(ref_c), 0.5)).all()
        g = outs[r]["grad"].reshape(-1, D * H * W)
        rg = ref_g[r * 3:(r + 1) * 3].reshape(-1, D * H * W)
        assert (np.abs(g - rg) <= 1e-5 * np.abs(rg).max(axis=1, keepdims=True) + 1e-12).all()
    X, Y = synth.procrustes_pairs(101, seed=4)
    rpa, rraw = P.pa_mpjpe(X, Y)
    assert abs(outs[0]["pa"] - rpa) < 1e-9 and abs(outs[1]["raw"] - rraw) < 1e-9


def _worker_graphed(rank, world, port, B, out_dir, geom):
    import torch.distributed as dist
    import hipe_b200
    from hipe_b200 import parallel as PL
    from tests import synth
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        D, H, W = geom
        lo, hi = PL.shard_bounds(B, world, rank)
        batches = []
        for seed in (42, 43, 44, 45, 46):
            hm = synth.heatmaps(B, J, D, H, W, "peaked", seed=seed)
            gt, vis = synth.targets(B, J, seed=seed, zero_vis_frac=0.1)
            batches.append((torch.as_tensor(hm[lo:hi]).cuda(), torch.as_tensor(gt[lo:hi]).cuda(), torch.as_tensor(vis[lo:hi]).cuda()))
        # eager sharded criterion: the reference of this test
        crit = PL.ShardedJointLocationLoss(global_batch=B)
        want = []
        for x, g, v in batches:
            xr = x.clone().requires_grad_(True)
            loss = crit(xr, g, v)
            loss.backward()
            want.append((loss.item(), xr.grad.clone(), crit.coords.clone()))
        # the graphed step, back to back without reading anything in between (replays run ahead of the
        # side stream), results read afterwards from the per-step copies
        step = hipe_b200.graphed_sharded(batches[0][0], batches[0][1], batches[0][2], global_batch=B)
        got = []
        for i, (x, g, v) in enumerate(batches):
            loss, grad = step(x, g, v)
            got.append([loss, grad.clone(), step.coords.clone()])
            if i == 3:
                # a loss value lives in a ring of four buffers: read the first four before the fifth step
                for rec in got:
                    rec[0] = (type(rec[0]).__name__, rec[0].item())
        got[4][0] = (type(got[4][0]).__name__, got[4][0].item())
        step.check()
        ok = True
        msgs = []
        for i, ((wl, wg, wc), (l, g_, c)) in enumerate(zip(want, got)):
            if l[0] != "DeferredLoss":
                ok = False; msgs.append("step %d: loss is a %s" % (i, l[0]))
            if l[1] != wl:
                ok = False; msgs.append("step %d: loss %r != %r" % (i, l[1], wl))
            if not torch.equal(g_, wg):
                ok = False; msgs.append("step %d: gradient differs by %g" % (i, (g_ - wg).abs().max().item()))
            if not torch.equal(c, wc):
                ok = False; msgs.append("step %d: joints differ" % i)
        np.savez(os.path.join(out_dir, "g%d.npz" % rank), ok=np.array(ok), msgs=np.array("; ".join(msgs)),
                 losses=np.array([w[0] for w in want]))
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_graphed_sharded_step_matches_eager_world2(tmp_path):
    """hipe_b200.graphed_sharded: local launch replayed from a CUDA graph, exchange launches outside it --
    same loss bits, gradient and gathered joints as the eager sharded criterion, five steps in a row."""
    import torch.multiprocessing as mp
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    B, world = 6, 2
    mp.spawn(_worker_graphed, args=(world, port, B, str(tmp_path), (32, 64, 64)), nprocs=world, join=True)
    outs = [np.load(os.path.join(str(tmp_path), "g%d.npz" % r)) for r in range(world)]
    for o in outs:
        assert bool(o["ok"]), str(o["msgs"])
    np.testing.assert_array_equal(outs[0]["losses"], outs[1]["losses"])
    assert len(set(outs[0]["losses"].tolist())) == 5          # five different batches went through
