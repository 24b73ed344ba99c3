"""CPU, world_size 2 over gloo: the exchange logic of the batch-sharded criterion
(hipe_b200/parallel.py) with an oracle-backed local function standing in for the kernels.
Checks the value every rank returns, the gathered joints, and the local gradients against the
single-process semantics of DataParallelCriterion (common/nets/balanced_parallel.py:117-127)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from hipe_b200 import parallel as PL
from oracle import integral_torch as OT
from tests import synth

J, D = 21, 6


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _oracle_local(heatmap, gt, vis):
    w, h = heatmap.shape[-1], heatmap.shape[-2]
    d = heatmap.shape[-3] // J
    if heatmap.shape[0] == 0:        # a rank whose shard is empty still takes part in the exchange
        return heatmap.sum(), heatmap.new_zeros((0, 3 * J))
    c = OT.integral_coords(heatmap, J, w, h, d)
    return (torch.abs(c - gt) * vis).sum(), c


def _worker(rank, world, port, B, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        hm = torch.as_tensor(synth.heatmaps(B, J, D, D, D, "peaked", seed=42)).double()
        gt, vis = (torch.as_tensor(a) for a in synth.targets(B, J, seed=42, zero_vis_frac=0.1))
        lo, hi = PL.shard_bounds(B, world, rank)
        x = hm[lo:hi].clone().requires_grad_(True)
        crit = PL.ShardedJointLocationLoss(local_fn=_oracle_local, global_batch=(B if B != 5 else None))
        loss = crit(x, gt[lo:hi], vis[lo:hi])
        if hi > lo:
            loss.backward()
        np.savez(os.path.join(out_dir, "r%d.npz" % rank), loss=loss.detach().numpy(), coords=crit.coords.numpy(),
                 grad=(x.grad.numpy() if x.grad is not None else np.zeros(0)), lo=lo, hi=hi)
        # eval-side reduction: [sum_pa, sum_raw, count] records summed in rank order
        rec = torch.tensor([1.5 * (rank + 1), 2.0 * (rank + 1), 10.0 * (rank + 1)], dtype=torch.float64)
        pa, raw = PL.reduce_error_sums(PL.gather_records(rec))
        assert abs(pa.item() - 4.5 / 30) < 1e-15 and abs(raw.item() - 6.0 / 30) < 1e-15
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("B", [4, 5, 1])
def test_sharded_criterion_world2(tmp_path, B):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), B, str(tmp_path)), nprocs=world, join=True)
    hm = torch.as_tensor(synth.heatmaps(B, J, D, D, D, "peaked", seed=42)).double()
    gt, vis = (torch.as_tensor(a) for a in synth.targets(B, J, seed=42, zero_vis_frac=0.1))
    # single-process semantics: per-replica sum/len, mean over the replicas that got rows
    xs, losses = [], []
    for r in range(world):
        lo, hi = PL.shard_bounds(B, world, r)
        if hi == lo:
            continue
        x = hm[lo:hi].clone().requires_grad_(True)
        xs.append(x)
        losses.append(OT.joint_location_loss(x, gt[lo:hi], vis[lo:hi]))
    ref = sum(losses) / len(losses)
    ref.backward()
    ref_coords = OT.integral_coords(hm, J, D, D, D).numpy()
    outs = [np.load(os.path.join(str(tmp_path), "r%d.npz" % r)) for r in range(world)]
    assert outs[0]["loss"] == outs[1]["loss"]                      # bit-identical on every rank
    assert abs(outs[0]["loss"] - ref.item()) <= 1e-15 * max(1.0, abs(ref.item()))
    i = 0
    for r in range(world):
        np.testing.assert_array_equal(outs[r]["coords"], ref_coords)
        if outs[r]["hi"] > outs[r]["lo"]:
            np.testing.assert_allclose(outs[r]["grad"], xs[i].grad.numpy(), rtol=1e-14, atol=1e-18)
            i += 1
    if B % world == 0:
        # equal shards: the replica mean IS the global sum / B_global (loss.py:88-89)
        full = OT.joint_location_loss(hm, gt, vis)
        assert abs(outs[0]["loss"] - full.item()) <= 1e-14


def test_shard_bounds_and_records():
    assert [PL.shard_bounds(10, 4, r) for r in range(4)] == [(0, 3), (3, 6), (6, 9), (9, 10)]
    assert [PL.shard_bounds(2, 4, r) for r in range(4)] == [(0, 1), (1, 2), (2, 2), (2, 2)]
    assert PL.shard_bounds(512, 8, 7) == (448, 512)
    # matches torch.chunk, which is what scatter uses
    for n, w in [(10, 4), (64, 8), (7, 3), (2, 4)]:
        chunks = torch.arange(n).chunk(w)
        for r in range(w):
            lo, hi = PL.shard_bounds(n, w, r)
            assert hi - lo == (len(chunks[r]) if r < len(chunks) else 0)
    with pytest.raises(ValueError):
        PL.shard_bounds(4, 2, 2)
    c = torch.arange(6, dtype=torch.float32).reshape(2, 3)
    rec = PL.pack_record(torch.tensor(2.5), c, 4)
    assert rec.shape == (2 + 12,) and rec.dtype == torch.float64
    sums, counts, allc = PL.unpack_records([rec, PL.pack_record(torch.tensor(1.0), c[:1] + 10, 4)], 3)
    assert counts == [2, 1] and allc.shape == (3, 3) and float(allc[2, 0]) == 10.0
    v, live = PL.combine_replica_losses(sums, counts)
    assert live == 2 and abs(v.item() - (2.5 / 2 + 1.0 / 1) / 2) < 1e-15
