"""Full-size (BASELINE.json configs[1]: 64 samples x 21 joints x 64^3 fp32 = 1344 volumes, 1.41 GB)
checks through size-independent properties, where the fp64 oracle would take minutes; and the
threaded calling convention of the reference's DataParallelCriterion (one Python thread per GPU,
common/nets/balanced_parallel.py:149-173)."""
import threading

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
J, D = 21, 64


def _logits(B, seed, device="cuda"):
    g = torch.Generator(device=device).manual_seed(seed)
    x = torch.randn(B, J * D, D, D, generator=g, device=device)
    # one-voxel spikes of height 14 at known positions: the expectation sits within ~1e-3 voxel of them
    idx = torch.randint(0, D, (B * J, 3), generator=g, device=device)
    flat = x.view(B * J, D, D, D)
    flat[torch.arange(B * J, device=device), idx[:, 2], idx[:, 1], idx[:, 0]] += 30.0
    return x, idx


@pytest.mark.parametrize("fused", [None, False])
def test_full_size_properties(fused):
    from nets import loss as L
    B = 64
    x, idx = _logits(B, 7)
    g = torch.Generator(device="cuda").manual_seed(8)
    gt = torch.rand(B, 3 * J, generator=g, device="cuda", dtype=torch.float64) - 0.5
    vis = torch.ones_like(gt)
    crit = L.JointLocationLoss(fused=fused)
    xa = x.clone().requires_grad_(True)
    la = crit(xa, gt, vis)
    la.backward()
    ga = xa.grad.view(B * J, -1)
    # a spike 30 above N(0,1) noise carries all but ~e^-30 * 262144 ~ 2e-8 of the mass:
    # coords = idx/dim - 0.5 (common/nets/loss.py:54-56; voxel indices start at 0)
    with torch.no_grad():
        c = L.softmax_integral_tensor(x, J, D, D, D).view(B * J, 3).double()
    want = idx.double() / D - 0.5
    assert (c - want).abs().max().item() < 1e-5
    # the gradient of a softmax expectation sums to zero over every volume
    scale = ga.abs().amax(dim=1)
    assert (ga.double().sum(dim=1).abs() <= 1e-4 * scale.double() + 1e-12).all()
    # shift invariance: adding a constant per volume changes neither the loss nor the gradient
    shift = torch.linspace(-3, 3, B * J, device="cuda").view(B, J, 1, 1, 1)
    xb = (x.view(B, J, D, D, D) + shift).view_as(x).clone().requires_grad_(True)
    lb = crit(xb, gt, vis)
    lb.backward()
    assert abs(la.item() - lb.item()) <= 1e-5 * abs(la.item())
    gb = xb.grad.view(B * J, -1)
    assert ((ga - gb).abs() <= 2e-5 * scale[:, None] + 1e-12).all()
    # the loss is what the coordinates say (fp64 sum of 64 x 63 terms)
    ref = ((c.view(B, 3 * J) - gt).abs() * vis).sum().item() / B
    assert abs(la.item() - ref) <= 1e-6 * abs(ref)


def test_single_launch_and_two_launch_plans_agree_at_full_size():
    from nets import loss as L
    B = 64
    x, _ = _logits(B, 17)
    g = torch.Generator(device="cuda").manual_seed(18)
    gt = torch.rand(B, 3 * J, generator=g, device="cuda", dtype=torch.float64) - 0.5
    vis = (torch.rand(B, 3 * J, generator=g, device="cuda") > 0.1).double()
    outs = []
    for fused in (None, False):
        xx = x.clone().requires_grad_(True)
        l = L.JointLocationLoss(fused=fused)(xx, gt, vis)
        l.backward()
        outs.append((l.item(), xx.grad))
    assert abs(outs[0][0] - outs[1][0]) <= 1e-6 * abs(outs[1][0])
    ga, gb = outs[0][1].view(B * J, -1), outs[1][1].view(B * J, -1)
    scale = gb.abs().amax(dim=1, keepdim=True)
    assert ((ga - gb).abs() <= 2e-5 * scale + 1e-12).all()


def test_one_python_thread_per_device_like_data_parallel_criterion():
    """Concurrent calls from Python threads (ctypes releases the GIL), each under its own
    torch.cuda.device, on every visible GPU -- and two threads on side streams of GPU 0."""
    from nets import loss as L
    B = 16
    n_dev = torch.cuda.device_count()
    jobs = [("cuda:%d" % d, None) for d in range(n_dev)] + [("cuda:0", "side"), ("cuda:0", "side")]
    results, errors = {}, []

    def work(k, dev, stream_kind):
        try:
            with torch.cuda.device(dev):
                stream = torch.cuda.Stream() if stream_kind else torch.cuda.current_stream()
                with torch.cuda.stream(stream):
                    x, _ = _logits(B, 100, device=dev)           # same seed everywhere: same answer everywhere
                    g = torch.Generator(device=dev).manual_seed(101)
                    gt = torch.rand(B, 3 * J, generator=g, device=dev, dtype=torch.float64) - 0.5
                    vis = torch.ones_like(gt)
                    vals = []
                    for _ in range(4):
                        xx = x.clone().requires_grad_(True)
                        l = L.JointLocationLoss()(xx, gt, vis)
                        l.backward()
                        vals.append((l.item(), float(xx.grad.double().abs().sum().item())))
                    stream.synchronize()
                results[k] = vals
        except Exception as e:   # surfaced below
            errors.append((k, repr(e)))

    threads = [threading.Thread(target=work, args=(k, d, s)) for k, (d, s) in enumerate(jobs)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not errors, errors
    first = results[0][0]
    for k, vals in results.items():
        for v in vals:
            assert v == first, (k, v, first)              # bit-identical across threads, devices, repeats
