"""CUDA-graph replay of the criterion's forward + backward for fixed shapes.

The reference trains 56^3 heat maps at 32 samples per GPU (main/config.py:23-25,39): 0.2 ms of
kernels per step, less than the host needs to enqueue them through Python, autograd and ctypes.
Every launch of this library is capture-safe (stream-ordered memset + kernels, the cooperative and
cluster launches included, no host synchronisation), so the step can be captured once and replayed:

    step = hipe_b200.graphed(nets.loss.JointLocationLoss(), heatmaps, gt, vis)
    loss, dheatmaps = step(heatmaps, gt, vis)      # copies into the static inputs, one graph launch
    heatmaps.backward(dheatmaps)                   # hand the gradient to the producer's graph

`loss` and `dheatmaps` are the graph's static output buffers: valid until the next call."""
import torch


class GraphedCriterion:
    def __init__(self, crit, heatmaps, gt, vis, warmup=2):
        if not heatmaps.is_cuda:
            raise RuntimeError("hipe_b200 has no CPU path: heatmaps live on %s" % heatmaps.device)
        self.crit = crit
        self.x = heatmaps.detach().clone().requires_grad_(True)
        self.gt = gt.detach().to(heatmaps.device).clone()
        self.vis = vis.detach().to(heatmaps.device).clone()
        side = torch.cuda.Stream(device=heatmaps.device)
        side.wait_stream(torch.cuda.current_stream(heatmaps.device))
        with torch.cuda.stream(side):
            for _ in range(max(1, warmup)):              # allocator pools, plan caches, workspaces
                self.x.grad = None
                self._step()
        torch.cuda.current_stream(heatmaps.device).wait_stream(side)
        self.x.grad = None
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self.loss = self._step()
        self.grad = self.x.grad                           # the buffer the capture allocated; replay refills it

    def _step(self):
        loss = self.crit(self.x, self.gt, self.vis)
        loss.backward()
        return loss.detach()

    def __call__(self, heatmaps=None, gt=None, vis=None):
        """Replay.  Arguments that are given are copied into the static inputs first (device to
        device, stream-ordered); pass nothing to replay on what is already there."""
        with torch.no_grad():
            if heatmaps is not None and heatmaps.data_ptr() != self.x.data_ptr():
                self.x.copy_(heatmaps)
            if gt is not None:
                self.gt.copy_(gt)
            if vis is not None:
                self.vis.copy_(vis)
        self.graph.replay()
        return self.loss, self.grad

    @property
    def static_input(self):
        """The static heat-map buffer: a producer that writes its output here needs no copy."""
        return self.x


def graphed(crit, heatmaps, gt, vis, warmup=2):
    """Capture `loss = crit(heatmaps, gt, vis); loss.backward()` for these shapes and dtypes."""
    return GraphedCriterion(crit, heatmaps, gt, vis, warmup)
