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


class GraphedShardedStep:
    """The batch-sharded training step (parallel.ShardedJointLocationLoss, equal shards, peer-memory
    exchange) with its local part -- the single launch that produces loss terms, joints and the
    FINISHED heat-map gradient (norm = 1 / (B_local * world), balanced_parallel.py:52-55,127) --
    captured in a CUDA graph, and the two exchange launches enqueued on the side stream outside it
    (inside the graph their rendezvous with the slowest rank would end every replay):

        step = hipe_b200.graphed_sharded(heatmaps_local, gt_local, vis_local, global_batch=B)
        loss, dheatmaps = step(heatmaps_local, gt_local, vis_local)
        heatmaps_local.backward(dheatmaps)
        step.coords                                   # (B_global, 3J) joints of all ranks

    Per step the host enqueues one graph launch, two kernels and four event operations: what a
    small shard needs at the strong-scaling end (bf16 64^3 at 64 samples per rank is 0.28 ms of
    kernels against 0.28-0.35 ms of eager enqueue with eight processes on one host).
    `loss` is a parallel.DeferredLoss (value valid until four steps later); `dheatmaps` is the graph's
    static output, valid until the next call."""

    def __init__(self, heatmaps, gt, vis, global_batch=None, group=None, timeout_s=20.0, warmup=2):
        import torch.distributed as dist
        from . import ops, parallel
        if not heatmaps.is_cuda:
            raise RuntimeError("hipe_b200 has no CPU path: heatmaps live on %s" % heatmaps.device)
        dev = heatmaps.device
        world = dist.get_world_size(group)
        b, k = int(heatmaps.shape[0]), int(gt.shape[1])
        if global_batch is not None and global_batch != b * world:
            raise ValueError("graphed_sharded needs equal shards: %d rows per rank x %d ranks != global batch %d"
                             % (b, world, global_batch))
        from nets.loss import FreiHandConfig
        self.J = k // 3
        self.D = heatmaps.shape[-3] // FreiHandConfig.num_joints
        self.H, self.W = int(heatmaps.shape[-2]), int(heatmaps.shape[-1])
        self.n_vol = b * self.J
        if ops.fused_plan(heatmaps, self.n_vol, self.D, self.H, self.W) is None:
            raise ValueError("no single-launch plan for %s heat maps of %dx%dx%d: use ShardedJointLocationLoss"
                             % (heatmaps.dtype, self.D, self.H, self.W))
        self.world, self.b, self.k = world, b, k
        self.norm = 1.0 / (b * world)
        self.inv_denom = 1.0 / (b * world)
        self._crit = parallel.ShardedJointLocationLoss(group=group, global_batch=b * world, timeout_s=timeout_s)
        self.px = self._crit._peer_exchange(world, b, k, dev)
        if self.px is None:
            raise RuntimeError("graphed_sharded needs the NVLink peer-memory exchange (symmetric memory over NCCL): %s"
                               % getattr(self._crit, "_peer_error", "unavailable on this box"))
        self.x = heatmaps.detach().clone()
        self.gt = gt.detach().to(device=dev, dtype=torch.float64).clone()
        self.vis = vis.detach().to(device=dev, dtype=torch.float64).clone()
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):
            for _ in range(max(1, warmup)):              # allocator pools, plan caches, workspaces
                self._local()
        torch.cuda.current_stream(dev).wait_stream(side)
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self.coords_local, self.terms, self.grad = self._local()
        self.values = [torch.zeros((), dtype=torch.float64, device=dev) for _ in range(4)]
        self.pub_done = [torch.cuda.Event() for _ in range(2)]
        self.step = 0
        self._coords = None
        self._coords64 = None
        self._coords_event = None

    def _local(self):
        from . import ops
        coords, _, _, terms, dlogits = ops.joint_loss_fused(self.x, self.n_vol, self.D, self.H, self.W, self.gt, self.vis,
                                                            self.norm, want_aux=False)
        return coords.view(self.b, self.k), terms, dlogits

    def __call__(self, heatmaps=None, gt=None, vis=None):
        from . import parallel
        dev = self.x.device
        main = torch.cuda.current_stream(dev)
        if self.step > 0:
            # the previous step's publish reads the static terms / joints the replay is about to overwrite
            main.wait_event(self.pub_done[(self.step - 1) & 1])
        with torch.no_grad():
            if heatmaps is not None and heatmaps.data_ptr() != self.x.data_ptr():
                self.x.copy_(heatmaps)
            if gt is not None:
                self.gt.copy_(gt)
            if vis is not None:
                self.vis.copy_(vis)
        self.graph.replay()
        value = self.values[self.step & 3]
        _, allc, event = self.px.run_deferred(self.terms, 1.0, self.coords_local, self.inv_denom, torch.float64,
                                              into=value, pub_done=self.pub_done[self.step & 1])
        self.step += 1
        self._coords, self._coords64, self._coords_event = None, allc.view(self.world * self.b, self.k), event
        loss = value.as_subclass(parallel.DeferredLoss)
        loss._hipe_event = event
        return loss, self.grad

    @property
    def coords(self):
        """(B_global, 3J) joints of all ranks (fp32); the current stream first waits for the collect."""
        if self._coords is None and self._coords64 is not None:
            torch.cuda.current_stream(self._coords64.device).wait_event(self._coords_event)
            self._coords = self._coords64.to(torch.float32)
        return self._coords

    def check(self):
        """Raise if an earlier collect gave up waiting for a rank (no synchronisation)."""
        self.px.check()

    @property
    def static_input(self):
        return self.x


def graphed_sharded(heatmaps, gt, vis, global_batch=None, group=None, timeout_s=20.0, warmup=2):
    """Capture the local part of the batch-sharded step for these shapes and dtypes (see GraphedShardedStep)."""
    return GraphedShardedStep(heatmaps, gt, vis, global_batch, group, timeout_s, warmup)
