"""Batch-sharded form of the path: one process per GPU (torchrun), `torch.distributed` for the
plumbing.  Replaces the reference's single-process DataParallelCriterion + Reduce
(common/nets/balanced_parallel.py:97-127,46-55) and the eval scripts' gather of whole heatmaps to
GPU 0 (main/evaluate.py:62-63, main/test.py:91-92).

Every (sample, joint) volume and every Procrustes sample is independent, so rank r keeps the
samples `shard_bounds(B, world, r)` -- the dim-0 chunking `scatter` does
(balanced_parallel.py:122) -- and no volume, probability or heatmap gradient ever crosses a GPU.
The only exchange per step is one packed record per rank,

    [ sum_r |c - gt| * vis ,  B_r ,  coords_r (B_max * 3J, zero padded) ]      (fp64)

after which every rank holds all predicted joints and combines the per-replica losses in fixed
rank order (bit-identical on every rank).  Backward needs no collective: each rank scales its own
gradient by 1/world, exactly what `Reduce.backward` broadcasts (balanced_parallel.py:52-55).

How the record travels:
  * equal shards known a priori on CUDA (the training loop): NVLink peer memory, no NCCL launch --
    `hipe_peer_publish` (stores + flag, never waits) and `hipe_peer_collect` (waits for the peers,
    combines) both on a side stream that waits for the step's loss kernel; they run on the SMs the
    persistent loss launch of the next step leaves idle.  The stream of the loss kernels only
    records an event per step and waits for the collect of step t-2 (which bounds the lifetime of
    the buffers the side stream touches), or for the collect of step t when the loss value / the
    gathered joints are consumed (the returned loss is a `DeferredLoss`: any use of its value first
    makes the current stream wait), so a slow rank or a host hiccup costs the others nothing until
    two steps later.  `HIPE_PUBLISH_STREAM=main` keeps the publish in stream order.  `defer=False` keeps the one-kernel form
    (`hipe_peer_exchange`) in stream order.
  * anything else (unequal / unknown shard sizes, CPU / gloo, no symmetric memory): one
    `all_gather` of the record.
"""
import torch
import torch.distributed as dist
import torch.nn as nn


def shard_bounds(n, world, rank):
    """[lo, hi) of rank's rows when n rows are split like torch.chunk / scatter: ceil(n/world) per
    rank, the tail ranks short or empty."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError("bad rank %d of world %d" % (rank, world))
    per = -(-n // world)
    lo = min(rank * per, n)
    return lo, min(lo + per, n)


def max_shard(n, world):
    return -(-n // world)


def pack_record(loss_sum, coords, b_max):
    """-> 1-D fp64 tensor [loss_sum, B_local, coords padded to b_max rows]."""
    b, k = coords.shape
    rec = torch.zeros(2 + b_max * k, dtype=torch.float64, device=coords.device)
    rec[0] = loss_sum.to(torch.float64)
    rec[1] = float(b)
    rec[2:2 + b * k] = coords.reshape(-1).to(torch.float64)
    return rec


def unpack_records(recs, k, counts=None):
    """list of per-rank records -> (loss_sums (world,), counts list[int], coords (B_global, k) fp64).
    `counts` known a priori (from `shard_bounds`) avoids reading them back from the device."""
    sums = torch.stack([r[0] for r in recs])
    if counts is None:
        counts = [int(round(float(r[1]))) for r in recs]
    rows = [r[2:2 + c * k].reshape(c, k) for r, c in zip(recs, counts)]
    return sums, counts, torch.cat(rows, dim=0)


def gather_records(rec, group=None):
    world = dist.get_world_size(group)
    out = [torch.empty_like(rec) for _ in range(world)]
    dist.all_gather(out, rec, group=group)
    return out


class _ValueWithGradOf(torch.autograd.Function):
    """forward: `value` (a detached global loss); backward: the incoming gradient goes to `carrier`
    unchanged.  The same thing as value + (carrier - carrier.detach()) without its three launches:
    at 0.5 ms of kernels per step the step is bound by the host's launch rate, not the GPU.
    `touch=False`: return a view of `value` without launching anything on the current stream (the
    value may still be in flight on the exchange's side stream)."""

    @staticmethod
    def forward(ctx, carrier, value, touch=True):
        if not touch and value.dtype == carrier.dtype:
            return value.view_as(carrier)
        return value.to(carrier.dtype).view_as(carrier).clone()

    @staticmethod
    def backward(ctx, g):
        return g, None, None


# attribute reads / calls on a DeferredLoss that do not look at its value
_NO_VALUE = frozenset(("backward", "grad_fn", "requires_grad", "is_leaf", "shape", "dtype", "device", "size", "dim",
                       "ndim", "numel", "is_cuda", "layout", "register_hook", "retain_grad", "retains_grad",
                       "_backward_hooks", "_version", "names", "output_nr", "_base", "is_sparse", "is_quantized",
                       "is_meta", "is_nested", "grad", "element_size", "nelement", "ndimension", "is_contiguous",
                       "stride", "storage_offset", "data_ptr", "__len__", "is_floating_point", "is_complex",
                       "requires_grad_", "_is_view", "is_mkldnn", "is_xla", "is_cpu", "is_ipu", "is_xpu",
                       "is_mps", "is_maia", "is_mtia", "is_vulkan", "is_inference", "_is_zerotensor", "is_neg",
                       "is_conj", "get_device", "type", "untyped_storage"))


class DeferredLoss(torch.Tensor):
    """The loss of a sharded step whose VALUE is produced by the exchange's collect launch on a side
    stream.  It behaves like the plain 0-dim tensor it wraps; anything that reads the value
    (`.item()`, `.detach()`, arithmetic, printing, ...) first makes the CURRENT stream wait for the
    collect's event -- a stream-side wait, not a host synchronisation -- while `.backward()` and
    metadata reads do not, which is what lets the rendezvous with the slowest rank slide under the
    next step."""

    _hipe_event = None

    def join(self):
        ev = self._hipe_event
        if ev is not None:
            with torch._C.DisableTorchFunctionSubclass():
                dev = torch.Tensor.device.__get__(self)
            torch.cuda.current_stream(dev).wait_event(ev)
        return self

    @classmethod
    def __torch_function__(cls, func, types, args=(), kwargs=None):
        kwargs = kwargs or {}
        name = getattr(getattr(func, "__self__", None), "__name__", None) or getattr(func, "__name__", "")
        if name not in _NO_VALUE:
            for a in args:
                if isinstance(a, DeferredLoss):
                    a.join()
            for a in kwargs.values():
                if isinstance(a, DeferredLoss):
                    a.join()
        with torch._C.DisableTorchFunctionSubclass():
            return func(*args, **kwargs)


class _Exchange:
    """Pre-allocated send / receive records of the step's one all-gather (fast path: every rank's
    shard size known a priori and equal), so a step issues: 2 copies into the record, 1
    all_gather_into_tensor, 1 reduction of the `world` loss sums."""

    def __init__(self, world, b, k, device):
        self.world, self.b, self.k = world, b, k
        self.send = torch.zeros(2 + b * k, dtype=torch.float64, device=device)
        self.send[1] = float(b)
        self.recv_flat = torch.empty(world * (2 + b * k), dtype=torch.float64, device=device)
        self.recv = self.recv_flat.view(world, 2 + b * k)

    def run(self, loss_sum, coords, group):
        self.send[0:1].copy_(loss_sum.reshape(1))
        self.send[2:].view(self.b, self.k).copy_(coords)
        dist.all_gather_into_tensor(self.recv_flat, self.send, group=group)
        return self.recv[:, 0], self.recv[:, 2:]


def combine_replica_losses(sums, counts, size_average=True):
    """What DataParallelCriterion returns (balanced_parallel.py:127): the mean over replicas of
    each replica's `sum / len(coord_out)` (loss.py:88-89); summed in rank order.  Replicas with an
    empty shard do not exist in the reference (scatter yields fewer chunks) and are skipped."""
    total = torch.zeros((), dtype=torch.float64, device=sums.device)
    live = 0
    for s, c in zip(sums, counts):
        if c == 0:
            continue
        total = total + (s / c if size_average else s)
        live += 1
    return total / max(live, 1), live


def _local_cuda_loss(heatmap_out, gt_coord, gt_vis):
    """(loss_sum_local [differentiable, fp64], coords (B_local, 3J) fp32) on this rank's GPU."""
    from nets.loss import FreiHandConfig, softmax_integral_tensor
    from . import ops
    joint_num = int(gt_coord.shape[1] / 3)
    w, h = heatmap_out.shape[-1], heatmap_out.shape[-2]
    d = heatmap_out.shape[-3] // FreiHandConfig.num_joints
    coords = softmax_integral_tensor(heatmap_out, joint_num, w, h, d)
    loss_sum = ops.IntegralL1Function.apply(coords, gt_coord.to(torch.float64), gt_vis.to(torch.float64), 1.0)
    return loss_sum, coords


class _PeerExchange:
    """The step's exchange over NVLink peer memory (no NCCL launch, no pack / unpack launches): every
    rank stores its record into all peers' symmetric buffers and publishes a sequence number; a
    collect launch waits for the peers' and combines.  `run` does both in ONE kernel in stream order
    (hipe_peer_exchange); `run_deferred` splits them: publish on the caller's stream, collect on a
    side stream (hipe_peer_publish / hipe_peer_collect).  The buffers come from
    torch.distributed._symmetric_memory (plumbing: allocation + rendezvous)."""

    def __init__(self, world, rank, b, k, device, group, timeout_s=20.0):
        import torch.distributed._symmetric_memory as symm
        from . import _lib
        lib = _lib.load()
        self.world, self.rank, self.b, self.k, self.group = world, rank, b, k, group
        self.timeout_s = float(timeout_s)
        nbytes = lib.hipe_peer_exchange_bytes(world, b, k)
        if nbytes <= 0:
            raise RuntimeError("peer exchange: unsupported world / shard size")
        self.buf = symm.empty(nbytes // 8, dtype=torch.float64, device=device)
        self.buf.zero_()
        self.hdl = symm.rendezvous(self.buf, dist.group.WORLD if group is None else group)
        self.peers_dev = int(self.hdl.buffer_ptrs_dev)
        self.err = torch.zeros(1, dtype=torch.int32, device=device)
        # split form: one word per rank in pinned host memory the collect kernel writes straight into
        # (UVA), so the host can see a timeout -- and WHICH rank is missing -- without synchronising
        self.err_host = torch.zeros(world, dtype=torch.int32).pin_memory()
        self.call_ring = [torch.empty((world * b, k), dtype=torch.float64, device=device) for _ in range(4)]
        self.keep = [None] * 4
        self.err_np = self.err_host.numpy()          # host view: reading it costs no torch dispatch
        self.err_ptr = self.err_host.data_ptr()
        self._lib = lib
        self.scratch = torch.zeros(world + 1, dtype=torch.float64, device=device)
        self.scratch_ptr = self.scratch.data_ptr()
        self.value = torch.empty(2, dtype=torch.float64, device=device)      # alternating outputs
        self.coords_all = [torch.empty((world * b, k), dtype=torch.float64, device=device) for _ in range(2)]
        self.seq = 0
        self.side = torch.cuda.Stream(device=device)
        self.ev_pub = [torch.cuda.Event() for _ in range(4)]
        self.ev_col = [torch.cuda.Event() for _ in range(4)]
        self.last_col = None
        # HIPE_PUBLISH_STREAM=main keeps the publish in stream order behind the loss kernels (18 us per
        # step on their stream); default: on the side stream with the collect
        import os
        self.publish_on_side = os.environ.get("HIPE_PUBLISH_STREAM", "side") != "main"
        torch.cuda.synchronize(device)
        dist.barrier(group)                    # every buffer is zeroed before anyone's first store

    def run(self, loss_scalar, in_scale, coords, value_scale):
        from . import _lib, ops
        self._join_side(coords.device)
        self.seq += 1
        slot = self.seq & 1
        out = self.value[slot:slot + 1]
        call = self.coords_all[slot]
        with ops._launch("peer_exchange", coords):
            _lib.check(_lib.load().hipe_peer_exchange(
                loss_scalar.data_ptr(), loss_scalar.numel(), float(in_scale), coords.data_ptr(), self.b, self.k,
                self.peers_dev,
                self.world, self.rank, self.seq, float(value_scale), out.data_ptr(), None, call.data_ptr(),
                self.err.data_ptr(), self.scratch.data_ptr(), self.timeout_s, torch.cuda.current_stream(coords.device).cuda_stream))
        return out.view(()), call

    def _join_side(self, device):
        if self.last_col is not None:
            torch.cuda.current_stream(device).wait_event(self.last_col)
            self.last_col = None

    def run_deferred(self, loss_scalar, in_scale, coords, value_scale, out_dtype=torch.float64, into=None, pub_done=None):
        """-> (value 0-dim `out_dtype`, coords_all (world*b, k) fp64, event).  The two tensors are
        written on the side stream; `event` fires when they are complete.  `into`: a 0-dim fp64 tensor
        that receives the value (the unwritten loss scalar of the single-launch autograd function),
        else a fresh one.  coords_all is one of four ring buffers: valid until four steps later.
        `pub_done`: an event to record once the publish launch has read `loss_scalar` and `coords` (for
        callers that overwrite those buffers in place, like the graphed step)."""
        from . import _lib, ops
        lib = self._lib
        dev = coords.device
        if self.err_np.any():
            self.check()
        main = torch.cuda.current_stream(dev)
        self.seq += 1
        i = self.seq & 3
        ev_pub, ev_col, side = self.ev_pub[i], self.ev_col[i], self.side
        timed = ops.KERNEL_EVENTS is not None
        value = into.detach() if into is not None else torch.empty((), dtype=torch.float64, device=dev)
        call = self.call_ring[i]
        if self.publish_on_side:
            # Publish AND collect on the side stream: the stream of the loss kernels only records an event.
            # Flow control (see hipe_exchange.cu) -- publish(t) after this rank's collect(t-1) -- is the
            # side stream's own order.  The loss kernels' stream may run at most two steps ahead of it:
            # that bounds the lifetime of what the side stream reads and writes (kept for four steps).
            if self.seq > 2:
                main.wait_event(self.ev_col[(self.seq - 2) & 3])
            ev_pub.record(main)
            side.wait_event(ev_pub)
            self.keep[i] = (value, loss_scalar, coords)
            ops.LAUNCHES += 1
            if timed:
                p0, p1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                p0.record(side)
            _lib.check(lib.hipe_peer_publish(
                loss_scalar.data_ptr(), loss_scalar.numel(), float(in_scale), coords.data_ptr(), self.b, self.k,
                self.peers_dev, self.world, self.rank, self.seq, side.cuda_stream))
            if timed:
                p1.record(side)
                ops.KERNEL_EVENTS.setdefault("peer_publish", []).append((p0, p1))
            if pub_done is not None:
                pub_done.record(side)
        else:
            # publish in stream order behind the loss kernels (HIPE_PUBLISH_STREAM=main); flow control:
            # publish(t) only after this rank's collect(t-1)
            if self.last_col is not None:
                main.wait_event(self.last_col)
            with ops._launch("peer_publish", coords):
                _lib.check(lib.hipe_peer_publish(
                    loss_scalar.data_ptr(), loss_scalar.numel(), float(in_scale), coords.data_ptr(), self.b, self.k,
                    self.peers_dev, self.world, self.rank, self.seq, main.cuda_stream))
            ev_pub.record(main)
            if pub_done is not None:
                pub_done.record(main)
            # the value's storage must outlive the side-stream write even if the caller drops the loss at
            # once: at most two collects are outstanding (flow control above), four are kept
            self.keep[i] = value
            side.wait_event(ev_pub)
        ops.LAUNCHES += 1
        if timed:
            t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            t0.record(side)
        _lib.check(lib.hipe_peer_collect(
            self.peers_dev, self.world, self.rank, self.seq, self.b, self.k, float(value_scale),
            value.data_ptr(), None, call.data_ptr(), self.err_ptr, self.scratch_ptr, self.timeout_s, side.cuda_stream))
        if timed:
            t1.record(side)
            ops.KERNEL_EVENTS.setdefault("peer_collect", []).append((t0, t1))
        out = value
        if out_dtype != torch.float64:
            with torch.cuda.stream(side):
                out = value.to(out_dtype)
            out.record_stream(side)
        ev_col.record(side)
        self.last_col = ev_col
        return out, call, ev_col

    def check(self):
        """Raise if an earlier collect gave up waiting (no synchronisation: the kernel writes the
        words into pinned host memory)."""
        if self.err_np.any():
            missing = [r for r in range(self.world) if int(self.err_np[r]) != 0]
            raise RuntimeError("hipe_b200 peer exchange: rank(s) %s did not publish their record within %.0f s "
                               "(a rank skipped a step, crashed or is stalled); the global loss of that step is NaN"
                               % (missing, self.timeout_s))

    def timed_out(self):
        torch.cuda.synchronize(self.buf.device)
        return bool(self.err.item() != 0) or bool(self.err_np.any())


def _peer_exchange_wanted(group, device):
    """HIPE_EXCHANGE = peer | nccl | auto (default: peer memory on CUDA with the NCCL backend)."""
    import os
    mode = os.environ.get("HIPE_EXCHANGE", "auto")
    if mode == "nccl" or device.type != "cuda":
        return False
    try:
        backend = dist.get_backend(group)
    except Exception:
        return False
    return mode == "peer" or "nccl" in str(backend)


def _local_cuda_loss_fused(heatmap_out, gt_coord, gt_vis, norm, reduce_loss=True):
    """(norm * loss_sum_local [differentiable], coords, loss terms) from ONE launch that also stores
    the heatmap gradient already scaled by `norm` (ops.FusedJointLossFunction), so backward is a
    hand-over.  reduce_loss=False: the scalar is only a gradient carrier (its value is unset)."""
    from nets.loss import FreiHandConfig
    from . import ops
    joint_num = int(gt_coord.shape[1] / 3)
    w, h = heatmap_out.shape[-1], heatmap_out.shape[-2]
    d = heatmap_out.shape[-3] // FreiHandConfig.num_joints
    scaled, coords, terms = ops.FusedJointLossFunction.apply(heatmap_out, gt_coord.to(torch.float64),
                                                             gt_vis.to(torch.float64), joint_num, d, h, w, norm,
                                                             reduce_loss)
    return scaled, coords, terms


def _fused_plan(heatmap_out, gt_coord):
    from nets.loss import FreiHandConfig
    from . import ops
    if not (heatmap_out.is_cuda and torch.is_grad_enabled() and heatmap_out.requires_grad):
        return None
    d = heatmap_out.shape[-3] // FreiHandConfig.num_joints
    n_vol = heatmap_out.shape[0] * int(gt_coord.shape[1] / 3)
    return ops.fused_plan(heatmap_out, n_vol, d, heatmap_out.shape[-2], heatmap_out.shape[-1])


class ShardedJointLocationLoss(nn.Module):
    """JointLocationLoss over a batch sharded across the ranks of `group`.

        crit = ShardedJointLocationLoss(global_batch=B)
        loss = crit(heatmap_local, gt_local, vis_local)     # same value on every rank
        loss.backward()                                      # local kernels only
        crit.coords                                          # (B_global, 3J) on every rank

    `local_fn(heatmap, gt, vis) -> (loss_sum, coords)` is the per-rank computation; the default
    runs the sm_100a kernels.  (The CPU tests inject an oracle-backed function to exercise the
    exchange logic over gloo.)

    `defer` (peer-memory exchange only): True (default) = publish on the caller's stream, collect on
    a side stream, the returned loss a `DeferredLoss`; False = the whole exchange as one kernel in
    stream order.  `timeout_s`: how long a collect waits for a missing rank before it gives up,
    poisons the loss with NaN and records the rank; the next `forward` (or `check()`) raises."""

    def __init__(self, group=None, local_fn=None, global_batch=None, defer=True, timeout_s=20.0):
        """`global_batch`: total rows over all ranks, when every rank holds its `shard_bounds`
        share; the exchange then needs no size negotiation and no host synchronisation."""
        super().__init__()
        self.size_average = True
        self.group = group
        self.local_fn = local_fn or _local_cuda_loss
        self.global_batch = global_batch
        self.defer = defer
        self.timeout_s = timeout_s
        self._coords = None
        self._coords64 = None
        self._coords_event = None
        self._counts_key = None
        self._counts = None
        self._fast = {}              # (heatmap shape, dtype) -> cached decisions of the steady-state step
        self._xch = None
        self._peers = {}             # (world, b, k) -> _PeerExchange: train / val / last short batch each keep theirs
        self._peer_off = False       # True = unavailable on this box (NCCL is used)
        self._peer = None            # the exchange object the last step used
        self.exchange = None         # "peer" | "nccl": what the last step used

    def _peer_exchange(self, world, b, k, device):
        """The peer-memory exchange object for this geometry, or None.  Every rank must reach the same
        decision: success is agreed on with one MIN all-reduce at set-up time."""
        if self._peer_off:
            return None
        p = self._peers.get((world, b, k))
        if p is not None:
            return p
        if not _peer_exchange_wanted(self.group, device):
            self._peer_off = True
            return None
        ok, obj = 1, None
        try:
            obj = _PeerExchange(world, dist.get_rank(self.group), b, k, device, self.group, self.timeout_s)
        except Exception as e:                           # no symmetric memory / P2P on this box
            ok, self._peer_error = 0, repr(e)
        t = torch.tensor([ok], dtype=torch.int32, device=device)
        dist.all_reduce(t, op=dist.ReduceOp.MIN, group=self.group)
        if int(t.item()) == 0:
            self._peer_off = True
            return None
        self._peers[(world, b, k)] = obj
        return obj

    def check(self):
        """Raise if a peer-memory exchange of an earlier step timed out (names the missing ranks)."""
        for p in self._peers.values():
            p.check()

    @property
    def coords(self):
        """(B_global, 3J) predicted joints of ALL ranks, in the local coordinates' dtype (converted
        from the fp64 exchange record on first access; the current stream first waits for the
        collect launch that produces them)."""
        if self._coords is None and self._coords64 is not None:
            c64, dt = self._coords64
            if self._coords_event is not None:
                torch.cuda.current_stream(c64.device).wait_event(self._coords_event)
            self._coords = c64.reshape(-1, c64.shape[-1]).to(dt)
        return self._coords

    @coords.setter
    def coords(self, v):
        self._coords, self._coords64, self._coords_event = v, None, None

    def _fast_step(self, fp, heatmap_out, gt_coord, gt_vis):
        """The steady-state training step (equal shards known a priori, single-launch plan, peer-memory
        exchange, deferred collect) with the decisions of the first step cached: one autograd node,
        the loss scalar it returns IS the DeferredLoss the collect fills."""
        from . import ops
        px, joint_num, d, h, w, norm, inv_denom, world, b_local, k = fp
        if gt_coord.dtype != torch.float64:
            gt_coord = gt_coord.to(torch.float64)
        if gt_vis.dtype != torch.float64:
            gt_vis = gt_vis.to(torch.float64)
        loss, coords, terms = ops.FusedJointLossFunction.apply(heatmap_out, gt_coord, gt_vis, joint_num, d, h, w, norm,
                                                               False, DeferredLoss)
        _, allc, event = px.run_deferred(terms, 1.0, coords, inv_denom, torch.float64, into=loss)
        self._coords, self._coords64, self._coords_event = None, (allc.view(world, b_local, k), coords.dtype), event
        loss._hipe_event = event
        return loss

    def forward(self, heatmap_out, gt_coord, gt_vis):
        if self._fast:
            fp = self._fast.get((heatmap_out.shape, heatmap_out.dtype))
            if fp is not None and heatmap_out.requires_grad and torch.is_grad_enabled() and heatmap_out.is_contiguous():
                return self._fast_step(fp, heatmap_out, gt_coord, gt_vis)
        world = dist.get_world_size(self.group)
        counts = None
        scaled = terms = None
        px = None
        if self.global_batch is not None:
            ck = (self.global_batch, world)
            if self._counts_key != ck:
                self._counts = [hi - lo for lo, hi in (shard_bounds(self.global_batch, world, r) for r in range(world))]
                self._counts_key = ck
            counts = self._counts
            b_local = heatmap_out.shape[0]
            if counts[dist.get_rank(self.group)] != b_local:
                raise ValueError("this rank holds %d rows, shard_bounds(%d, %d) gives it %d"
                                 % (b_local, self.global_batch, world, counts[dist.get_rank(self.group)]))
            b_max = max_shard(self.global_batch, world)
            if self.local_fn is _local_cuda_loss and b_local > 0 and _fused_plan(heatmap_out, gt_coord):
                # shard sizes known a priori => so is this replica's weight in the global loss: the
                # single-launch kernel can store the finished gradient (no scaling pass in backward)
                live = sum(1 for c in counts if c > 0)
                norm = 1.0 / (b_local * live) if self.size_average else 1.0 / live
                equal = all(c == b_local for c in counts)
                px = self._peer_exchange(world, b_local, int(gt_coord.shape[1]), heatmap_out.device) if equal else None
                # with the peer exchange the loss terms are summed inside the exchange kernel
                scaled, coords, terms = _local_cuda_loss_fused(heatmap_out, gt_coord, gt_vis, norm,
                                                               reduce_loss=px is None)
                loss_sum = scaled.detach() / norm if px is None else None
        if scaled is None:
            loss_sum, coords = self.local_fn(heatmap_out, gt_coord, gt_vis)
        b_local, k = coords.shape
        if self.global_batch is not None:
            pass
        else:
            # one extra tiny collective only when the caller did not promise the shard sizes
            t = torch.tensor([b_local], dtype=torch.int64, device=coords.device)
            dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self.group)
            b_max = int(t.item())
        if counts is not None and b_local > 0 and all(c == b_local for c in counts):
            # fast path: equal shards known a priori (the training loop's case)
            denom = float(world * b_local if self.size_average else world)
            if px is None and coords.is_cuda and coords.dtype == torch.float32 and self.local_fn is _local_cuda_loss:
                px = self._peer_exchange(world, b_local, k, coords.device)
            event = None
            direct = False
            with torch.no_grad():
                if px is not None:
                    # NVLink peer memory: (sum the loss terms,) store to all peers, flag; wait, combine
                    src = terms if terms is not None else loss_sum.detach()
                    if src.dtype != torch.float64:
                        src = src.to(torch.float64)
                    carrier_dtype = scaled.dtype if scaled is not None else loss_sum.dtype
                    if self.defer:
                        # the collect writes the global value straight into the (so far unwritten) scalar
                        # the single-launch function returned: no second autograd node around it
                        into = scaled if (scaled is not None and scaled.dtype == torch.float64) else None
                        value, allc, event = px.run_deferred(src.reshape(-1), 1.0, coords.detach().contiguous(),
                                                             1.0 / denom, carrier_dtype, into=into)
                        direct = into is not None
                    else:
                        value, allc = px.run(src.reshape(-1), 1.0, coords.detach().contiguous(), 1.0 / denom)
                    self.exchange = "peer"
                    self._peer = px
                else:
                    x = self._xch
                    if x is None or (x.world, x.b, x.k) != (world, b_local, k) or x.send.device != coords.device:
                        x = self._xch = _Exchange(world, b_local, k, coords.device)
                    sums, allc = x.run(loss_sum.detach(), coords.detach(), self.group)
                    # mean over replicas of sum / len (balanced_parallel.py:127, loss.py:88-89), summed in
                    # rank order by one deterministic reduction of `world` values
                    value = sums.sum() / denom
                    self.exchange = "nccl"
            self._coords, self._coords64 = None, (allc.reshape(world, b_local, k), coords.dtype)
            self._coords_event = event
            if direct:
                out = scaled.as_subclass(DeferredLoss)       # outside no_grad: stays attached to the graph
                out._hipe_event = event
                if self.size_average and self.global_batch is not None:
                    # the next steps of this shape skip every decision taken above
                    from nets.loss import FreiHandConfig
                    self._fast[(heatmap_out.shape, heatmap_out.dtype)] = (
                        px, int(gt_coord.shape[1] / 3), heatmap_out.shape[-3] // FreiHandConfig.num_joints,
                        heatmap_out.shape[-2], heatmap_out.shape[-1], 1.0 / (b_local * world), 1.0 / denom, world,
                        b_local, k)
                return out
            if scaled is not None:
                carrier = scaled
            else:
                own = loss_sum / b_local if self.size_average else loss_sum
                carrier = own / world
            out = _ValueWithGradOf.apply(carrier, value, event is None)
            if event is not None:
                out = out.as_subclass(DeferredLoss)
                out._hipe_event = event
            return out
        with torch.no_grad():
            recs = gather_records(pack_record(loss_sum.detach(), coords.detach(), b_max), self.group)
            sums, counts, allc = unpack_records(recs, k, counts)
            value, live = combine_replica_losses(sums, counts, self.size_average)
        self.coords = allc.to(coords.dtype)
        if b_local == 0:
            return value
        if scaled is not None:
            return value + (scaled - scaled.detach())
        # value of the global loss, gradient of this replica's share (1/live of its own loss)
        own = loss_sum / b_local if self.size_average else loss_sum
        return value + (own - own.detach()) / live


def sharded_pa_mpjpe(gt_local, pred_local, group=None, mode=0):
    """(PA-MPJPE, MPJPE) over all ranks' samples (data/FreiHand/FreiHand.py:742,766-767,804-807):
    local batched Procrustes, then one all-gather of [sum_pa, sum_raw, n_local*J]."""
    from . import evaluate, ops
    X, Y = evaluate._as_cuda64(gt_local), evaluate._as_cuda64(pred_local)
    o = ops.procrustes(X, Y, mode, True, want=("pa_dist", "raw_dist"))
    rec = torch.stack([ops.sum_f64(o["pa_dist"]), ops.sum_f64(o["raw_dist"]),
                       torch.tensor(float(o["pa_dist"].numel()), dtype=torch.float64, device=X.device)])
    return reduce_error_sums(gather_records(rec, group))


def reduce_error_sums(recs):
    tot = torch.zeros(3, dtype=torch.float64, device=recs[0].device)
    for r in recs:                      # fixed rank order
        tot = tot + r
    return tot[0] / tot[2], tot[1] / tot[2]
