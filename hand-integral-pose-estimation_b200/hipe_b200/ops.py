"""Torch-facing wrappers of the C ABI: tensors in, tensors out, autograd wired up.

PyTorch is used for device memory, streams and autograd plumbing only; every number is
produced by the sm_100a kernels in csrc/ (called through ctypes, which releases the GIL, so
the reference's one-thread-per-GPU criterion, common/nets/balanced_parallel.py:149-173,
really overlaps).  Non-CUDA tensors raise: there is no CPU path.
"""
import torch

from . import _lib
from ._lib import BF16, F32, F64, check

_DT = {torch.float32: F32, torch.bfloat16: BF16}
_TGT = {torch.float32: F32, torch.float64: F64}

# tuning knobs a caller may set (0 = library default); tests and bench.py sweep them
DEFAULT_CLUSTER = 0
DEFAULT_VARIANT = 0


# bench.py's accounting: kernels launched through this module, and (when a dict is installed)
# CUDA-event pairs recorded on the launching stream around each named launch.
LAUNCHES = 0
KERNEL_EVENTS = None


class _launch:
    """Counts one kernel launch; with KERNEL_EVENTS set, brackets it with events on `stream`."""
    __slots__ = ("name", "tensor", "ev")

    def __init__(self, name, tensor):
        self.name, self.tensor = name, tensor

    def __enter__(self):
        global LAUNCHES
        LAUNCHES += 1
        if KERNEL_EVENTS is not None:
            self.ev = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
            self.ev[0].record(torch.cuda.current_stream(self.tensor.device))
        return self

    def __exit__(self, *exc):
        if KERNEL_EVENTS is not None and exc[0] is None:
            self.ev[1].record(torch.cuda.current_stream(self.tensor.device))
            KERNEL_EVENTS.setdefault(self.name, []).append(self.ev)
        return False


class _on:
    """`with _on(dev)` only when dev is not already current (the context manager costs
    ~10 us of host time per launch, a tenth of the step's launch budget)."""

    def __init__(self, dev):
        self.ctx = None if dev.index == torch.cuda.current_device() else torch.cuda.device(dev)

    def __enter__(self):
        if self.ctx is not None:
            self.ctx.__enter__()

    def __exit__(self, *exc):
        if self.ctx is not None:
            return self.ctx.__exit__(*exc)
        return False


def _need_cuda(t, name):
    if not isinstance(t, torch.Tensor):
        raise TypeError("%s must be a torch.Tensor, got %s" % (name, type(t).__name__))
    if not t.is_cuda:
        raise RuntimeError("hipe_b200 has no CPU path: %s lives on %s" % (name, t.device))


def _logits_code(t):
    try:
        return _DT[t.dtype]
    except KeyError:
        raise TypeError("heatmaps must be float32 or bfloat16, got %s" % t.dtype) from None


def _stream(t):
    return torch.cuda.current_stream(t.device).cuda_stream


def _ptr(t):
    return None if t is None else t.data_ptr()


def softargmax_forward(logits, n_vol, D, H, W, apply_softmax=True, cluster=None, variant=None):
    """-> coords (n_vol,3) fp32 [x/W-.5, y/H-.5, z/D-.5], expect (n_vol,3) fp64 voxel units,
    stats (n_vol,2) fp32 (max, sum exp)."""
    _need_cuda(logits, "heatmaps")
    code = _logits_code(logits)
    if not logits.is_contiguous():
        logits = logits.contiguous()
    if logits.numel() != n_vol * D * H * W:
        raise ValueError("heatmaps have %d elements, expected %d x %dx%dx%d" % (logits.numel(), n_vol, D, H, W))
    dev = logits.device
    coords = torch.empty((n_vol, 3), dtype=torch.float32, device=dev)
    expect = torch.empty((n_vol, 3), dtype=torch.float64, device=dev)
    stats = torch.empty((n_vol, 2), dtype=torch.float32, device=dev)
    with _on(dev), _launch("softargmax_fwd", logits):
        check(_lib.load().hipe_softargmax_fwd(
            logits.data_ptr(), code, n_vol, D, H, W, 1 if apply_softmax else 0,
            coords.data_ptr(), expect.data_ptr(), stats.data_ptr(),
            DEFAULT_CLUSTER if cluster is None else cluster,
            DEFAULT_VARIANT if variant is None else variant, _stream(logits)))
    return coords, expect, stats


def softargmax_backward(logits, n_vol, D, H, W, stats, expect, gcoords, gscale=None, cluster=None,
                        variant=None):
    """-> dlogits, same shape/dtype as `logits`; gcoords is (n_vol,3) fp32 w.r.t. coords."""
    _need_cuda(logits, "heatmaps")
    code = _logits_code(logits)
    if not logits.is_contiguous():
        logits = logits.contiguous()
    gcoords = gcoords.to(device=logits.device, dtype=torch.float32).contiguous()
    expect = expect.to(device=logits.device, dtype=torch.float64).contiguous()
    if gscale is not None:
        gscale = gscale.to(device=logits.device, dtype=torch.float32).reshape(1).contiguous()
    out = torch.empty_like(logits)
    with _on(logits.device), _launch("softargmax_bwd", logits):
        check(_lib.load().hipe_softargmax_bwd(
            logits.data_ptr(), code, n_vol, D, H, W, stats.data_ptr(), expect.data_ptr(),
            gcoords.data_ptr(), _ptr(gscale), out.data_ptr(),
            DEFAULT_CLUSTER if cluster is None else cluster,
            DEFAULT_VARIANT if variant is None else variant, _stream(logits)))
    return out


def volume_softmax(logits, n_vol, D, H, W, stats=None, cluster=None, variant=None):
    """-> probabilities with the shape/dtype of `logits` (softmax over each D*H*W volume)."""
    _need_cuda(logits, "heatmaps")
    code = _logits_code(logits)
    if not logits.is_contiguous():
        logits = logits.contiguous()
    if stats is None:
        _, _, stats = softargmax_forward(logits, n_vol, D, H, W, True, cluster, variant)
    out = torch.empty_like(logits)
    with _on(logits.device), _launch("volume_softmax", logits):
        check(_lib.load().hipe_volume_softmax(
            logits.data_ptr(), code, n_vol, D, H, W, stats.data_ptr(), out.data_ptr(),
            DEFAULT_CLUSTER if cluster is None else cluster,
            DEFAULT_VARIANT if variant is None else variant, _stream(logits)))
    return out


def volume_entropy(logits, n_vol, D, H, W, stats=None):
    """-> (n_vol,) fp64 entropies (nats) of the per-volume softmax distributions, without the
    probabilities ever existing (hipe_volume_entropy)."""
    _need_cuda(logits, "heatmaps")
    code = _logits_code(logits)
    if not logits.is_contiguous():
        logits = logits.contiguous()
    if stats is None:
        _, _, stats = softargmax_forward(logits, n_vol, D, H, W, True)
    out = torch.empty((n_vol,), dtype=torch.float64, device=logits.device)
    with _on(logits.device), _launch("volume_entropy", logits):
        check(_lib.load().hipe_volume_entropy(logits.data_ptr(), code, n_vol, D, H, W, stats.data_ptr(),
                                              out.data_ptr(), _stream(logits)))
    return out


def _targets(coords, gt, vis):
    if (gt.dtype is vis.dtype and gt.dtype in _TGT and gt.device == coords.device and vis.device == coords.device
            and gt.is_contiguous() and vis.is_contiguous()):
        return gt, vis, _TGT[gt.dtype], gt.dtype          # the training loop's case: nothing to convert
    tdt = torch.promote_types(gt.dtype, vis.dtype)
    if tdt not in _TGT:
        tdt = torch.float64 if tdt == torch.float64 else torch.float32
    gt = gt.to(device=coords.device, dtype=tdt).contiguous()
    vis = vis.to(device=coords.device, dtype=tdt).contiguous()
    return gt, vis, _TGT[tdt], tdt


def l1_loss(coords, gt, vis, norm, want_grad=True):
    """sum(|coords-gt|*vis)*norm as a 0-dim fp64 tensor, and sign(coords-gt)*vis*norm (fp32)."""
    _need_cuda(coords, "coords")
    coords = coords.to(torch.float32).contiguous()
    gt, vis, code, _ = _targets(coords, gt, vis)
    if gt.shape != coords.shape or vis.shape != coords.shape:
        raise ValueError("coords %s, gt %s, vis %s must have the same shape"
                         % (tuple(coords.shape), tuple(gt.shape), tuple(vis.shape)))
    B, K = coords.shape
    loss = torch.empty((), dtype=torch.float64, device=coords.device)
    g = torch.empty_like(coords) if want_grad else None
    with _on(coords.device), _launch("l1_loss", coords):
        check(_lib.load().hipe_l1_loss_fwd(coords.data_ptr(), gt.data_ptr(), vis.data_ptr(), code, B, K,
                                           float(norm), loss.data_ptr(), _ptr(g), _stream(coords)))
    return loss, g


def sum_f64(x, scale=1.0):
    _need_cuda(x, "x")
    x = x.to(torch.float64).contiguous()
    out = torch.empty((), dtype=torch.float64, device=x.device)
    with _on(x.device), _launch("sum_f64", x):
        check(_lib.load().hipe_sum_f64(x.data_ptr(), x.numel(), float(scale), out.data_ptr(), _stream(x)))
    return out


# "auto": the launch measured fastest for the geometry (fused_plan); "resident" / "stream" force one
FUSED_MODE = "auto"          # "auto" | "cluster" | "resident" | "stream"


_PLAN_CACHE = {}


def fused_plan(logits, n_vol, D, H, W):
    key = (logits.dtype, n_vol, D, H, W)
    if key not in _PLAN_CACHE:
        _PLAN_CACHE[key] = _fused_plan(logits, n_vol, D, H, W)
    return _PLAN_CACHE[key]


def _fused_plan(logits, n_vol, D, H, W):
    """Which launch computes loss + gradient fastest for this geometry, from same-box B200 A/Bs
    (profiles/r2_configs.md):
      "cluster"  the cluster launch (hipe_joint_loss_cluster: one HBM read + one write per volume) for
                 the geometries its loops are specialised for.  Resident form (the volume stays in the
                 cluster's shared memory, nothing in L2) -- 4096- or 3136-byte visits with every stage
                 starting at y = 0: bf16 64^3 x 10752 2.09 ms vs 2.40 stream, x 1344 0.273 vs 0.331;
                 fp32 56^3 x 672 0.177 vs 0.202, x 2016 0.530 vs 0.576; fp32 32^3 x 5376 0.244 vs 0.282;
                 bf16 32^3 (64 KiB volumes) x 5376 0.162 vs 0.224 stream / 0.248 pair.  Streamed form (two
                 rings, second read through L2 a bounded lead behind): fp32 64^3 x 1344 0.473-0.495 vs
                 0.500-0.518 resident form / 0.540 resident launch / 0.556-0.572 stream, x 672 0.257 vs
                 0.269 / 0.301; fp32 96^3 (fits no cluster; wrapped 12 KiB stages on 16 CTAs) x 672
                 0.81-0.85 vs 0.99-1.02 stream;
      "stream"   one launch whose second read goes through L2 unbounded (hipe_joint_loss_fused) for
                 every other fp32 / bf16 volume >= 128 KiB (fp32 48^3 x 2016: 0.362 vs 0.42 cluster; the
                 general-geometry streamed cluster form is untuned);
      None       the two-launch forward + recompute backward, which still wins on tiny volumes (fp32
                 16^3: 0.366 ms vs 0.433 fused; bf16 32^3 is a tie).
    "resident" (hipe_joint_loss_resident, the L2-window launch that was the fp32 64^3 default before
    the cluster launch: 0.540 ms) stays reachable through FUSED_MODE / mode="resident"."""
    esz = logits.element_size()
    vol_bytes = D * H * W * esz
    if logits.dtype not in (torch.float32, torch.bfloat16):
        return None
    if vol_bytes >= (1 << 16) and n_vol >= 8:
        cp = cluster_plan(logits, n_vol, D, H, W)
        if cp is not None:
            stage_bytes, lead, row_bytes = cp[3], cp[4], W * esz
            sr = stage_bytes // row_bytes                       # rows per stage
            # the specialised loops; 2-byte 56^3 (12.25 KB stages on 2 CTAs) is the one case among them
            # where the stream launch is still ahead (bf16 56^3 x 2016: 0.328 ms vs 0.366)
            if sr % H == 0 and (stage_bytes == 16384 or (stage_bytes == 12544 and esz == 4)) and \
                    (lead == 0 or (esz == 4 and lead * stage_bytes * 120 <= 46e6)):      # streamed: the L2 must hold it
                return "cluster"
            # wrapped visits (fp32 rows of 24 vectors): 3 x 4096-byte visits = 32 whole rows
            if lead > 0 and esz == 4 and stage_bytes == 12288 and row_bytes == 384:
                return "cluster"
    if (W * esz) % 16 == 0 and vol_bytes >= (1 << 17):
        return "stream"
    return None


_WORKSPACES = {}             # (device index, stream) -> uint8 scratch tensor, grown on demand
_NEED = {}                   # (kind, dtype code, n_vol, D, H, W) -> workspace bytes (0 = unsupported)


def _need_bytes(kind, code, n_vol, D, H, W):
    key = (kind, code, n_vol, D, H, W)
    v = _NEED.get(key)
    if v is None:
        lib = _lib.load()
        fn = lib.hipe_resident_workspace_bytes if kind == "resident" else lib.hipe_cluster_workspace_bytes
        v = _NEED[key] = int(fn(code, n_vol, D, H, W))
    return v


def resident_supported(logits, n_vol, D, H, W):
    return _need_bytes("resident", _logits_code(logits), n_vol, D, H, W) > 0


def _workspace(dev, stream_handle, nbytes):
    key = (dev.index, stream_handle)
    ws = _WORKSPACES.get(key)
    if ws is None or ws.numel() < nbytes:
        ws = torch.empty((nbytes + 255) // 256 * 256, dtype=torch.uint8, device=dev)
        _WORKSPACES[key] = ws
    return ws


def resident_timed_out(device, n_vol):
    """True if an inter-CTA wait of the last resident launch on this device/stream hit its spin
    limit (word [n_vol] of the workspace; its outputs are NaN then).  Synchronises; for tests and
    diagnostics."""
    dev = torch.device(device)
    ws = _WORKSPACES.get((dev.index, torch.cuda.current_stream(dev).cuda_stream))
    if ws is None:
        return False
    return bool(ws[4 * n_vol:4 * n_vol + 4].view(torch.int32).item() == 1)


def cluster_plan(logits, n_vol, D, H, W):
    """(cluster size, ring slots per CTA, stages per CTA share, bytes per stage, lead) of the cluster
    launch for this geometry, or None when it cannot take it.  lead = 0: the volume stays resident in
    the cluster's shared memory; lead > 0: streamed shares, the second read comes through L2 at most
    that many stages behind the first."""
    import ctypes
    c, ns, sh, sb, ld = (ctypes.c_int() for _ in range(5))
    rc = _lib.load().hipe_cluster_plan(_logits_code(logits), n_vol, D, H, W, ctypes.byref(c), ctypes.byref(ns),
                                       ctypes.byref(sh), ctypes.byref(sb), ctypes.byref(ld))
    return None if rc != 0 else (c.value, ns.value, sh.value, sb.value, ld.value)


def cluster_supported(logits, n_vol, D, H, W):
    return _lib.load().hipe_cluster_workspace_bytes(_logits_code(logits), n_vol, D, H, W) > 0


def _single_launch(kind, logits, n_vol, D, H, W, gt, vis, norm, loss_out, want_aux):
    """hipe_joint_loss_cluster / hipe_joint_loss_resident.  want_aux=False leaves `expect` and `stats`
    unallocated (the kernels take NULL): the autograd path needs neither."""
    _need_cuda(logits, "heatmaps")
    code = _logits_code(logits)
    if not logits.is_contiguous():
        logits = logits.contiguous()
    dev = logits.device
    lib = _lib.load()
    need = _need_bytes(kind, code, n_vol, D, H, W)
    if need <= 0:
        raise _lib.HipeError("libhipe_b200: error %d: geometry not supported by the %s launch"
                             % (_lib.E_UNSUPPORTED, kind))
    coords = torch.empty((n_vol, 3), dtype=torch.float32, device=dev)
    gt, vis, tcode, _ = _targets(coords, gt, vis)
    if gt.numel() != n_vol * 3 or vis.numel() != n_vol * 3:
        raise ValueError("targets must hold 3 values per volume")
    expect = torch.empty((n_vol, 3), dtype=torch.float64, device=dev) if want_aux else None
    stats = torch.empty((n_vol, 2), dtype=torch.float32, device=dev) if want_aux else None
    terms = torch.empty((n_vol * 3,), dtype=torch.float64, device=dev)
    dlogits = torch.empty_like(logits)
    fn = lib.hipe_joint_loss_cluster if kind == "cluster" else lib.hipe_joint_loss_resident
    with _on(dev):
        st = torch.cuda.current_stream(dev).cuda_stream
        ws = _workspace(dev, st, need)
        with _launch("joint_loss_" + kind, logits):
            check(fn(logits.data_ptr(), code, n_vol, D, H, W, gt.data_ptr(), vis.data_ptr(), tcode, float(norm),
                     coords.data_ptr(), _ptr(expect), _ptr(stats), terms.data_ptr(), _ptr(loss_out),
                     dlogits.data_ptr(), ws.data_ptr(), ws.numel(), st))
    return coords, expect, stats, terms, dlogits


def joint_loss_cluster(logits, n_vol, D, H, W, gt, vis, norm, loss_out=None, want_aux=True):
    """hipe_joint_loss_cluster: same outputs as joint_loss_fused; each volume lives in the shared
    memory of one thread-block cluster between the reduction and the gradient write.
    `loss_out`: optional 0-dim fp64 CUDA tensor that receives norm * sum(loss terms) from the same
    launch (the last cluster to finish sums them)."""
    return _single_launch("cluster", logits, n_vol, D, H, W, gt, vis, norm, loss_out, want_aux)


def joint_loss_resident(logits, n_vol, D, H, W, gt, vis, norm, loss_out=None, want_aux=True):
    """hipe_joint_loss_resident: same outputs as joint_loss_fused, each volume read from HBM once
    (second read served by the L2 inside a bounded window).  `loss_out` as for joint_loss_cluster."""
    return _single_launch("resident", logits, n_vol, D, H, W, gt, vis, norm, loss_out, want_aux)


def joint_loss_fused(logits, n_vol, D, H, W, gt, vis, norm, cluster=None, variant=None, mode=None, loss_out=None,
                     want_aux=True):
    """One launch: coords, expect, stats, per-coordinate loss terms (n_vol*3 fp64), dlogits.
    `loss_out` (0-dim fp64 CUDA tensor, optional) is filled with the loss scalar: by the cluster /
    resident launch itself, else by hipe_sum_f64.  mode: None = FUSED_MODE; "auto" = fused_plan's
    choice; "cluster" | "resident" | "stream" force a launch."""
    _need_cuda(logits, "heatmaps")
    mode = FUSED_MODE if mode is None else mode
    plan = fused_plan(logits, n_vol, D, H, W) if (mode == "auto" and cluster is None and variant is None) else None
    if mode == "cluster" or plan == "cluster":
        try:
            return joint_loss_cluster(logits, n_vol, D, H, W, gt, vis, norm, loss_out, want_aux)
        except _lib.HipeError as e:
            # the plan is made from the geometry alone; a device that cannot co-schedule the cluster shape
            # (a partition with fewer SMs per GPC than the cluster) answers UNSUPPORTED at launch time:
            # the streaming launch below takes every geometry
            if mode == "cluster" or "error %d:" % _lib.E_UNSUPPORTED not in str(e):
                raise
    if mode == "resident" or plan == "resident":
        try:
            return joint_loss_resident(logits, n_vol, D, H, W, gt, vis, norm, loss_out, want_aux)
        except _lib.HipeError as e:
            if mode == "resident" or "error %d:" % _lib.E_UNSUPPORTED not in str(e):
                raise
    code = _logits_code(logits)
    if not logits.is_contiguous():
        logits = logits.contiguous()
    dev = logits.device
    coords = torch.empty((n_vol, 3), dtype=torch.float32, device=dev)
    gt, vis, tcode, _ = _targets(coords, gt, vis)
    if gt.numel() != n_vol * 3 or vis.numel() != n_vol * 3:
        raise ValueError("targets must hold 3 values per volume")
    expect = torch.empty((n_vol, 3), dtype=torch.float64, device=dev) if want_aux else None
    stats = torch.empty((n_vol, 2), dtype=torch.float32, device=dev) if want_aux else None
    terms = torch.empty((n_vol * 3,), dtype=torch.float64, device=dev)
    dlogits = torch.empty_like(logits)
    with _on(dev):
        st = torch.cuda.current_stream(dev).cuda_stream
        with _launch("joint_loss_fused", logits):
            check(_lib.load().hipe_joint_loss_fused(
                logits.data_ptr(), code, n_vol, D, H, W, gt.data_ptr(), vis.data_ptr(), tcode, float(norm),
                coords.data_ptr(), _ptr(expect), _ptr(stats), terms.data_ptr(), dlogits.data_ptr(),
                DEFAULT_CLUSTER if cluster is None else cluster,
                DEFAULT_VARIANT if variant is None else variant, st))
        if loss_out is not None:
            with _launch("sum_f64", terms):
                check(_lib.load().hipe_sum_f64(terms.data_ptr(), terms.numel(), float(norm), loss_out.data_ptr(), st))
    return coords, expect, stats, terms, dlogits


def procrustes(X, Y, mode=0, optimal_scale=True, want=("Z", "pa_dist", "raw_dist")):
    """Batched alignment of Y onto X, both (N, J, 3) fp64 CUDA tensors.  Returns a dict with the
    requested outputs among d (N,), Z (N,J,3), T (N,3,3), b (N,), c (N,3), pa_dist (N,J),
    raw_dist (N,J)."""
    _need_cuda(X, "X")
    _need_cuda(Y, "Y")
    X = X.to(torch.float64).contiguous()
    Y = Y.to(device=X.device, dtype=torch.float64).contiguous()
    if X.shape != Y.shape or X.dim() != 3 or X.shape[2] != 3:
        raise ValueError("X and Y must both be (N, J, 3); got %s and %s" % (tuple(X.shape), tuple(Y.shape)))
    N, J, _ = X.shape
    shapes = {"d": (N,), "Z": (N, J, 3), "T": (N, 3, 3), "b": (N,), "c": (N, 3), "pa_dist": (N, J),
              "raw_dist": (N, J)}
    out = {k: torch.empty(shapes[k], dtype=torch.float64, device=X.device) for k in want}
    with _on(X.device), _launch("procrustes", X):
        check(_lib.load().hipe_procrustes_batched(
            X.data_ptr(), Y.data_ptr(), N, J, int(mode), 1 if optimal_scale else 0,
            _ptr(out.get("d")), _ptr(out.get("Z")), _ptr(out.get("T")), _ptr(out.get("b")),
            _ptr(out.get("c")), _ptr(out.get("pa_dist")), _ptr(out.get("raw_dist")), _stream(X)))
    return out


def keypoint_measures(dist, val_min, val_max, steps, vis=None):
    """hipe_keypoint_measures on a CUDA (N, K) distance matrix; see evaluate.keypoint_measures."""
    _need_cuda(dist, "dist")
    d = dist.to(torch.float64).contiguous()
    if d.dim() != 2:
        raise ValueError("dist must be (N, K)")
    N, K = d.shape
    v = None
    if vis is not None:
        if not isinstance(vis, torch.Tensor):
            import numpy as np
            vis = torch.as_tensor(np.ascontiguousarray(vis))
        v = (vis.to(d.device) != 0).to(torch.uint8).contiguous()
        if tuple(v.shape) != (N, K):
            raise ValueError("vis must be (N, K)")
    lib = _lib.load()
    steps = int(steps)
    need = lib.hipe_keypoint_measures_workspace_bytes(K, steps)
    ws = torch.empty((max(need, 8) // 8,), dtype=torch.float64, device=d.device)
    summary = torch.empty((3,), dtype=torch.float64, device=d.device)
    curve = torch.empty((max(steps, 0),), dtype=torch.float64, device=d.device)
    with _on(d.device), _launch("keypoint_measures", d):
        check(lib.hipe_keypoint_measures(d.data_ptr(), _ptr(v), N, K, float(val_min), float(val_max), steps,
                                         summary.data_ptr(), curve.data_ptr(), ws.data_ptr(), ws.numel() * 8,
                                         _stream(d)))
    per = {"count": ws[0:K], "mean": ws[K:2 * K], "median": ws[2 * K:3 * K], "auc": ws[3 * K:4 * K],
           "pck": ws[4 * K:4 * K + K * steps].view(K, steps)}
    return {"summary": summary, "pck_curve": curve, "per_keypoint": per}


def _f64(t, dev, shape=None):
    if t is None:
        return None
    if not isinstance(t, torch.Tensor):
        import numpy as np
        t = torch.as_tensor(np.ascontiguousarray(t, dtype=np.float64))
    t = t.to(device=dev, dtype=torch.float64).contiguous()
    if shape is not None and tuple(t.shape) != tuple(shape):
        raise ValueError("expected shape %s, got %s" % (tuple(shape), tuple(t.shape)))
    return t


def patch_to_camera(coords, N, J, tprime, K, trans=None, bbox=None, R=None, bone_len=None, bone=(9, 10),
                    patch=(224.0, 224.0), dst=(224.0, 224.0), box_scale=1.0, gt=None, mode=0, optimal_scale=True,
                    want=("cam",)):
    """The coordinate chain of data/FreiHand/FreiHand.py:683-737 / :853-883 on the device, optionally
    fused with the alignment against `gt` (hipe_eval_chain_procrustes).
      coords : CUDA float32 (N, 3J) soft-argmax output, or float64 (N, J, 3|4) patch pixels
      want   : any of "img", "cam" (chain only) or "cam", "Z", "pa_dist", "raw_dist" (with gt)
    Returns a dict of CUDA fp64 tensors."""
    _need_cuda(coords, "coords")
    dev = coords.device
    if coords.dtype == torch.float32 and coords.dim() == 2:
        kind, stride = 0, 3
        if tuple(coords.shape) != (N, 3 * J):
            raise ValueError("soft-argmax coordinates must be (N, 3J)")
        coords = coords.contiguous()
    else:
        coords = coords.to(torch.float64).contiguous()
        if coords.dim() != 3 or coords.shape[0] != N or coords.shape[1] != J or coords.shape[2] < 3:
            raise ValueError("patch coordinates must be (N, J, >=3)")
        kind, stride = 1, int(coords.shape[2])
    if trans is None and bbox is None:
        raise ValueError("one of trans (N,3,3) or bbox (N,4) is required")
    tprime, K = _f64(tprime, dev, (N,)), _f64(K, dev, (N, 3, 3))
    trans, bbox = _f64(trans, dev, (N, 3, 3)), _f64(bbox, dev, (N, 4))
    R, bone_len = _f64(R, dev, (N, 3, 3)), _f64(bone_len, dev, (N,))
    shapes = {"img": (N, J, 3), "cam": (N, J, 3), "Z": (N, J, 3), "pa_dist": (N, J), "raw_dist": (N, J)}
    out = {k: torch.empty(shapes[k], dtype=torch.float64, device=dev) for k in want}
    lib = _lib.load()
    common = (coords.data_ptr(), kind, stride, N, J, float(patch[0]), float(patch[1]), _ptr(trans), _ptr(bbox),
              float(box_scale), float(dst[0]), float(dst[1]), tprime.data_ptr(), K.data_ptr(), _ptr(R),
              _ptr(bone_len), int(bone[0]), int(bone[1]))
    with _on(dev):
        if gt is None:
            if set(want) - {"img", "cam"}:
                raise ValueError("Z / pa_dist / raw_dist need gt")
            with _launch("patch_to_camera", coords):
                check(lib.hipe_patch_to_camera(*common, _ptr(out.get("img")), _ptr(out.get("cam")), _stream(coords)))
        else:
            if "img" in want:
                raise ValueError("img is only produced by the chain-only launch")
            gt = _f64(gt, dev, (N, J, 3))
            with _launch("eval_chain_procrustes", coords):
                check(lib.hipe_eval_chain_procrustes(
                    *common, gt.data_ptr(), int(mode), 1 if optimal_scale else 0, _ptr(out.get("cam")),
                    _ptr(out.get("Z")), _ptr(out.get("pa_dist")), _ptr(out.get("raw_dist")), _stream(coords)))
    return out


# ----------------------------------------------------------------------------------------
# autograd
# ----------------------------------------------------------------------------------------
class SoftArgmaxFunction(torch.autograd.Function):
    """softmax_integral_tensor (common/nets/loss.py:46-59) as one differentiable op.  Saves the
    logits and 5 floats per volume -- never the probabilities."""

    @staticmethod
    def forward(ctx, preds, num_joints, hm_width, hm_height, hm_depth):
        B = preds.shape[0]
        n_vol = B * num_joints
        logits = preds.contiguous()
        coords, expect, stats = softargmax_forward(logits, n_vol, hm_depth, hm_height, hm_width)
        ctx.save_for_backward(logits, expect, stats)
        ctx.geom = (n_vol, hm_depth, hm_height, hm_width)
        ctx.in_shape = preds.shape
        return coords.view(B, num_joints * 3)

    @staticmethod
    def backward(ctx, grad_coords):
        logits, expect, stats = ctx.saved_tensors
        n_vol, D, H, W = ctx.geom
        g = grad_coords.reshape(n_vol, 3)
        d = softargmax_backward(logits, n_vol, D, H, W, stats, expect, g)
        return d.view(ctx.in_shape), None, None, None, None


class IntegralL1Function(torch.autograd.Function):
    """sum(|coords - gt| * vis) * norm (common/nets/loss.py:86-89) with the fp64 accumulation
    and result dtype torch's promotion gives the reference."""

    @staticmethod
    def forward(ctx, coords, gt, vis, norm):
        loss, g = l1_loss(coords, gt, vis, norm, want_grad=True)
        ctx.save_for_backward(g)
        ctx.mark_non_differentiable()
        out_dtype = torch.promote_types(coords.dtype, torch.promote_types(gt.dtype, vis.dtype))
        return loss.to(out_dtype)

    @staticmethod
    def backward(ctx, grad_loss):
        (g,) = ctx.saved_tensors
        return g * grad_loss.to(g.dtype), None, None, None


class FusedJointLossFunction(torch.autograd.Function):
    """JointLocationLoss forward with the gradient produced in the same launch
    (hipe_joint_loss_cluster / _resident / _fused): backward only scales the stored dlogits when the
    upstream gradient is not 1, and hands the buffer over.  The stored gradient is CONSUMED by that
    hand-over (scaled in place, possibly adopted by autograd as `.grad`): a second backward through
    the same forward (retain_graph=True) raises instead of silently compounding the scale -- use
    JointLocationLoss(fused=False) for graphs that are back-propagated more than once."""

    @staticmethod
    def forward(ctx, heatmap_out, gt, vis, num_joints, D, H, W, norm, reduce_loss=True, loss_cls=None):
        """-> (loss, coords (B, 3J), loss_terms (B*3J,) fp64).  reduce_loss=False leaves the scalar
        unwritten (a carrier for the gradient only): the sharded path sums the terms inside its
        exchange kernel instead of a separate launch."""
        B = heatmap_out.shape[0]
        n_vol = B * num_joints
        out_dtype = torch.promote_types(torch.float32, torch.promote_types(gt.dtype, vis.dtype))
        loss64 = torch.empty((), dtype=torch.float64, device=heatmap_out.device)
        coords, _, _, terms, dlogits = joint_loss_fused(heatmap_out, n_vol, D, H, W, gt, vis, norm,
                                                        loss_out=loss64 if reduce_loss else None, want_aux=False)
        loss = loss64 if out_dtype == torch.float64 else loss64.to(out_dtype)
        if loss_cls is not None:                 # e.g. parallel.DeferredLoss: the value arrives later, on a side stream
            loss = loss.as_subclass(loss_cls)
        ctx.dlogits = dlogits                    # not save_for_backward: it is mutated, once, by backward
        ctx.in_shape = heatmap_out.shape
        coords = coords.view(B, num_joints * 3)
        ctx.mark_non_differentiable(coords, terms)
        return loss, coords, terms

    @staticmethod
    def backward(ctx, grad_loss, _grad_coords, _grad_terms):
        dlogits = ctx.dlogits
        if dlogits is None:
            raise RuntimeError(
                "hipe_b200: second backward through the single-launch JointLocationLoss: its heatmap "
                "gradient was stored by the forward launch and handed over (scaled in place) by the first "
                "backward.  Build the criterion with fused=False for retain_graph / repeated backward.")
        ctx.dlogits = None
        # in place, on device, no host sync; the kernel returns at once when the scalar is 1
        gs = grad_loss
        if gs.device != dlogits.device or gs.dtype not in (torch.float32, torch.float64) or not gs.is_contiguous():
            gs = gs.to(device=dlogits.device, dtype=torch.float32).reshape(1).contiguous()
        with _on(dlogits.device), _launch("scale_inplace", dlogits):
            check(_lib.load().hipe_scale_inplace_s(dlogits.data_ptr(), _logits_code(dlogits), dlogits.numel(),
                                                   gs.data_ptr(), _lib.F32 if gs.dtype == torch.float32 else _lib.F64,
                                                   _stream(dlogits)))
        return dlogits.view(ctx.in_shape), None, None, None, None, None, None, None, None, None


class Orthonormalize3Function(torch.autograd.Function):
    """make_orthonormal (procrustes_encoding/nrsfm/nrsfm_modules.py:46-67) for (..., 3, 3) CUDA
    matrices, fp32 or fp64: U diag(1, 1, det(U V^T)) V^T with its exact derivative, one thread per
    matrix (hipe_orthonormalize3_fwd / _bwd).  Saves 21 doubles per matrix."""

    @staticmethod
    def forward(ctx, mats):
        _need_cuda(mats, "input_mat")
        if mats.shape[-2:] != (3, 3):
            raise ValueError("expected (..., 3, 3) matrices, got %s" % (tuple(mats.shape),))
        if mats.dtype not in (torch.float32, torch.float64):
            raise TypeError("orthonormalize3 takes float32 or float64 matrices")
        a = mats.contiguous()
        n = a.numel() // 9
        code = _lib.F32 if a.dtype == torch.float32 else _lib.F64
        out = torch.empty_like(a)
        aux = torch.empty((n, 21), dtype=torch.float64, device=a.device)
        with _on(a.device), _launch("orthonormalize3_fwd", a):
            check(_lib.load().hipe_orthonormalize3_fwd(a.data_ptr(), code, n, out.data_ptr(), aux.data_ptr(), _stream(a)))
        ctx.save_for_backward(aux)
        ctx.code = code
        return out

    @staticmethod
    def backward(ctx, grad):
        (aux,) = ctx.saved_tensors
        g = grad.contiguous()
        n = g.numel() // 9
        d = torch.empty_like(g)
        with _on(g.device), _launch("orthonormalize3_bwd", g):
            check(_lib.load().hipe_orthonormalize3_bwd(g.data_ptr(), ctx.code, n, aux.data_ptr(), d.data_ptr(), _stream(g)))
        return d
