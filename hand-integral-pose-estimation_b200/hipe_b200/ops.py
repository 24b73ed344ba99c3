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
    with torch.cuda.device(dev), _launch("softargmax_fwd", logits):
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
    with torch.cuda.device(logits.device), _launch("softargmax_bwd", logits):
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
    with torch.cuda.device(logits.device), _launch("volume_softmax", logits):
        check(_lib.load().hipe_volume_softmax(
            logits.data_ptr(), code, n_vol, D, H, W, stats.data_ptr(), out.data_ptr(),
            DEFAULT_CLUSTER if cluster is None else cluster,
            DEFAULT_VARIANT if variant is None else variant, _stream(logits)))
    return out


def _targets(coords, gt, vis):
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
    with torch.cuda.device(coords.device), _launch("l1_loss", coords):
        check(_lib.load().hipe_l1_loss_fwd(coords.data_ptr(), gt.data_ptr(), vis.data_ptr(), code, B, K,
                                           float(norm), loss.data_ptr(), _ptr(g), _stream(coords)))
    return loss, g


def sum_f64(x, scale=1.0):
    _need_cuda(x, "x")
    x = x.to(torch.float64).contiguous()
    out = torch.empty((), dtype=torch.float64, device=x.device)
    with torch.cuda.device(x.device), _launch("sum_f64", x):
        check(_lib.load().hipe_sum_f64(x.data_ptr(), x.numel(), float(scale), out.data_ptr(), _stream(x)))
    return out


# "auto": the persistent on-chip launch whenever the geometry allows it, else the L2 re-read launch
FUSED_MODE = "auto"          # "auto" | "resident" | "stream"
_WORKSPACES = {}             # (device index, stream) -> uint8 scratch tensor, grown on demand


def resident_supported(logits, n_vol, D, H, W):
    return _lib.load().hipe_resident_workspace_bytes(_logits_code(logits), n_vol, D, H, W) > 0


def _workspace(dev, nbytes):
    key = (dev.index, torch.cuda.current_stream(dev).cuda_stream)
    ws = _WORKSPACES.get(key)
    if ws is None or ws.numel() < nbytes:
        ws = torch.empty((nbytes + 255) // 256 * 256, dtype=torch.uint8, device=dev)
        _WORKSPACES[key] = ws
    return ws


def resident_timed_out(device, n_vol):
    """True if an inter-CTA wait of the last resident launch on this device/stream hit its spin
    limit (word [n_vol] of the workspace).  Synchronises; for tests and diagnostics."""
    dev = torch.device(device)
    ws = _WORKSPACES.get((dev.index, torch.cuda.current_stream(dev).cuda_stream))
    if ws is None:
        return False
    return bool(ws[4 * n_vol:4 * n_vol + 4].view(torch.int32).item() != 0)


def joint_loss_resident(logits, n_vol, D, H, W, gt, vis, norm):
    """hipe_joint_loss_resident: same outputs as joint_loss_fused, each volume read from HBM once."""
    _need_cuda(logits, "heatmaps")
    code = _logits_code(logits)
    if not logits.is_contiguous():
        logits = logits.contiguous()
    dev = logits.device
    lib = _lib.load()
    need = lib.hipe_resident_workspace_bytes(code, n_vol, D, H, W)
    if need <= 0:
        raise _lib.HipeError("libhipe_b200: error %d: geometry not supported by the resident launch"
                             % _lib.E_UNSUPPORTED)
    coords = torch.empty((n_vol, 3), dtype=torch.float32, device=dev)
    gt, vis, tcode, _ = _targets(coords, gt, vis)
    if gt.numel() != n_vol * 3 or vis.numel() != n_vol * 3:
        raise ValueError("targets must hold 3 values per volume")
    expect = torch.empty((n_vol, 3), dtype=torch.float64, device=dev)
    stats = torch.empty((n_vol, 2), dtype=torch.float32, device=dev)
    terms = torch.empty((n_vol * 3,), dtype=torch.float64, device=dev)
    dlogits = torch.empty_like(logits)
    with torch.cuda.device(dev):
        ws = _workspace(dev, need)
        with _launch("joint_loss_resident", logits):
            check(lib.hipe_joint_loss_resident(
                logits.data_ptr(), code, n_vol, D, H, W, gt.data_ptr(), vis.data_ptr(), tcode, float(norm),
                coords.data_ptr(), expect.data_ptr(), stats.data_ptr(), terms.data_ptr(), dlogits.data_ptr(),
                ws.data_ptr(), ws.numel(), _stream(logits)))
    return coords, expect, stats, terms, dlogits


def joint_loss_fused(logits, n_vol, D, H, W, gt, vis, norm, cluster=None, variant=None, mode=None):
    """One launch: coords, expect, stats, per-coordinate loss terms (n_vol*3 fp64), dlogits."""
    _need_cuda(logits, "heatmaps")
    mode = FUSED_MODE if mode is None else mode
    if mode == "resident" or (mode == "auto" and cluster is None and variant is None
                              and resident_supported(logits, n_vol, D, H, W)):
        try:
            return joint_loss_resident(logits, n_vol, D, H, W, gt, vis, norm)
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
    expect = torch.empty((n_vol, 3), dtype=torch.float64, device=dev)
    stats = torch.empty((n_vol, 2), dtype=torch.float32, device=dev)
    terms = torch.empty((n_vol * 3,), dtype=torch.float64, device=dev)
    dlogits = torch.empty_like(logits)
    with torch.cuda.device(dev), _launch("joint_loss_fused", logits):
        check(_lib.load().hipe_joint_loss_fused(
            logits.data_ptr(), code, n_vol, D, H, W, gt.data_ptr(), vis.data_ptr(), tcode, float(norm),
            coords.data_ptr(), expect.data_ptr(), stats.data_ptr(), terms.data_ptr(), dlogits.data_ptr(),
            DEFAULT_CLUSTER if cluster is None else cluster,
            DEFAULT_VARIANT if variant is None else variant, _stream(logits)))
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
    with torch.cuda.device(X.device), _launch("procrustes", X):
        check(_lib.load().hipe_procrustes_batched(
            X.data_ptr(), Y.data_ptr(), N, J, int(mode), 1 if optimal_scale else 0,
            _ptr(out.get("d")), _ptr(out.get("Z")), _ptr(out.get("T")), _ptr(out.get("b")),
            _ptr(out.get("c")), _ptr(out.get("pa_dist")), _ptr(out.get("raw_dist")), _stream(X)))
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
    (hipe_joint_loss_fused): backward only scales the stored dlogits when the upstream
    gradient is not 1."""

    @staticmethod
    def forward(ctx, heatmap_out, gt, vis, num_joints, D, H, W, norm):
        B = heatmap_out.shape[0]
        n_vol = B * num_joints
        coords, _, _, terms, dlogits = joint_loss_fused(heatmap_out, n_vol, D, H, W, gt, vis, norm)
        loss = sum_f64(terms, norm)
        ctx.save_for_backward(dlogits)
        ctx.in_shape = heatmap_out.shape
        out_dtype = torch.promote_types(torch.float32, torch.promote_types(gt.dtype, vis.dtype))
        ctx.mark_non_differentiable(coords)
        return loss.to(out_dtype), coords.view(B, num_joints * 3)

    @staticmethod
    def backward(ctx, grad_loss, _grad_coords):
        (dlogits,) = ctx.saved_tensors
        # in place, on device, no host sync; the kernel returns at once when the scalar is 1
        gs = grad_loss.to(device=dlogits.device, dtype=torch.float32).reshape(1).contiguous()
        with torch.cuda.device(dlogits.device), _launch("scale_inplace", dlogits):
            check(_lib.load().hipe_scale_inplace(dlogits.data_ptr(), _logits_code(dlogits), dlogits.numel(),
                                                 gs.data_ptr(), _stream(dlogits)))
        return dlogits.view(ctx.in_shape), None, None, None, None, None, None, None
