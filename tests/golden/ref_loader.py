"""Load the LIVE reference functions from /root/reference (this container only).

Used by `make_golden.py` to generate the committed fixtures and by the
`test_oracle_vs_live_reference` tests (skipped when /root/reference is absent,
i.e. on the GPU box).  Nothing under `hand-integral-pose-estimation_b200/`,
`bench.py` or `__graft_entry__.py` imports this file.

Shims (none touch arithmetic):
  * empty stub modules `config` (only `cfg` is read), `augment`, `config_panet`
    so `common/nets/loss.py` imports without mkdir-ing into the read-only tree;
  * `torch.cuda.comm.broadcast` returns `[t]` when `devices == [None]`, i.e. on
    CPU tensors (`common/nets/loss.py:35-37` would raise otherwise);
  * `compute_similarity_transform` (`common/augment.py:18-77`) and
    `align_w_scale` (`main/eval.py:142-165`), `EvalUtil` (`main/eval.py:15-106`)
    are exec'd from source slices because their modules need matplotlib.
"""
import importlib.util
import os
import sys
import types

REF = os.environ.get("HIPE_REFERENCE_ROOT", "/root/reference")


def available() -> bool:
    return os.path.isfile(os.path.join(REF, "common", "nets", "loss.py"))


def _slice(path, first, last):
    with open(os.path.join(REF, path)) as f:
        lines = f.readlines()
    return "".join(lines[first - 1:last])


_cache = {}


def load_loss_module():
    """The unmodified `common/nets/loss.py` as a module object."""
    if "loss" in _cache:
        return _cache["loss"]
    import torch
    import torch.cuda.comm  # noqa: F401

    cfg_mod = types.ModuleType("config")
    cfg_mod.cfg = types.SimpleNamespace(_lambda=0, patch_width=224, patch_height=224)
    saved = {k: sys.modules.get(k) for k in ("config", "augment", "config_panet", "FreiHand_config")}
    sys.modules["config"] = cfg_mod
    sys.modules["augment"] = types.ModuleType("augment")
    sys.modules["config_panet"] = types.ModuleType("config_panet")
    fh = os.path.join(REF, "data", "FreiHand")
    sys.path.insert(0, fh)

    orig = torch.cuda.comm.broadcast

    def broadcast(tensor, devices=None, *, out=None):
        if devices == [None]:
            return [tensor]
        return orig(tensor, devices=devices, out=out)

    torch.cuda.comm.broadcast = broadcast
    try:
        spec = importlib.util.spec_from_file_location(
            "_hipe_reference_loss", os.path.join(REF, "common", "nets", "loss.py"))
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
    finally:
        sys.path.remove(fh)
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v
    _cache["loss"] = mod
    return mod


def load_compute_similarity_transform():
    if "cst" not in _cache:
        import numpy as np
        ns = {"np": np}
        exec(_slice("common/augment.py", 18, 77), ns)
        _cache["cst"] = ns["compute_similarity_transform"]
    return _cache["cst"]


def load_eval_module():
    """`align_w_scale` and `EvalUtil` from main/eval.py."""
    if "eval" not in _cache:
        import numpy as np
        from scipy.linalg import orthogonal_procrustes
        if not hasattr(np, "trapz"):
            np.trapz = np.trapezoid
        ns = {"np": np, "orthogonal_procrustes": orthogonal_procrustes}
        exec(_slice("main/eval.py", 15, 106), ns)
        exec(_slice("main/eval.py", 142, 165), ns)
        _cache["eval"] = types.SimpleNamespace(align_w_scale=ns["align_w_scale"], EvalUtil=ns["EvalUtil"])
    return _cache["eval"]


def load_chain_functions():
    """The coordinate chain between soft-argmax and Procrustes: source slices
    of common/augment.py (pixel2cam :89-103, trans_coords_from_patch_to_org{,_3d} :186-217,
    gen_trans_from_patch_cv + trans_point2d :457-511) and FreiHand.scale_result
    (data/FreiHand/FreiHand.py:632-668, a method: dedented and called with self=None)."""
    if "chain" not in _cache:
        import textwrap
        import cv2
        import numpy as np
        ns = {"np": np, "cv2": cv2}
        exec(_slice("common/augment.py", 89, 103), ns)
        exec(_slice("common/augment.py", 186, 217), ns)
        exec(_slice("common/augment.py", 457, 511), ns)
        exec(textwrap.dedent(_slice("data/FreiHand/FreiHand.py", 632, 668)), ns)
        _cache["chain"] = types.SimpleNamespace(
            pixel2cam=ns["pixel2cam"], trans_coords_from_patch_to_org_3d=ns["trans_coords_from_patch_to_org_3d"],
            gen_trans_from_patch_cv=ns["gen_trans_from_patch_cv"], scale_result=ns["scale_result"])
    return _cache["chain"]


def load_nrsfm_modules():
    """The unmodified procrustes_encoding/nrsfm/nrsfm_modules.py (make_orthonormal :46-67,
    batch_det_3x3 :24-43).  Its `from .batch_svd import batch_svd` needs the external
    torch_batch_svd CUDA extension (absent): a stand-in package provides
    `batch_svd(x) -> (U, S, V)` with x = U S V^T, descending S -- that extension's contract --
    from torch.linalg.svd."""
    if "nrsfm" not in _cache:
        import torch
        pkg = types.ModuleType("_hipe_ref_nrsfm")
        pkg.__path__ = []
        sub = types.ModuleType("_hipe_ref_nrsfm.batch_svd")

        def batch_svd(x):
            u, s, vh = torch.linalg.svd(x)
            return u, s, vh.transpose(1, 2)

        sub.batch_svd = batch_svd
        sys.modules["_hipe_ref_nrsfm"] = pkg
        sys.modules["_hipe_ref_nrsfm.batch_svd"] = sub
        spec = importlib.util.spec_from_file_location(
            "_hipe_ref_nrsfm.nrsfm_modules", os.path.join(REF, "procrustes_encoding", "nrsfm", "nrsfm_modules.py"))
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        _cache["nrsfm"] = mod
    return _cache["nrsfm"]


def load_loss2_class():
    """JointLocationLoss2 of the unmodified common/nets/loss.py (:207-277), runnable on CPU: its
    module-level `augment` (a stub here) gets the reference's own test_get_joint_loc_res
    (common/augment.py:186-194, source slice, with the same cfg), trans_coords_from_patch_to_org_3d
    and pixel2cam (source slices, see load_chain_functions), and the one `.cuda()` of its result
    (:268) is made an identity for the duration of a call."""
    if "loss2" in _cache:
        return _cache["loss2"]
    import numpy as np
    import torch
    L = load_loss_module()
    ch = load_chain_functions()
    ns = {"np": np, "cfg": types.SimpleNamespace(patch_width=224, patch_height=224)}
    exec(_slice("common/augment.py", 186, 194), ns)
    L.augment.test_get_joint_loc_res = ns["test_get_joint_loc_res"]
    L.augment.trans_coords_from_patch_to_org_3d = ch.trans_coords_from_patch_to_org_3d
    L.augment.pixel2cam = ch.pixel2cam

    def run(*args):
        orig = torch.Tensor.cuda
        torch.Tensor.cuda = lambda self, *a, **k: self
        try:
            return L.JointLocationLoss2()(*args)
        finally:
            torch.Tensor.cuda = orig

    _cache["loss2"] = run
    return run


def fixture_arrays():
    """The two joint-set arrays the reference ships."""
    import numpy as np
    return (np.load(os.path.join(REF, "main", "evaluation_predictions.npy")),
            np.load(os.path.join(REF, "procrustes_encoding", "local_data", "hand_test.npy")))
