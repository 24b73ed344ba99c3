"""The drop-in package imported the way the reference's own call sites import it
(common/base.py:15-16,92-95,182-185, main/train.py:12, main/test.py:10, main/model.py:2,
data/dataset.py, data/FreiHand/FreiHand.py:745,865,888, procrustes_encoding/nrsfm/nrsfmnet.py:5),
with this package's directory in FRONT of the reference's directories on sys.path as
INTEGRATION.md section 1 prescribes.  Runs in a subprocess (a clean sys.modules); needs the
reference tree, so it is skipped on the GPU box."""
import os
import subprocess
import sys
import textwrap

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "hand-integral-pose-estimation_b200")
REF = "/root/reference"

SCRIPT = textwrap.dedent("""
    import sys, types
    REF, PKG = %(ref)r, %(pkg)r
    # what main/config.py:93-100 builds, minus the mkdirs under the read-only tree
    for d in ("data/FreiHand", "data", "main", "procrustes_encoding", "lib", "common"):
        sys.path.insert(0, REF + "/" + d)
    sys.path.insert(0, PKG)                         # INTEGRATION.md section 1
    cfgmod = types.ModuleType("config")
    class _Cfg:
        patch_width = 224; patch_height = 224; _lambda = 0; input_shape = (224, 224); loss = "L_combined"
        num_gpus = 1
    cfgmod.cfg = _Cfg()
    sys.modules["config"] = cfgmod
    for name in ("matplotlib", "matplotlib.pyplot", "config_panet", "pycrayon"):   # not installed here
        sys.modules.setdefault(name, types.ModuleType(name))
    sys.modules["matplotlib.pyplot"].switch_backend = lambda *a, **k: None          # common/augment.py:14
    import torchvision.models.resnet as _r
    if not hasattr(_r, "model_urls"):
        _r.model_urls = {}                          # name removed upstream; common/nets/resnet.py:5

    # common/base.py:15-16
    from nets.balanced_parallel import DataParallelModel, DataParallelCriterion
    from nets import loss
    # main/model.py:2
    from nets.resnet import ResNetBackbone
    import nets
    assert any(p.startswith(PKG) for p in nets.__path__) and any(p.startswith(REF) for p in nets.__path__), list(nets.__path__)
    assert loss.__file__.startswith(PKG), loss.__file__
    # common/base.py:92-95,182-185
    crits = [DataParallelCriterion(loss.CombinedLoss()), DataParallelCriterion(loss.JointLocationLoss()),
             DataParallelCriterion(loss.JointLocationLoss2())]
    assert loss.CombinedLoss().lambda_panet == 0
    _ = loss.softmax_integral_tensor                # common/base.py:181
    # main/train.py:12, main/test.py:10
    from nets.loss import softmax_integral_tensor, JointLocationLoss
    # the augment namespace: hot-path names are ours, everything else the reference's own
    import augment
    assert augment.__file__.startswith(PKG)
    assert augment.REFERENCE_AUGMENT == REF + "/common/augment.py" and augment.REFERENCE_AUGMENT_ERROR is None, \\
        (augment.REFERENCE_AUGMENT, augment.REFERENCE_AUGMENT_ERROR)
    for name in ("get_aug_config", "generate_patch_image", "trans_point2d", "generate_joint_location_label",
                 "generate_input_unlabelled", "projectPoints", "gen_trans_from_patch_cv", "prepare_panet_input",
                 "test_get_joint_loc_res", "pixel2cam_torch", "find_bb", "sample_rotation_matrix"):
        f = getattr(augment, name)
        assert f.__code__.co_filename == REF + "/common/augment.py", (name, f.__code__.co_filename)
    for name in ("compute_similarity_transform", "get_joint_location_result", "pixel2cam",
                 "trans_coords_from_patch_to_org_3d", "compute_similarity_transform_batched"):
        f = getattr(augment, name)
        assert f.__code__.co_filename.startswith(PKG), (name, f.__code__.co_filename)
    import numpy as np
    assert augment.get_aug_config()[0] == 1.0       # the reference's own body runs (augment.py:235-250)
    pts = augment.trans_point2d(np.array([1.0, 2.0]), np.eye(3))
    assert pts.tolist() == [1.0, 2.0]
    # the nrsfm namespace: sparse-coding layers are the reference's, the SVD-backed functions ours
    from nrsfm.nrsfmnet import PANet
    from nrsfm import nrsfm_modules as M
    assert M.__file__.startswith(PKG) and M.REFERENCE_MODULE_ERROR is None, M.REFERENCE_MODULE_ERROR
    assert M.SfMSparseCodingLayer.__module__ == "nrsfm.nrsfm_modules"
    assert M.make_orthonormal.__code__.co_filename.startswith(PKG)
    assert M.relu_threshold.__code__.co_filename == REF + "/procrustes_encoding/nrsfm/nrsfm_modules.py"
    print("DROPIN-OK")
""")


@pytest.mark.skipif(not os.path.isdir(REF), reason="needs the reference tree (not on the GPU box)")
def test_reference_call_sites_import_the_dropin():
    env = dict(os.environ, HIPE_NO_BUILD="1")
    r = subprocess.run([sys.executable, "-c", SCRIPT % {"ref": REF, "pkg": PKG}], capture_output=True, text=True,
                       env=env, timeout=600)
    assert r.returncode == 0 and "DROPIN-OK" in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]


def test_dropin_alone_has_the_hot_path_names():
    """No reference tree on the path: the modules still import and carry the hot-path surface."""
    code = textwrap.dedent("""
        import sys
        sys.path.insert(0, %r)
        from nets import loss
        import augment
        from nrsfm import nrsfm_modules as M
        assert augment.REFERENCE_AUGMENT is None and M.REFERENCE_MODULE is None
        for n in ("softmax_integral_tensor", "generate_3d_integral_preds_tensor", "softmax_integral_tensor2",
                  "JointLocationLoss", "JointLocationLoss2", "CombinedLoss", "computeMPJPE"):
            assert hasattr(loss, n), n
        for n in ("compute_similarity_transform", "get_joint_location_result", "pixel2cam",
                  "trans_coords_from_patch_to_org_3d"):
            assert hasattr(augment, n), n
        loss.JointLocationLoss(); loss.JointLocationLoss2(); loss.CombinedLoss()
        print("ALONE-OK")
    """ % PKG)
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True,
                       env=dict(os.environ, HIPE_NO_BUILD="1"), timeout=600)
    assert r.returncode == 0 and "ALONE-OK" in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]
