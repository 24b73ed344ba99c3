"""ctypes binding of libhipe_b200.so (the C ABI declared in include/hipe_b200.h).

There is no CPU fallback and no alternative backend: if the library cannot be loaded, or a
tensor is not on a CUDA device, the call raises."""
import ctypes
import os
import threading

from . import _build

F32, BF16, F64 = 0, 1, 3
E_UNSUPPORTED = -7
E_LIMIT = -8
VARIANT_AUTO, VARIANT_TMA, VARIANT_LDG = 0, 1, 2

_c = ctypes
_vp, _i, _ll, _d = _c.c_void_p, _c.c_int, _c.c_longlong, _c.c_double

# name -> (restype, argtypes); mirrors include/hipe_b200.h one to one
SIGNATURES = {
    "hipe_version": (_i, []),
    "hipe_strerror": (_c.c_char_p, [_i]),
    "hipe_softargmax_fwd": (_i, [_vp, _i, _ll, _i, _i, _i, _i, _vp, _vp, _vp, _i, _i, _vp]),
    "hipe_softargmax_bwd": (_i, [_vp, _i, _ll, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _i, _i, _vp]),
    "hipe_volume_softmax": (_i, [_vp, _i, _ll, _i, _i, _i, _vp, _vp, _i, _i, _vp]),
    "hipe_volume_entropy": (_i, [_vp, _i, _ll, _i, _i, _i, _vp, _vp, _vp]),
    "hipe_l1_loss_fwd": (_i, [_vp, _vp, _vp, _i, _ll, _i, _d, _vp, _vp, _vp]),
    "hipe_joint_loss_fused": (_i, [_vp, _i, _ll, _i, _i, _i, _vp, _vp, _i, _d, _vp, _vp, _vp, _vp, _vp,
                                   _i, _i, _vp]),
    "hipe_resident_workspace_bytes": (_ll, [_i, _ll, _i, _i, _i]),
    "hipe_joint_loss_resident": (_i, [_vp, _i, _ll, _i, _i, _i, _vp, _vp, _i, _d, _vp, _vp, _vp, _vp, _vp, _vp,
                                      _vp, _ll, _vp]),
    "hipe_cluster_workspace_bytes": (_ll, [_i, _ll, _i, _i, _i]),
    "hipe_cluster_plan": (_i, [_i, _ll, _i, _i, _i, _vp, _vp, _vp, _vp, _vp]),
    "hipe_joint_loss_cluster": (_i, [_vp, _i, _ll, _i, _i, _i, _vp, _vp, _i, _d, _vp, _vp, _vp, _vp, _vp, _vp,
                                     _vp, _ll, _vp]),
    "hipe_scale_inplace": (_i, [_vp, _i, _ll, _vp, _vp]),
    "hipe_scale_inplace_s": (_i, [_vp, _i, _ll, _vp, _i, _vp]),
    "hipe_sum_f64": (_i, [_vp, _ll, _d, _vp, _vp]),
    "hipe_procrustes_batched": (_i, [_vp, _vp, _ll, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "hipe_patch_to_camera": (_i, [_vp, _i, _i, _ll, _i, _d, _d, _vp, _vp, _d, _d, _d, _vp, _vp, _vp, _vp, _i, _i,
                                  _vp, _vp, _vp]),
    "hipe_eval_chain_procrustes": (_i, [_vp, _i, _i, _ll, _i, _d, _d, _vp, _vp, _d, _d, _d, _vp, _vp, _vp, _vp, _i,
                                        _i, _vp, _i, _i, _vp, _vp, _vp, _vp, _vp]),
    "hipe_keypoint_measures_workspace_bytes": (_ll, [_i, _i]),
    "hipe_keypoint_measures": (_i, [_vp, _vp, _ll, _i, _d, _d, _i, _vp, _vp, _vp, _ll, _vp]),
    "hipe_keypoint_measures_noalloc": (_i, [_vp, _vp, _ll, _i, _d, _d, _i, _vp, _vp, _vp]),
    "hipe_orthonormalize3_fwd": (_i, [_vp, _i, _ll, _vp, _vp, _vp]),
    "hipe_orthonormalize3_bwd": (_i, [_vp, _i, _ll, _vp, _vp, _vp]),
    "hipe_peer_exchange_reclen": (_ll, [_ll, _i]),
    "hipe_peer_exchange_bytes": (_ll, [_i, _ll, _i]),
    "hipe_peer_exchange": (_i, [_vp, _ll, _d, _vp, _ll, _i, _vp, _i, _i, _c.c_ulonglong, _d, _vp, _vp, _vp, _vp, _vp, _d, _vp]),
    "hipe_peer_publish": (_i, [_vp, _ll, _d, _vp, _ll, _i, _vp, _i, _i, _c.c_ulonglong, _vp]),
    "hipe_peer_collect": (_i, [_vp, _i, _i, _c.c_ulonglong, _ll, _i, _d, _vp, _vp, _vp, _vp, _vp, _d, _vp]),
    "hipe_joint_location_loss_host": (_i, [_vp, _i, _ll, _i, _i, _i, _i, _vp, _vp, _i, _i, _vp, _vp, _vp,
                                           _vp, _ll]),
    "hipe_procrustes_host": (_i, [_vp, _vp, _ll, _i, _i, _i, _vp, _vp, _vp]),
}

_lock = threading.Lock()
_lib = None


def lib_path():
    return _build.LIB_PATH


def load():
    """Load (building first if the sources are newer) and type the library."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        if os.environ.get("HIPE_NO_BUILD") != "1":
            try:
                _build.build()
            except RuntimeError:
                if not os.path.exists(_build.LIB_PATH):
                    raise
        if not os.path.exists(_build.LIB_PATH):
            raise RuntimeError("libhipe_b200.so is missing (%s): run __graft_entry__.build()" % _build.LIB_PATH)
        lib = ctypes.CDLL(_build.LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


class HipeError(RuntimeError):
    pass


def check(rc):
    if rc != 0:
        msg = load().hipe_strerror(int(rc)).decode()
        raise HipeError("libhipe_b200: error %d: %s" % (rc, msg))
