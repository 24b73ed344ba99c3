"""hipe_b200: B200-native (sm_100a) kernels for the integral-regression hot path of
mqadri9/hand-integral-pose-estimation, behind the reference's own operator names.

    from hipe_b200 import ops            # tensors in, tensors out (torch, CUDA only)
    from nets import loss                # drop-in for the reference's common/nets/loss.py
    import augment                       # drop-in for the Procrustes / joint-result helpers
"""
from . import _lib  # noqa: F401
from ._lib import HipeError, load, lib_path  # noqa: F401
from .graphed import GraphedCriterion, GraphedShardedStep, graphed, graphed_sharded  # noqa: F401

__all__ = ["HipeError", "load", "lib_path", "graphed", "GraphedCriterion", "graphed_sharded", "GraphedShardedStep"]
