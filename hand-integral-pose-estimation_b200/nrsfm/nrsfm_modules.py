"""Drop-in for the SVD-backed functions of procrustes_encoding/nrsfm/nrsfm_modules.py:
`batch_det_3x3` (:24-43) and `make_orthonormal` (:46-67).  The reference takes its batched SVD
from the external, un-vendored torch_batch_svd extension (nrsfm/batch_svd.py:5-35); here both
directions are one hand-written sm_100a kernel each (one thread per 3x3 matrix, Jacobi SVD in
fp64 registers).  The sparse-coding layers of that file are dense linear algebra outside the
hot path and are not rebuilt.  CUDA tensors only."""
import torch

from hipe_b200 import ops


def batch_det_3x3(x):
    """nrsfm_modules.py:24-43: determinant of (..., 3, 3) matrices by the rule of Sarrus."""
    return (x[..., 0, 0] * x[..., 1, 1] * x[..., 2, 2] + x[..., 1, 0] * x[..., 2, 1] * x[..., 0, 2]
            + x[..., 2, 0] * x[..., 1, 2] * x[..., 0, 1] - x[..., 0, 2] * x[..., 1, 1] * x[..., 2, 0]
            - x[..., 1, 2] * x[..., 2, 1] * x[..., 0, 0] - x[..., 2, 2] * x[..., 1, 0] * x[..., 0, 1])


def make_orthonormal(input_mat):
    """nrsfm_modules.py:46-67: [-1 x 3 x 3] -> the nearest rotation U diag(1, 1, det(U V^T)) V^T,
    differentiable."""
    if input_mat.dim() != 3 or input_mat.shape[-1] != 3 or input_mat.shape[-2] != 3:
        raise ValueError("make_orthonormal here covers the 3x3 case the reference's camera uses "
                         "(nrsfm_modules.py:58-65); got %s" % (tuple(input_mat.shape),))
    return ops.Orthonormalize3Function.apply(input_mat)
