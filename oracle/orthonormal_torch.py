"""ORACLE (test infrastructure, not product code): make_orthonormal
(procrustes_encoding/nrsfm/nrsfm_modules.py:46-67) restated on torch.linalg.svd, CPU, with
autograd supplying the backward.  See `oracle/integral_np.py` for the import rules.

Parity status: pinned against the live reference FUNCTION (nrsfm_modules.py is imported
unmodified by tests/golden/ref_loader.py) with its external SVD dependency substituted: the
reference calls torch_batch_svd (KinglittleQ/torch-batch-svd, a cuSOLVER gesvdjBatched wrapper,
un-vendored, unpinned, not installed -- SURVEY.md 8c), whose published contract is x = U S V^T
with descending singular values; torch.linalg.svd provides that contract on CPU.  Goldens:
tests/golden/orthonormal.npz.
"""
import torch


def batch_det_3x3(x):
    """nrsfm_modules.py:24-43."""
    return torch.linalg.det(x)


def make_orthonormal(input_mat):
    """nrsfm_modules.py:46-67: U V^T with the last column of U multiplied by det(U V^T)."""
    u, s, vh = torch.linalg.svd(input_mat)
    orth = u @ vh
    det = batch_det_3x3(orth)
    u_flip = torch.cat((u[..., :2], u[..., 2:3] * det.view(-1, 1, 1)), 2)
    return u_flip @ vh


def make_orthonormal_with_grad(mats, upstream):
    """(R, dL/dA) for L = sum(R * upstream), float64."""
    a = torch.as_tensor(mats, dtype=torch.float64).clone().requires_grad_(True)
    r = make_orthonormal(a)
    (r * torch.as_tensor(upstream, dtype=torch.float64)).sum().backward()
    return r.detach().numpy(), a.grad.numpy()
