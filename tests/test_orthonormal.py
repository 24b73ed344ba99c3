"""Batched 3x3 orthonormalisation with backward: the torch oracle against
the golden vectors frozen from the live make_orthonormal, and the CUDA kernels against both.
Tolerances: fp64 I/O 1e-10 absolute on R, 1e-9 relative to max|dA| on the gradient; fp32 I/O
2e-6 / 1e-5 (one fp32 rounding of an fp64-internal result)."""
import numpy as np
import pytest
import torch

from oracle import orthonormal_torch as OT
from tests import synth
from tests.golden import ref_loader
from tests.test_oracle_golden import gold


def test_orthonormal_oracle_golden():
    g = gold("orthonormal")
    r, dA = OT.make_orthonormal_with_grad(g["A"], g["G"])
    assert np.abs(r - g["R"]).max() < 1e-13
    assert np.abs(dA - g["dA"]).max() < 1e-11 * np.abs(g["dA"]).max()
    assert np.allclose(np.linalg.det(r), 1.0, atol=1e-12)                       # always a proper rotation
    assert np.abs(np.einsum("nji,njk->nik", r, r) - np.eye(3)).max() < 1e-13


@pytest.mark.skipif(not ref_loader.available(), reason="/root/reference is not mounted here")
def test_orthonormal_oracle_vs_live_reference():
    M = ref_loader.load_nrsfm_modules()
    A = synth.matrices3(40, seed=9)
    G = np.random.default_rng(10).standard_normal(A.shape)
    a = torch.as_tensor(A).clone().requires_grad_(True)
    r = M.make_orthonormal(a)
    (r * torch.as_tensor(G)).sum().backward()
    r2, g2 = OT.make_orthonormal_with_grad(A, G)
    assert np.abs(r.detach().numpy() - r2).max() < 1e-14
    assert np.abs(a.grad.numpy() - g2).max() < 1e-11 * np.abs(g2).max()
    assert np.allclose(M.batch_det_3x3(torch.as_tensor(A)).numpy(), np.linalg.det(A), rtol=1e-10, atol=1e-12)


@pytest.mark.gpu
@pytest.mark.parametrize("dtype,tol_r,tol_g", [(torch.float64, 1e-10, 1e-9), (torch.float32, 2e-6, 1e-5)])
def test_make_orthonormal_kernel_vs_golden(dtype, tol_r, tol_g):
    from nrsfm import nrsfm_modules as N
    g = gold("orthonormal")
    A = torch.as_tensor(g["A"]).to(dtype).cuda()
    if dtype == torch.float32:      # the oracle sees the rounded input
        ref_r, ref_g = OT.make_orthonormal_with_grad(A.cpu().double().numpy(), g["G"].astype(np.float32))
    else:
        ref_r, ref_g = g["R"], g["dA"]
    a = A.clone().requires_grad_(True)
    r = N.make_orthonormal(a)
    assert r.dtype == dtype and r.shape == a.shape
    (r * torch.as_tensor(g["G"]).to(dtype).cuda()).sum().backward()
    assert np.abs(r.detach().cpu().double().numpy() - ref_r).max() < tol_r
    assert np.abs(a.grad.cpu().double().numpy() - ref_g).max() < tol_g * np.abs(ref_g).max()
    det = N.batch_det_3x3(A)
    assert np.allclose(det.cpu().double().numpy(), np.linalg.det(A.cpu().double().numpy()), rtol=1e-4, atol=1e-5)


@pytest.mark.gpu
@pytest.mark.parametrize("n", [0, 1, 127, 128, 100000])
def test_make_orthonormal_properties(n):
    """Size-independent properties: R^T R = I, det R = +1, idempotence (a rotation maps to itself),
    invariance to positive scaling, gradient orthogonal to the scaling direction."""
    from nrsfm import nrsfm_modules as N
    if n == 0:
        out = N.make_orthonormal(torch.empty((0, 3, 3), device="cuda"))
        assert out.shape == (0, 3, 3)
        return
    A = torch.as_tensor(synth.matrices3(n, seed=n)).cuda().requires_grad_(True)
    r = N.make_orthonormal(A)
    eye = torch.eye(3, dtype=torch.float64, device="cuda")
    assert (r.transpose(1, 2) @ r - eye).abs().max().item() < 1e-12
    assert (torch.linalg.det(r) - 1).abs().max().item() < 1e-12
    assert (N.make_orthonormal(r.detach()) - r.detach()).abs().max().item() < 1e-12
    assert (N.make_orthonormal(A.detach() * 3.7) - r.detach()).abs().max().item() < 1e-12
    G = torch.randn_like(r)
    (r * G).sum().backward()
    scale_dir = (A.grad * A.detach()).sum(dim=(1, 2))            # d/dt R(t A) = 0
    assert scale_dir.abs().max().item() < 1e-9 * A.grad.abs().max().item()


@pytest.mark.gpu
def test_make_orthonormal_errors():
    from nrsfm import nrsfm_modules as N
    with pytest.raises(RuntimeError):
        N.make_orthonormal(torch.zeros(2, 3, 3))                 # CPU tensor: no CPU path
    with pytest.raises(ValueError):
        N.make_orthonormal(torch.zeros(2, 4, 4, device="cuda"))
