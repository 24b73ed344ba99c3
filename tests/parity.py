"""Shared parity checks of the loss + gradient launches against the fp64 oracle (tests only).

Gates: coordinates |ours - ref64| <= 1e-5 * max(|ref64|, 0.5); loss terms 1e-5
relative (+1e-10 absolute); fp32 gradients <= 1e-5 * max|grad_ref64| per volume (+1e-12); bf16
gradients within 1 bf16 ulp of the fp64 reference rounded to bf16.

sign(c - gt) is undecidable on an axis where |c_ref - gt| is below the coordinate tolerance: on
exactly those axes the reference gradient is built with OUR sign (either choice is a correct
answer there); every other axis, and the rest of the volume, is still held to the oracle."""
import numpy as np

from oracle import integral_np as O

J = O.NUM_JOINTS
GRAD_GATE = 1e-5
SIGN_BAND = 1e-5


def oracle_volumes(vols64, gt, vis, norm, our_coords=None):
    """fp64 oracle on n volumes (n, D, H, W): coords (n,3), loss terms (n,3), gradient (n, V)
    of norm * sum |c - gt| vis.  `our_coords` (n,3): used for the sign on undecidable axes."""
    n, D, H, W = vols64.shape
    pad = (-n) % J
    v = np.concatenate([vols64, np.zeros((pad, D, H, W))], axis=0) if pad else vols64
    b = v.shape[0] // J
    hm = v.reshape(b, J * D, H, W)
    c = O.integral_coords(hm, J, W, H, D).reshape(-1, 3)
    gtp = np.concatenate([gt, np.zeros((pad, 3))], axis=0) if pad else gt
    visp = np.concatenate([vis, np.zeros((pad, 3))], axis=0) if pad else vis
    sgn = np.sign(c - gtp)
    band = np.abs(c - gtp) < SIGN_BAND
    if our_coords is not None:
        ours = np.concatenate([our_coords, np.zeros((pad, 3))], axis=0) if pad else our_coords
        sgn = np.where(band, np.sign(ours - gtp), sgn)
    g = (sgn * visp * norm).reshape(b, 3 * J)
    grad = O.integral_coords_vjp(hm, J, W, H, D, g).reshape(b * J, -1)
    terms = np.abs(c - gtp) * visp
    return c[:n], terms[:n], grad[:n], band[:n]


def check_volumes(vols, gt, vis, norm, coords, terms, dlogits, bf16=False, label=""):
    """vols (n, D, H, W) logits as the kernel saw them (bf16 already rounded, upcast); coords (n,3),
    terms (n,3), dlogits (n, V) from the launch.  Returns the worst gradient error relative to the
    gate (<= 1 passes) so callers can report the margin."""
    v64 = np.asarray(vols, dtype=np.float64)
    c = np.asarray(coords, dtype=np.float64)
    rc, rt, rg, band = oracle_volumes(v64, np.asarray(gt, np.float64), np.asarray(vis, np.float64), norm, c)
    cerr = np.abs(c - rc)
    assert (cerr <= 1e-5 * np.maximum(np.abs(rc), 0.5)).all(), "%s coords: max err %.3e" % (label, cerr.max())
    if terms is not None:
        t = np.asarray(terms, dtype=np.float64)
        # a term inherits the coordinate's tolerance (|c - gt| * vis with gt, vis exact)
        assert (np.abs(t - rt) <= 1e-5 * np.maximum(np.abs(rc), 0.5) * np.asarray(vis, np.float64) + 1e-12).all(), \
            "%s loss terms: max err %.3e" % (label, np.abs(t - rt).max())
    d = np.asarray(dlogits, dtype=np.float64).reshape(len(v64), -1)
    if bf16:
        import torch
        want = torch.as_tensor(rg).to(torch.bfloat16).double().numpy()
        ulp = np.maximum(np.abs(want), 1e-30) * 2.0 ** -7
        err = np.abs(d - want)
        worst = float((err / (ulp + 1e-12)).max())
        assert (err <= ulp + 1e-12).all(), "%s bf16 gradient: worst %.2f ulp" % (label, worst)
        return worst
    scale = np.abs(rg).max(axis=1, keepdims=True)
    err = np.abs(d - rg)
    worst = float((err / (GRAD_GATE * scale + 1e-12)).max())
    assert (err <= GRAD_GATE * scale + 1e-12).all(), \
        "%s gradient: worst error %.3e of the volume maximum (gate %.0e)" % (label, worst * GRAD_GATE, GRAD_GATE)
    return worst
