"""Seeded synthetic inputs shared by tests, smoke() and bench.py."""
import numpy as np

REGIMES = {"noise": (0.0, 2.0), "peaked": (12.0, 2.0), "sharp": (20.0, 1.5)}


def heatmaps(B, J, D, H, W, regime="peaked", seed=0, dtype=np.float32):
    """(B, J*D, H, W) logits = A*exp(-|v-c|^2/(2 s^2)) + N(0,1), centre c uniform per volume."""
    A, sig = REGIMES[regime]
    rng = np.random.default_rng(seed)
    out = rng.standard_normal((B, J, D, H, W)).astype(np.float32)
    if A != 0.0:
        c = rng.uniform(0, 1, size=(B, J, 3)) * np.array([W - 1, H - 1, D - 1])
        z = np.arange(D, dtype=np.float32)[None, None, :, None, None]
        y = np.arange(H, dtype=np.float32)[None, None, None, :, None]
        x = np.arange(W, dtype=np.float32)[None, None, None, None, :]
        cx = c[..., 0].astype(np.float32)[:, :, None, None, None]
        cy = c[..., 1].astype(np.float32)[:, :, None, None, None]
        cz = c[..., 2].astype(np.float32)[:, :, None, None, None]
        r2 = (x - cx) ** 2 + (y - cy) ** 2 + (z - cz) ** 2
        out += (A * np.exp(-r2 / (2 * sig * sig))).astype(np.float32)
    return out.reshape(B, J * D, H, W).astype(dtype)


def targets(B, J, seed=0, zero_vis_frac=0.0):
    rng = np.random.default_rng(seed + 7919)
    gt = rng.uniform(-0.5, 0.5, size=(B, 3 * J))
    vis = np.ones((B, 3 * J))
    if zero_vis_frac > 0:
        vis[rng.uniform(size=vis.shape) < zero_vis_frac] = 0.0
    return gt, vis


def random_rotations(n, rng):
    q, r = np.linalg.qr(rng.standard_normal((n, 3, 3)))
    q = q * np.sign(np.diagonal(r, axis1=1, axis2=2))[:, None, :]
    q[:, :, 0] *= np.sign(np.linalg.det(q))[:, None]
    return q


def hand_like_joints(n, J=21, seed=0):
    """Synthetic hand-sized joint sets in metres (the reference's own arrays are not shipped to
    the GPU box): an anisotropic blob ~ 8 cm across, 0.5-1 m from the camera."""
    rng = np.random.default_rng(seed)
    base = rng.standard_normal((n, J, 3)) * np.array([0.035, 0.03, 0.015])
    return base + rng.uniform([-0.1, -0.1, 0.5], [0.1, 0.1, 1.0], size=(n, 1, 3))


def procrustes_pairs(n, J=21, seed=0, noise=0.005, mirror_frac=0.0):
    """(X, Y): X = 1.1 * Y R + 0.3 + N(0, noise); a fraction of X mirrored."""
    rng = np.random.default_rng(seed)
    Y = hand_like_joints(n, J, seed + 1)
    R = random_rotations(n, rng)
    X = 1.1 * np.einsum("njk,nkl->njl", Y, R) + 0.3 + rng.standard_normal((n, J, 3)) * noise
    if mirror_frac > 0:
        m = rng.uniform(size=n) < mirror_frac
        X[m, :, 0] *= -1.0
    return X, Y


def chain_inputs(n, J=21, seed=0, patch=224.0):
    """Per-sample inputs of the patch -> camera chain (data/FreiHand/FreiHand.py:683-742) with
    FreiHAND-like magnitudes: a hand 0.5-1 m away, 224x224 images, focal length ~ 480 px.
    Ground truth in camera space is projected into a box crop to make consistent patch
    coordinates, which are then perturbed (the 'prediction')."""
    rng = np.random.default_rng(seed)
    gt_cam = hand_like_joints(n, J, seed + 3)
    R = random_rotations(n, rng)
    K = np.zeros((n, 3, 3))
    K[:, 0, 0] = rng.uniform(440, 520, n)
    K[:, 1, 1] = K[:, 0, 0] * rng.uniform(0.98, 1.02, n)
    K[:, 0, 2] = rng.uniform(100, 124, n)
    K[:, 1, 2] = rng.uniform(100, 124, n)
    K[:, 2, 2] = 1.0
    K[:, 0, 1] = rng.uniform(-0.5, 0.5, n)              # a little skew: exercises the full inverse
    rot = np.einsum("nik,njk->nji", R, gt_cam)          # (R x^T)^T
    uvw = np.einsum("nik,njk->nji", K, rot)
    uv = uvw[..., :2] / uvw[..., 2:3]
    z_mm = rot[..., 2] * 1000.0
    lo, hi = uv.min(axis=1), uv.max(axis=1)
    ctr = 0.5 * (lo + hi)
    size = (hi - lo).max(axis=1) * rng.uniform(1.2, 1.6, n)
    bbox = np.stack([ctr[:, 0], ctr[:, 1], size, size * rng.uniform(0.9, 1.1, n)], axis=1)
    tprime = z_mm[:, 9] + rng.uniform(-5, 5, n)
    x = (uv[..., 0] - (bbox[:, None, 0] - 0.5 * bbox[:, None, 2])) * patch / bbox[:, None, 2]
    y = (uv[..., 1] - (bbox[:, None, 1] - 0.5 * bbox[:, None, 3])) * patch / bbox[:, None, 3]
    z = z_mm - tprime[:, None]
    pt = np.stack([x, y, z, np.ones_like(x)], axis=2)
    pt[..., 0:3] += rng.standard_normal((n, J, 3)) * np.array([2.0, 2.0, 4.0])
    bone = np.sqrt(((gt_cam[:, 9] - gt_cam[:, 10]) ** 2).sum(axis=1))
    return {"patch": pt, "bbox": bbox, "tprime": tprime, "K": K, "R": R, "bone_len": bone, "gt_cam": gt_cam}


def matrices3(n, seed=0):
    """3x3 matrices for the orthonormalisation: generic, with negative determinant (the flip must
    act), one small singular value, and near-rotations (the PANet camera's usual input)."""
    rng = np.random.default_rng(seed)
    A = rng.standard_normal((n, 3, 3))
    A[1::4] = random_rotations(len(A[1::4]), rng) + rng.standard_normal(A[1::4].shape) * 0.05
    A[2::4, :, 0] *= -1.0
    q = random_rotations(len(A[3::4]), rng)
    sv = np.stack([rng.uniform(1, 2, len(q)), rng.uniform(0.3, 0.9, len(q)), rng.uniform(0.02, 0.1, len(q))], axis=1)
    A[3::4] = np.einsum("nij,nj,nkj->nik", q, sv, random_rotations(len(q), rng))
    return A
