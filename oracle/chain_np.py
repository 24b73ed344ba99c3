"""ORACLE (test infrastructure, not product code): numpy restatement of the coordinate chain the
reference runs between the soft-argmax and the Procrustes alignment.
See `oracle/integral_np.py` for the import rules.  Parity status: pinned against the live
reference functions (tests/golden/chain.npz, made by tests/golden/make_golden.py from source
slices of common/augment.py and data/FreiHand/FreiHand.py).

All functions are batched over a leading sample axis N and evaluate in float64, the dtype the
reference runs them in.  Citations are into /root/reference.
"""
import numpy as np


def joint_location_result(coords, patch_width, patch_height):
    """common/augment.py:170-184: (B, 3J) soft-argmax output -> (B, J, 4) patch pixels + score 1."""
    c = np.asarray(coords).astype(float)
    c = c.reshape(c.shape[0], c.shape[1] // 3, 3).copy()
    c[:, :, 0] = (c[:, :, 0] + 0.5) * patch_width
    c[:, :, 1] = (c[:, :, 1] + 0.5) * patch_height
    c[:, :, 2] = c[:, :, 2] * patch_width
    return np.concatenate((c, np.ones(c.shape[:2] + (1,))), axis=2)


def gen_trans_from_patch(c_x, c_y, src_width, src_height, dst_width, dst_height, scale, inv=False):
    """common/augment.py:457-506 for batches of axis-aligned boxes, in closed form.

    The reference hands four float32 corner pairs to cv2.getPerspectiveTransform; for two
    axis-aligned rectangles that homography is a per-axis scale and offset with last row
    (0, 0, 1).  The float32 rounding of the corners (:473-476) is kept."""
    c_x, c_y = np.asarray(c_x, dtype=np.float64), np.asarray(c_y, dtype=np.float64)
    src_w = np.asarray(src_width, dtype=np.float64) * scale
    src_h = np.asarray(src_height, dtype=np.float64) * scale
    xl = (c_x - src_w * 0.5).astype(np.float32).astype(np.float64)
    xr = (c_x + src_w * 0.5).astype(np.float32).astype(np.float64)
    yt = (c_y - src_h * 0.5).astype(np.float32).astype(np.float64)
    yb = (c_y + src_h * 0.5).astype(np.float32).astype(np.float64)
    dw = float(np.float32(dst_width))
    dh = float(np.float32(dst_height))
    T = np.zeros(c_x.shape + (3, 3))
    T[..., 2, 2] = 1.0
    if inv:      # patch -> original image
        T[..., 0, 0] = (xr - xl) / dw
        T[..., 0, 2] = xl
        T[..., 1, 1] = (yb - yt) / dh
        T[..., 1, 2] = yt
    else:        # original image -> patch
        T[..., 0, 0] = dw / (xr - xl)
        T[..., 0, 2] = -xl * dw / (xr - xl)
        T[..., 1, 1] = dh / (yb - yt)
        T[..., 1, 2] = -yt * dh / (yb - yt)
    return T


def patch_to_org_3d(coords_in_patch, trans, tprime):
    """common/augment.py:186-217 with trans_point2d (:508-511): (u, v) = first two rows of
    trans @ (x, y, 1) -- NO division by the third row, as in the reference -- and z + tprime.
    coords_in_patch (N, J, C>=3); further columns (the score) pass through."""
    c = np.asarray(coords_in_patch, dtype=np.float64)
    trans = np.asarray(trans, dtype=np.float64)
    out = c.copy()
    x, y = c[..., 0], c[..., 1]
    out[..., 0] = trans[:, None, 0, 0] * x + trans[:, None, 0, 1] * y + trans[:, None, 0, 2]
    out[..., 1] = trans[:, None, 1, 0] * x + trans[:, None, 1, 1] * y + trans[:, None, 1, 2]
    out[..., 2] = c[..., 2] + np.asarray(tprime, dtype=np.float64)[:, None]
    return out


def pixel2cam(pixel_coord, K):
    """common/augment.py:89-103: K^-1 (u, v, 1) * z / 1000."""
    p = np.asarray(pixel_coord, dtype=np.float64)
    Kinv = np.linalg.inv(np.asarray(K, dtype=np.float64))
    uv1 = np.concatenate((p[..., 0:2], np.ones(p.shape[:-1] + (1,))), axis=-1)
    xyz = np.einsum("nik,njk->nji", Kinv, uv1)
    return xyz * (p[..., 2:3] / 1000.0)


def rotate_back(cam, R):
    """data/FreiHand/FreiHand.py:722: (R^T cam^T)^T = cam @ R."""
    return np.einsum("njk,nkl->njl", np.asarray(cam, dtype=np.float64), np.asarray(R, dtype=np.float64))


def scale_by_bone(pred, bone_length, a=9, b=10):
    """data/FreiHand/FreiHand.py:632-653 (method='scale'): alpha = bone_length / |p_a - p_b|."""
    pred = np.asarray(pred, dtype=np.float64)
    d = np.sqrt(((pred[:, a] - pred[:, b]) ** 2).sum(axis=1))
    alpha = np.asarray(bone_length, dtype=np.float64) / d
    return pred * alpha[:, None, None], alpha


def camera_chain(coords_in_patch, trans, tprime, K, R=None, bone_length=None, a=9, b=10):
    """The per-sample sequence of data/FreiHand/FreiHand.py:683-737 (R given) or :853-883
    (R None): patch -> image -> camera -> un-rotated -> bone-length rescaled.
    Returns (in_img (N,J,3), cam (N,J,3))."""
    img = patch_to_org_3d(np.asarray(coords_in_patch)[..., 0:3], trans, tprime)
    cam = pixel2cam(img, K)
    if R is not None:
        cam = rotate_back(cam, R)
    if bone_length is not None:
        cam, _ = scale_by_bone(cam, bone_length, a, b)
    return img, cam
