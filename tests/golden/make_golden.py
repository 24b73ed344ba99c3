"""Freeze golden vectors from the LIVE reference (run in the build container only):

    python tests/golden/make_golden.py

imports the unmodified /root/reference/common/nets/loss.py and the source slices of
common/augment.py:18-77 and main/eval.py:15-165 through `ref_loader`, evaluates them on seeded
inputs and writes tests/golden/*.npz.  The reference has no golden vectors of its own for this
path, so these files are what pins the oracle and, through it, the CUDA
path.  Inputs small enough are stored; the 64^3 case stores a SHA-256 of its regenerated input.
"""
import hashlib
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from tests import synth  # noqa: E402
from tests.golden import ref_loader  # noqa: E402

J = 21

# name -> (B, D, H, W, regime, seed, store_input)
INTEGRAL_CASES = {
    "cube8": (2, 8, 8, 8, "peaked", 101, True),
    "ragged567": (3, 5, 6, 7, "sharp", 102, True),
    "box6x10x12": (1, 6, 10, 12, "noise", 103, True),
    "cube64": (1, 64, 64, 64, "peaked", 104, False),
}
N_GRAD_SAMPLES = 512


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def integral_case(L, B, D, H, W, regime, seed, store_input):
    hm = synth.heatmaps(B, J, D, H, W, regime, seed=seed)
    gt, vis = synth.targets(B, J, seed=seed, zero_vis_frac=0.1)
    out = {"shape": np.array([B, D, H, W]), "seed": np.array(seed), "regime": np.array(regime),
           "input_sha256": np.array(sha(hm)), "gt": gt, "vis": vis}
    if store_input:
        out["logits"] = hm
    t_gt, t_vis = torch.as_tensor(gt), torch.as_tensor(vis)
    # fp64 evaluation of the reference code on the upcast input = ground truth
    x = torch.as_tensor(hm).double().requires_grad_(True)
    coords = L.softmax_integral_tensor(x, J, W, H, D)
    out["coords64"] = coords.detach().numpy()
    crit = L.JointLocationLoss()
    loss = crit(x, t_gt, t_vis)
    loss.backward()
    out["loss64"] = np.array(loss.item())
    grad = x.grad.numpy().reshape(B * J, -1)
    # vjp with a seeded dense upstream gradient (exercises all three axes with non-+-1 weights)
    rng = np.random.default_rng(seed + 1)
    g = rng.standard_normal((B, 3 * J))
    x2 = torch.as_tensor(hm).double().requires_grad_(True)
    (L.softmax_integral_tensor(x2, J, W, H, D) * torch.as_tensor(g)).sum().backward()
    vjp = x2.grad.numpy().reshape(B * J, -1)
    out["upstream"] = g
    if store_input:
        out["loss_grad64"] = grad
        out["vjp64"] = vjp
    else:
        idx = np.random.default_rng(seed + 2).integers(0, grad.shape[1], size=N_GRAD_SAMPLES)
        out["grad_idx"] = idx
        out["loss_grad64_sampled"] = grad[:, idx]
        out["vjp64_sampled"] = vjp[:, idx]
        out["loss_grad64_abs_sum"] = np.abs(grad).sum(axis=1)
        out["vjp64_abs_sum"] = np.abs(vjp).sum(axis=1)
        out["vjp64_abs_max"] = np.abs(vjp).max(axis=1)
        out["loss_grad64_abs_max"] = np.abs(grad).max(axis=1)
    # the reference in its own fp32 arithmetic ("the reference's own error")
    with torch.no_grad():
        x32 = torch.as_tensor(hm)
        out["coords32"] = L.softmax_integral_tensor(x32, J, W, H, D).numpy()
        out["loss32"] = np.array(crit(x32, t_gt, t_vis).item())
        if store_input:
            p = L.softmax_integral_tensor2(torch.as_tensor(hm).double(), J, W, H, D)
            out["probs64"] = p.numpy()
            ex, ey, ez = L.generate_3d_integral_preds_tensor(p, J, W, H, D)
            out["expect64"] = torch.cat((ex, ey, ez), dim=2).numpy()
            out["mpjpe64"] = np.array(L.computeMPJPE(coords.detach(), t_gt).item())
    return out


def combined_case(L):
    """CombinedLoss (loss.py:99-203) with cfg._lambda = 0 and pass-through PANet stubs."""
    B, D, H, W = 4, 8, 8, 8
    hm = synth.heatmaps(B, J, D, H, W, "peaked", seed=201)
    gt, vis = synth.targets(B, J, seed=201, zero_vis_frac=0.1)
    rng = np.random.default_rng(202)
    teacher = rng.uniform(-0.5, 0.5, size=(1, B, 3 * J))
    labelled = np.array([True, False, True, False])
    L.augment.prepare_panet_input = lambda pts, *a, **k: pts

    class Tester:
        def forward(self, pts):
            return pts, None, None, None

    x = torch.as_tensor(hm).double().requires_grad_(True)
    z = torch.zeros(B)
    loss, s_mpjpe, t_mpjpe, sup, unsup = L.CombinedLoss()(
        x, torch.as_tensor(teacher), torch.as_tensor(gt), torch.as_tensor(vis), torch.as_tensor(labelled),
        z, z, z, z, z, z, z, Tester())
    loss.backward()
    return {"logits": hm, "gt": gt, "vis": vis, "teacher": teacher, "labelled": labelled,
            "loss64": np.array(loss.item()), "student_mpjpe": np.array(s_mpjpe.item()),
            "teacher_mpjpe": np.array(t_mpjpe.item()), "sup": np.array(sup.item()),
            "unsup": np.array(unsup.item()), "grad64": x.grad.numpy()}


def procrustes_cases():
    cst = ref_loader.load_compute_similarity_transform()
    ev = ref_loader.load_eval_module()
    preds, hands = ref_loader.fixture_arrays()
    rng = np.random.default_rng(0)
    # every 31st FreiHAND evaluation prediction (metres) as Y; SURVEY.md 8d recipe for X
    Y = preds[::31][:128].copy()
    n = len(Y)
    R = synth.random_rotations(n, rng)
    X = 1.1 * np.einsum("njk,nkl->njl", Y, R) + 0.3 + rng.standard_normal(Y.shape) * 0.005
    X[::16, :, 0] *= -1.0                    # mirrored targets: the reflection repair must act
    Y2 = hands[::101][:32].copy()            # the PANet test hands: a second shape distribution
    X2 = 0.8 * np.einsum("njk,nkl->njl", Y2, synth.random_rotations(len(Y2), rng)) - 0.1 \
        + rng.standard_normal(Y2.shape) * 0.01
    out = {}
    for tag, (A, Bm) in {"freihand": (X, Y), "panet": (X2, Y2)}.items():
        out[tag + "_X"], out[tag + "_Y"] = A, Bm
        for scale in (True, False):
            res = [cst(A[i], Bm[i], compute_optimal_scale=scale) for i in range(len(A))]
            k = "%s_s%d_" % (tag, int(scale))
            out[k + "d"] = np.array([r[0] for r in res])
            out[k + "Z"] = np.array([r[1] for r in res])
            out[k + "T"] = np.array([r[2] for r in res])
            out[k + "b"] = np.array([float(r[3]) for r in res])
            out[k + "c"] = np.array([r[4] for r in res])
        out[tag + "_aws_Z"] = np.array([ev.align_w_scale(A[i], Bm[i]) for i in range(len(A))])
        # data/FreiHand/FreiHand.py:766-767,804-807
        Z = out[tag + "_s1_Z"]
        out[tag + "_pa_mpjpe"] = np.array(np.mean(np.sqrt(np.sum((Z - A) ** 2, axis=2))))
        out[tag + "_mpjpe"] = np.array(np.mean(np.sqrt(np.sum((Bm - A) ** 2, axis=2))))
        # main/eval.py:185-206: EvalUtil on raw and aligned predictions
        for nm, P in (("raw", Bm), ("aligned", out[tag + "_aws_Z"])):
            u = ev.EvalUtil(num_kp=21)
            for i in range(len(A)):
                u.feed(A[i], np.ones(21), P[i])
            # raw predictions are ~0.3 m off: widen the range so the PCK curve is not all zero
            vmax = 0.05 if nm == "aligned" else 1.0
            m = u.get_measures(0.0, vmax, 100)
            out["%s_%s_measures" % (tag, nm)] = np.array([m[0], m[1], m[2]])
            out["%s_%s_pck" % (tag, nm)] = np.asarray(m[3])
            out["%s_%s_thr" % (tag, nm)] = np.asarray(m[4])
    return out


def chain_case():
    """data/FreiHand/FreiHand.py:683-767 (labelled split: R, Procrustes) and :853-883 (evaluation
    split: trans from the box, no R) per sample through the live reference functions."""
    ch = ref_loader.load_chain_functions()
    cst = ref_loader.load_compute_similarity_transform()
    c = synth.chain_inputs(48, seed=301)
    N = len(c["tprime"])
    out = dict(c)
    trans = np.array([ch.gen_trans_from_patch_cv(*c["bbox"][i], 224, 224, 1.0, inv=True) for i in range(N)])
    out["trans"] = trans
    img, cam, cam_norot, Z = [], [], [], []
    for i in range(N):
        pim = ch.trans_coords_from_patch_to_org_3d(c["patch"][i].copy(), *c["bbox"][i], 224, 224, 1.0,
                                                   trans[i], c["tprime"][i])
        img.append(pim.copy())
        p3 = ch.pixel2cam(pim[:, 0:3].copy(), c["K"][i])
        e, _ = ch.scale_result(None, p3, method="scale", bone_length=float(c["bone_len"][i]))
        cam_norot.append(e)                                     # FreiHand.py:879-883
        p3r = np.matmul(c["R"][i].T, p3.T).T                    # FreiHand.py:722
        p3r, _ = ch.scale_result(None, p3r, method="scale", bone_length=float(c["bone_len"][i]))
        cam.append(p3r)
        Z.append(cst(c["gt_cam"][i], p3r, compute_optimal_scale=True)[1])
    out["img"], out["cam"], out["cam_norot"], out["Z"] = map(np.array, (img, cam, cam_norot, Z))
    out["pa_mpjpe"] = np.array(np.mean(np.sqrt(np.sum((out["Z"] - c["gt_cam"]) ** 2, axis=2))))
    out["mpjpe"] = np.array(np.mean(np.sqrt(np.sum((out["cam"] - c["gt_cam"]) ** 2, axis=2))))
    return out


def orthonormal_case():
    """make_orthonormal (nrsfm_modules.py:46-67) forward and its autograd backward in fp64 on
    random, reflected, nearly singular and nearly rotation matrices."""
    M = ref_loader.load_nrsfm_modules()
    A = synth.matrices3(96, seed=401)
    G = np.random.default_rng(402).standard_normal(A.shape)
    a = torch.as_tensor(A).clone().requires_grad_(True)
    r = M.make_orthonormal(a)
    (r * torch.as_tensor(G)).sum().backward()
    return {"A": A, "G": G, "R": r.detach().numpy(), "dA": a.grad.numpy(),
            "det": M.batch_det_3x3(torch.as_tensor(A)).numpy()}


def main():
    if not ref_loader.available():
        raise SystemExit("the reference tree is not mounted; goldens can only be made in the build container")
    L = ref_loader.load_loss_module()
    for name, spec in INTEGRAL_CASES.items():
        d = integral_case(L, *spec)
        np.savez_compressed(os.path.join(HERE, "integral_%s.npz" % name), **d)
        print(name, "loss64 %.12f" % d["loss64"], "|c32-c64| max %.2e" % np.abs(d["coords32"] - d["coords64"]).max())
    np.savez_compressed(os.path.join(HERE, "combined_cube8.npz"), **combined_case(L))
    ch = chain_case()
    np.savez_compressed(os.path.join(HERE, "chain.npz"), **ch)
    print("chain PA-MPJPE %.6f m, MPJPE %.6f m" % (ch["pa_mpjpe"], ch["mpjpe"]))
    o = orthonormal_case()
    np.savez_compressed(os.path.join(HERE, "orthonormal.npz"), **o)
    print("orthonormal: max |R^T R - I| %.2e" % np.abs(np.einsum("nji,njk->nik", o["R"], o["R"]) - np.eye(3)).max())
    if "--chain-only" in sys.argv:
        return
    p = procrustes_cases()
    np.savez_compressed(os.path.join(HERE, "procrustes.npz"), **p)
    print("freihand PA-MPJPE %.6f m, MPJPE %.6f m" % (p["freihand_pa_mpjpe"], p["freihand_mpjpe"]))
    for f in sorted(os.listdir(HERE)):
        if f.endswith(".npz"):
            print(f, os.path.getsize(os.path.join(HERE, f)))


if __name__ == "__main__":
    main()
