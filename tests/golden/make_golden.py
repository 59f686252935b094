"""Generate tests/golden/*.npz by running the REFERENCE's own Python wrapper.

Run in the authoring container only (needs /root/reference; the GPU box has
neither it nor any need for this script):

    python tests/golden/make_golden.py

What runs: /root/reference/lib/ransac_voting_gpu_layer/ransac_voting_gpu.py is
imported unmodified (by path).  Its native module `ransac_voting` (CUDA, cannot
run here -- no GPU) is replaced by a shim over oracle/pvv_oracle.c, and three
torch-2.11 compatibility patches documented in SURVEY.md F3 are applied from
the outside:
  * Tensor.masked_select accepts the uint8 masks the wrapper passes (:550),
  * torch.gesv (removed from torch) is provided on top of torch.linalg.solve
    so that b_inv (:503-518) does not fall into its bare-except identity path,
  * Tensor.random_ / Tensor.uniform_ are wrapped to *record* what they drew, so
    that the same idxs / selection can be injected into the oracle and product.
Everything else -- compaction order, argmax tie rule, best-point update, the
refinement and covariance algebra -- is the reference's code executing on CPU
tensors.  The fixtures therefore pin oracle/oracle.py's control flow against
the reference wrapper; the kernel arithmetic itself is pinned on the GPU box
against oracle/_ref (the reference's CUDA build).
"""
from __future__ import annotations

import importlib.util
import os
import sys
import types

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import oracle as orc  # noqa: E402
from pvnet_b200 import synth  # noqa: E402

REF = "/root/reference"
OUT = os.path.dirname(os.path.abspath(__file__))
LOG = {"idxs": [], "selection": [], "hyp": [], "counts": []}


def _install_shim():
    shim = types.ModuleType("lib.ransac_voting_gpu_layer.ransac_voting")

    def generate_hypothesis(direct, coords, idxs):
        hyp = orc.generate_hypothesis(direct.numpy(), coords.numpy(), idxs.numpy())
        LOG["hyp"].append(hyp.copy())
        return torch.from_numpy(hyp)

    def voting_for_hypothesis(direct, coords, hyp, inliers, thresh):
        assert inliers.dtype == torch.uint8 and inliers.is_contiguous()
        orc.voting_for_hypothesis(direct.numpy(), coords.numpy(), hyp.numpy(), inliers.numpy(), thresh)
        LOG["counts"].append(inliers.numpy().astype(np.int64).sum(2))

    shim.generate_hypothesis = generate_hypothesis
    shim.voting_for_hypothesis = voting_for_hypothesis
    for name in ("lib", "lib.ransac_voting_gpu_layer"):
        m = types.ModuleType(name)
        m.__path__ = []
        sys.modules[name] = m
    sys.modules["lib.ransac_voting_gpu_layer.ransac_voting"] = shim
    sys.modules["lib.ransac_voting_gpu_layer"].ransac_voting = shim

    _ms = torch.Tensor.masked_select
    torch.Tensor.masked_select = lambda self, m: _ms(self, m.bool() if m.dtype == torch.uint8 else m)
    torch.gesv = lambda B, A: (torch.linalg.solve(A, B), None)
    _rand, _unif = torch.Tensor.random_, torch.Tensor.uniform_

    def random_(self, *a, **k):
        r = _rand(self, *a, **k)
        LOG["idxs"].append(r.numpy().copy())
        return r

    def uniform_(self, *a, **k):
        r = _unif(self, *a, **k)
        LOG["selection"].append(r.numpy().copy())
        return r

    torch.Tensor.random_, torch.Tensor.uniform_ = random_, uniform_


def _load_reference():
    spec = importlib.util.spec_from_file_location(
        "ref_ransac_voting_gpu", os.path.join(REF, "lib/ransac_voting_gpu_layer/ransac_voting_gpu.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def _clear():
    for v in LOG.values():
        v.clear()


def make_cat_fixture():
    from PIL import Image

    m = np.array(Image.open(os.path.join(REF, "data/demo/cat_mask.png")))[..., 0] != 0
    pts3d = np.loadtxt(os.path.join(REF, "data/demo/cat_points_3d.txt"))
    pose = np.load(os.path.join(REF, "data/demo/cat_pose.npy"))
    K = np.array([[572.4114, 0.0, 325.2611], [0.0, 573.57043, 242.04899], [0.0, 0.0, 1.0]])  # tools/demo.py:176-178
    p = (pts3d @ pose[:, :3].T + pose[:, 3]) @ K.T
    kp2d = p[:, :2] / p[:, 2:]
    np.savez_compressed(os.path.join(OUT, "cat_fixture.npz"), mask_bits=np.packbits(m.reshape(-1)), kp2d=kp2d)
    print("cat fixture: fg", int(m.sum()), "kp2d\n", kp2d)


def run_v3(ref, mask, vertex, hn, thresh, min_num, max_num, seed):
    """Runs the reference v3; returns its output plus the recorded draws."""
    _clear()
    torch.manual_seed(seed)
    out = ref.ransac_voting_layer_v3(torch.from_numpy(mask), torch.from_numpy(vertex), hn, inlier_thresh=thresh,
                                     min_num=min_num, max_num=max_num).numpy()
    b, h, w, vn, _ = vertex.shape
    # re-associate the recorded draws with images: skipped images draw nothing,
    # subsampled ones draw selection then idxs
    idxs = np.zeros((b, hn, vn, 2), np.int32)
    sel = np.zeros((b, h, w), np.float32)
    first_counts = np.full((b, hn, vn), -1, np.int64)
    it_i, it_s, it_c = iter(LOG["idxs"]), iter(LOG["selection"]), 0
    for bi in range(b):
        fg = int(torch.from_numpy(mask[bi]).byte().sum())
        if fg < min_num:
            continue
        if fg > max_num:
            sel[bi] = next(it_s)
        idxs[bi] = next(it_i)
        first_counts[bi] = LOG["counts"][it_c]
        # every further round re-votes the same idxs: identical counts
        while LOG["counts"][it_c].shape[0] != 1:
            assert np.array_equal(LOG["counts"][it_c], first_counts[bi])
            it_c += 1
        it_c += 1  # the [1,vn,tn] winner re-vote that ends the image
    return out, idxs, sel, first_counts


def run_v5(ref, mask, vertex, hn, thresh, min_num, max_num, seed):
    """Runs the reference's ransac_voting_layer_v5 (:769-864); returns (keypoints, confidence) plus the recorded draws."""
    _clear()
    torch.manual_seed(seed)
    out, conf = ref.ransac_voting_layer_v5(torch.from_numpy(mask), torch.from_numpy(vertex), hn, inlier_thresh=thresh,
                                           min_num=min_num, max_num=max_num)
    b, h, w, vn, _ = vertex.shape
    idxs = np.zeros((b, hn, vn, 2), np.int32)
    sel = np.zeros((b, h, w), np.float32)
    it_i, it_s = iter(LOG["idxs"]), iter(LOG["selection"])
    for bi in range(b):
        fg = int(torch.from_numpy(mask[bi]).byte().sum())
        if fg < min_num:
            continue
        if fg > max_num:
            sel[bi] = next(it_s)
        idxs[bi] = next(it_i)
    return out.numpy(), conf.numpy(), idxs, sel


def run_cov(ref, mask, vertex, mean, round_hyp, min_hyp, thresh, min_num, max_num, seed):
    _clear()
    torch.manual_seed(seed)
    m, cov = ref.estimate_voting_distribution_with_mean(
        torch.from_numpy(mask), torch.from_numpy(vertex), torch.from_numpy(mean), round_hyp_num=round_hyp,
        min_hyp_num=min_hyp, inlier_thresh=thresh, min_num=min_num, max_num=max_num)
    b, h, w, vn, _ = vertex.shape
    rounds = int(np.ceil(min_hyp / round_hyp))
    idxs = np.zeros((b, rounds, round_hyp, vn, 2), np.int32)
    sel = np.zeros((b, h, w), np.float32)
    it_i, it_s = iter(LOG["idxs"]), iter(LOG["selection"])
    for bi in range(b):
        fg = int((mask[bi] == 1).sum())
        if fg < min_num:
            continue
        if fg > max_num:
            sel[bi] = next(it_s)
        for r in range(rounds):
            idxs[bi, r] = next(it_i)
    return m.numpy(), cov.numpy(), idxs, sel


def main():
    _install_shim()
    ref = _load_reference()
    make_cat_fixture()
    cat_mask, cat_kp = synth.cat_fixture()

    # ---- G1/G2: configs[0] cat fixture, hn=128 --------------------------------
    for name, sigma, thresh, seed in (("cat_exact", 0.0, 0.99, 0), ("cat_noisy", 0.05, 0.999, 1)):
        rng = np.random.default_rng(100 + seed)
        vertex = synth.unit_field(cat_mask, cat_kp, sigma, rng, junk_background=False)[None]
        mask = cat_mask[None]
        out, idxs, sel, counts = run_v3(ref, mask, vertex, 128, thresh, 100, 30000, seed)
        ys, xs = np.nonzero(cat_mask)
        np.savez_compressed(os.path.join(OUT, f"v3_{name}.npz"), fg_vertex=vertex[0][ys, xs], idxs=idxs,
                            counts=counts, out=out, thresh=np.float32(thresh), hn=128, min_num=100, max_num=30000)
        print(name, "kp err px", np.abs(out[0] - cat_kp).max(), "best count", counts[0].max(0))

    # ---- G3: small batch with a skipped and a subsampled image ----------------
    h, w, vn, b = 60, 80, 9, 4
    rng = np.random.default_rng(7)
    mask = np.zeros((b, h, w), np.int64)
    vertex = np.zeros((b, h, w, vn, 2), np.float32)
    fracs = [0.10, 0.002, 0.45, 0.2]           # image 1 falls below min_num, image 2 exceeds max_num
    for i in range(b):
        mask[i] = synth.blob_mask(rng, fracs[i], h, w)
        kp = synth.keypoints_in_bbox(rng, mask[i], vn)
        vertex[i] = synth.unit_field(mask[i], kp, 0.08, rng)
    mask[3][mask[3] == 1] = rng.integers(1, 4, size=int((mask[3] == 1).sum()))  # labels 1..3: v3 and ==1 disagree
    out, idxs, sel, counts = run_v3(ref, mask, vertex, 64, 0.99, 50, 1500, 11)
    mean = out.copy()
    m2, cov, cidxs, csel = run_cov(ref, mask, vertex, mean, 32, 128, 0.99, 20, 1500, 12)
    np.savez_compressed(os.path.join(OUT, "small_batch.npz"), mask=mask.astype(np.uint8), vertex=vertex,
                        v3_idxs=idxs, v3_sel=sel, v3_counts=counts, v3_out=out,
                        cov_idxs=cidxs, cov_sel=csel, cov=cov, cov_mean=m2,
                        v3_args=np.array([64, 50, 1500]), cov_args=np.array([32, 128, 20, 1500]))
    # ---- G4: ransac_voting_layer_v5 the way EvalWrapper(use_uncertainty) calls it (tools/train_linemod.py:122-123):
    # 128 hypotheses, thresh 0.99, max_num=100 (every image is subsampled), min_num=5 ----
    mask5 = mask.copy()
    mask5[1] = 0
    mask5[1, 10, 10:13] = 1                       # 3 px < min_num = 5: skipped image -> zeros / zero confidence
    k5, c5, i5, s5 = run_v5(ref, mask5, vertex, 128, 0.99, 5, 100, 13)
    np.savez_compressed(os.path.join(OUT, "v5_small_batch.npz"), mask=mask5.astype(np.uint8), idxs=i5, sel=s5, out=k5,
                        conf=c5, args=np.array([128, 5, 100]))
    print("v5 conf\n", c5)
    # ---- G5: ransac_motion_voting (:966-987) on the same batch (image 1 of mask5 has 3 fg px, an empty one is added) ----
    mask6 = mask5.copy()
    mask6[1] = 0
    import warnings
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")      # uint8 mask indexing (:983) is deprecated in torch 2.x but still a mask
        mv = ref.ransac_motion_voting(torch.from_numpy(mask6), torch.from_numpy(vertex)).numpy()
    np.savez_compressed(os.path.join(OUT, "motion_small_batch.npz"), mask=mask6.astype(np.uint8), out=mv)
    print("motion voting\n", mv[:, :2])
    print("small batch v3 out\n", out[:, :2], "\ncov[0,0]\n", cov[0, 0], "\nfg", [(int((mask[i] != 0).sum())) for i in range(b)])


if __name__ == "__main__":
    main()
