"""BASELINE.json configs[2], configs[3] (one rank's share) and the configs[4] sweep, on one GPU or -- under torchrun -- on
N GPUs of one box (weak scaling: every rank runs the same cases on data of its own; a case's time is the MAXIMUM over the
ranks, its images/s the sum; rank 0 prints).

    python tools/sweep.py [--quick] > gpurun_out/sweep.jsonl
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/sweep.py --quick

One JSON line per case: stage times from CUDA events inside the C ABI, hyp.pixel tests/s of the fused counting kernel
and its fraction of the FP32 peak measured in the same process (12 FLOP per test).  Parity at these
sizes is checked through a size-independent property: the guard-banded filter's counts equal the exact-only kernel's.
"""
import argparse
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pvnet_b200 import ransac_voting_gpu as pv  # noqa: E402
from pvnet_b200 import synth  # noqa: E402

WORLD, RANK, LOCAL = (int(os.environ.get(k, d)) for k, d in (("WORLD_SIZE", "1"), ("RANK", "0"), ("LOCAL_RANK", "0")))
DEV = f"cuda:{LOCAL}"
FLOP = 12.0
if WORLD > 1:
    import torch.distributed as dist
    torch.cuda.set_device(LOCAL)
    dist.init_process_group("nccl", device_id=torch.device(DEV))


def over_ranks(st):
    """stage times -> their maximum over the ranks (the same dict when there is one rank)"""
    if WORLD == 1:
        return st
    keys = sorted(st)
    t = torch.tensor([st[k] for k in keys], device=DEV, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return dict(zip(keys, (float(x) for x in t.cpu())))


def emit(rec):
    if RANK == 0:
        rec["n_gpus"] = WORLD
        print(json.dumps(rec), flush=True)


def timed(fn, reps):
    st = []
    for _ in range(reps + 2):
        fn(st)
    keys = st[0].keys()
    return {k: float(np.median([s[k] for s in st[2:]])) for k in keys}


def case(name, mask, vertex, hn, thresh, peak, reps=10, check=True, **kw):
    tm, tv = torch.from_numpy(mask).to(DEV), synth.as_net_output(vertex, device=DEV)
    st = over_ranks(timed(lambda s: pv.ransac_voting_layer_v3(tm, tv, hn, thresh, seed=1, stage_ms=s, **kw), reps))
    out, dbg = pv.ransac_voting_layer_v3(tm, tv, hn, thresh, seed=1, return_debug=True, **kw)
    tn = dbg["tn"].cpu().numpy().astype(np.int64)
    tests = float(tn.sum()) * vertex.shape[3] * hn
    same = None
    if check:
        _, d2 = pv.ransac_voting_layer_v3(tm, tv, hn, thresh, seed=1, return_debug=True, force_exact=True, **kw)
        same = bool(torch.equal(dbg["counts"], d2["counts"]))
    tps = tests / (st["count"] * 1e-3) if st["count"] > 0 else 0.0
    rec = {"case": name, "b": int(mask.shape[0]), "vn": int(vertex.shape[3]), "hn": hn, "thresh": thresh,
           "fg_px_total": int((mask != 0).sum()), "tn_total": int(tn.sum()), "subsampled_images": int((tn < (mask != 0).sum((1, 2))).sum()),
           "stage_ms": st, "images_per_s": WORLD * mask.shape[0] / (st["total"] * 1e-3), "k3_tests_per_s": tps,
           "k3_frac_of_fp32_peak": tps * FLOP / 1e12 / peak, "amb_frac": float(int(dbg["amb_count"]) / max(tests, 1.0)),
           "filter_equals_exact": same}
    emit(rec)
    del tm, tv
    torch.cuda.empty_cache()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--quick", action="store_true")
    a = ap.parse_args()
    peak = pv.fp32_peak_tflops(4096)
    emit({"fp32_peak_tflops_measured": peak, "gpu": torch.cuda.get_device_name(0)})

    # ---- configs[2]: uncertainty path, batch 64: v3(512, 0.99) -> estimate_voting_distribution_with_mean defaults ----
    b = 16 if a.quick else 64
    mask, vertex, _ = synth.make_batch(b, vn=9, frac=0.03, sigma=0.05, seed=4321 + 100 * RANK)
    kp = case("configs[2] v3(512, 0.99) feeding the distribution", mask, vertex, 512, 0.99, peak)
    tm, tv = torch.from_numpy(mask).to(DEV), synth.as_net_output(vertex, device=DEV)
    st = over_ranks(timed(lambda s: pv.estimate_voting_distribution_with_mean(tm, tv, kp, seed=2, stage_ms=s), 5))
    _, cov, dbg = pv.estimate_voting_distribution_with_mean(tm, tv, kp, seed=2, return_debug=True)
    _, cov2, d2 = pv.estimate_voting_distribution_with_mean(tm, tv, kp, seed=2, return_debug=True, force_exact=True)
    tests = float(dbg["tn"].sum()) * 9 * 4096
    tps = tests / (st["count"] * 1e-3)
    emit({"case": "configs[2] estimate_voting_distribution_with_mean (16 x 256 hyp, 0.99)", "b": b, "vn": 9,
          "hn": 4096, "stage_ms": st, "images_per_s": WORLD * b / (st["total"] * 1e-3), "k3_tests_per_s": tps,
          "k3_frac_of_fp32_peak": tps * FLOP / 1e12 / peak,
          "filter_equals_exact": bool(torch.equal(dbg["counts"], d2["counts"]) and torch.equal(cov, cov2)),
          "cov_finite": bool(torch.isfinite(cov).all())})
    del tm, tv
    torch.cuda.empty_cache()

    # ---- configs[3]: Occlusion-LINEMOD shape, one rank's share: 16 images x 8 instance masks = 128 units ----
    units = 32 if a.quick else 128
    rng = np.random.default_rng(99)
    masks = np.zeros((units, synth.H, synth.W), np.int64)
    verts = np.zeros((units, synth.H, synth.W, 9, 2), np.float32)
    for i in range(units):
        r = np.random.default_rng(5000 + i + 1000 * RANK)
        masks[i] = synth.blob_mask(r, float(rng.uniform(0.005, 0.03)))
        verts[i] = synth.unit_field(masks[i], synth.keypoints_in_bbox(r, masks[i], 9), 0.05, r)
    case("configs[3] 16 images x 8 instance masks (one of 8 ranks)", masks, verts, 512, 0.999, peak)
    del masks, verts

    # ---- configs[4]: sweep foreground fraction x hypotheses x keypoints ----
    fracs = [0.01, 0.10, 0.60] if a.quick else [0.01, 0.03, 0.10, 0.30, 0.60]
    hns = [128, 2048] if a.quick else [128, 256, 512, 2048]
    vns = [9, 17] if a.quick else [9, 13, 17]
    b = 8
    for vn in vns:
        for frac in fracs:
            mask, vertex, _ = synth.make_batch(b, vn=vn, frac=frac, sigma=0.05, seed=int(1000 * frac) + vn + 100 * RANK)
            for hn in hns:
                case(f"configs[4] fg={frac:.0%} hn={hn} vn={vn}", mask, vertex, hn, 0.999, peak, reps=6, check=(hn <= 512))


if __name__ == "__main__":
    main()
    if WORLD > 1:
        dist.destroy_process_group()
