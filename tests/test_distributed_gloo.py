"""N > 1 host logic on CPU: two gloo ranks shard a batch by image, vote on their slice, and one packed all-gather
assembles keypoints (+ covariances) on every rank.  The CUDA entry points cannot run here, so the per-rank vote is a
stand-in built on the CPU oracle (tests may use oracle/); what is under test is pvnet_b200.distributed: the shard
bounds, the global image ids handed to the draw stream, padding of uneven shards and the single collective."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from pvnet_b200 import distributed as pd
from pvnet_b200 import synth

H, W, VN, HN = 60, 80, 4, 32


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _idxs_for(image_id, tn, hn, seed):
    # a draw stream keyed by (seed, GLOBAL image id) like the library's Philox stream (pvv_common.cuh: pvv_draw_pair)
    return np.random.default_rng([seed, image_id]).integers(0, tn, size=(hn, VN, 2)).astype(np.int32)


def _vote_fn(calls):
    from oracle import oracle as orc

    def vote(mask, vertex, hn, seed=0, image_id_base=0, **kw):
        calls.append(("v3", int(image_id_base), int(mask.shape[0])))
        m, v = mask.numpy(), vertex.numpy()
        idxs = np.stack([_idxs_for(image_id_base + i, max(int((m[i] != 0).sum()), 1), hn, seed) for i in range(m.shape[0])])
        return torch.from_numpy(orc.ransac_voting_layer_v3(m, v, hn, 0.99, min_num=10, idxs=idxs))
    return vote


def _cov_fn(calls):
    from oracle import oracle as orc

    def cov(mask, vertex, mean, seed=0, image_id_base=0, **kw):
        calls.append(("cov", int(image_id_base), int(mask.shape[0])))
        m, v = mask.numpy(), vertex.numpy()
        idxs = np.stack([_idxs_for(image_id_base + i, max(int((m[i] == 1).sum()), 1), 2 * HN, seed).reshape(2, HN, VN, 2)
                         for i in range(m.shape[0])])
        _, c = orc.estimate_voting_distribution_with_mean(m, v, mean.numpy(), HN, 2 * HN, min_num=10, idxs=idxs)
        return mean, torch.from_numpy(c)
    return cov


def _worker(rank, world, port, global_b, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        mask, vertex, _ = synth.make_batch(global_b, vn=VN, frac=0.1, sigma=0.05, seed=7, h=H, w=W)
        lo, hi = pd.shard_bounds(global_b, world, rank)
        calls = []
        kp, cov = pd.ransac_voting_layer_v3_sharded(torch.from_numpy(mask[lo:hi]), torch.from_numpy(vertex[lo:hi]), HN,
                                                    global_b, seed=3, with_cov=True, vote_fn=_vote_fn(calls),
                                                    cov_fn=_cov_fn(calls))
        kp_only = pd.ransac_voting_layer_v3_sharded(torch.from_numpy(mask[lo:hi]), torch.from_numpy(vertex[lo:hi]), HN,
                                                    global_b, seed=3, vote_fn=_vote_fn(calls))
        with pytest.raises(ValueError):
            pd.ransac_voting_layer_v3_sharded(torch.from_numpy(mask[:1]), torch.from_numpy(vertex[:1]), HN, global_b + 4,
                                              vote_fn=_vote_fn([]))
        # the asynchronous gather (bench.py's N > 1 path): three steps through two rotating receive buffers
        async_ok = None
        if global_b % world == 0:
            g = pd.AsyncGather(global_b, (VN, 6), torch.device("cpu"))
            slots = [pd.ransac_voting_layer_v3_sharded(torch.from_numpy(mask[lo:hi]), torch.from_numpy(vertex[lo:hi]), HN,
                                                       global_b, seed=3, with_cov=True, vote_fn=_vote_fn([]),
                                                       cov_fn=_cov_fn([]), gather=g) for _ in range(3)]
            assert slots == [0, 1, 0]
            g.drain()
            full = g.wait(slots[-1])
            async_ok = bool(torch.equal(full[:, :, :2], kp) and torch.equal(full[:, :, 2:].reshape(global_b, VN, 2, 2), cov)
                            and torch.equal(g.wait(1), full))
            with pytest.raises(ValueError):
                pd.AsyncGather(global_b + 1, (VN, 2), torch.device("cpu"))
        q.put((rank, kp.numpy(), cov.numpy(), kp_only.numpy(), calls, (lo, hi), async_ok))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("global_b", [4, 5])
def test_two_ranks_equal_single_process(global_b):
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, global_b, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted((q.get(timeout=240) for _ in range(world)), key=lambda t: t[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    # single-process answer over the whole batch with the same global-id-keyed draws
    mask, vertex, _ = synth.make_batch(global_b, vn=VN, frac=0.1, sigma=0.05, seed=7, h=H, w=W)
    calls = []
    kp1 = _vote_fn(calls)(torch.from_numpy(mask), torch.from_numpy(vertex), HN, seed=3, image_id_base=0)
    _, cov1 = _cov_fn(calls)(torch.from_numpy(mask), torch.from_numpy(vertex), kp1, seed=4, image_id_base=0)
    covered = []
    for rank, kp, cov, kp_only, rcalls, (lo, hi), async_ok in res:
        assert async_ok is (True if global_b % world == 0 else None)
        assert kp.shape == (global_b, VN, 2) and cov.shape == (global_b, VN, 2, 2)
        assert np.array_equal(kp, kp1.numpy())            # every rank holds the full, rank-count-invariant result
        assert np.array_equal(cov, cov1.numpy())
        assert np.array_equal(kp_only, kp1.numpy())
        assert rcalls[0] == ("v3", lo, hi - lo) and rcalls[1] == ("cov", lo, hi - lo)
        covered += list(range(lo, hi))
    assert covered == list(range(global_b))                # contiguous partition, nothing dropped or doubled


def test_shard_bounds_partition():
    for b in (1, 7, 32, 128, 1024):
        for world in (1, 2, 3, 4, 8):
            spans = [pd.shard_bounds(b, world, r) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == b
            assert all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1


def test_world_size_one_needs_no_process_group():
    mask, vertex, _ = synth.make_batch(2, vn=VN, frac=0.1, sigma=0.05, seed=7, h=H, w=W)
    kp = pd.ransac_voting_layer_v3_sharded(torch.from_numpy(mask), torch.from_numpy(vertex), HN, 2, seed=3,
                                           vote_fn=_vote_fn([]))
    assert kp.shape == (2, VN, 2) and not dist.is_initialized()
