"""pvv_v3_noalloc (include/pvvote.h): the entry point that allocates nothing -- scratch in a static __device__ array.
Collected last on purpose.  Called through the C ABI (ctypes) and compared with the
existing entry point pvv_v3 on the same inputs: bit-identical keypoints (same kernels, same seed), on two calls in a row
with different inputs (the shared static scratch keeps the previous call's contents)."""
import ctypes as C

import numpy as np
import pytest
import torch

from pvnet_b200 import _lib, synth
from pvnet_b200 import ransac_voting_gpu as pv

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _job(tm, tv, hn, thresh, seed):
    job = pv._make_job(tm, tv, hn, thresh, 100, 30000, _lib.FG_BYTE, None, None, seed, 0, 0)
    job.stream = C.c_void_p(torch.cuda.current_stream(torch.device(DEV)).cuda_stream)
    return job


def test_v3_noalloc_matches_pvv_v3_twice_in_a_row():
    L = _lib.lib()
    limit = int(L.pvv_static_workspace_bytes())
    assert limit == 32 << 20
    with torch.cuda.device(torch.device(DEV)):
        for call, (seed, frac, hn, thresh) in enumerate([(3, 0.03, 512, 0.999), (4, 0.06, 128, 0.99)]):
            mask, vertex, _ = synth.make_batch(2, vn=9, frac=frac, sigma=0.05, seed=900 + call)
            tm, tv = torch.from_numpy(mask).to(DEV), synth.as_net_output(vertex, device=DEV)
            # existing entry point, caller-owned workspace
            job = _job(tm, tv, hn, thresh, seed)
            nbytes = int(L.pvv_workspace_bytes(C.byref(job)))
            assert 0 < nbytes <= limit
            ws = torch.empty(nbytes, dtype=torch.uint8, device=DEV)
            job.workspace, job.workspace_bytes = ws.data_ptr(), nbytes
            want = torch.empty((2, 9, 2), dtype=torch.float32, device=DEV)
            assert L.pvv_v3(C.byref(job), 0.99, 100, want.data_ptr(), None) == 0
            # the no-allocation entry point: no workspace passed at all
            job2 = _job(tm, tv, hn, thresh, seed)
            got = torch.full((2, 9, 2), float("nan"), dtype=torch.float32, device=DEV)
            before = torch.cuda.memory_allocated(DEV)
            assert L.pvv_v3_noalloc(C.byref(job2), 0.99, 100, got.data_ptr()) == 0
            assert torch.cuda.memory_allocated(DEV) == before
            torch.cuda.synchronize()
            assert torch.equal(got, want), call                      # same kernels, same draws: bit-identical
            assert np.isfinite(got.cpu().numpy()).all() and float(got.abs().sum()) > 0


def test_v3_noalloc_rejects_jobs_beyond_the_static_maximum():
    L = _lib.lib()
    with torch.cuda.device(torch.device(DEV)):
        mask, vertex, _ = synth.make_batch(8, vn=9, frac=0.03, sigma=0.05, seed=7)
        tm, tv = torch.from_numpy(mask).to(DEV), synth.as_net_output(vertex, device=DEV)
        job = _job(tm, tv, 512, 0.99, 1)
        assert int(L.pvv_workspace_bytes(C.byref(job))) > int(L.pvv_static_workspace_bytes())
        out = torch.zeros((8, 9, 2), dtype=torch.float32, device=DEV)
        assert L.pvv_v3_noalloc(C.byref(job), 0.99, 100, out.data_ptr()) == -2      # PVV_ERR_WORKSPACE, nothing launched
        torch.cuda.synchronize()
        assert not out.any()
        assert L.pvv_v3_noalloc(None, 0.99, 100, out.data_ptr()) == -1               # PVV_ERR_INVALID
