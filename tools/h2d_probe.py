"""What limits the host-buffer path when several ranks share one box?  Every rank copies a pinned 78.6 MB buffer (the
int64 mask of configs[1]) to its own GPU, all ranks at once; reported per N: the slowest rank's GB/s and the aggregate.
Run once per N under torchrun (N = 1: plain python).  Also times the host-side narrowing (8 workers) of the same buffer
through the host entry point with a trivially small voting job, i.e. the CPU side of the e2e path.
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/h2d_probe.py"""
import json, os, sys, time
import numpy as np, torch
import torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
world, rank, local = (int(os.environ.get(k, d)) for k, d in (("WORLD_SIZE", "1"), ("RANK", "0"), ("LOCAL_RANK", "0")))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
nbytes = 32 * 480 * 640 * 8
h = torch.zeros(nbytes, dtype=torch.uint8).pin_memory()
d = torch.empty(nbytes, dtype=torch.uint8, device=dev)
for _ in range(3):
    d.copy_(h, non_blocking=True)
torch.cuda.synchronize()
if world > 1:
    dist.barrier()
    torch.cuda.synchronize()
reps = 30
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(reps):
    d.copy_(h, non_blocking=True)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / reps
t = torch.tensor([ms], device=dev)
if world > 1:
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
worst = float(t.item())
# CPU read bandwidth over the same pinned buffer from this process (numpy sum, one thread): a lower bound on what a
# narrowing worker sees
a = h.numpy().view(np.int64)
t0 = time.perf_counter()
for _ in range(5):
    a.sum()
cpu_gbs = 5 * nbytes / (time.perf_counter() - t0) / 1e9
t2 = torch.tensor([cpu_gbs], device=dev)
if world > 1:
    dist.all_reduce(t2, op=dist.ReduceOp.MIN)
if rank == 0:
    print(json.dumps({"n_gpus": world, "h2d_mb": nbytes / 1e6, "ms_per_copy_slowest_rank": worst, "gbs_per_rank_slowest": nbytes / worst / 1e6,
                      "gbs_aggregate": world * nbytes / worst / 1e6, "cpu_read_gbs_one_thread_slowest_rank": float(t2.item()),
                      "host_threads": os.cpu_count()}), flush=True)
if world > 1:
    dist.destroy_process_group()
