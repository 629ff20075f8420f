"""time a 64 MB fp32 sum all-reduce (the packed gradient payload at 1 M Gaussians) under the current NCCL environment
usage: torchrun --nproc-per-node N tools/allreduce_probe.py"""
import os
import torch
import torch.distributed as dist
rank = int(os.environ["RANK"]); local = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
x = torch.ones(16 * 1_000_000, device="cuda")
for _ in range(5):
    dist.all_reduce(x)
torch.cuda.synchronize(); dist.barrier()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(20):
    dist.all_reduce(x)
e1.record(); e1.synchronize()
if rank == 0:
    ms = e0.elapsed_time(e1) / 20
    print(f"NCCL env {{{', '.join(k + '=' + v for k, v in os.environ.items() if k.startswith('NCCL'))}}}: {ms * 1e3:.0f} us per 64 MB all-reduce, "
          f"algbw {64e6 / ms / 1e6:.0f} GB/s", flush=True)

# ---- the same payload through gs_dp_allreduce_grads (own kernels over NVLink peer memory, csrc/peer_reduce.cu)
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gaussiansplatting_b200 import capi as gs
world = dist.get_world_size()
n = 1_000_000
ctx = gs.Context(local)
stream = torch.cuda.ExternalStream(ctx.stream, device=local)
D = gs.Grads(ctx, n); D.zero(n)
pr = gs.PeerReduce(ctx, rank, world)
blobs = [None] * world
dist.all_gather_object(blobs, pr.export(D))
pr.connect(blobs)
for _ in range(5):
    pr.allreduce(D, n)
ctx.sync(); dist.barrier()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(stream)
for _ in range(20):
    pr.allreduce(D, n)
e1.record(stream); e1.synchronize()
pr.check()
if rank == 0:
    ms = e0.elapsed_time(e1) / 20
    print(f"peer kernels: {ms * 1e3:.0f} us per 64 MB all-reduce, algbw {64e6 / ms / 1e6:.0f} GB/s", flush=True)
dist.barrier()

# ---- the in-switch (NVLS) sum on symmetric multicast memory, swept over the launch shape of its kernel
from gaussiansplatting_b200 import dp
try:
    SG = dp.SymmetricGrads(gs, ctx, dist, n, local)
except Exception as e:
    SG = None
    if rank == 0:
        print("no symmetric memory:", repr(e)[:200], flush=True)
if SG is not None and SG.multicast:
    def time_nvls(reps=20):
        for _ in range(3):
            SG.reduce.allreduce(SG.grads, n)
        ctx.sync(); dist.barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        for _ in range(reps):
            SG.reduce.allreduce(SG.grads, n)
        b.record(stream); b.synchronize()
        t = torch.tensor([a.elapsed_time(b) / reps], device="cuda"); dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())
    rows = []
    for threads in (256, 512):
        for unroll in (1, 2, 4, 8):
            for grid in (16, 32, 64, 148, 296, 592, 1184):
                gs.tune("dp_mc_threads", threads); gs.tune("dp_mc_unroll", unroll); gs.tune("dp_mc_grid", grid)
                rows.append((time_nvls(10), threads, unroll, grid))
    SG.reduce.check()
    if rank == 0:
        rows.sort()
        for ms, threads, unroll, grid in rows[:12] + rows[-3:]:
            print(f"nvls threads {threads} unroll {unroll} grid {grid}: {ms * 1e3:.0f} us, algbw {64e6 / ms / 1e6:.0f} GB/s", flush=True)
    dist.barrier()
    SG.close()
pr.close(); D.close(); ctx.close()
dist.destroy_process_group()
