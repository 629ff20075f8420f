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
pr.close(); D.close(); ctx.close()
dist.destroy_process_group()
