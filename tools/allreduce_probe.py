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
dist.destroy_process_group()
