"""probe: does torch symmetric memory give a multicast (NVLS) pointer on this box?  torchrun --nproc-per-node N tools/symm_probe.py"""
import os, sys, traceback
import torch, torch.distributed as dist
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
try:
    import torch.distributed._symmetric_memory as symm_mem
    print(rank, "symm_mem api:", [n for n in dir(symm_mem) if not n.startswith("_")][:40], flush=True)
    t = symm_mem.empty(16 * 1024 * 1024, dtype=torch.float32, device=f"cuda:{local}")
    hdl = symm_mem.rendezvous(t, dist.group.WORLD.group_name)
    print(rank, "handle attrs:", [n for n in dir(hdl) if not n.startswith("_")], flush=True)
    print(rank, "multicast_ptr", hex(getattr(hdl, "multicast_ptr", 0) or 0), "buffer_ptrs", [hex(p) for p in hdl.buffer_ptrs],
          "signal_pad_ptrs", [hex(p) for p in hdl.signal_pad_ptrs][:2], "local", hex(t.data_ptr()), flush=True)
    t.fill_(float(rank + 1))
    dist.barrier(); torch.cuda.synchronize()
    try:
        out = torch.ops.symm_mem.multimem_all_reduce_(t, "sum", dist.group.WORLD.group_name)
        torch.cuda.synchronize()
        print(rank, "multimem_all_reduce_ ->", float(t[0]), float(t[-1]), flush=True)
    except Exception as e:
        print(rank, "multimem_all_reduce_ failed:", repr(e)[:300], flush=True)
except Exception:
    traceback.print_exc()
dist.barrier()
dist.destroy_process_group()
