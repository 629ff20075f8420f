"""Stand-alone GPURadixSort64 / 32 throughput (Gkeys/s) on full-entropy keys.  usage: python tools/sort_bench.py [reps]"""
import os
import sys
import time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gaussiansplatting_b200 import capi as gs, synth
import torch

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 20
ctx = gs.Context(0)
stream = torch.cuda.ExternalStream(ctx.stream, device=0)
sizes = [int(x) for x in sys.argv[2].split(',')] if len(sys.argv) > 2 else (20, 23, 24, 26)
for ballot, logn in [(b, l) for l in sizes for b in (0, 1)]:
    gs.tune("sort_api_ballot", ballot)
    m = 1 << logn
    keys = synth.splitmix64(1234 + logn, m, 5)
    dk, dv = ctx.alloc(m * 8), ctx.alloc(m * 4)
    ctx.upload(dk, keys); ctx.upload(dv, np.arange(m, dtype=np.uint32))
    S = gs.Sort64(ctx, m)
    for _ in range(3):
        S.sort(dk, dv, m)
    ctx.sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(reps):
        S.sort(dk, dv, m)
    e1.record(stream); e1.synchronize()
    ms = e0.elapsed_time(e1) / reps
    print(f"hybrid={S.last_plan_hybrid} ballot={ballot} u64+u32 n=2^{logn}: {ms:.3f} ms  {m / ms / 1e6:.2f} Gkeys/s  {m * 200 / ms / 1e6:.0f} GB/s algorithmic (8 passes)")
    S.close(); ctx.free(dk); ctx.free(dv)
