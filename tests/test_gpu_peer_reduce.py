"""gs_dp_*: the gradient sum of view-parallel training over NVLink peer memory (csrc/peer_reduce.cu) against numpy.
The sum is taken in rank order by the owner of each slice, so every rank must hold exactly float32(g0 + g1 (+ g2 ...)).
Needs two GPUs for the exchange itself (skipped on a one-GPU box; run with `gpurun --gpus 2`)."""
import os
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _grads_for(rank, n):
    from gaussiansplatting_b200.layouts import GRADIENTS
    rng = np.random.default_rng(100 + rank)
    g = np.zeros(n, GRADIENTS)
    for f, w in (("position", 3), ("scale", 3), ("rotation", 4)):
        g[f] = rng.standard_normal((n, w)).astype(np.float32)
    g["opacity"] = rng.standard_normal(n).astype(np.float32)
    g["sh"][:, [0, 4, 8]] = rng.standard_normal((n, 3)).astype(np.float32)
    g["viewspace"] = rng.standard_normal((n, 2)).astype(np.float32)
    return g


def _worker(rank, world, port, n, steps, out_dir):
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.cuda.set_device(rank)
    from gaussiansplatting_b200 import capi as gs
    ctx = gs.Context(rank)
    D = gs.Grads(ctx, n)
    pr = gs.PeerReduce(ctx, rank, world)
    blobs = [None] * world
    dist.all_gather_object(blobs, pr.export(D))
    pr.connect(blobs)
    ok = True
    for s in range(steps):
        mine = _grads_for(rank + 10 * s, n)
        D.upload(mine)
        dist.barrier()                          # nobody starts before everyone's upload is enqueued AND done (upload synchronises)
        pr.allreduce(D, n)
        pr.check()
        got = D.download(n)
        want = {f: _grads_for(0 + 10 * s, n)[f].astype(np.float32) for f in ("position", "opacity", "scale", "rotation", "sh", "viewspace")}
        for q in range(1, world):
            gq = _grads_for(q + 10 * s, n)
            for f in want:
                want[f] = (want[f] + gq[f]).astype(np.float32)
        for f in want:
            ok &= bool(np.array_equal(got[f], want[f]))
    np.save(os.path.join(out_dir, f"ok_{rank}.npy"), np.array([ok]))
    pr.close(); D.close(); ctx.close()
    dist.destroy_process_group()


@pytest.mark.parametrize("n", [1000, 250007])
def test_peer_allreduce_two_gpus(tmp_path, n):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    import torch.multiprocessing as mp
    port = 29600 + (n % 97)
    mp.spawn(_worker, args=(2, port, n, 3, str(tmp_path)), nprocs=2, join=True)
    for r in range(2):
        assert bool(np.load(tmp_path / f"ok_{r}.npy")[0]), f"rank {r}"


def test_peer_allreduce_single_rank_is_a_no_op_and_misuse_is_loud(gs, ctx):
    n = 500
    D = gs.Grads(ctx, n)
    g = _grads_for(7, n)
    D.upload(g)
    one = gs.PeerReduce(ctx, 0, 1)
    two = gs.PeerReduce(ctx, 0, 2)
    try:
        one.allreduce(D, n); one.check()
        assert np.array_equal(D.download(n)["position"], g["position"])
        with pytest.raises(gs.GsError):
            two.allreduce(D, n)                  # not connected
        with pytest.raises(gs.GsError):
            gs.PeerReduce(ctx, 3, 2)             # rank outside the world
    finally:
        two.close(); one.close(); D.close()
