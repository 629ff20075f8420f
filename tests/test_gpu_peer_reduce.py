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


def _lost_peer_worker(rank, world, port, n, out_dir):
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
    D.upload(_grads_for(rank, n))
    dist.barrier()
    pr.allreduce(D, n); pr.check()                 # a healthy step first
    res = {}
    if rank == 0:                                  # rank 1 never enters the second step
        D.upload(_grads_for(5, n))
        pr.allreduce(D, n)                         # enqueues; barrier A gives up after 5 s on the device
        ctx.sync()
        got = D.download(n)
        res["poisoned"] = bool(np.isnan(got["position"][:, 0]).all())
        res["others_untouched"] = bool(np.array_equal(got["opacity"], _grads_for(5, n)["opacity"]))
        for name, call in (("check", pr.check), ("next", lambda: pr.allreduce(D, n))):
            try:
                call(); res[name] = "no error"
            except gs.GsError as e:
                res[name] = str(e)
    dist.barrier()
    np.save(os.path.join(out_dir, f"res_{rank}.npy"), np.array([repr(res)]))
    pr.close(); D.close(); ctx.close()
    dist.destroy_process_group()


def test_a_lost_peer_poisons_the_step_and_is_reported(tmp_path):
    """a barrier that times out must not let a partial sum reach Adam.  The reduce kernel withholds the sum, turns
    this rank's position-x gradients into NaN (adamStep's guard then skips every Gaussian, shaders.metal:567-570), and both
    gs_dp_check and the NEXT gs_dp_allreduce_grads return GS_ERR_NCCL (mapped pinned error word, no stream sync needed)"""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    import torch.multiprocessing as mp
    mp.spawn(_lost_peer_worker, args=(2, 29711, 4000, str(tmp_path)), nprocs=2, join=True)
    res = eval(str(np.load(tmp_path / "res_0.npy")[0]))
    assert res["poisoned"] and res["others_untouched"], res
    assert "did not arrive" in res["check"] and "did not arrive" in res["next"], res


def _step_worker(rank, world, port, out_dir):
    """one view-parallel training step on `world` GPUs: each rank renders its view, gradients are summed over NVLink peer
    memory, every rank takes the same Adam step"""
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.cuda.set_device(rank)
    from gaussiansplatting_b200 import capi as gs, synth
    n, W, H = 3000, 160, 128
    ctx = gs.Context(rank)
    g = synth.make_scene(n, 61, 0.04); gt_g = synth.make_scene(n, 62, 0.04)
    G = gs.Gaussians(ctx, n); R = gs.Rasterizer(ctx, n); D = gs.Grads(ctx, n); A = gs.Adam(ctx, n)
    img, gti = gs.Image(ctx, W, H), gs.Image(ctx, W, H)
    pr = gs.PeerReduce(ctx, rank, world)
    blobs = [None] * world
    dist.all_gather_object(blobs, pr.export(D))
    pr.connect(blobs)
    u = synth.make_camera(rank, world, W, H, n)
    G.upload(gt_g); R.forward(G, u, gti)
    G.upload(g)
    R.forward(G, u, img); R.backward(G, D, u, img, gti)
    ctx.sync(); dist.barrier()
    pr.allreduce(D, n); pr.check()
    grads = D.download(n)
    A.step(G, D, [1.6e-4, 5e-3, 1e-3, 2.5e-2, 2.5e-3])
    np.save(os.path.join(out_dir, f"grads_{rank}.npy"), grads)
    np.save(os.path.join(out_dir, f"params_{rank}.npy"), G.download())
    for o in (pr, gti, img, A, D, R, G):
        o.close()
    ctx.close()
    dist.destroy_process_group()


def test_two_rank_training_step_against_the_oracle_and_bit_equal_replicas(tmp_path, orc):
    """the summed gradients of a 2-rank step against the oracle's sum over the same two views (1e-3 relative bar of
    oracle/parity.py), both ranks holding bit-identical sums and bit-identical parameters after Adam"""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    import torch.multiprocessing as mp
    from gaussiansplatting_b200 import synth
    from oracle import parity
    mp.spawn(_step_worker, args=(2, 29713, str(tmp_path)), nprocs=2, join=True)
    g0, g1 = np.load(tmp_path / "grads_0.npy"), np.load(tmp_path / "grads_1.npy")
    p0, p1 = np.load(tmp_path / "params_0.npy"), np.load(tmp_path / "params_1.npy")
    assert g0.tobytes() == g1.tobytes() and p0.tobytes() == p1.tobytes()
    n, W, H = 3000, 160, 128
    g = synth.make_scene(n, 61, 0.04); gt_g = synth.make_scene(n, 62, 0.04)
    tot = tabs = tcnt = None
    for v in range(2):
        u = synth.make_camera(v, 2, W, H, n)
        r = orc.render(g, u); gt = orc.render(gt_g, u)["image"]
        gr, ab, cn = orc.backward(g, r["proj"], r["svals"], r["ranges"], u, r["last"], r["image"], gt, want_abs=True)
        if tot is None:
            tot, tabs, tcnt = gr.copy(), ab.copy(), cn.copy()
        else:
            for f in parity.GROUPS:
                tot[f] = tot[f] + gr[f]; tabs[f] = tabs[f] + ab[f]
            tcnt = tcnt + cn
    parity.assert_grads(g0, tot, tabs, tcnt)
    assert np.abs(tot["opacity"]).max() > 0 and not np.array_equal(p0["position"], g["position"])


def test_one_process_two_devices_sort_on_both(gs):
    """the >48 KB dynamic shared memory attribute of the onesweep / local-sort kernels is per device"""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    n = 1 << 20
    keys = np.random.default_rng(3).integers(0, 1 << 63, n, dtype=np.uint64)
    vals = np.arange(n, dtype=np.uint32)
    o = np.argsort(keys, kind="stable")
    for dev in (0, 1, 0):
        c = gs.Context(dev)
        S = gs.Sort64(c, n)
        try:
            k, v = S.sort_host(keys, vals)
        finally:
            S.close(); c.close()
        assert np.array_equal(k, keys[o]) and np.array_equal(v, vals[o]), dev


def _nvls_worker(rank, world, port, n, steps, out_dir):
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("cpu:gloo,cuda:nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    from gaussiansplatting_b200 import capi as gs, dp
    ctx = gs.Context(rank)
    res = {"multicast": False, "ok": True}
    try:
        SG = dp.SymmetricGrads(gs, ctx, dist, n, rank)
        res["multicast"] = SG.multicast
        for s in range(steps):
            SG.grads.upload(_grads_for(rank + 10 * s, n))
            dist.barrier()
            SG.reduce.allreduce(SG.grads, n)
            SG.reduce.check()
            got = SG.grads.download(n)
            for f in ("position", "opacity", "scale", "rotation", "sh", "viewspace"):
                want = _grads_for(0 + 10 * s, n)[f].astype(np.float64)
                for q in range(1, world):
                    want = want + _grads_for(q + 10 * s, n)[f]
                w32 = want.astype(np.float32)
                # two ranks: one rounding, so the switch's sum IS float32(a + b); more ranks: its order of addition is unspecified
                if world == 2:
                    res["ok"] &= bool(np.array_equal(got[f], w32))
                else:
                    res["ok"] &= bool(np.all(np.abs(got[f] - want) <= 4e-7 * np.maximum(np.abs(want), 1.0) * world))
        # a second buffer size through the same objects is not possible (external memory cannot grow): must be loud
        try:
            SG.grads.upload(_grads_for(0, SG.capacity + 64)); res["grow"] = "no error"
        except gs.GsError as e:
            res["grow"] = str(e)
        SG.close()
    except Exception as e:
        res["error"] = repr(e)
    np.save(os.path.join(out_dir, f"nvls_{rank}.npy"), np.array([repr(res)]))
    ctx.close()
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("n", [1000, 250007])
def test_nvls_allreduce_in_the_switch(tmp_path, n):
    """gs_grads_create_external + gs_dp_connect_ptrs: the gradient sum by multimem.ld_reduce / multimem.st on a symmetric,
    multicast-mapped buffer (plumbing: torch.distributed._symmetric_memory); every rank ends with the same bits"""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    import torch.multiprocessing as mp
    mp.spawn(_nvls_worker, args=(2, 29720 + (n % 7), n, 3, str(tmp_path)), nprocs=2, join=True)
    res = [eval(str(np.load(tmp_path / f"nvls_{r}.npy")[0])) for r in range(2)]
    if not res[0].get("multicast") and "error" not in res[0]:
        pytest.skip("no multicast (NVLS) memory on this box")
    for r in res:
        assert "error" not in r, r
        assert r["ok"] and "cannot grow" in r["grow"], r
